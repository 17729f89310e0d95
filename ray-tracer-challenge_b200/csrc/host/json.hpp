// Minimal JSON DOM (objects keep key order) for the scene schema of src/parsing/scene.zig:28-210.
// Numbers are parsed with strtod (correctly rounded, like Zig's std.fmt.parseFloat).
#pragma once
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace rtc {

struct JsonError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

struct Json {
    enum Type { Null, Bool, Number, String, Array, Object } type = Null;
    bool b = false;
    double num = 0;
    std::string str;
    std::vector<Json> arr;
    std::vector<std::pair<std::string, Json>> obj;

    const Json* find(const std::string& k) const {
        for (auto& kv : obj)
            if (kv.first == k) return &kv.second;
        return nullptr;
    }
    const Json& at(const std::string& k) const {
        const Json* j = find(k);
        if (!j) throw JsonError("missing field '" + k + "'");
        return *j;
    }
    double number() const {
        if (type != Number) throw JsonError("expected number");
        return num;
    }
    bool boolean() const {
        if (type != Bool) throw JsonError("expected bool");
        return b;
    }
    const std::string& string() const {
        if (type != String) throw JsonError("expected string");
        return str;
    }
    const std::vector<Json>& array() const {
        if (type != Array) throw JsonError("expected array");
        return arr;
    }
    // A Zig union(enum) is encoded as an object with exactly one key.
    const std::pair<std::string, Json>& single() const {
        if (type != Object || obj.size() != 1) throw JsonError("expected single-key object (tagged union)");
        return obj[0];
    }
};

class JsonParser {
   public:
    explicit JsonParser(const std::string& s) : p_(s.c_str()), end_(s.c_str() + s.size()) {}
    Json parse() {
        Json j = value();
        ws();
        if (p_ != end_) fail("trailing characters");
        return j;
    }

   private:
    const char* p_;
    const char* end_;
    [[noreturn]] void fail(const char* m) { throw JsonError(std::string("json: ") + m); }
    void ws() {
        while (p_ < end_ && (*p_ == ' ' || *p_ == '\n' || *p_ == '\t' || *p_ == '\r')) p_++;
    }
    bool lit(const char* s) {
        size_t n = std::strlen(s);
        if ((size_t)(end_ - p_) >= n && std::memcmp(p_, s, n) == 0) {
            p_ += n;
            return true;
        }
        return false;
    }
    Json value() {
        ws();
        if (p_ >= end_) fail("unexpected end");
        Json j;
        char c = *p_;
        if (c == '{') {
            p_++;
            j.type = Json::Object;
            ws();
            if (p_ < end_ && *p_ == '}') {
                p_++;
                return j;
            }
            for (;;) {
                ws();
                if (p_ >= end_ || *p_ != '"') fail("expected key");
                std::string k = str();
                ws();
                if (p_ >= end_ || *p_ != ':') fail("expected ':'");
                p_++;
                j.obj.emplace_back(std::move(k), value());
                ws();
                if (p_ < end_ && *p_ == ',') {
                    p_++;
                    continue;
                }
                if (p_ < end_ && *p_ == '}') {
                    p_++;
                    break;
                }
                fail("expected ',' or '}'");
            }
        } else if (c == '[') {
            p_++;
            j.type = Json::Array;
            ws();
            if (p_ < end_ && *p_ == ']') {
                p_++;
                return j;
            }
            for (;;) {
                j.arr.push_back(value());
                ws();
                if (p_ < end_ && *p_ == ',') {
                    p_++;
                    continue;
                }
                if (p_ < end_ && *p_ == ']') {
                    p_++;
                    break;
                }
                fail("expected ',' or ']'");
            }
        } else if (c == '"') {
            j.type = Json::String;
            j.str = str();
        } else if (lit("true")) {
            j.type = Json::Bool;
            j.b = true;
        } else if (lit("false")) {
            j.type = Json::Bool;
            j.b = false;
        } else if (lit("null")) {
            j.type = Json::Null;
        } else {
            char* e = nullptr;
            j.num = std::strtod(p_, &e);
            if (e == p_) fail("bad value");
            j.type = Json::Number;
            p_ = e;
        }
        return j;
    }
    std::string str() {
        std::string out;
        p_++;  // opening quote
        while (p_ < end_ && *p_ != '"') {
            if (*p_ == '\\') {
                p_++;
                if (p_ >= end_) fail("bad escape");
                switch (*p_) {
                    case 'n': out += '\n'; break;
                    case 't': out += '\t'; break;
                    case 'r': out += '\r'; break;
                    case 'b': out += '\b'; break;
                    case 'f': out += '\f'; break;
                    case 'u': {
                        if (end_ - p_ < 5) fail("bad \\u");
                        unsigned v = (unsigned)std::strtoul(std::string(p_ + 1, 4).c_str(), nullptr, 16);
                        p_ += 4;
                        if (v < 0x80) out += (char)v;
                        else if (v < 0x800) { out += (char)(0xC0 | (v >> 6)); out += (char)(0x80 | (v & 0x3F)); }
                        else { out += (char)(0xE0 | (v >> 12)); out += (char)(0x80 | ((v >> 6) & 0x3F)); out += (char)(0x80 | (v & 0x3F)); }
                        break;
                    }
                    default: out += *p_;
                }
                p_++;
            } else {
                out += *p_++;
            }
        }
        if (p_ >= end_) fail("unterminated string");
        p_++;
        return out;
    }
};

}  // namespace rtc
