// OBJ subset loader, host mirror of src/parsing/obj.zig:10-283:
//   v / vn / f (fan triangulation, "i", "i/t", "i//n", "i/t/n") / g ; everything else is counted in
//   lines_ignored; optional normalisation of vertices to [-1,1] around the bbox centre.
// Lines are split on '\n' and tokens on single spaces, exactly like std.mem.tokenizeScalar.
#include <cstdlib>

#include "loader.hpp"

namespace rtc {

namespace {
std::vector<std::string> tokenize(const std::string& s, char sep) {
    std::vector<std::string> out;
    size_t i = 0;
    while (i < s.size()) {
        while (i < s.size() && s[i] == sep) i++;
        size_t j = i;
        while (j < s.size() && s[j] != sep) j++;
        if (j > i) out.emplace_back(s, i, j - i);
        i = j;
    }
    return out;
}
// std.fmt.parseFloat: the whole token must be a number.
bool parse_float(const std::string& t, double& out) {
    if (t.empty()) return false;
    char* e = nullptr;
    out = std::strtod(t.c_str(), &e);
    return e == t.c_str() + t.size();
}
bool parse_index(const std::string& t, size_t& out) {
    if (t.empty()) return false;
    size_t i = 0;
    if (t[0] == '+') i = 1;
    if (i >= t.size()) return false;
    size_t v = 0;
    for (; i < t.size(); i++) {
        if (t[i] == '_') continue;  // Zig's parseInt accepts '_' separators
        if (t[i] < '0' || t[i] > '9') return false;
        v = v * 10 + (size_t)(t[i] - '0');
    }
    out = v;
    return true;
}
struct FaceRef {
    size_t v = 0;
    bool has_n = false;
    size_t n = 0;
};
// handleFaceHelper (obj.zig:83-99)
bool parse_face_ref(const std::string& tok, FaceRef& r) {
    std::vector<std::string> parts;  // splitScalar keeps empty fields
    size_t start = 0;
    for (;;) {
        size_t k = tok.find('/', start);
        if (k == std::string::npos) {
            parts.push_back(tok.substr(start));
            break;
        }
        parts.push_back(tok.substr(start, k - start));
        start = k + 1;
    }
    if (!parse_index(parts[0], r.v)) return false;
    if (parts.size() < 3) return true;
    if (!parse_index(parts[2], r.n)) return false;
    r.has_n = true;
    return true;
}
}  // namespace

Shape& ObjParser::active() {
    return active_path.empty() ? default_group : default_group.children[active_path[0]];
}

const Shape* ObjParser::named_group(const std::string& name) const {
    for (auto& kv : named_groups)
        if (kv.first == name) return &default_group.children[kv.second];
    return nullptr;
}

bool ObjParser::handle_line(const std::string& line, const ObjInherited& st) {
    std::vector<std::string> tok = tokenize(line, ' ');
    if (tok.empty()) return false;
    const std::string& first = tok[0];
    if (first == "v") {
        double x, y, z;
        if (tok.size() < 4 || !parse_float(tok[1], x) || !parse_float(tok[2], y) || !parse_float(tok[3], z)) return false;
        // point - offset then / scale : w ends up 1/scale, harmless (SURVEY §8a notes)
        vertices.push_back(Tuple::point(x, y, z).sub(offset).div(scale));
        return true;
    }
    if (first == "vn") {
        double x, y, z;
        if (tok.size() < 4 || !parse_float(tok[1], x) || !parse_float(tok[2], y) || !parse_float(tok[3], z)) return false;
        normals.push_back(Tuple::vec3(x, y, z));
        return true;
    }
    if (first == "f") {
        if (tok.size() < 4) return false;
        FaceRef f0, last;
        if (!parse_face_ref(tok[1], f0) || !parse_face_ref(tok[2], last)) return false;
        for (size_t i = 3; i < tok.size(); i++) {
            FaceRef cur;
            if (!parse_face_ref(tok[i], cur)) return false;  // triangles already emitted stay (as in Zig)
            if (f0.v < 1 || last.v < 1 || cur.v < 1 || f0.v > vertices.size() || last.v > vertices.size() ||
                cur.v > vertices.size())
                return false;
            const Tuple &p1 = vertices[f0.v - 1], &p2 = vertices[last.v - 1], &p3 = vertices[cur.v - 1];
            bool smooth = f0.has_n && last.has_n && cur.has_n;
            if (smooth && (f0.n < 1 || last.n < 1 || cur.n < 1 || f0.n > normals.size() || last.n > normals.size() ||
                           cur.n > normals.size()))
                return false;
            Shape tri = smooth ? Shape::smooth_triangle(p1, p2, p3, normals[f0.n - 1], normals[last.n - 1],
                                                        normals[cur.n - 1])
                               : Shape::triangle(p1, p2, p3);
            tri.material = st.has_material ? st.material : Material();
            tri.casts_shadow = st.casts_shadow < 0 ? true : (st.casts_shadow != 0);
            active().add_child(std::move(tri));
            last = cur;
        }
        return true;
    }
    if (first == "g") {
        if (tok.size() < 2) return false;
        default_group.add_child(Shape::group());
        size_t idx = default_group.children.size() - 1;
        named_groups.emplace_back(tok[1], idx);
        active_path.assign(1, idx);
        return true;
    }
    return false;
}

void ObjParser::load_obj(const std::string& text, const ObjInherited& st, bool normalize) {
    // lines = the non-empty pieces between '\n' (std.mem.tokenizeScalar); walked in place, one reused buffer
    auto for_each_line = [&](auto&& fn) {
        std::string line;
        size_t i = 0;
        while (i < text.size()) {
            while (i < text.size() && text[i] == '\n') i++;
            size_t j = i;
            while (j < text.size() && text[j] != '\n') j++;
            if (j > i) {
                line.assign(text, i, j - i);
                fn(line);
            }
            i = j;
        }
    };
    if (normalize) {  // obj.zig:198-268
        double mn[3] = {kInf, kInf, kInf}, mx[3] = {-kInf, -kInf, -kInf};
        for_each_line([&](const std::string& line) {
            // only lines whose first token is exactly "v" count: look before splitting the whole line
            size_t k = 0;
            while (k < line.size() && line[k] == ' ') k++;
            if (k >= line.size() || line[k] != 'v' || (k + 1 < line.size() && line[k + 1] != ' ')) return;
            std::vector<std::string> tok = tokenize(line, ' ');
            if (tok.empty() || tok[0] != "v") return;
            for (int c = 0; c < 3; c++) {
                double v;
                if ((size_t)(c + 1) < tok.size() && parse_float(tok[c + 1], v)) {
                    if (v < mn[c]) mn[c] = v;
                    if (v > mx[c]) mx[c] = v;
                }
            }
        });
        double sx = mx[0] - mn[0], sy = mx[1] - mn[1], sz = mx[2] - mn[2];
        offset = Tuple::vec3(mn[0] + 0.5 * sx, mn[1] + 0.5 * sy, mn[2] + 0.5 * sz);
        scale = 0.5 * std::fmax(sx, std::fmax(sy, sz));
    }
    for_each_line([&](const std::string& line) {
        if (!handle_line(line, st)) lines_ignored++;
    });
}

}  // namespace rtc
