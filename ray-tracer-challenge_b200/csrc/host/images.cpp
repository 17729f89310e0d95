// Image decoders standing in for zigimg (PNG, un-vendored dependency pinned at
// SinclaM/zigimg@3c4f73a, build.zig.zon:5-8) and Canvas.fromPpm (src/raytracer/canvas.zig:48-121).
#include <zlib.h>

#include <cstring>
#include <fstream>
#include <sstream>

#include "loader.hpp"

namespace rtc {

std::string read_file(const std::string& path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("FileNotFound: " + path);
    std::ostringstream ss;
    ss << f.rdbuf();
    return ss.str();
}

FileLoader directory_loader(const std::string& dir) {
    return [dir](const std::string& name) { return read_file(dir.empty() ? name : dir + "/" + name); };
}

static uint32_t be32(const unsigned char* p) {
    return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3];
}

std::shared_ptr<Image> decode_png(const std::string& bytes, const std::string& name) {
    static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    if (bytes.size() < 8 || std::memcmp(bytes.data(), sig, 8) != 0) throw std::runtime_error("png: bad signature");
    const unsigned char* p = (const unsigned char*)bytes.data() + 8;
    const unsigned char* end = (const unsigned char*)bytes.data() + bytes.size();
    uint32_t w = 0, h = 0;
    int depth = 0, ctype = 0, interlace = 0;
    bool have_ihdr = false;
    std::string idat;
    std::vector<unsigned char> palette;
    while (p + 12 <= end) {
        uint32_t len = be32(p);
        const char* type = (const char*)p + 4;
        const unsigned char* data = p + 8;
        if (data + len + 4 > end) throw std::runtime_error("png: truncated chunk");
        if (!have_ihdr && std::memcmp(type, "IHDR", 4)) throw std::runtime_error("png: IHDR is not the first chunk");
        if (!std::memcmp(type, "IHDR", 4)) {
            if (len != 13 || have_ihdr) throw std::runtime_error("png: bad IHDR");
            have_ihdr = true;
            w = be32(data);
            h = be32(data + 4);
            depth = data[8];
            ctype = data[9];
            interlace = data[12];
        } else if (!std::memcmp(type, "PLTE", 4)) {
            palette.assign(data, data + len);
        } else if (!std::memcmp(type, "IDAT", 4)) {
            idat.append((const char*)data, len);
        } else if (!std::memcmp(type, "IEND", 4)) {
            break;
        }
        p = data + len + 4;
    }
    if (!have_ihdr) throw std::runtime_error("png: no IHDR");
    // 16-bit samples and Adam7 interlacing are not decoded (the reference's decoder, zigimg, is not vendored: what it
    // accepts beyond the shipped 8-bit RGB, non-interlaced textures was not checked)
    if (depth != 8 || interlace != 0) throw std::runtime_error("png: only 8-bit non-interlaced supported");
    // sizes come from the file: nothing below may wrap (1 <= w, h and w * h * channels <= 2^31)
    if (w == 0 || h == 0 || (uint64_t)w * h > (uint64_t(1) << 29)) throw std::runtime_error("png: image size out of range");
    int ch = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : ctype == 6 ? 4 : 0;
    if (!ch) throw std::runtime_error("png: bad colour type");
    size_t stride = (size_t)w * ch;
    std::vector<unsigned char> raw((stride + 1) * h);
    uLongf out_len = raw.size();
    if (uncompress(raw.data(), &out_len, (const Bytef*)idat.data(), idat.size()) != Z_OK || out_len != raw.size())
        throw std::runtime_error("png: inflate failed");
    // undo the per-scanline filters (PNG spec 9.2)
    std::vector<unsigned char> pix(stride * h);
    for (uint32_t y = 0; y < h; y++) {
        int ft = raw[(stride + 1) * y];
        const unsigned char* src = &raw[(stride + 1) * y + 1];
        unsigned char* dst = &pix[stride * y];
        const unsigned char* up = y ? &pix[stride * (y - 1)] : nullptr;
        for (size_t i = 0; i < stride; i++) {
            int a = i >= (size_t)ch ? dst[i - ch] : 0;
            int b = up ? up[i] : 0;
            int c = (up && i >= (size_t)ch) ? up[i - ch] : 0;
            int v = src[i];
            switch (ft) {
                case 0: break;
                case 1: v += a; break;
                case 2: v += b; break;
                case 3: v += (a + b) / 2; break;
                case 4: {
                    int pa = std::abs(b - c), pb = std::abs(a - c), pc = std::abs(a + b - 2 * c);
                    v += (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
                    break;
                }
                default: throw std::runtime_error("png: bad filter");
            }
            dst[i] = (unsigned char)v;
        }
    }
    auto img = std::make_shared<Image>();
    img->width = (int)w;
    img->height = (int)h;
    img->name = name;
    img->rgb8.resize((size_t)w * h * 3);
    for (size_t i = 0; i < (size_t)w * h; i++) {
        unsigned char r, g, b;
        if (ctype == 2 || ctype == 6) {
            r = pix[i * ch];
            g = pix[i * ch + 1];
            b = pix[i * ch + 2];
        } else if (ctype == 3) {
            size_t k = (size_t)pix[i] * 3;
            if (k + 3 > palette.size()) throw std::runtime_error("png: palette index");
            r = palette[k];
            g = palette[k + 1];
            b = palette[k + 2];
        } else {
            r = g = b = pix[i * ch];
        }
        img->rgb8[i * 3] = r;
        img->rgb8[i * 3 + 1] = g;
        img->rgb8[i * 3 + 2] = b;
    }
    return img;
}

// Canvas.fromPpm (canvas.zig:48-121): "P3", optional # comment lines before the dimensions, scale,
// then whitespace-separated samples; colour = sample / scale.
std::shared_ptr<Image> decode_ppm(const std::string& text) {
    std::vector<std::string> lines;
    {
        std::string cur;
        for (char c : text) {
            if (c == '\n') {
                if (!cur.empty()) lines.push_back(cur);
                cur.clear();
            } else
                cur += c;
        }
        if (!cur.empty()) lines.push_back(cur);
    }
    size_t li = 0;
    if (lines.empty() || lines[li++] != "P3") throw std::runtime_error("InvalidMagicNumber");
    auto next_noncomment = [&](const char* err) -> std::string {
        while (li < lines.size() && lines[li].rfind("#", 0) == 0) li++;
        if (li >= lines.size()) throw std::runtime_error(err);
        return lines[li++];
    };
    std::string dims = next_noncomment("InvalidDimensions");
    long w = 0, h = 0;
    {
        std::istringstream ss(dims);
        if (!(ss >> w >> h) || w <= 0 || h <= 0) throw std::runtime_error("InvalidDimensions");
    }
    std::string sc = next_noncomment("InvalidScale");
    double scale = 0;
    {
        std::istringstream ss(sc);
        if (!(ss >> scale) || scale <= 0) throw std::runtime_error("InvalidScale");
    }
    auto img = std::make_shared<Image>();
    img->width = (int)w;
    img->height = (int)h;
    img->name = "<ppm>";
    for (; li < lines.size(); li++) {
        if (lines[li].rfind("#", 0) == 0) continue;
        std::istringstream ss(lines[li]);
        double v;
        while (ss >> v) img->rgbf.push_back(v / scale);
    }
    if (img->rgbf.size() != (size_t)w * h * 3) throw std::runtime_error("InvalidPixelData");
    return img;
}

std::shared_ptr<Image> decode_image(const std::string& bytes, const std::string& name) {
    if (bytes.size() >= 2 && bytes[0] == 'P' && bytes[1] == '3') return decode_ppm(bytes);
    return decode_png(bytes, name);
}

std::shared_ptr<Image> standin_image(const std::vector<std::shared_ptr<const Image>>& others,
                                     const std::string& name) {
    auto img = std::make_shared<Image>();
    img->width = img->height = 1;
    img->name = name + " (stand-in: mean of sibling images)";
    double sum[3] = {0, 0, 0};
    double n = 0;
    for (auto& o : others) {
        if (o->rgb8.empty()) continue;
        size_t px = (size_t)o->width * o->height;
        unsigned long long s[3] = {0, 0, 0};
        for (size_t i = 0; i < px; i++)
            for (int c = 0; c < 3; c++) s[c] += o->rgb8[i * 3 + c];
        for (int c = 0; c < 3; c++) sum[c] += (double)s[c];
        n += (double)px;
    }
    img->rgb8.resize(3);
    for (int c = 0; c < 3; c++) img->rgb8[c] = n > 0 ? (uint8_t)std::lround(sum[c] / n) : 128;
    return img;
}

}  // namespace rtc
