// The step right after the render: Canvas.ppm (canvas.zig:181-254) and the rgb24 PNG that
// main.zig:95-99 writes from Canvas.toImage (canvas.zig:150-176).  Host code, no GPU involved.
#include <zlib.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../../include/rtc_b200.h"

namespace {

// color.zig:61-71 clamp(): round half away from zero of c * 255, clipped to [0, 255]
inline unsigned clamp255(double c) {
    double v = std::round(c * 255.0);
    if (!(v > 0.0)) return 0;  // also NaN
    return v > 255.0 ? 255u : (unsigned)v;
}

void put_be32(std::vector<unsigned char>& o, uint32_t v) {
    o.push_back((unsigned char)(v >> 24));
    o.push_back((unsigned char)(v >> 16));
    o.push_back((unsigned char)(v >> 8));
    o.push_back((unsigned char)v);
}

void chunk(std::vector<unsigned char>& o, const char type[4], const unsigned char* data, size_t n) {
    put_be32(o, (uint32_t)n);
    size_t at = o.size();
    o.insert(o.end(), type, type + 4);
    if (n) o.insert(o.end(), data, data + n);
    put_be32(o, (uint32_t)crc32(0L, o.data() + at, (uInt)(n + 4)));
}

}  // namespace

extern "C" {

int rtc_canvas_ppm(const double* rgb, uint64_t width, uint64_t height, char** text_out, size_t* nbytes_out) {
    if (!rgb || !text_out || !nbytes_out || width == 0 || height == 0) return RTC_ERR_INVALID;
    std::string s;
    s.reserve((size_t)(width * height * 12));
    s += "P3\n" + std::to_string(width) + " " + std::to_string(height) + "\n255\n";
    size_t col = 0;
    char num[8];
    for (uint64_t i = 0; i < width * height; i++) {
        for (int ch = 0; ch < 3; ch++) {
            int len = std::snprintf(num, sizeof num, "%u", clamp255(rgb[3 * i + ch]));
            if (col + (size_t)len >= 70) {  // the value does not fit on the line: break it (canvas.zig:199-203)
                s += '\n';
                col = 0;
            } else if (ch != 0 || i % width != 0) {  // a space, except before the first value of a row
                s += ' ';
                col += 1;
            }
            s.append(num, (size_t)len);
            col += (size_t)len;
        }
        if ((i + 1) % width == 0) {
            s += '\n';
            col = 0;
        }
    }
    char* p = static_cast<char*>(std::malloc(s.size() + 1));
    if (!p) return RTC_ERR_INVALID;
    std::memcpy(p, s.c_str(), s.size() + 1);
    *text_out = p;
    *nbytes_out = s.size();
    return RTC_OK;
}

int rtc_write_png(const char* path, const uint8_t* rgba, uint64_t width, uint64_t height) {
    if (!path || !rgba || width == 0 || height == 0 || width > 0x7fffffffu || height > 0x7fffffffu) return RTC_ERR_INVALID;
    // rgb24, filter 0 on every scanline (the alpha byte of the RGBA8 frame is dropped, like toImage's rgb24)
    std::vector<unsigned char> raw((size_t)height * (1 + 3 * (size_t)width));
    size_t k = 0;
    for (uint64_t y = 0; y < height; y++) {
        raw[k++] = 0;
        const uint8_t* row = rgba + 4 * y * width;
        for (uint64_t x = 0; x < width; x++) {
            raw[k++] = row[4 * x];
            raw[k++] = row[4 * x + 1];
            raw[k++] = row[4 * x + 2];
        }
    }
    uLongf zn = compressBound((uLong)raw.size());
    std::vector<unsigned char> z(zn);
    if (compress2(z.data(), &zn, raw.data(), (uLong)raw.size(), 6) != Z_OK) return RTC_ERR_INVALID;
    std::vector<unsigned char> out = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    std::vector<unsigned char> ihdr;
    put_be32(ihdr, (uint32_t)width);
    put_be32(ihdr, (uint32_t)height);
    const unsigned char tail[5] = {8, 2, 0, 0, 0};  // 8 bits, colour type 2 (RGB), deflate, adaptive filtering, no interlace
    ihdr.insert(ihdr.end(), tail, tail + 5);
    chunk(out, "IHDR", ihdr.data(), ihdr.size());
    chunk(out, "IDAT", z.data(), zn);
    chunk(out, "IEND", nullptr, 0);
    FILE* f = std::fopen(path, "wb");
    if (!f) return RTC_ERR_INVALID;
    bool ok = std::fwrite(out.data(), 1, out.size(), f) == out.size();
    ok = (std::fclose(f) == 0) && ok;
    return ok ? RTC_OK : RTC_ERR_INVALID;
}

}  // extern "C"
