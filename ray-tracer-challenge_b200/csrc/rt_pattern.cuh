// Pattern evaluation on the device: Pattern.patternAt (patterns/pattern.zig:112-124) and every
// variant, UV maps and image lookups (patterns/texture_map.zig), Perlin noise (noise.zig).
//
// The reference recurses through child patterns; every combining pattern is LINEAR in its
// children's colours (blend: 0.5/0.5, gradient: 1-f / f), and the selecting ones (stripes, rings,
// checkers, uv checkers, align-check) forward to exactly one child, so the recursion becomes a
// small explicit stack of (pattern, object point, weight) and the colour is the weighted sum of
// the solid / texel leaves reached.
#pragma once
#include "rt_math.cuh"

namespace rtc {

static __device__ __constant__ unsigned char c_perm[256] = {
    151, 160, 137, 91,  90,  15,  131, 13,  201, 95,  96,  53,  194, 233, 7,   225, 140, 36,  103, 30,  69,  142,
    8,   99,  37,  240, 21,  10,  23,  190, 6,   148, 247, 120, 234, 75,  0,   26,  197, 62,  94,  252, 219, 203,
    117, 35,  11,  32,  57,  177, 33,  88,  237, 149, 56,  87,  174, 20,  125, 136, 171, 168, 68,  175, 74,  165,
    71,  134, 139, 48,  27,  166, 77,  146, 158, 231, 83,  111, 229, 122, 60,  211, 133, 230, 220, 105, 92,  41,
    55,  46,  245, 40,  244, 102, 143, 54,  65,  25,  63,  161, 1,   216, 80,  73,  209, 76,  132, 187, 208, 89,
    18,  169, 200, 196, 135, 130, 116, 188, 159, 86,  164, 100, 109, 198, 173, 186, 3,   64,  52,  217, 226, 250,
    124, 123, 5,   202, 38,  147, 118, 126, 255, 82,  85,  212, 207, 206, 59,  227, 47,  16,  58,  17,  182, 189,
    28,  42,  223, 183, 170, 213, 119, 248, 152, 2,   44,  154, 163, 70,  221, 153, 101, 155, 167, 43,  172, 9,
    129, 22,  39,  253, 19,  98,  108, 110, 79,  113, 224, 232, 178, 185, 112, 104, 218, 246, 97,  228, 251, 34,
    242, 193, 238, 210, 144, 12,  191, 179, 162, 241, 81,  51,  145, 235, 249, 14,  239, 107, 49,  192, 214, 31,
    181, 199, 106, 157, 184, 84,  204, 176, 115, 121, 50,  45,  127, 4,   150, 254, 138, 236, 205, 93,  222, 114,
    67,  29,  24,  72,  243, 141, 128, 195, 78,  66,  215, 61,  156, 180};

__device__ __forceinline__ int perm(int i) { return c_perm[i & 255]; }  // noise.zig:25-33 (table twice)
__device__ __forceinline__ double nfade(double t) { return t * t * t * (t * (t * 6.0 - 15.0) + 10.0); }
__device__ __forceinline__ double nlerp(double t, double a, double b) { return a + t * (b - a); }
__device__ __forceinline__ double ngrad(int hash, double x, double y, double z) {  // noise.zig:98-104
    int h = hash & 15;
    double u = h < 8 ? x : y;
    double v = h < 4 ? y : ((h == 12 || h == 14) ? x : z);
    return ((h & 1) == 0 ? u : -u) + ((h & 2) == 0 ? v : -v);
}
static __device__ __noinline__ double perlin(double x, double y, double z) {  // noise.zig:52-88
    double fx = floor(x), fy = floor(y), fz = floor(z);
    int X = (int)((long long)fx & 255), Y = (int)((long long)fy & 255), Z = (int)((long long)fz & 255);
    x -= fx;
    y -= fy;
    z -= fz;
    double u = nfade(x), v = nfade(y), w = nfade(z);
    int A = perm(X) + Y, AA = perm(A) + Z, AB = perm(A + 1) + Z;
    int B = perm(X + 1) + Y, BA = perm(B) + Z, BB = perm(B + 1) + Z;
    return nlerp(w,
                 nlerp(v, nlerp(u, ngrad(perm(AA), x, y, z), ngrad(perm(BA), x - 1, y, z)),
                       nlerp(u, ngrad(perm(AB), x, y - 1, z), ngrad(perm(BB), x - 1, y - 1, z))),
                 nlerp(v, nlerp(u, ngrad(perm(AA + 1), x, y, z - 1), ngrad(perm(BA + 1), x - 1, y, z - 1)),
                       nlerp(u, ngrad(perm(AB + 1), x, y - 1, z - 1), ngrad(perm(BB + 1), x - 1, y - 1, z - 1))));
}
__device__ __forceinline__ double octave_noise(double x, double y, double z, int octaves, double persistence) {
    double total = 0, frequency = 1, amplitude = 1, max_value = 0;  // noise.zig:35-50
    for (int i = 0; i < octaves; i++) {
        total += perlin(x * frequency, y * frequency, z * frequency) * amplitude;
        max_value += amplitude;
        amplitude *= persistence;
        frequency *= 2.0;
    }
    return total / max_value;
}

// Canvas.getPixelPointer on a texture (canvas.zig:139-147).  RGBA8 texels become colours the way
// zigimg + Canvas.fromImage do: f32(u8)/255 widened to f64 (canvas.zig:39-42; DESIGN.md "third-party").
__device__ __forceinline__ V3 texel(const DScene& S, const rtc_image& im, long ix, long iy, int* err) {
    if (ix < 0 || iy < 0 || ix >= im.width || iy >= im.height) {
        *err = 1;  // the reference unwraps a null pointer here (panic); we flag the frame instead
        return v3(0, 0, 0);
    }
    size_t i = (size_t)iy * im.width + ix;
    if (im.format == RTC_IMG_RGBA8) {
        uchar4 c = __ldg(reinterpret_cast<const uchar4*>(S.texels + im.offset) + i);
        return v3((double)((float)c.x / 255.0f), (double)((float)c.y / 255.0f), (double)((float)c.z / 255.0f));
    }
    const double* p = reinterpret_cast<const double*>(S.texels + im.offset) + i * 3;
    return v3(__ldg(p), __ldg(p + 1), __ldg(p + 2));
}

// LEAN scenes (render_kernel.cuh): every material's pattern is a solid colour or ONE selecting / mixing pattern over two
// solid colours (stripes, rings, checkers, gradient, radial gradient, blend) -- most shipped scenes.  Same operations in
// the same order as pattern_color below performs for such a tree (weight 1: x * 1.0 and 0.0 + x are exact), without the
// explicit stack, the texture code and the call: bit-identical, tested against the general instantiation.
__device__ __forceinline__ V3 simple_pattern_color(const DScene& S, int root, V3 op) {
    const rtc_pattern& P = S.patterns[root];
    const int kind = P.kind;
    if (kind == RTC_PAT_SOLID) return v3(__ldg(P.color), __ldg(P.color + 1), __ldg(P.color + 2));
    M34 inv = load_m34(P.inverse);
    V3 pp = xf_point(inv, op);
    const double* ca = S.patterns[P.a].color;
    const double* cb = S.patterns[P.b].color;
    bool first;
    if (kind == RTC_PAT_STRIPES) first = zmod(pp.x, 2.0) < 1.0;                                          // stripes.zig:27-33
    else if (kind == RTC_PAT_RINGS) first = zmod(floor(sqrt(pp.x * pp.x + pp.z * pp.z)), 2.0) < 1.0;    // rings.zig:27-33
    else if (kind == RTC_PAT_CHECKERS) first = zmod(floor(pp.x) + floor(pp.y) + floor(pp.z), 2.0) < 1.0;  // checkers.zig:23-29
    else {
        double f = 0.5;  // blend.zig:24-27
        if (kind == RTC_PAT_GRADIENT) f = pp.x - floor(pp.x);  // gradient.zig:27-32
        if (kind == RTC_PAT_RADIAL_GRADIENT) {                 // gradient.zig:49-55
            double mag = sqrt(pp.x * pp.x + pp.z * pp.z);
            f = mag - floor(mag);
        }
        const double wa = 1.0 - f;
        return v3(__ldg(ca), __ldg(ca + 1), __ldg(ca + 2)) * wa + v3(__ldg(cb), __ldg(cb + 1), __ldg(cb + 2)) * f;
    }
    const double* c = first ? ca : cb;
    return v3(__ldg(c), __ldg(c + 1), __ldg(c + 2));
}

struct PatTask {
    int pat;
    V3 p;      // object-space point handed to this pattern
    double w;  // weight of its colour in the result
};
#define RTC_PAT_STACK 8

// Returns the pattern colour at object point `op`.  *err is set on a texture lookup out of range
// or when the explicit stack overflows.
static __device__ __noinline__ V3 pattern_color(const DScene& S, int root, V3 op, int* err) {
    PatTask stack[RTC_PAT_STACK];
    int sp = 0;
    stack[sp++] = {root, op, 1.0};
    V3 acc = v3(0, 0, 0);
    while (sp > 0) {
        PatTask tk = stack[--sp];
        for (;;) {  // follow single-child forwards without touching the stack
            const rtc_pattern& P = S.patterns[tk.pat];
            M34 inv = load_m34(P.inverse);
            V3 pp = xf_point(inv, tk.p);
            int kind = P.kind;
            int next = -1;
            if (kind == RTC_PAT_SOLID) {
                acc = acc + v3(P.color[0], P.color[1], P.color[2]) * tk.w;
            } else if (kind == RTC_PAT_TEST) {
                acc = acc + pp * tk.w;
            } else if (kind == RTC_PAT_STRIPES) {  // stripes.zig:27-33
                next = zmod(pp.x, 2.0) < 1.0 ? P.a : P.b;
            } else if (kind == RTC_PAT_RINGS) {  // rings.zig:27-33
                next = zmod(floor(sqrt(pp.x * pp.x + pp.z * pp.z)), 2.0) < 1.0 ? P.a : P.b;
            } else if (kind == RTC_PAT_CHECKERS) {  // checkers.zig:23-29
                next = zmod(floor(pp.x) + floor(pp.y) + floor(pp.z), 2.0) < 1.0 ? P.a : P.b;
            } else if (kind == RTC_PAT_BLEND || kind == RTC_PAT_GRADIENT || kind == RTC_PAT_RADIAL_GRADIENT) {
                double f = 0.5;  // blend.zig:24-27 : (a + b) * 0.5
                if (kind == RTC_PAT_GRADIENT) f = pp.x - floor(pp.x);  // gradient.zig:27-32 : a + (b - a) f
                if (kind == RTC_PAT_RADIAL_GRADIENT) {                 // gradient.zig:49-55
                    double mag = sqrt(pp.x * pp.x + pp.z * pp.z);
                    f = mag - floor(mag);
                }
                if (sp >= RTC_PAT_STACK) {
                    *err = 2;
                } else {
                    stack[sp++] = {P.b, tk.p, tk.w * f};
                }
                tk.w = tk.w * (1.0 - f);
                next = P.a;
            } else if (kind == RTC_PAT_PERTURB) {  // perturb.zig:31-47
                double ox = octave_noise(tk.p.x, tk.p.y, tk.p.z, P.octaves, P.perturb_persistence);
                double oy = octave_noise(tk.p.x, tk.p.y, tk.p.z + 1.0, P.octaves, P.perturb_persistence);
                double oz = octave_noise(tk.p.x, tk.p.y, tk.p.z + 2.0, P.octaves, P.perturb_persistence);
                tk.p = tk.p + v3(ox, oy, oz) * P.perturb_scale;
                next = P.a;
            } else if (kind == RTC_PAT_TEXTURE_MAP) {
                const double pi = 3.14159265358979323846;
                double u, v;
                int face = 0;
                if (P.mapping == RTC_MAP_SPHERICAL) {  // texture_map.zig:171-190
                    double theta = atan2(pp.x, pp.z);
                    double radius = sqrt(pp.x * pp.x + pp.y * pp.y + pp.z * pp.z);
                    double phi = acos(pp.y / radius);
                    double raw_u = theta / (2.0 * pi);
                    u = 1 - (raw_u + 0.5);
                    v = 1 - phi / pi;
                } else if (P.mapping == RTC_MAP_PLANAR) {  // :192-200
                    u = zmod(pp.x, 1.0);
                    v = zmod(pp.z, 1.0);
                } else if (P.mapping == RTC_MAP_CYLINDRICAL) {  // :202-213
                    double theta = atan2(pp.x, pp.z);
                    double raw_u = theta / (2.0 * pi);
                    u = 1.0 - (raw_u + 0.5);
                    v = zmod(pp.y, 1.0);
                } else {  // cubic :215-290 ; record order front,back,left,right,up,down
                    double coord = fmax(fabs(pp.x), fmax(fabs(pp.y), fabs(pp.z)));
                    if (coord == pp.x) face = 3;
                    else if (coord == -pp.x) face = 2;
                    else if (coord == pp.y) face = 4;
                    else if (coord == -pp.y) face = 5;
                    else if (coord == pp.z) face = 0;
                    else face = 1;
                    double a, b;
                    switch (face) {
                        case 0: a = pp.x + 1.0; b = pp.y + 1.0; break;
                        case 1: a = 1.0 - pp.x; b = pp.y + 1.0; break;
                        case 2: a = pp.z + 1.0; b = pp.y + 1.0; break;
                        case 3: a = 1.0 - pp.z; b = pp.y + 1.0; break;
                        case 4: a = pp.x + 1.0; b = 1.0 - pp.z; break;
                        default: a = pp.x + 1.0; b = pp.z + 1.0; break;
                    }
                    u = zmod(a, 2.0) / 2.0;
                    v = zmod(b, 2.0) / 2.0;
                }
                const rtc_uv_pattern& U = S.uvs[P.uv[face]];
                if (U.kind == RTC_UV_TEST) {  // texture_map.zig:9-19
                    acc = acc + v3(u, v, 0.0) * tk.w;
                } else if (U.kind == RTC_UV_ALIGN_CHECK) {  // :21-43
                    int k = 0;
                    if (v > 0.8) {
                        if (u < 0.2) k = 1;
                        else if (u > 0.8) k = 2;
                    } else if (v < 0.2) {
                        if (u < 0.2) k = 3;
                        else if (u > 0.8) k = 4;
                    }
                    next = U.sub[k];
                } else if (U.kind == RTC_UV_CHECKERS) {  // :45-63
                    double ua = floor(u * U.width), va = floor(v * U.height);
                    next = zmod(ua + va, 2.0) < 1.0 ? U.sub[0] : U.sub[1];
                } else {  // image :65-105 : no wrap, no half-texel offset
                    const rtc_image& im = S.images[U.image];
                    double x = u * (double)(im.width - 1);
                    double y = (1.0 - v) * (double)(im.height - 1);
                    V3 c;
                    if (!U.bilinear) {
                        c = texel(S, im, (long)round(x), (long)round(y), err);
                    } else {
                        double x1 = floor(x), x2 = ceil(x), y1 = floor(y), y2 = ceil(y);
                        V3 c11 = texel(S, im, (long)x1, (long)y1, err), c21 = texel(S, im, (long)x2, (long)y1, err);
                        V3 c12 = texel(S, im, (long)x1, (long)y2, err), c22 = texel(S, im, (long)x2, (long)y2, err);
                        V3 cx1 = c11 * (x2 - x) + c21 * (x - x1);
                        V3 cx2 = c12 * (x2 - x) + c22 * (x - x1);
                        c = cx1 * (y2 - y) + cx2 * (y - y1);
                    }
                    acc = acc + c * tk.w;
                }
            }
            if (next < 0) break;
            tk.pat = next;
        }
    }
    return acc;
}

}  // namespace rtc
