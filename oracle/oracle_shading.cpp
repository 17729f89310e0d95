// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.hpp).  Patterns, noise and Phong lighting.
#include "oracle.hpp"

namespace oracle {

// ---- noise.zig (Ken Perlin's improved noise) ---------------------------------------------------
static const uint8_t kPerm[256] = {
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
static inline int perm(int i) { return kPerm[i & 255]; }  // p[512] is the table twice (noise.zig:25-33)
static inline double fade(double t) { return t * t * t * (t * (t * 6.0 - 15.0) + 10.0); }
static inline double lerp(double t, double a, double b) { return a + t * (b - a); }
static inline double grad(int hash, double x, double y, double z) {  // noise.zig:98-104
    int h = hash & 15;
    double u = h < 8 ? x : y;
    double v = h < 4 ? y : ((h == 12 || h == 14) ? x : z);
    return ((h & 1) == 0 ? u : -u) + ((h & 2) == 0 ? v : -v);
}

double noise(double x, double y, double z) {  // noise.zig:52-88
    // Cell index = floor(coordinate) & 255 (pinned by the negative-coordinate KAT, noise.zig:108).
    int X = (int)((long long)std::floor(x) & 255), Y = (int)((long long)std::floor(y) & 255),
        Z = (int)((long long)std::floor(z) & 255);
    x -= std::floor(x);
    y -= std::floor(y);
    z -= std::floor(z);
    double u = fade(x), v = fade(y), w = fade(z);
    // u8 arithmetic in the reference wraps; the table has period 256 so masking is equivalent.
    int A = perm(X) + Y, AA = perm(A) + Z, AB = perm(A + 1) + Z;
    int B = perm(X + 1) + Y, BA = perm(B) + Z, BB = perm(B + 1) + Z;
    return lerp(w,
                lerp(v, lerp(u, grad(perm(AA), x, y, z), grad(perm(BA), x - 1, y, z)),
                     lerp(u, grad(perm(AB), x, y - 1, z), grad(perm(BB), x - 1, y - 1, z))),
                lerp(v, lerp(u, grad(perm(AA + 1), x, y, z - 1), grad(perm(BA + 1), x - 1, y, z - 1)),
                     lerp(u, grad(perm(AB + 1), x, y - 1, z - 1), grad(perm(BB + 1), x - 1, y - 1, z - 1))));
}

double octave_noise(double x, double y, double z, int octaves, double persistence) {  // noise.zig:35-50
    double total = 0, frequency = 1, amplitude = 1, max_value = 0;
    for (int i = 0; i < octaves; i++) {
        total += noise(x * frequency, y * frequency, z * frequency) * amplitude;
        max_value += amplitude;
        amplitude *= persistence;
        frequency *= 2.0;
    }
    return total / max_value;
}

// ---- texture_map.zig ---------------------------------------------------------------------------
Color uv_pattern_at(const UvPattern& uv, double u, double v, const Tuple& op) {
    switch (uv.kind) {
        case UvPattern::Test: return {u, v, 0.0};  // texture_map.zig:9-19
        case UvPattern::AlignCheck:                // texture_map.zig:21-43
            if (v > 0.8) {
                if (u < 0.2) return pattern_at(*uv.sub[1], op);
                if (u > 0.8) return pattern_at(*uv.sub[2], op);
            } else if (v < 0.2) {
                if (u < 0.2) return pattern_at(*uv.sub[3], op);
                if (u > 0.8) return pattern_at(*uv.sub[4], op);
            }
            return pattern_at(*uv.sub[0], op);
        case UvPattern::Checkers: {  // texture_map.zig:45-63
            double ua = std::floor(u * uv.width), va = std::floor(v * uv.height);
            return floored_mod(ua + va, 2.0) < 1.0 ? pattern_at(*uv.sub[0], op) : pattern_at(*uv.sub[1], op);
        }
        case UvPattern::ImageMap: {  // texture_map.zig:65-105 : no wrap, no half-texel offset
            const Image& im = *uv.image;
            double x = u * (double)(im.width - 1);
            double y = (1.0 - v) * (double)(im.height - 1);
            auto px = [&](double fx, double fy) {
                long ix = (long)fx, iy = (long)fy;
                if (ix < 0 || iy < 0 || ix >= im.width || iy >= im.height)
                    throw std::runtime_error("texture lookup out of bounds (the reference would panic)");
                return im.texel((int)ix, (int)iy);
            };
            if (!uv.bilinear) return px(std::round(x), std::round(y));
            double x1 = std::floor(x), x2 = std::ceil(x), y1 = std::floor(y), y2 = std::ceil(y);
            Color c11 = px(x1, y1), c21 = px(x2, y1), c12 = px(x1, y2), c22 = px(x2, y2);
            Color cx1 = c11.mul(x2 - x).add(c21.mul(x - x1));
            Color cx2 = c12.mul(x2 - x).add(c22.mul(x - x1));
            return cx1.mul(y2 - y).add(cx2.mul(y - y1));
        }
    }
    return {};
}

static Color texture_map_at(const Pattern& p, const Tuple& pp, const Tuple& op) {
    const double pi = 3.14159265358979323846;
    switch (p.mapping) {
        case Pattern::Spherical: {  // texture_map.zig:171-190
            double theta = std::atan2(pp.x, pp.z);
            double radius = Tuple::vec3(pp.x, pp.y, pp.z).magnitude();
            double phi = std::acos(pp.y / radius);
            double raw_u = theta / (2.0 * pi);
            double u = 1 - (raw_u + 0.5);
            double v = 1 - phi / pi;
            return uv_pattern_at(p.uv[0], u, v, op);
        }
        case Pattern::Planar:  // texture_map.zig:192-200
            return uv_pattern_at(p.uv[0], floored_mod(pp.x, 1.0), floored_mod(pp.z, 1.0), op);
        case Pattern::Cylindrical: {  // texture_map.zig:202-213
            double theta = std::atan2(pp.x, pp.z);
            double raw_u = theta / (2.0 * pi);
            double u = 1.0 - (raw_u + 0.5);
            return uv_pattern_at(p.uv[0], u, floored_mod(pp.y, 1.0), op);
        }
        case Pattern::Cubic: {  // texture_map.zig:215-290 ; face order front,back,left,right,up,down
            double ax = std::fabs(pp.x), ay = std::fabs(pp.y), az = std::fabs(pp.z);
            double coord = std::fmax(ax, std::fmax(ay, az));
            int face;
            if (coord == pp.x) face = 3;        // right
            else if (coord == -pp.x) face = 2;  // left
            else if (coord == pp.y) face = 4;   // up
            else if (coord == -pp.y) face = 5;  // down
            else if (coord == pp.z) face = 0;   // front
            else face = 1;                      // back
            double u, v;
            auto m2 = [](double a) { return floored_mod(a, 2.0) / 2.0; };
            switch (face) {
                case 0: u = m2(pp.x + 1.0); v = m2(pp.y + 1.0); break;
                case 1: u = m2(1.0 - pp.x); v = m2(pp.y + 1.0); break;
                case 2: u = m2(pp.z + 1.0); v = m2(pp.y + 1.0); break;
                case 3: u = m2(1.0 - pp.z); v = m2(pp.y + 1.0); break;
                case 4: u = m2(pp.x + 1.0); v = m2(1.0 - pp.z); break;
                default: u = m2(pp.x + 1.0); v = m2(pp.z + 1.0); break;
            }
            return uv_pattern_at(p.uv[face], u, v, op);
        }
    }
    return {};
}

// pattern.zig:112-124 : pattern point = own inverse * OBJECT point; children restart from op.
Color pattern_at(const Pattern& p, const Tuple& op) {
    Tuple pp = p.inverse.tuple_mul(op);
    switch (p.kind) {
        case Pattern::Test: return {pp.x, pp.y, pp.z};
        case Pattern::Solid: return p.color;
        case Pattern::Blend: return pattern_at(*p.a, op).add(pattern_at(*p.b, op)).mul(0.5);  // blend.zig:24-27
        case Pattern::Perturb: {  // perturb.zig:31-47
            Tuple offset = Tuple::vec3(
                octave_noise(op.x, op.y, op.z, p.perturb_octaves, p.perturb_persistence),
                octave_noise(op.x, op.y, op.z + 1.0, p.perturb_octaves, p.perturb_persistence),
                octave_noise(op.x, op.y, op.z + 2.0, p.perturb_octaves, p.perturb_persistence));
            return pattern_at(*p.a, op.add(offset.mul(p.perturb_scale)));
        }
        case Pattern::Stripes:  // stripes.zig:27-33
            return floored_mod(pp.x, 2.0) < 1.0 ? pattern_at(*p.a, op) : pattern_at(*p.b, op);
        case Pattern::Gradient: {  // gradient.zig:27-32
            double f = pp.x - std::floor(pp.x);
            Color a = pattern_at(*p.a, op), b = pattern_at(*p.b, op);
            return a.add(b.sub(a).mul(f));
        }
        case Pattern::RadialGradient: {  // gradient.zig:49-55
            double mag = std::sqrt(pp.x * pp.x + pp.z * pp.z);
            double f = mag - std::floor(mag);
            Color a = pattern_at(*p.a, op), b = pattern_at(*p.b, op);
            return a.add(b.sub(a).mul(f));
        }
        case Pattern::Rings:  // rings.zig:27-33
            return floored_mod(std::floor(std::sqrt(pp.x * pp.x + pp.z * pp.z)), 2.0) < 1.0 ? pattern_at(*p.a, op)
                                                                                             : pattern_at(*p.b, op);
        case Pattern::Checkers:  // checkers.zig:23-29
            return floored_mod(std::floor(pp.x) + std::floor(pp.y) + std::floor(pp.z), 2.0) < 1.0
                       ? pattern_at(*p.a, op)
                       : pattern_at(*p.b, op);
        case Pattern::TextureMap: return texture_map_at(p, pp, op);
    }
    return {};
}

Color pattern_at_shape(const Pattern& p, const Shape& s, const Tuple& world_point) {  // pattern.zig:128-131
    return pattern_at(p, s.world_to_object(world_point));
}

// material.zig:40-74
Color lighting(const Material& m, const Light& light, const Shape& object, const Tuple& point,
               const Tuple& eyev, const Tuple& normal, bool in_shadow) {
    Color color = pattern_at_shape(*m.pattern, object, point);
    Color effective = color.hadamard(light.intensity);
    Tuple lightv = light.position.sub(point).normalized();
    Color ambient = effective.mul(m.ambient);
    if (in_shadow) return ambient;
    Color diffuse{0, 0, 0}, specular{0, 0, 0};
    double light_dot_normal = lightv.dot(normal);
    if (light_dot_normal >= 0.0) {
        diffuse = effective.mul(m.diffuse * light_dot_normal);
        Tuple reflectv = lightv.reflect(normal);
        double reflect_dot_eye = reflectv.negate().dot(eyev);
        if (reflect_dot_eye > 0.0) specular = light.intensity.mul(m.specular * zig_pow(reflect_dot_eye, m.shininess));
    }
    return ambient.add(diffuse).add(specular);
}

}  // namespace oracle
