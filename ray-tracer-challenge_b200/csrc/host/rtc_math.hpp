// Host-side f64 math used by the scene loader, the flattener and (read-only) by the oracle.
//
// Mirrors the arithmetic of the reference's L0 layer so that matrices baked into the
// device scene are the ones the reference would have computed:
//   Tuple   -> src/raytracer/tuple.zig:11-141   (dot / magnitude INCLUDE w)
//   Matrix  -> src/raytracer/matrix.zig:9-325   (cofactor inverse, |det| < 1e-5 rejects)
//   Color   -> src/raytracer/color.zig:7-71     (clamp = round-half-away(c*255) in [0,255])
// The shipped reference binary renders in f64 (src/main.zig:71), so only double exists here.
#pragma once
#include <cmath>
#include <limits>
#include <cstdint>
#include <stdexcept>

namespace rtc {

struct Tuple {
    double x = 0, y = 0, z = 0, w = 0;
    static Tuple point(double x, double y, double z) { return {x, y, z, 1.0}; }
    static Tuple vec3(double x, double y, double z) { return {x, y, z, 0.0}; }
    Tuple add(const Tuple& o) const { return {x + o.x, y + o.y, z + o.z, w + o.w}; }
    Tuple sub(const Tuple& o) const { return {x - o.x, y - o.y, z - o.z, w - o.w}; }
    Tuple negate() const { return {-x, -y, -z, -w}; }
    Tuple mul(double v) const { return {x * v, y * v, z * v, w * v}; }
    Tuple div(double v) const { return {x / v, y / v, z / v, w / v}; }
    double dot(const Tuple& o) const { return x * o.x + y * o.y + z * o.z + w * o.w; }
    double magnitude() const { return std::sqrt(x * x + y * y + z * z + w * w); }
    Tuple normalized() const {
        double m = magnitude();
        return m == 0.0 ? *this : div(m);
    }
    // left-handed cross product, w dropped (tuple.zig:127-133)
    Tuple cross(const Tuple& o) const {
        return vec3(y * o.z - z * o.y, z * o.x - x * o.z, x * o.y - y * o.x);
    }
    Tuple reflect(const Tuple& n) const { return sub(n.mul(2.0 * dot(n))); }
    bool approx(const Tuple& o, double tol = 1e-5) const {
        return std::fabs(x - o.x) < tol && std::fabs(y - o.y) < tol && std::fabs(z - o.z) < tol &&
               std::fabs(w - o.w) < tol;
    }
};

struct Color {
    double r = 0, g = 0, b = 0;
    Color add(const Color& o) const { return {r + o.r, g + o.g, b + o.b}; }
    Color sub(const Color& o) const { return {r - o.r, g - o.g, b - o.b}; }
    Color mul(double v) const { return {r * v, g * v, b * v}; }
    Color hadamard(const Color& o) const { return {r * o.r, g * o.g, b * o.b}; }
    bool approx(const Color& o, double tol = 1e-5) const {
        return std::fabs(r - o.r) < tol && std::fabs(g - o.g) < tol && std::fabs(b - o.b) < tol;
    }
};

// color.zig:61-71 : @round is half-away-from-zero, same as std::round.
inline uint8_t clamp_channel(double c) {
    double t = std::round(c * 255.0);
    if (!(t > 0.0)) return 0;  // negatives and NaN -> 0
    if (t > 255.0) return 255;
    return (uint8_t)t;
}

struct NotInvertible : std::runtime_error {
    NotInvertible() : std::runtime_error("NotInvertible") {}
};

template <int N>
struct MatN {
    double d[N][N];
    double det() const;
    MatN<N - 1> submatrix(int row, int col) const {
        MatN<N - 1> s;
        for (int r = 0; r < N; r++) {
            if (r == row) continue;
            for (int c = 0; c < N; c++) {
                if (c == col) continue;
                s.d[r - (r > row)][c - (c > col)] = d[r][c];
            }
        }
        return s;
    }
    double minor_(int row, int col) const { return submatrix(row, col).det(); }
    double cofactor(int row, int col) const {
        return ((row + col) % 2 == 0) ? minor_(row, col) : -minor_(row, col);
    }
};
template <>
struct MatN<1> {
    double d[1][1];
    double det() const { return d[0][0]; }
};
template <int N>
inline double MatN<N>::det() const {
    if (N == 2) return d[0][0] * d[1][1] - d[0][1] * d[1][0];
    double s = 0.0;
    for (int c = 0; c < N; c++) s += d[0][c] * cofactor(0, c);
    return s;
}

struct Mat4 : MatN<4> {
    static Mat4 zero() {
        Mat4 m;
        for (auto& r : m.d)
            for (auto& v : r) v = 0.0;
        return m;
    }
    static Mat4 identity() {
        Mat4 m = zero();
        for (int i = 0; i < 4; i++) m.d[i][i] = 1.0;
        return m;
    }
    static Mat4 rows(const double (&a)[4][4]) {
        Mat4 m;
        for (int i = 0; i < 4; i++)
            for (int j = 0; j < 4; j++) m.d[i][j] = a[i][j];
        return m;
    }
    Mat4 mul(const Mat4& o) const {  // matrix.zig:106-120
        Mat4 r;
        for (int i = 0; i < 4; i++)
            for (int j = 0; j < 4; j++) {
                double s = 0;
                for (int k = 0; k < 4; k++) s += d[i][k] * o.d[k][j];
                r.d[i][j] = s;
            }
        return r;
    }
    Mat4 add(const Mat4& o) const {
        Mat4 r;
        for (int i = 0; i < 4; i++)
            for (int j = 0; j < 4; j++) r.d[i][j] = d[i][j] + o.d[i][j];
        return r;
    }
    Mat4 scalar_mul(double v) const {
        Mat4 r;
        for (int i = 0; i < 4; i++)
            for (int j = 0; j < 4; j++) r.d[i][j] = d[i][j] * v;
        return r;
    }
    Tuple tuple_mul(const Tuple& t) const {  // matrix.zig:124-140 : row.dot(tup)
        auto row = [&](int i) { return d[i][0] * t.x + d[i][1] * t.y + d[i][2] * t.z + d[i][3] * t.w; };
        return {row(0), row(1), row(2), row(3)};
    }
    Mat4 transpose() const {
        Mat4 r;
        for (int i = 0; i < 4; i++)
            for (int j = 0; j < 4; j++) r.d[i][j] = d[j][i];
        return r;
    }
    Mat4 inverse() const {  // matrix.zig:202-217
        double dt = det();
        if (std::fabs(dt) < 1e-5) throw NotInvertible();
        Mat4 inv;
        for (int r = 0; r < 4; r++)
            for (int c = 0; c < 4; c++) inv.d[c][r] = cofactor(r, c) / dt;
        return inv;
    }
    // Fluent builders: each returns  (new op) * self   (matrix.zig:222-325)
    Mat4 translate(double x, double y, double z) const {
        return rows({{1, 0, 0, x}, {0, 1, 0, y}, {0, 0, 1, z}, {0, 0, 0, 1}}).mul(*this);
    }
    Mat4 scale(double x, double y, double z) const {
        return rows({{x, 0, 0, 0}, {0, y, 0, 0}, {0, 0, z, 0}, {0, 0, 0, 1}}).mul(*this);
    }
    Mat4 rotate_x(double a) const {
        double c = std::cos(a), s = std::sin(a);
        return rows({{1, 0, 0, 0}, {0, c, -s, 0}, {0, s, c, 0}, {0, 0, 0, 1}}).mul(*this);
    }
    Mat4 rotate_y(double a) const {
        double c = std::cos(a), s = std::sin(a);
        return rows({{c, 0, s, 0}, {0, 1, 0, 0}, {-s, 0, c, 0}, {0, 0, 0, 1}}).mul(*this);
    }
    Mat4 rotate_z(double a) const {
        double c = std::cos(a), s = std::sin(a);
        return rows({{c, -s, 0, 0}, {s, c, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}}).mul(*this);
    }
    // Rodrigues rotation about a unit axis (matrix.zig:292-303), used by lib.zig's camera orbit.
    Mat4 rotate(const Tuple& axis, double a) const {
        Mat4 C = rows({{0, -axis.z, axis.y, 0}, {axis.z, 0, -axis.x, 0}, {-axis.y, axis.x, 0, 0}, {0, 0, 0, 0}});
        Mat4 R = identity().add(C.scalar_mul(std::sin(a))).add(C.mul(C).scalar_mul(1.0 - std::cos(a)));
        R.d[3][3] = 1.0;
        return R.mul(*this);
    }
    Mat4 shear(double xy, double xz, double yx, double yz, double zx, double zy) const {
        return rows({{1, xy, xz, 0}, {yx, 1, yz, 0}, {zx, zy, 1, 0}, {0, 0, 0, 1}}).mul(*this);
    }
    static Mat4 view_transform(const Tuple& from, const Tuple& to, const Tuple& up) {  // matrix.zig:54-67
        Tuple forward = to.sub(from).normalized();
        Tuple left = forward.cross(up.normalized());
        Tuple true_up = left.cross(forward);
        Mat4 o = rows({{left.x, left.y, left.z, 0},
                       {true_up.x, true_up.y, true_up.z, 0},
                       {-forward.x, -forward.y, -forward.z, 0},
                       {0, 0, 0, 1}});
        return o.mul(identity().translate(-from.x, -from.y, -from.z));
    }
    bool approx(const Mat4& o, double tol = 1e-5) const {
        for (int i = 0; i < 4; i++)
            for (int j = 0; j < 4; j++)
                if (std::fabs(d[i][j] - o.d[i][j]) > tol) return false;
        return true;
    }
};

// Zig's @mod for floats: floored modulo, result has the sign of the divisor.
// Restated the way Zig lowers it for floats (fmod; for a negative numerator fmod(rem + b, b)),
// so that rem + b rounding up to b folds back to 0 exactly as in the reference binary.
inline double floored_mod(double a, double b) {
    double r = std::fmod(a, b);
    return (a < 0.0) ? std::fmod(r + b, b) : r;
}

// std.math.pow for f64 as the Zig standard library computes it (lib/std/math/pow.zig, a port of
// Go's math.Pow; Zig 0.12-dev per build.zig.zon -- the std lib is not vendored in the reference, so
// this is a restatement of the published algorithm): special cases, then x^yi by squaring on the
// frexp mantissa with the exponent carried separately, times exp(yf*log(x)) for a fractional part.
// Call sites: material.zig:69 (specular, y = shininess) and world.zig:288 (schlick, y = 5).
// For integral y every operation is an exact IEEE multiply / frexp / scalbn, so the device copy
// (rt_math.cuh zig_pow) returns the same bits.
inline double zig_pow(double x, double y) {
    if (y == 0.0 || x == 1.0) return 1.0;
    if (std::isnan(x) || std::isnan(y)) return std::numeric_limits<double>::quiet_NaN();
    if (y == 1.0) return x;
    if (x == 0.0) {
        bool odd = std::fabs(y) < 9007199254740992.0 && std::fmod(std::fabs(y), 2.0) == 1.0;
        if (y < 0.0) return odd ? std::copysign(INFINITY, x) : INFINITY;
        return odd ? x : 0.0;
    }
    if (std::isinf(y)) {
        if (x == -1.0) return 1.0;
        if ((std::fabs(x) < 1.0) == (y > 0.0)) return 0.0;
        return INFINITY;
    }
    if (std::isinf(x)) {
        if (x < 0.0) {
            bool odd = std::fabs(y) < 9007199254740992.0 && std::fmod(std::fabs(y), 2.0) == 1.0;
            if (y < 0.0) return odd ? -0.0 : 0.0;
            return odd ? -INFINITY : INFINITY;
        }
        return y < 0.0 ? 0.0 : INFINITY;
    }
    if (y == 0.5) return std::sqrt(x);
    if (y == -0.5) return 1.0 / std::sqrt(x);
    double yi = std::trunc(std::fabs(y));
    double yf = std::fabs(y) - yi;
    if (yf != 0.0 && x < 0.0) return std::numeric_limits<double>::quiet_NaN();
    if (yi >= 9.223372036854775808e18) return std::exp(y * std::log(x));
    double a1 = 1.0;
    int ae = 0;
    if (yf != 0.0) {
        if (yf > 0.5) {
            yf -= 1.0;
            yi += 1.0;
        }
        a1 = std::exp(yf * std::log(x));
    }
    int xe;
    double x1 = std::frexp(x, &xe);
    for (long long i = (long long)yi; i != 0; i >>= 1) {
        if (xe < -4096 || 4096 < xe) {  // catastrophic over/underflow: the result is 0 or inf either way
            ae += xe;
            break;
        }
        if (i & 1) {
            a1 *= x1;
            ae += xe;
        }
        x1 *= x1;
        xe <<= 1;
        if (x1 < 0.5) {
            x1 += x1;
            xe -= 1;
        }
    }
    if (y < 0.0) {
        a1 = 1.0 / a1;
        ae = -ae;
    }
    return std::scalbn(a1, ae);
}

}  // namespace rtc
