// Device-side f64 vector helpers and the per-leaf closed-form intersection / normal routines.
// Each routine cites the reference function it computes the same values as, in the reference's operation
// order; the library is compiled with -fmad=false (Makefile), so a*b+c stays an IEEE multiply and an IEEE add as
// in the Zig build and the results are bit-identical to the oracle's.
#pragma once
#include <cuda_runtime.h>

#include "device_scene.h"

namespace rtc {

struct V3 {
    double x, y, z;
};
__device__ __forceinline__ V3 v3(double x, double y, double z) { return {x, y, z}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 operator*(V3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ V3 neg(V3 a) { return {-a.x, -a.y, -a.z}; }
__device__ __forceinline__ double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 cross(V3 a, V3 b) {  // tuple.zig:127-133
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
__device__ __forceinline__ V3 normalized(V3 a) {  // tuple.zig:118-123 (w = 0 for every vector here)
    double m = sqrt(dot(a, a));
    return {a.x / m, a.y / m, a.z / m};
}

// 16-byte loads of read-only scene data
__device__ __forceinline__ double2 ld2(const double* p) { return __ldg(reinterpret_cast<const double2*>(p)); }

struct M34 {
    double m[12];
};
__device__ __forceinline__ M34 load_m34(const double* p) {
    M34 r;
#pragma unroll
    for (int i = 0; i < 6; i++) {
        double2 v = ld2(p + 2 * i);
        r.m[2 * i] = v.x;
        r.m[2 * i + 1] = v.y;
    }
    return r;
}
// Matrix.tupleMul on a point (w = 1) / a vector (w = 0)  (matrix.zig:124-140, ray.zig:30-32)
__device__ __forceinline__ V3 xf_point(const M34& a, V3 p) {
    return {a.m[0] * p.x + a.m[1] * p.y + a.m[2] * p.z + a.m[3], a.m[4] * p.x + a.m[5] * p.y + a.m[6] * p.z + a.m[7],
            a.m[8] * p.x + a.m[9] * p.y + a.m[10] * p.z + a.m[11]};
}
__device__ __forceinline__ V3 xf_vec(const M34& a, V3 v) {
    return {a.m[0] * v.x + a.m[1] * v.y + a.m[2] * v.z, a.m[4] * v.x + a.m[5] * v.y + a.m[6] * v.z,
            a.m[8] * v.x + a.m[9] * v.y + a.m[10] * v.z};
}
// Shape.normalToWorld (shape.zig:139-145): inverse-transpose * n, w forced to 0, normalised.
__device__ __forceinline__ V3 xf_normal(const M34& a, V3 n) {
    V3 r = {a.m[0] * n.x + a.m[4] * n.y + a.m[8] * n.z, a.m[1] * n.x + a.m[5] * n.y + a.m[9] * n.z,
            a.m[2] * n.x + a.m[6] * n.y + a.m[10] * n.z};
    return normalized(r);
}

// Zig @mod on floats: floored modulo (see host rtc_math.hpp floored_mod)
__device__ __forceinline__ double zmod(double a, double b) {
    double r = fmod(a, b);
    return (a < 0.0) ? fmod(r + b, b) : r;
}

#define RTC_INF (__longlong_as_double(0x7ff0000000000000LL))

// std.math.pow(f64) as Zig computes it -- device copy of host/rtc_math.hpp zig_pow (same operations in
// the same order; for integral y all of them are exact IEEE operations, so the bits match the oracle).
// Replaces CUDA's general pow(): ~40 instructions on the hot path instead of ~300.
// Special cases that cannot occur at the two call sites (x is a finite number >= 0) are folded.
__device__ __forceinline__ double zig_pow(double x, double y) {
    if (y == 0.0 || x == 1.0) return 1.0;
    if (isnan(x) || isnan(y)) return __longlong_as_double(0x7ff8000000000000LL);
    if (y == 1.0) return x;
    if (x == 0.0) return y < 0.0 ? RTC_INF : 0.0;
    if (isinf(y)) return ((fabs(x) < 1.0) == (y > 0.0)) ? 0.0 : RTC_INF;
    if (isinf(x)) return y < 0.0 ? 0.0 : RTC_INF;
    if (y == 0.5) return sqrt(x);
    if (y == -0.5) return 1.0 / sqrt(x);
    double yi = trunc(fabs(y));
    double yf = fabs(y) - yi;
    if (yf != 0.0 && x < 0.0) return __longlong_as_double(0x7ff8000000000000LL);
    if (yi >= 9.223372036854775808e18) return exp(y * log(x));
    double a1 = 1.0;
    int ae = 0;
    if (yf != 0.0) {
        if (yf > 0.5) {
            yf -= 1.0;
            yi += 1.0;
        }
        a1 = exp(yf * log(x));
    }
    int xe;
    double x1 = frexp(x, &xe);
#pragma unroll 1
    for (long long i = (long long)yi; i != 0; i >>= 1) {
        if (xe < -4096 || 4096 < xe) {
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
    return scalbn(a1, ae);
}

// cube.zig:24-48 == bounding_box.zig:112-136
__device__ __forceinline__ void check_axis(double origin, double direction, double mn, double mx, double& tmin,
                                           double& tmax) {
    double a = mn - origin, b = mx - origin;
    if (fabs(direction) >= 1e-5) {
        a = a / direction;
        b = b / direction;
    } else {
        a = a * RTC_INF;
        b = b * RTC_INF;
    }
    if (a > b) {
        double s = a;
        a = b;
        b = s;
    }
    tmin = a;
    tmax = b;
}
// The reference slab test (cube.zig:50-79, bounding_box.zig:138-165).  fmax/fmin drop NaN like @max/@min.
__device__ __forceinline__ bool ref_slabs(V3 o, V3 d, const double mn[3], const double mx[3], double& tmin,
                                          double& tmax) {
    double xa, xb, ya, yb, za, zb;
    check_axis(o.x, d.x, mn[0], mx[0], xa, xb);
    check_axis(o.y, d.y, mn[1], mx[1], ya, yb);
    check_axis(o.z, d.z, mn[2], mx[2], za, zb);
    tmin = fmax(xa, fmax(ya, za));
    tmax = fmin(xb, fmin(yb, zb));
    return !(tmin > tmax);
}

// Möller–Trumbore exactly as triangle.zig:29-63 / :225-259.
__device__ __forceinline__ bool tri_intersect(V3 p1, V3 e1, V3 e2, V3 o, V3 d, double& t, double& u, double& v) {
    V3 dir_cross_e2 = cross(d, e2);
    double det = dot(e1, dir_cross_e2);
    if (fabs(det) < 1e-5) return false;
    double f = 1.0 / det;
    V3 p1_to_origin = o - p1;
    u = f * dot(p1_to_origin, dir_cross_e2);
    if (u < 0.0 || u > 1.0) return false;
    V3 origin_cross_e1 = cross(p1_to_origin, e1);
    v = f * dot(d, origin_cross_e1);
    if (v < 0.0 || (u + v) > 1.0) return false;
    t = f * dot(e2, origin_cross_e1);
    return true;
}

// Intersections of one analytic leaf with the OBJECT-space ray (o,d), in the reference's append
// order.  Returns the count (0..4).
__device__ __forceinline__ int leaf_intersect(int kind, uint32_t flags, double mn, double mx, V3 o, V3 d,
                                              double ts[4]) {
    int n = 0;
    switch (kind) {
        case RTC_SPHERE: {  // sphere.zig:24-46
            double a = dot(d, d);
            double b = 2.0 * dot(o, d);
            double c = dot(o, o) - 1.0;
            double disc = b * b - 4.0 * a * c;
            if (disc >= 0.0) {
                double s = sqrt(disc);
                ts[0] = (-b - s) / (2.0 * a);
                ts[1] = (-b + s) / (2.0 * a);
                n = 2;
            }
            break;
        }
        case RTC_PLANE: {  // plane.zig:25-36
            if (fabs(d.y) > 1e-5) {
                ts[0] = -o.y / d.y;
                n = 1;
            }
            break;
        }
        case RTC_CUBE: {  // cube.zig:50-79
            const double lo[3] = {-1.0, -1.0, -1.0}, hi[3] = {1.0, 1.0, 1.0};
            double t0, t1;
            if (ref_slabs(o, d, lo, hi, t0, t1)) {
                ts[0] = t0;
                ts[1] = t1;
                n = 2;
            }
            break;
        }
        case RTC_CYLINDER: {  // cylinder.zig:30-98
            double a = d.x * d.x + d.z * d.z;
            bool wall = !(fabs(a) < 1e-5);
            bool keep = true;
            if (wall) {
                double b = 2.0 * o.x * d.x + 2.0 * o.z * d.z;
                double c = o.x * o.x + o.z * o.z - 1.0;
                double disc = b * b - 4.0 * a * c;
                if (disc < 0.0) {
                    keep = false;  // "Ray does not intersect": returns before the caps
                } else {
                    double s = sqrt(disc);
                    double t0 = (-b - s) / (2.0 * a), t1 = (-b + s) / (2.0 * a);
                    if (t0 > t1) {
                        double w = t0;
                        t0 = t1;
                        t1 = w;
                    }
                    double y0 = o.y + t0 * d.y;
                    if (mn < y0 && y0 < mx) ts[n++] = t0;
                    double y1 = o.y + t1 * d.y;
                    if (mn < y1 && y1 < mx) ts[n++] = t1;
                }
            }
            if (keep && (flags & RTC_CLOSED) && !(fabs(d.y) < 1e-5)) {  // intersectCaps :37-51
                double t = (mn - o.y) / d.y;
                double x = o.x + t * d.x, z = o.z + t * d.z;
                if (x * x + z * z <= 1.0) ts[n++] = t;
                t = (mx - o.y) / d.y;
                x = o.x + t * d.x, z = o.z + t * d.z;
                if (x * x + z * z <= 1.0) ts[n++] = t;
            }
            break;
        }
        case RTC_CONE: {  // cone.zig:30-113, tolerance 1e-4
            double a = d.x * d.x - d.y * d.y + d.z * d.z;
            double b = 2.0 * o.x * d.x - 2.0 * o.y * d.y + 2.0 * o.z * d.z;
            bool caps = true;
            if (!(fabs(a) < 1e-4 && fabs(b) < 1e-4)) {
                double c = o.x * o.x - o.y * o.y + o.z * o.z;
                if (fabs(a) < 1e-4) {
                    ts[n++] = -c / (2.0 * b);
                } else {
                    double disc = b * b - 4.0 * a * c;
                    if (disc < 0.0) {
                        caps = false;
                    } else {
                        double s = sqrt(disc);
                        double t0 = (-b - s) / (2.0 * a), t1 = (-b + s) / (2.0 * a);
                        if (t0 > t1) {
                            double w = t0;
                            t0 = t1;
                            t1 = w;
                        }
                        double y0 = o.y + t0 * d.y;
                        if (mn < y0 && y0 < mx) ts[n++] = t0;
                        double y1 = o.y + t1 * d.y;
                        if (mn < y1 && y1 < mx) ts[n++] = t1;
                    }
                }
            }
            if (caps && (flags & RTC_CLOSED) && !(fabs(d.y) < 1e-4)) {  // intersectCaps :37-51
                double t = (mn - o.y) / d.y;
                double x = o.x + t * d.x, z = o.z + t * d.z;
                if (x * x + z * z <= mn * mn) ts[n++] = t;
                t = (mx - o.y) / d.y;
                x = o.x + t * d.x, z = o.z + t * d.z;
                if (x * x + z * z <= mx * mx) ts[n++] = t;
            }
            break;
        }
        default: break;  // triangles are handled by the callers; test shapes never intersect
    }
    return n;
}

// localNormalAt of every leaf kind (object space, un-normalised like the reference).
__device__ __forceinline__ V3 leaf_local_normal(int kind, double mn, double mx, V3 p, const double* nrm, double u,
                                                double v) {
    switch (kind) {
        case RTC_SPHERE: return p;                       // sphere.zig:48-53
        case RTC_PLANE: return v3(0.0, 1.0, 0.0);        // plane.zig:38-43
        case RTC_CUBE: {                                 // cube.zig:81-97
            double ax = fabs(p.x), ay = fabs(p.y), az = fabs(p.z);
            double m = fmax(ax, fmax(ay, az));
            if (m == ax) return v3(p.x, 0.0, 0.0);
            if (m == ay) return v3(0.0, p.y, 0.0);
            return v3(0.0, 0.0, p.z);
        }
        case RTC_CYLINDER: {  // cylinder.zig:100-112
            double dist = p.x * p.x + p.z * p.z;
            if (dist < 1.0 && p.y >= mx - 1e-5) return v3(0.0, 1.0, 0.0);
            if (dist < 1.0 && p.y <= mn + 1e-5) return v3(0.0, -1.0, 0.0);
            return v3(p.x, 0.0, p.z);
        }
        case RTC_CONE: {  // cone.zig:115-127
            double dist = p.x * p.x + p.z * p.z;
            if (dist < mx * mx && p.y >= mx - 1e-4) return v3(0.0, 1.0, 0.0);
            if (dist < mn * mn && p.y <= mn + 1e-4) return v3(0.0, -1.0, 0.0);
            double sgn = p.y > 0.0 ? 1.0 : (p.y < 0.0 ? -1.0 : 0.0);
            return v3(p.x, -sgn * sqrt(dist), p.z);
        }
        case RTC_TRIANGLE: return v3(__ldg(nrm), __ldg(nrm + 1), __ldg(nrm + 2));  // triangle.zig:65-70
        case RTC_SMOOTH_TRIANGLE: {                                                // triangle.zig:261-265
            double w = 1.0 - u - v;
            return v3(__ldg(nrm + 3) * u + __ldg(nrm + 6) * v + __ldg(nrm + 0) * w,
                      __ldg(nrm + 4) * u + __ldg(nrm + 7) * v + __ldg(nrm + 1) * w,
                      __ldg(nrm + 5) * u + __ldg(nrm + 8) * v + __ldg(nrm + 2) * w);
        }
        default: return v3(0.0, 0.0, 0.0);
    }
}

}  // namespace rtc
