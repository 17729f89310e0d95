// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.hpp).  Shape kernels of the reference, restated.
#include <algorithm>

#include "oracle.hpp"

namespace oracle {

// shape.zig:64-66 : stable ascending sort by t (std.mem.sort is stable).  Applied to the sub-range
// a shape appended, which is what sorting its own freshly built list amounts to.
void sort_intersections(Intersections& xs, size_t from) {
    size_t n = xs.size() - from;
    if (n < 2) return;
    if (n <= 24) {  // insertion sort: stable, no allocation
        for (size_t i = from + 1; i < xs.size(); i++) {
            Intersection key = xs[i];
            size_t j = i;
            while (j > from && key.t < xs[j - 1].t) {
                xs[j] = xs[j - 1];
                j--;
            }
            xs[j] = key;
        }
    } else {
        std::stable_sort(xs.begin() + (long)from, xs.end(),
                         [](const Intersection& a, const Intersection& b) { return a.t < b.t; });
    }
}

// shape.zig:71-80 : first intersection with t >= 0 in a sorted list
long hit_index(const Intersections& xs, size_t from) {
    for (size_t i = from; i < xs.size(); i++)
        if (xs[i].t >= 0.0) return (long)i;
    return -1;
}

// cube.zig:24-48 and bounding_box.zig:112-136 share this slab rule.
static inline void check_axis(double origin, double direction, double mn, double mx, double& tmin, double& tmax) {
    double tmin_num = mn - origin, tmax_num = mx - origin;
    if (std::fabs(direction) >= 1e-5) {
        tmin = tmin_num / direction;
        tmax = tmax_num / direction;
    } else {
        tmin = tmin_num * kInf;
        tmax = tmax_num * kInf;
    }
    if (tmin > tmax) std::swap(tmin, tmax);
}

static inline bool slabs(const Tuple& mn, const Tuple& mx, const Ray& r, double& tmin, double& tmax) {
    double xa, xb, ya, yb, za, zb;
    check_axis(r.origin.x, r.direction.x, mn.x, mx.x, xa, xb);
    check_axis(r.origin.y, r.direction.y, mn.y, mx.y, ya, yb);
    check_axis(r.origin.z, r.direction.z, mn.z, mx.z, za, zb);
    tmin = std::fmax(xa, std::fmax(ya, za));  // @max/@min: non-NaN operand wins
    tmax = std::fmin(xb, std::fmin(yb, zb));
    return !(tmin > tmax);
}

// bounding_box.zig:138-165 (the box's own transform is always identity)
bool bbox_hit(const BBox& b, const Ray& ray) {
    double t0, t1;
    return slabs(b.min, b.max, ray, t0, t1);
}
void bbox_intersect(const BBox& b, const Ray& ray, Intersections& out, const Shape* tag) {
    double t0, t1;
    if (!slabs(b.min, b.max, ray, t0, t1)) return;
    out.push_back({t0, tag});
    out.push_back({t1, tag});
}

// sphere.zig:24-46
static void sphere_intersect(const Shape& s, const Ray& ray, Intersections& out) {
    Tuple sphere_to_ray = ray.origin.sub(Tuple::point(0, 0, 0));
    double a = ray.direction.dot(ray.direction);
    double b = 2.0 * sphere_to_ray.dot(ray.direction);
    double c = sphere_to_ray.dot(sphere_to_ray) - 1.0;
    double disc = b * b - 4.0 * a * c;
    if (disc >= 0.0) {
        size_t from = out.size();
        out.push_back({(-b - std::sqrt(disc)) / (2.0 * a), &s});
        out.push_back({(-b + std::sqrt(disc)) / (2.0 * a), &s});
        sort_intersections(out, from);
    }
}

// plane.zig:25-36
static void plane_intersect(const Shape& s, const Ray& ray, Intersections& out) {
    if (std::fabs(ray.direction.y) > 1e-5) out.push_back({-ray.origin.y / ray.direction.y, &s});
}

// cube.zig:50-79
static void cube_intersect(const Shape& s, const Ray& ray, Intersections& out) {
    double t0, t1;
    if (!slabs(Tuple::point(-1, -1, -1), Tuple::point(1, 1, 1), ray, t0, t1)) return;
    out.push_back({t0, &s});
    out.push_back({t1, &s});
}

// cylinder.zig:30-98
static void cylinder_caps(const Shape& s, const Ray& ray, Intersections& out) {
    if (!s.closed || std::fabs(ray.direction.y) < 1e-5) return;
    auto check = [&](double t) {
        double x = ray.origin.x + t * ray.direction.x, z = ray.origin.z + t * ray.direction.z;
        return x * x + z * z <= 1.0;
    };
    double t = (s.minimum - ray.origin.y) / ray.direction.y;
    if (check(t)) out.push_back({t, &s});
    t = (s.maximum - ray.origin.y) / ray.direction.y;
    if (check(t)) out.push_back({t, &s});
}
static void cylinder_intersect(const Shape& s, const Ray& ray, Intersections& out) {
    const Tuple &o = ray.origin, &d = ray.direction;
    double a = d.x * d.x + d.z * d.z;
    if (std::fabs(a) < 1e-5) {
        cylinder_caps(s, ray, out);
        return;
    }
    double b = 2.0 * o.x * d.x + 2.0 * o.z * d.z;
    double c = o.x * o.x + o.z * o.z - 1.0;
    double disc = b * b - 4.0 * a * c;
    if (disc < 0.0) return;
    double t0 = (-b - std::sqrt(disc)) / (2.0 * a), t1 = (-b + std::sqrt(disc)) / (2.0 * a);
    if (t0 > t1) std::swap(t0, t1);
    double y0 = o.y + t0 * d.y;
    if (s.minimum < y0 && y0 < s.maximum) out.push_back({t0, &s});
    double y1 = o.y + t1 * d.y;
    if (s.minimum < y1 && y1 < s.maximum) out.push_back({t1, &s});
    cylinder_caps(s, ray, out);
}

// cone.zig:30-113 (tolerance 1e-4)
static void cone_caps(const Shape& s, const Ray& ray, Intersections& out) {
    if (!s.closed || std::fabs(ray.direction.y) < 1e-4) return;
    auto check = [&](double t, double radius) {
        double x = ray.origin.x + t * ray.direction.x, z = ray.origin.z + t * ray.direction.z;
        return x * x + z * z <= radius * radius;
    };
    double t = (s.minimum - ray.origin.y) / ray.direction.y;
    if (check(t, s.minimum)) out.push_back({t, &s});
    t = (s.maximum - ray.origin.y) / ray.direction.y;
    if (check(t, s.maximum)) out.push_back({t, &s});
}
static void cone_intersect(const Shape& s, const Ray& ray, Intersections& out) {
    const Tuple &o = ray.origin, &d = ray.direction;
    double a = d.x * d.x - d.y * d.y + d.z * d.z;
    double b = 2.0 * o.x * d.x - 2.0 * o.y * d.y + 2.0 * o.z * d.z;
    if (std::fabs(a) < 1e-4 && std::fabs(b) < 1e-4) {
        cone_caps(s, ray, out);
        return;
    }
    double c = o.x * o.x - o.y * o.y + o.z * o.z;
    if (std::fabs(a) < 1e-4) {
        out.push_back({-c / (2.0 * b), &s});
        cone_caps(s, ray, out);
        return;
    }
    double disc = b * b - 4.0 * a * c;
    if (disc < 0.0) return;
    double t0 = (-b - std::sqrt(disc)) / (2.0 * a), t1 = (-b + std::sqrt(disc)) / (2.0 * a);
    if (t0 > t1) std::swap(t0, t1);
    double y0 = o.y + t0 * d.y;
    if (s.minimum < y0 && y0 < s.maximum) out.push_back({t0, &s});
    double y1 = o.y + t1 * d.y;
    if (s.minimum < y1 && y1 < s.maximum) out.push_back({t1, &s});
    cone_caps(s, ray, out);
}

// triangle.zig:29-63 and :225-259 (Möller–Trumbore; smooth triangles keep u,v)
static void triangle_intersect(const Shape& s, const Ray& ray, Intersections& out) {
    Tuple dir_cross_e2 = ray.direction.cross(s.e2);
    double det = s.e1.dot(dir_cross_e2);
    if (std::fabs(det) < 1e-5) return;
    double f = 1.0 / det;
    Tuple p1_to_origin = ray.origin.sub(s.p1);
    double u = f * p1_to_origin.dot(dir_cross_e2);
    if (u < 0.0 || u > 1.0) return;
    Tuple origin_cross_e1 = p1_to_origin.cross(s.e1);
    double v = f * ray.direction.dot(origin_cross_e1);
    if (v < 0.0 || (u + v) > 1.0) return;
    double t = f * s.e2.dot(origin_cross_e1);
    if (s.kind == ShapeKind::SmoothTriangle)
        out.push_back({t, &s, u, v});
    else
        out.push_back({t, &s});
}

// csg.zig:115-121
bool csg_intersection_allowed(CsgOp op, bool lhit, bool inl, bool inr) {
    switch (op) {
        case CsgOp::Union: return (lhit && !inr) || !(lhit || inl);
        case CsgOp::Intersection: return (lhit && inr) || (!lhit && inl);
        case CsgOp::Difference: return (lhit && !inr) || (!lhit && inl);
    }
    return false;
}

// csg.zig:123-140
bool csg_includes(const Shape& a, const Shape& b) {
    if (a.kind == ShapeKind::Group) {
        for (const Shape& c : a.children)
            if (csg_includes(c, b)) return true;
        return false;
    }
    if (a.kind == ShapeKind::Csg) return csg_includes(*a.left, b) || csg_includes(*a.right, b);
    return a.id == b.id;
}

// Csg.filterIntersections (csg.zig:51-73) on the sorted tail xs[from..]
void csg_filter(const Shape& csg, Intersections& xs, size_t from) {
    bool inl = false, inr = false;
    size_t w = from;
    for (size_t i = from; i < xs.size(); i++) {
        bool lhit = csg_includes(*csg.left, *xs[i].object);
        if (csg_intersection_allowed(csg.op, lhit, inl, inr)) xs[w++] = xs[i];
        if (lhit) inl = !inl;
        else inr = !inr;
    }
    xs.resize(w);
}

// shape.zig:313-335 : leaves see the ray in object space, groups / CSG pass it through.
void intersect(const Shape& s, const Ray& ray, Intersections& out) {
    switch (s.kind) {
        case ShapeKind::Group: {  // group.zig:39-62
            if (!bbox_hit(s.bbox, ray)) return;
            size_t from = out.size();
            for (const Shape& c : s.children) intersect(c, ray, out);
            sort_intersections(out, from);
            return;
        }
        case ShapeKind::Csg: {  // csg.zig:75-96 + filterIntersections :51-73
            if (!bbox_hit(s.bbox, ray)) return;
            size_t from = out.size();
            intersect(*s.left, ray, out);
            intersect(*s.right, ray, out);
            sort_intersections(out, from);
            csg_filter(s, out, from);
            return;
        }
        default: break;
    }
    Ray local = ray.transform(s.inverse);
    switch (s.kind) {
        case ShapeKind::Test: break;  // TestShape records the ray only (shape.zig:404-439)
        case ShapeKind::Sphere: sphere_intersect(s, local, out); break;
        case ShapeKind::Plane: plane_intersect(s, local, out); break;
        case ShapeKind::Cube: cube_intersect(s, local, out); break;
        case ShapeKind::Cylinder: cylinder_intersect(s, local, out); break;
        case ShapeKind::Cone: cone_intersect(s, local, out); break;
        case ShapeKind::Triangle:
        case ShapeKind::SmoothTriangle: triangle_intersect(s, local, out); break;
        default: break;
    }
}

static inline double zig_sign(double v) { return v > 0 ? 1.0 : (v < 0 ? -1.0 : 0.0); }

Tuple local_normal_at(const Shape& s, const Tuple& p, const Intersection& hit) {
    switch (s.kind) {
        case ShapeKind::Test: return Tuple::point(0, 0, 0);  // shape.zig:420-426
        case ShapeKind::Sphere: return p.sub(Tuple::point(0, 0, 0));  // sphere.zig:48-53
        case ShapeKind::Plane: return Tuple::vec3(0, 1, 0);           // plane.zig:38-43
        case ShapeKind::Cube: {                                       // cube.zig:81-97
            double ax = std::fabs(p.x), ay = std::fabs(p.y), az = std::fabs(p.z);
            double m = std::fmax(ax, std::fmax(ay, az));
            if (m == ax) return Tuple::vec3(p.x, 0, 0);
            if (m == ay) return Tuple::vec3(0, p.y, 0);
            return Tuple::vec3(0, 0, p.z);
        }
        case ShapeKind::Cylinder: {  // cylinder.zig:100-112
            double dist = p.x * p.x + p.z * p.z;
            if (dist < 1.0 && p.y >= s.maximum - 1e-5) return Tuple::vec3(0, 1, 0);
            if (dist < 1.0 && p.y <= s.minimum + 1e-5) return Tuple::vec3(0, -1, 0);
            return Tuple::vec3(p.x, 0, p.z);
        }
        case ShapeKind::Cone: {  // cone.zig:115-127
            double dist = p.x * p.x + p.z * p.z;
            if (dist < s.maximum * s.maximum && p.y >= s.maximum - 1e-4) return Tuple::vec3(0, 1, 0);
            if (dist < s.minimum * s.minimum && p.y <= s.minimum + 1e-4) return Tuple::vec3(0, -1, 0);
            double y = -zig_sign(p.y) * std::sqrt(p.x * p.x + p.z * p.z);
            return Tuple::vec3(p.x, y, p.z);
        }
        case ShapeKind::Triangle: return s.normal;  // triangle.zig:65-70
        case ShapeKind::SmoothTriangle:             // triangle.zig:261-265
            return s.n2.mul(hit.u).add(s.n3.mul(hit.v)).add(s.n1.mul(1.0 - hit.u - hit.v));
        default: throw std::runtime_error("localNormalAt not implemented for groups / CSG");
    }
}

// shape.zig:338-350
Tuple normal_at(const Shape& s, const Tuple& world_point, const Intersection& hit) {
    Tuple local_point = s.world_to_object(world_point);
    return s.normal_to_world(local_normal_at(s, local_point, hit));
}

}  // namespace oracle
