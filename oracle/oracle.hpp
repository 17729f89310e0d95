// ORACLE — TEST INFRASTRUCTURE ONLY.
// CPU restatement (f64, C++17) of the reference's per-pixel hot path, Camera.render -> World.colorAt
// (src/raytracer/camera.zig:64-125, src/raytracer/world.zig:71-290) and everything below it
// (shapes/*.zig, material.zig, patterns/*.zig, noise.zig).  Each function cites the reference
// lines it follows.  It walks the SAME pointer-based World tree the reference walks (all-hit
// intersection lists, stable sort by t at every level, per-leaf ray transform, the reference's own
// divide(8) group tree), NOT the flattened device scene.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// build, link, load or call anything in this directory.  The product (librtc_b200.so) never does.
//
// Pinning: the reference is Zig and cannot be compiled here (no zig toolchain), so
// oracle/_ref does not exist.  The oracle is pinned by the reference's own known-answer tests
// (SURVEY §8c), transcribed in oracle/kat_tests.cpp.  Full-frame images are NOT pinned by any
// reference artefact (the reference ships no golden image): "parity unpinned" for whole scenes.
#pragma once
#include <cstdint>
#include <vector>

#include "../ray-tracer-challenge_b200/csrc/host/scene.hpp"

namespace oracle {
using namespace rtc;

struct Ray {
    Tuple origin, direction;
    Tuple position(double t) const { return origin.add(direction.mul(t)); }       // ray.zig:25-27
    Ray transform(const Mat4& m) const { return {m.tuple_mul(origin), m.tuple_mul(direction)}; }  // ray.zig:30-32
};

struct Intersection {  // shape.zig:23-47
    double t;
    const Shape* object;
    double u = 0.0, v = 0.0;
};
using Intersections = std::vector<Intersection>;

struct RayCounts {
    uint64_t primary = 0, shadow = 0, secondary = 0;
};

struct Comps {  // PreComputations (world.zig:193-210)
    Intersection intersection;
    Tuple point, over_point, under_point, eyev, normal, reflectv;
    bool inside;
    double n1, n2;
};

// shapes
void intersect(const Shape& s, const Ray& ray, Intersections& out);   // appends, sub-range sorted
void bbox_intersect(const BBox& b, const Ray& ray, Intersections& out, const Shape* tag);
bool bbox_hit(const BBox& b, const Ray& ray);
Tuple local_normal_at(const Shape& s, const Tuple& local_point, const Intersection& hit);
Tuple normal_at(const Shape& s, const Tuple& world_point, const Intersection& hit);
void sort_intersections(Intersections& xs, size_t from);
long hit_index(const Intersections& xs, size_t from = 0);
bool csg_intersection_allowed(CsgOp op, bool lhit, bool inl, bool inr);
bool csg_includes(const Shape& a, const Shape& b);
void csg_filter(const Shape& csg, Intersections& xs, size_t from);

// patterns / shading
double noise(double x, double y, double z);
double octave_noise(double x, double y, double z, int octaves, double persistence);
Color pattern_at(const Pattern& p, const Tuple& object_point);
Color pattern_at_shape(const Pattern& p, const Shape& s, const Tuple& world_point);
Color uv_pattern_at(const UvPattern& uv, double u, double v, const Tuple& object_point);
Color lighting(const Material& m, const Light& light, const Shape& object, const Tuple& point,
               const Tuple& eyev, const Tuple& normal, bool in_shadow);

// world
void world_intersect(const World& w, const Ray& ray, Intersections& out);
Comps prepare_computations(const Intersection& hit, const Ray& ray, const Intersections& xs);
double schlick(const Comps& c);
bool is_shadowed(const World& w, const Tuple& point, const Light& light, RayCounts* rc = nullptr);
Color shade_hit(const World& w, const Comps& c, size_t remaining, RayCounts* rc = nullptr);
Color color_at(const World& w, const Ray& ray, size_t remaining, RayCounts* rc = nullptr);
Color reflected_color(const World& w, const Comps& c, size_t remaining, RayCounts* rc = nullptr);
Color refracted_color(const World& w, const Comps& c, size_t remaining, RayCounts* rc = nullptr);

// camera
Ray ray_for_pixel(const Camera& c, size_t x, size_t y);
// Camera.render: one task per scanline on `threads` workers (camera.zig:88-97); rows [y0,y1).
// out = H*W*3 doubles, row-major, y down (Canvas.pixels).  Returns ray counts.
RayCounts render(const Camera& c, const World& w, double* out, int threads, size_t y0 = 0,
                 size_t y1 = (size_t)-1, size_t max_depth = 5);

}  // namespace oracle
