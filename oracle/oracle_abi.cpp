// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.hpp).  ctypes-friendly entry points used by
// tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
#include <cstdio>
#include <cstring>
#include <string>

#include "../ray-tracer-challenge_b200/csrc/host/loader.hpp"
#include "oracle.hpp"

using namespace oracle;

static void set_err(char* err, size_t n, const std::string& m) {
    if (err && n) {
        std::strncpy(err, m.c_str(), n - 1);
        err[n - 1] = 0;
    }
}

extern "C" {

void* oracle_scene_parse(const char* json_text, const char* asset_dir, long w, long h, char* err, size_t errlen) {
    try {
        auto* s = new rtc::SceneInfo(rtc::parse_scene(json_text, rtc::directory_loader(asset_dir ? asset_dir : ""), w, h));
        return s;
    } catch (const std::exception& e) {
        set_err(err, errlen, e.what());
        return nullptr;
    }
}

void* oracle_scene_load(const char* json_path, const char* asset_dir, long w, long h, char* err, size_t errlen) {
    try {
        std::string text = rtc::read_file(json_path);
        return oracle_scene_parse(text.c_str(), asset_dir, w, h, err, errlen);
    } catch (const std::exception& e) {
        set_err(err, errlen, e.what());
        return nullptr;
    }
}

void oracle_scene_dims(void* h, long* w, long* ht) {
    auto* s = (rtc::SceneInfo*)h;
    *w = (long)s->camera.hsize;
    *ht = (long)s->camera.vsize;
}

// Renders rows [y0,y1) into out (full-frame H*W*3 doubles; untouched rows are left alone).
int oracle_render(void* h, int threads, long y0, long y1, int max_depth, double* out, unsigned long long counts[3]) {
    auto* s = (rtc::SceneInfo*)h;
    try {
        RayCounts rc = render(s->camera, s->world, out, threads, (size_t)y0, y1 < 0 ? (size_t)-1 : (size_t)y1,
                              (size_t)max_depth);
        if (counts) {
            counts[0] = rc.primary;
            counts[1] = rc.shadow;
            counts[2] = rc.secondary;
        }
        return 0;
    } catch (const std::exception&) {
        return -1;
    }
}

// Camera orbit helpers of src/lib.zig:166-190 applied to the oracle's camera (for re-render tests).
void oracle_set_camera(void* h, const double from[3], const double to[3], const double up[3]) {
    auto* s = (rtc::SceneInfo*)h;
    s->camera.from = rtc::Tuple::point(from[0], from[1], from[2]);
    s->camera.to = rtc::Tuple::point(to[0], to[1], to[2]);
    s->camera.up = rtc::Tuple::vec3(up[0], up[1], up[2]);
    s->camera.set_transform(rtc::Mat4::view_transform(s->camera.from, s->camera.to, s->camera.up));
}

void oracle_scene_free(void* h) { delete (rtc::SceneInfo*)h; }

}  // extern "C"

// Debug aid for tests: prints the recursion tree of one pixel (hit kind, t, n1, n2) to stdout.
namespace {
void debug_color_at(const rtc::World& w, const Ray& ray, size_t remaining, int indent) {
    Intersections xs;
    world_intersect(w, ray, xs);
    long h = hit_index(xs);
    if (h < 0) {
        std::printf("%*smiss (remaining %zu, %zu intersections)\n", indent, "", remaining, xs.size());
        return;
    }
    Comps c = prepare_computations(xs[(size_t)h], ray, xs);
    const rtc::Shape& o = *c.intersection.object;
    std::printf("%*shit kind=%d id=%zu t=%.17g n1=%.6g n2=%.6g refl=%.3g transp=%.3g remaining=%zu nxs=%zu hit_index=%ld\n", indent, "",
                (int)o.kind, o.id, c.intersection.t, c.n1, c.n2, o.material.reflective, o.material.transparency, remaining,
                xs.size(), h);
    if (remaining > 0 && o.material.reflective != 0.0) {
        std::printf("%*sreflect:\n", indent, "");
        debug_color_at(w, Ray{c.over_point, c.reflectv}, remaining - 1, indent + 2);
    }
    double n_ratio = c.n1 / c.n2, cos_i = c.eyev.dot(c.normal), sin2_t = n_ratio * n_ratio * (1.0 - cos_i * cos_i);
    if (!(sin2_t > 1.0) && remaining > 0 && o.material.transparency != 0.0) {
        double cos_t = std::sqrt(1.0 - sin2_t);
        rtc::Tuple dir = c.normal.mul(n_ratio * cos_i - cos_t).sub(c.eyev.mul(n_ratio));
        std::printf("%*srefract:\n", indent, "");
        debug_color_at(w, Ray{c.under_point, dir}, remaining - 1, indent + 2);
    } else if (o.material.transparency != 0.0) {
        std::printf("%*sno refraction (sin2_t=%.6g)\n", indent, "", sin2_t);
    }
}
}  // namespace

extern "C" void oracle_debug_pixel(void* h, long x, long y, int max_depth) {
    auto* s = (rtc::SceneInfo*)h;
    debug_color_at(s->world, ray_for_pixel(s->camera, (size_t)x, (size_t)y), (size_t)max_depth, 0);
}
