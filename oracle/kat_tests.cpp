// ORACLE — TEST INFRASTRUCTURE ONLY.  Known-answer tests that PIN the oracle (and the host scene
// model it walks) to the reference: every block below restates a `test "..."` block of the
// reference with the same inputs and the same expected values (file:line cited per block), run
// here in f64 with the reference's own tolerance (1e-5, tuple.zig:14 / color.zig:11).
// Built and run by `make -C oracle kat` and by tests/test_oracle_kat.py.
#include <cmath>
#include <cstdio>
#include <string>

#include "../ray-tracer-challenge_b200/csrc/host/loader.hpp"
#include "oracle.hpp"

using namespace oracle;
using rtc::Mat4;

static int g_fail = 0, g_checks = 0;
static const char* g_block = "";
#define CHECK(cond)                                                                  \
    do {                                                                             \
        g_checks++;                                                                  \
        if (!(cond)) {                                                               \
            g_fail++;                                                                \
            std::printf("FAIL [%s] %s:%d  %s\n", g_block, __FILE__, __LINE__, #cond); \
        }                                                                            \
    } while (0)
#define BLOCK(name) g_block = name

static bool near(double a, double b, double tol = 1e-5) { return std::fabs(a - b) <= tol; }
static Tuple P(double x, double y, double z) { return Tuple::point(x, y, z); }
static Tuple V(double x, double y, double z) { return Tuple::vec3(x, y, z); }
static bool capprox(const Color& c, double r, double g, double b, double tol = 1e-5) {
    return near(c.r, r, tol) && near(c.g, g, tol) && near(c.b, b, tol);
}
static bool tapprox(const Tuple& t, const Tuple& o, double tol = 1e-5) { return t.approx(o, tol); }
static const double S2 = std::sqrt(2.0), S3 = std::sqrt(3.0), PI = 3.14159265358979323846;

static Comps comps_single(const Shape& s, double t, const Ray& r) {
    Intersections xs = {{t, &s}};
    return prepare_computations(xs[0], r, xs);
}

// ---- world.zig:293-892 ----------------------------------------------------------------------------
static void world_tests() {
    BLOCK("world.zig:293 Intersection");
    {
        World w = World::default_world();
        Intersections xs;
        world_intersect(w, Ray{P(0, 0, -5), V(0, 0, 1)}, xs);
        CHECK(xs.size() == 4);
        CHECK(near(xs[0].t, 4.0) && near(xs[1].t, 4.5) && near(xs[2].t, 5.5) && near(xs[3].t, 6.0));
    }
    BLOCK("world.zig:311 PreComputations");
    {
        Shape s = Shape::sphere();
        Comps c = comps_single(s, 4.0, Ray{P(0, 0, -5), V(0, 0, 1)});
        CHECK(tapprox(c.point, P(0, 0, -1)) && tapprox(c.eyev, V(0, 0, -1)) && tapprox(c.normal, V(0, 0, -1)));
        CHECK(!c.inside);
        c = comps_single(s, 1.0, Ray{P(0, 0, 0), V(0, 0, 1)});
        CHECK(tapprox(c.point, P(0, 0, 1)) && tapprox(c.eyev, V(0, 0, -1)) && tapprox(c.normal, V(0, 0, -1)));
        CHECK(c.inside);
        Shape t = Shape::sphere();
        t.set_transform(Mat4::identity().translate(0, 0, 1));
        c = comps_single(t, 5.0, Ray{P(0, 0, -5), V(0, 0, 1)});
        CHECK(c.over_point.z < -1e-5 / 2.0 && c.point.z > c.over_point.z);
        Shape g = Shape::glass_sphere();
        g.set_transform(Mat4::identity().translate(0, 0, 1));
        c = comps_single(g, 5.0, Ray{P(0, 0, -5), V(0, 0, 1)});
        CHECK(c.under_point.z > 1e-5 / 2.0 && c.point.z < c.under_point.z);
    }
    BLOCK("world.zig:379 Shading");
    {
        World w = World::default_world();
        Comps c = comps_single(w.objects[0], 4.0, Ray{P(0, 0, -5), V(0, 0, 1)});
        CHECK(capprox(shade_hit(w, c, 3), 0.38066, 0.47583, 0.2855));
        w.lights[0] = {P(0, 0.25, 0), {1, 1, 1}};
        c = comps_single(w.objects[1], 0.5, Ray{P(0, 0, 0), V(0, 0, 1)});
        CHECK(capprox(shade_hit(w, c, 3), 0.90498, 0.90498, 0.90498));
        World w2;
        w2.lights.push_back({P(0, 0, -10), {1, 1, 1}});
        w2.objects.push_back(Shape::sphere());
        Shape s2 = Shape::sphere();
        s2.set_transform(Mat4::identity().translate(0, 0, 10));
        w2.objects.push_back(s2);
        c = comps_single(w2.objects[1], 4.0, Ray{P(0, 0, 5), V(0, 0, 1)});
        CHECK(capprox(shade_hit(w2, c, 3), 0.1, 0.1, 0.1));
    }
    BLOCK("world.zig:445 Coloring");
    {
        World w = World::default_world();
        CHECK(capprox(color_at(w, Ray{P(0, 0, -5), V(0, 1, 0)}, 3), 0, 0, 0));
        CHECK(capprox(color_at(w, Ray{P(0, 0, -5), V(0, 0, 1)}, 3), 0.38066, 0.47583, 0.2855));
        w.objects[0].material.ambient = 1.0;
        w.objects[1].material.ambient = 1.0;
        Color inner = pattern_at_shape(*w.objects[1].material.pattern, w.objects[1], P(0, 0, 0));
        CHECK(capprox(color_at(w, Ray{P(0, 0, 0.75), V(0, 0, -1)}, 3), inner.r, inner.g, inner.b));
    }
    BLOCK("world.zig:493 isShadowed");
    {
        World w = World::default_world();
        CHECK(!is_shadowed(w, P(0, 10, 0), w.lights[0]));
        CHECK(is_shadowed(w, P(10, -10, 10), w.lights[0]));
        CHECK(!is_shadowed(w, P(-20, 20, -20), w.lights[0]));
        CHECK(!is_shadowed(w, P(-2, 2, -2), w.lights[0]));
        w.objects[0].casts_shadow = false;
        w.objects[1].casts_shadow = true;
        CHECK(is_shadowed(w, P(0, 0, 0), w.lights[0]));
        w.objects[0].casts_shadow = true;
        w.objects[1].casts_shadow = false;
        CHECK(is_shadowed(w, P(0, 0, 0), w.lights[0]));
        w.objects[0].casts_shadow = false;
        CHECK(!is_shadowed(w, P(0, 0, 0), w.lights[0]));
    }
    BLOCK("world.zig:527 Reflections");
    {
        Shape pl = Shape::plane();
        Comps c = comps_single(pl, S2, Ray{P(0, 1, -1), V(0, -1 / S2, 1 / S2)});
        CHECK(tapprox(c.reflectv, V(0, 1 / S2, 1 / S2)));
        World w = World::default_world();
        w.objects[1].material.ambient = 1.0;  // harmless; the reference checks a non-reflective hit is black
        c = comps_single(w.objects[1], 1.0, Ray{P(0, 0, 0), V(0, 0, 1)});
        CHECK(capprox(reflected_color(w, c, 3), 0, 0, 0));
        World w2 = World::default_world();
        Shape fl = Shape::plane();
        fl.material.reflective = 0.5;
        fl.set_transform(Mat4::identity().translate(0, -1, 0));
        w2.objects.push_back(fl);
        Ray r{P(0, 0, -3), V(0, -1 / S2, 1 / S2)};
        c = comps_single(w2.objects[2], S2, r);
        CHECK(capprox(reflected_color(w2, c, 3), 0.19033, 0.23791, 0.14275, 1e-4));
        CHECK(capprox(shade_hit(w2, c, 3), 0.87676, 0.92434, 0.82917, 1e-4));
        CHECK(capprox(reflected_color(w2, c, 0), 0, 0, 0));
        // mutually reflective planes terminate (world.zig:624-650)
        World w3;
        w3.lights.push_back({P(0, 0, 0), {1, 1, 1}});
        Shape lo = Shape::plane(), up = Shape::plane();
        lo.material.reflective = 1.0;
        lo.set_transform(Mat4::identity().translate(0, -1, 0));
        up.material.reflective = 1.0;
        up.set_transform(Mat4::identity().translate(0, 1, 0));
        w3.objects.push_back(lo);
        w3.objects.push_back(up);
        Color cc = color_at(w3, Ray{P(0, 0, 0), V(0, 1, 0)}, 3);
        CHECK(std::isfinite(cc.r));
    }
    BLOCK("world.zig:681 Refraction");
    {
        World w = World::default_world();
        Ray r{P(0, 0, -5), V(0, 0, 1)};
        Intersections xs = {{4.0, &w.objects[0]}, {6.0, &w.objects[0]}};
        CHECK(capprox(refracted_color(w, prepare_computations(xs[0], r, xs), 3), 0, 0, 0));
        w.objects[0].material.transparency = 1.0;
        w.objects[0].material.refractive_index = 1.5;
        CHECK(capprox(refracted_color(w, prepare_computations(xs[0], r, xs), 0), 0, 0, 0));
        // total internal reflection (world.zig:727-748)
        Ray r2{P(0, 0, 1 / S2), V(0, 1, 0)};
        Intersections xs2 = {{-1 / S2, &w.objects[0]}, {1 / S2, &w.objects[0]}};
        CHECK(capprox(refracted_color(w, prepare_computations(xs2[1], r2, xs2), 5), 0, 0, 0));
    }
    BLOCK("world.zig:750 Recursive refraction");
    {
        World w = World::default_world();
        w.objects[0].material.ambient = 1.0;
        rtc::Pattern tp;
        tp.kind = rtc::Pattern::Test;
        w.objects[0].material.pattern = rtc::share(tp);
        w.objects[1].material.transparency = 1.0;
        w.objects[1].material.refractive_index = 1.5;
        Ray r{P(0, 0, 0.1), V(0, 1, 0)};
        Intersections xs = {{-0.9899, &w.objects[0]}, {-0.4899, &w.objects[1]}, {0.4899, &w.objects[1]}, {0.9899, &w.objects[0]}};
        CHECK(capprox(refracted_color(w, prepare_computations(xs[2], r, xs), 5), 0.0, 0.99887, 0.04721, 1e-4));
        World w2 = World::default_world();
        Shape fl = Shape::plane();
        fl.set_transform(Mat4::identity().translate(0, -1, 0));
        fl.material.transparency = 0.5;
        fl.material.refractive_index = 1.5;
        Shape ball = Shape::sphere();
        ball.set_transform(Mat4::identity().translate(0, -3.5, -0.5));
        ball.material.pattern = rtc::share(rtc::Pattern::solid({1, 0, 0}));
        ball.material.ambient = 0.5;
        w2.objects.push_back(fl);
        w2.objects.push_back(ball);
        Ray r2{P(0, 0, -3), V(0, -1 / S2, 1 / S2)};
        CHECK(capprox(shade_hit(w2, comps_single(w2.objects[2], S2, r2), 5), 0.93642, 0.68642, 0.68642));
        // Schlick combined (world.zig:862-891)
        w2.objects[2].material.reflective = 0.5;
        CHECK(capprox(shade_hit(w2, comps_single(w2.objects[2], S2, r2), 5), 0.93391, 0.69643, 0.69243));
    }
    BLOCK("world.zig:809 Schlick");
    {
        Shape g = Shape::glass_sphere();
        Ray r{P(0, 0, 1 / S2), V(0, 1, 0)};
        Intersections xs = {{-1 / S2, &g}, {1 / S2, &g}};
        CHECK(schlick(prepare_computations(xs[1], r, xs)) == 1.0);
        Ray r2{P(0, 0, 0), V(0, 1, 0)};
        Intersections xs2 = {{-1.0, &g}, {1.0, &g}};
        CHECK(near(schlick(prepare_computations(xs2[1], r2, xs2)), 0.04));
        Ray r3{P(0, 0.99, -2), V(0, 0, 1)};
        Intersections xs3 = {{1.8589, &g}};
        CHECK(near(schlick(prepare_computations(xs3[0], r3, xs3)), 0.48873));
    }
}

static Intersections isect(const Shape& s, const Tuple& o, const Tuple& d) {
    Intersections xs;
    intersect(s, Ray{o, d}, xs);
    return xs;
}
static Intersection no_hit() { return {0.0, nullptr}; }

// ---- shapes/*.zig ---------------------------------------------------------------------------------
static void shape_tests() {
    BLOCK("shape.zig:459 Hit");
    {
        Shape s = Shape::sphere();
        Intersections a = {{1.0, &s}, {2.0, &s}};
        sort_intersections(a, 0);
        CHECK(a[(size_t)hit_index(a)].t == 1.0);
        Intersections b = {{-1.0, &s}, {1.0, &s}};
        sort_intersections(b, 0);
        CHECK(b[(size_t)hit_index(b)].t == 1.0);
        Intersections c = {{-2.0, &s}, {-1.0, &s}};
        sort_intersections(c, 0);
        CHECK(hit_index(c) < 0);
        Intersections d = {{5.0, &s}, {7.0, &s}, {-3.0, &s}, {2.0, &s}};
        sort_intersections(d, 0);
        CHECK(d[(size_t)hit_index(d)].t == 2.0);
    }
    BLOCK("shape.zig:520 Refraction n1/n2 table");
    {
        Shape a = Shape::glass_sphere(), b = Shape::glass_sphere(), c = Shape::glass_sphere();
        a.set_transform(Mat4::identity().scale(2, 2, 2));
        a.material.refractive_index = 1.5;
        b.set_transform(Mat4::identity().translate(0, 0, -0.25));
        b.material.refractive_index = 2.0;
        c.set_transform(Mat4::identity().translate(0, 0, 0.25));
        c.material.refractive_index = 2.5;
        Ray r{P(0, 0, -4), V(0, 0, 1)};
        Intersections xs = {{2.0, &a}, {2.75, &b}, {3.25, &c}, {4.75, &b}, {5.25, &c}, {6.0, &a}};
        const double want[6][2] = {{1.0, 1.5}, {1.5, 2.0}, {2.0, 2.5}, {2.5, 2.5}, {2.5, 1.5}, {1.5, 1.0}};
        for (int i = 0; i < 6; i++) {
            Comps k = prepare_computations(xs[(size_t)i], r, xs);
            CHECK(k.n1 == want[i][0] && k.n2 == want[i][1]);
        }
    }
    BLOCK("shape.zig:560 nested groups: world<->object, normals");
    {
        auto build = [](const Mat4& m2) {
            Shape g1 = Shape::group(), g2 = Shape::group(), s = Shape::sphere();
            s.set_transform(Mat4::identity().translate(5, 0, 0));
            g2.add_child(s);
            g1.add_child(g2);
            g1.children[0].set_transform(m2);
            g1.set_transform(Mat4::identity().rotate_y(PI / 2.0));
            return g1;
        };
        Shape g = build(Mat4::identity().scale(2, 2, 2));
        CHECK(tapprox(g.children[0].children[0].world_to_object(P(-2, 0, -10)), P(0, 0, -1)));
        Shape h = build(Mat4::identity().scale(1, 2, 3));
        const Shape& s = h.children[0].children[0];
        CHECK(tapprox(s.normal_to_world(V(1 / S3, 1 / S3, 1 / S3)), V(0.28571, 0.42857, -0.85714)));
        CHECK(tapprox(normal_at(s, P(1.7321, 1.1547, -5.5774), no_hit()), V(0.2857, 0.42854, -0.85716), 1e-4));
    }
    BLOCK("shape.zig:637 parent space bounds");
    {
        Shape s = Shape::sphere();
        s.set_transform(Mat4::identity().scale(0.5, 2, 4).translate(1, -3, 5));
        rtc::BBox b = s.parent_space_bounds();
        CHECK(tapprox(b.min, P(0.5, -5, 1)) && tapprox(b.max, P(1.5, -1, 9)));
    }
    BLOCK("sphere.zig:67-184");
    {
        Shape s = Shape::sphere();
        Intersections xs = isect(s, P(0, 0, -5), V(0, 0, 1));
        CHECK(xs.size() == 2 && near(xs[0].t, 4) && near(xs[1].t, 6));
        xs = isect(s, P(0, 1, -5), V(0, 0, 1));
        CHECK(xs.size() == 2 && near(xs[0].t, 5) && near(xs[1].t, 5));
        CHECK(isect(s, P(0, 2, -5), V(0, 0, 1)).empty());
        xs = isect(s, P(0, 0, 0), V(0, 0, 1));
        CHECK(xs.size() == 2 && near(xs[0].t, -1) && near(xs[1].t, 1));
        xs = isect(s, P(0, 0, 5), V(0, 0, 1));
        CHECK(xs.size() == 2 && near(xs[0].t, -6) && near(xs[1].t, -4));
        Shape sc = Shape::sphere();
        sc.set_transform(Mat4::identity().scale(2, 2, 2));
        xs = isect(sc, P(0, 0, -5), V(0, 0, 1));
        CHECK(xs.size() == 2 && near(xs[0].t, 3) && near(xs[1].t, 7));
        Shape tr = Shape::sphere();
        tr.set_transform(Mat4::identity().translate(5, 0, 0));
        CHECK(isect(tr, P(0, 0, -5), V(0, 0, 1)).empty());
        CHECK(tapprox(normal_at(s, P(1, 0, 0), no_hit()), V(1, 0, 0)));
        CHECK(tapprox(normal_at(s, P(S3 / 3, S3 / 3, S3 / 3), no_hit()), V(S3 / 3, S3 / 3, S3 / 3)));
        Shape t2 = Shape::sphere();
        t2.set_transform(Mat4::identity().translate(0, 1, 0));
        CHECK(tapprox(normal_at(t2, P(0, 1.70711, -0.70711), no_hit()), V(0, 0.70711, -0.70711)));
        Shape t3 = Shape::sphere();
        t3.set_transform(Mat4::identity().rotate_z(PI / 5.0).scale(1, 0.5, 1));
        CHECK(tapprox(normal_at(t3, P(0, S2 / 2, -S2 / 2), no_hit()), V(0, 0.97014, -0.24254)));
    }
    BLOCK("plane.zig:58-107");
    {
        Shape p = Shape::plane();
        CHECK(tapprox(normal_at(p, P(10, 0, -10), no_hit()), V(0, 1, 0)));
        CHECK(isect(p, P(0, 10, 0), V(0, 0, 1)).empty());
        CHECK(isect(p, P(0, 0, 0), V(0, 0, 1)).empty());
        Intersections xs = isect(p, P(0, 1, 0), V(0, -1, 0));
        CHECK(xs.size() == 1 && near(xs[0].t, 1));
        xs = isect(p, P(0, -1, 0), V(0, 1, 0));
        CHECK(xs.size() == 1 && near(xs[0].t, 1));
    }
    BLOCK("cube.zig:111-209");
    {
        Shape c = Shape::cube();
        const double hits[6][8] = {{5, 0.5, 0, -1, 0, 0, 4, 6},  {-5, 0.5, 0, 1, 0, 0, 4, 6}, {0.5, 5, 0, 0, -1, 0, 4, 6},
                                   {0.5, -5, 0, 0, 1, 0, 4, 6},  {0.5, 0, -5, 0, 0, 1, 4, 6}, {0, 0.5, 0, 0, 0, 1, -1, 1}};
        for (auto& h : hits) {
            Intersections xs = isect(c, P(h[0], h[1], h[2]), V(h[3], h[4], h[5]));
            CHECK(xs.size() == 2 && near(xs[0].t, h[6]) && near(xs[1].t, h[7]));
        }
        const double miss[6][6] = {{-2, 0, 0, 0.2673, 0.5345, 0.8018}, {0, -2, 0, 0.8018, 0.2673, 0.5345},
                                   {0, 0, -2, 0.5345, 0.8018, 0.2673}, {2, 0, 2, 0, 0, -1},
                                   {0, 2, 2, 0, -1, 0},                {2, 2, 0, -1, 0, 0}};
        for (auto& m : miss) CHECK(isect(c, P(m[0], m[1], m[2]), V(m[3], m[4], m[5])).empty());
        const double nrm[8][6] = {{1, 0.5, -0.8, 1, 0, 0},  {-1, -0.2, 0.9, -1, 0, 0}, {-0.4, 1, -0.1, 0, 1, 0},
                                  {0.3, -1, -0.7, 0, -1, 0}, {-0.6, 0.3, 1, 0, 0, 1},   {0.4, 0.4, -1, 0, 0, -1},
                                  {1, 1, 1, 1, 0, 0},        {-1, -1, -1, -1, 0, 0}};
        for (auto& n : nrm) CHECK(tapprox(normal_at(c, P(n[0], n[1], n[2]), no_hit()), V(n[3], n[4], n[5])));
    }
    BLOCK("cylinder.zig:124-351");
    {
        Shape c = Shape::cylinder();
        CHECK(isect(c, P(1, 0, 0), V(0, 1, 0).normalized()).empty());
        CHECK(isect(c, P(0, 0, 0), V(0, 1, 0).normalized()).empty());
        CHECK(isect(c, P(0, 0, -5), V(1, 1, 1).normalized()).empty());
        Intersections xs = isect(c, P(1, 0, -5), V(0, 0, 1));
        CHECK(xs.size() == 2 && near(xs[0].t, 5) && near(xs[1].t, 5));
        xs = isect(c, P(0, 0, -5), V(0, 0, 1));
        CHECK(xs.size() == 2 && near(xs[0].t, 4) && near(xs[1].t, 6));
        xs = isect(c, P(0.5, 0, -5), V(0.1, 1, 1).normalized());
        CHECK(xs.size() == 2 && near(xs[0].t, 6.80800, 1e-4) && near(xs[1].t, 7.08869, 1e-4));
        CHECK(tapprox(normal_at(c, P(1, 0, 0), no_hit()), V(1, 0, 0)));
        CHECK(tapprox(normal_at(c, P(0, 5, -1), no_hit()), V(0, 0, -1)));
        CHECK(tapprox(normal_at(c, P(0, -2, 1), no_hit()), V(0, 0, 1)));
        CHECK(tapprox(normal_at(c, P(-1, 1, 0), no_hit()), V(-1, 0, 0)));
        CHECK(c.minimum == -rtc::kInf && c.maximum == rtc::kInf && !c.closed);
        Shape t = Shape::cylinder();
        t.minimum = 1;
        t.maximum = 2;
        const double tr[6][7] = {{0, 1.5, 0, 0.1, 1, 0, 0}, {0, 3, -5, 0, 0, 1, 0}, {0, 0, -5, 0, 0, 1, 0},
                                 {0, 2, -5, 0, 0, 1, 0},    {0, 1, -5, 0, 0, 1, 0}, {0, 1.5, -2, 0, 0, 1, 2}};
        for (auto& k : tr) CHECK(isect(t, P(k[0], k[1], k[2]), V(k[3], k[4], k[5]).normalized()).size() == (size_t)k[6]);
        t.closed = true;
        const double cl[5][6] = {{0, 3, 0, 0, -1, 0}, {0, 3, -2, 0, -1, 2}, {0, 4, -2, 0, -1, 1}, {0, 0, -2, 0, 1, 2}, {0, -1, -2, 0, 1, 1}};
        for (auto& k : cl) CHECK(isect(t, P(k[0], k[1], k[2]), V(k[3], k[4], k[5]).normalized()).size() == 2);
        const double cn[6][4] = {{0, 1, 0, -1}, {0.5, 1, 0, -1}, {0, 1, 0.5, -1}, {0, 2, 0, 1}, {0.5, 2, 0, 1}, {0, 2, 0.5, 1}};
        for (auto& k : cn) CHECK(tapprox(normal_at(t, P(k[0], k[1], k[2]), no_hit()), V(0, k[3], 0)));
        rtc::BBox b = t.bounds();
        CHECK(b.min.y == 1 && b.max.y == 2 && b.min.x == -1 && b.max.z == 1);
    }
    BLOCK("cone.zig:141-258");
    {
        Shape c = Shape::cone();
        Intersections xs = isect(c, P(0, 0, -5), V(0, 0, 1).normalized());
        CHECK(xs.size() == 2 && near(xs[0].t, 5, 1e-4) && near(xs[1].t, 5, 1e-4));
        xs = isect(c, P(0, 0, -5), V(1, 1, 1).normalized());
        CHECK(xs.size() == 2 && near(xs[0].t, 8.66025, 1e-4) && near(xs[1].t, 8.66025, 1e-4));
        xs = isect(c, P(1, 1, -5), V(-0.5, -1, 1).normalized());
        CHECK(xs.size() == 2 && near(xs[0].t, 4.55006, 1e-4) && near(xs[1].t, 49.44994, 1e-4));
        xs = isect(c, P(0, 0, -1), V(0, 1, 1).normalized());
        CHECK(xs.size() == 1 && near(xs[0].t, 0.35355, 1e-4));
        Shape k = Shape::cone();
        k.minimum = -0.5;
        k.maximum = 0.5;
        k.closed = true;
        CHECK(isect(k, P(0, 0, -5), V(0, 1, 0).normalized()).size() == 0);
        CHECK(isect(k, P(0, 0, -0.25), V(0, 1, 1).normalized()).size() == 2);
        CHECK(isect(k, P(0, 0, -0.25), V(0, 1, 0).normalized()).size() == 4);
        CHECK(tapprox(normal_at(c, P(0, 0, 0), no_hit()), V(0, 0, 0)));
        CHECK(tapprox(normal_at(c, P(1, 1, 1), no_hit()), V(1, -S2, 1).normalized()));
        CHECK(tapprox(normal_at(c, P(-1, -1, 0), no_hit()), V(-1, 1, 0).normalized()));
        Shape bb = Shape::cone();
        bb.minimum = -5;
        bb.maximum = 3;
        rtc::BBox b = bb.bounds();
        CHECK(tapprox(b.min, P(-5, -5, -5)) && tapprox(b.max, P(5, 3, 5)));
    }
}

static void mesh_tests() {
    BLOCK("triangle.zig:83-196");
    {
        Shape t = Shape::triangle(P(0, 1, 0), P(-1, 0, 0), P(1, 0, 0));
        CHECK(tapprox(t.e1, V(-1, -1, 0)) && tapprox(t.e2, V(1, -1, 0)) && tapprox(t.normal, V(0, 0, -1)));
        CHECK(tapprox(local_normal_at(t, P(0, 0.5, 0), no_hit()), t.normal));
        CHECK(tapprox(local_normal_at(t, P(-0.5, 0.75, 0), no_hit()), t.normal));
        CHECK(isect(t, P(0, -1, -2), V(0, 1, 0)).empty());   // parallel
        CHECK(isect(t, P(1, 1, -2), V(0, 0, 1)).empty());    // p1-p3 edge
        CHECK(isect(t, P(-1, 1, -2), V(0, 0, 1)).empty());   // p1-p2 edge
        CHECK(isect(t, P(0, -1, -2), V(0, 0, 1)).empty());   // p2-p3 edge
        Intersections xs = isect(t, P(0, 0.5, -2), V(0, 0, 1));
        CHECK(xs.size() == 1 && xs[0].t == 2.0);
        Shape b = Shape::triangle(P(-3, 7, 2), P(6, 2, -4), P(2, -1, -1));
        rtc::BBox bb = b.bounds();
        CHECK(tapprox(bb.min, P(-3, -1, -4)) && tapprox(bb.max, P(6, 7, 2)));
    }
    BLOCK("triangle.zig:278-342 smooth");
    {
        Shape t = Shape::smooth_triangle(P(0, 1, 0), P(-1, 0, 0), P(1, 0, 0), V(0, 1, 0), V(-1, 0, 0), V(1, 0, 0));
        Intersections xs = isect(t, P(-0.2, 0.3, -2), V(0, 0, 1));
        CHECK(xs.size() == 1 && near(xs[0].u, 0.45) && near(xs[0].v, 0.25));
        Intersection i{1.0, &t, 0.45, 0.25};
        CHECK(tapprox(normal_at(t, P(0, 0, 0), i), V(-0.5547, 0.83205, 0)));
        Intersections one = {i};
        Comps c = prepare_computations(i, Ray{P(-0.2, 0.3, -2), V(0, 0, 1)}, one);
        CHECK(tapprox(c.normal, V(-0.5547, 0.83205, 0)));
    }
    BLOCK("bounding_box.zig:183-423");
    {
        rtc::BBox b;
        b.add(P(-5, 2, 0));
        b.add(P(7, 0, -3));
        CHECK(tapprox(b.min, P(-5, 0, -3)) && tapprox(b.max, P(7, 2, 0)));
        rtc::BBox c;
        c.min = P(5, -2, 0);
        c.max = P(11, 4, 7);
        CHECK(c.contains_point(P(5, -2, 0)) && c.contains_point(P(11, 4, 7)) && c.contains_point(P(8, 1, 3)));
        CHECK(!c.contains_point(P(3, 0, 3)) && !c.contains_point(P(8, -4, 3)) && !c.contains_point(P(8, 1, -1)));
        CHECK(!c.contains_point(P(13, 1, 3)) && !c.contains_point(P(8, 5, 3)) && !c.contains_point(P(8, 1, 8)));
        auto box = [](Tuple mn, Tuple mx) {
            rtc::BBox x;
            x.min = mn;
            x.max = mx;
            return x;
        };
        CHECK(c.contains_box(box(P(5, -2, 0), P(11, 4, 7))) && c.contains_box(box(P(6, -1, 1), P(10, 3, 6))));
        CHECK(!c.contains_box(box(P(4, -3, -1), P(10, 3, 6))) && !c.contains_box(box(P(6, -1, 1), P(12, 5, 8))));
        rtc::BBox u = box(P(-1, -1, -1), P(1, 1, 1));
        rtc::BBox ut = u.transform(Mat4::identity().rotate_y(PI / 4).rotate_x(PI / 4));
        CHECK(tapprox(ut.min, P(-1.41421, -1.7071, -1.7071), 1e-4) && tapprox(ut.max, P(1.41421, 1.7071, 1.7071), 1e-4));
        const double unit[13][7] = {{5, 0.5, 0, -1, 0, 0, 1}, {-5, 0.5, 0, 1, 0, 0, 1}, {0.5, -5, 0, 0, -1, 0, 1}, {0.5, -5, 0, 0, 1, 0, 1},
                                    {0.5, 0, 5, 0, 0, -1, 1}, {0.5, 0, -5, 0, 0, 1, 1}, {0, 0.5, 0, 0, 0, 1, 1},   {-2, 0, 0, 2, 4, 6, 0},
                                    {0, -2, 0, 6, 2, 4, 0},   {0, 0, -2, 4, 6, 2, 0},   {2, 0, 2, 0, 0, -1, 0},    {0, 2, 2, 0, -1, 0, 0},
                                    {2, 2, 0, -1, 0, 0, 0}};
        for (auto& k : unit) CHECK(bbox_hit(u, Ray{P(k[0], k[1], k[2]), V(k[3], k[4], k[5]).normalized()}) == (k[6] != 0));
        const double nc[13][7] = {{15, 1, 2, -1, 0, 0, 1}, {-5, -1, 4, 1, 0, 0, 1}, {7, 6, 5, 0, -1, 0, 1},  {9, -5, 6, 0, 1, 0, 1},
                                  {8, 2, 12, 0, 0, -1, 1}, {6, 0, -5, 0, 0, 1, 1},  {8, 1, 3.5, 0, 0, 1, 1}, {9, -1, -8, 2, 4, 6, 0},
                                  {8, 3, -4, 6, 2, 4, 0},  {9, -1, -2, 4, 6, 2, 0}, {4, 0, 9, 0, 0, -1, 0},  {8, 6, -1, 0, -1, 0, 0},
                                  {12, 5, 4, -1, 0, 0, 0}};
        for (auto& k : nc) CHECK(bbox_hit(c, Ray{P(k[0], k[1], k[2]), V(k[3], k[4], k[5]).normalized()}) == (k[6] != 0));
        rtc::BBox l, r;
        box(P(-1, -4, -5), P(9, 6, 5)).split(l, r);
        CHECK(tapprox(l.min, P(-1, -4, -5)) && tapprox(l.max, P(4, 6, 5)) && tapprox(r.min, P(4, -4, -5)) && tapprox(r.max, P(9, 6, 5)));
        box(P(-1, -2, -3), P(9, 5.5, 3)).split(l, r);
        CHECK(tapprox(l.max, P(4, 5.5, 3)) && tapprox(r.min, P(4, -2, -3)));
        box(P(-1, -2, -3), P(5, 8, 3)).split(l, r);
        CHECK(tapprox(l.max, P(5, 3, 3)) && tapprox(r.min, P(-1, 3, -3)));
        box(P(-1, -2, -3), P(5, 3, 7)).split(l, r);
        CHECK(tapprox(l.max, P(5, 3, 2)) && tapprox(r.min, P(-1, -2, 2)));
    }
}

static void group_csg_tests() {
    BLOCK("group.zig:163-215 intersect");
    {
        Shape e = Shape::group();
        CHECK(isect(e, P(0, 0, 0), V(0, 0, 1)).empty());
        Shape g = Shape::group();
        Shape s1 = Shape::sphere(), s2 = Shape::sphere(), s3 = Shape::sphere();
        s2.set_transform(Mat4::identity().translate(0, 0, -3));
        s3.set_transform(Mat4::identity().translate(5, 0, 0));
        g.add_child(s1);
        g.add_child(s2);
        g.add_child(s3);
        Intersections xs = isect(g, P(0, 0, -5), V(0, 0, 1));
        CHECK(xs.size() == 4);
        CHECK(xs[0].object == &g.children[1] && xs[1].object == &g.children[1] && xs[2].object == &g.children[0] &&
              xs[3].object == &g.children[0]);
        Shape t = Shape::group();
        Shape s = Shape::sphere();
        s.set_transform(Mat4::identity().translate(5, 0, 0));
        t.add_child(s);
        t.set_transform(Mat4::identity().scale(2, 2, 2));
        CHECK(isect(t, P(10, 0, -10), V(0, 0, 1)).size() == 2);
    }
    BLOCK("group.zig:217-236 bounds");
    {
        Shape s = Shape::sphere();
        s.set_transform(Mat4::identity().scale(2, 2, 2).translate(2, 5, -3));
        Shape c = Shape::cylinder();
        c.minimum = -2;
        c.maximum = 2;
        c.set_transform(Mat4::identity().scale(0.5, 1, 0.5).translate(-4, -1, 4));
        Shape g = Shape::group();
        g.add_child(s);
        g.add_child(c);
        CHECK(tapprox(g.bbox.min, P(-4.5, -3, -5)) && tapprox(g.bbox.max, P(4, 7, 4.5)));
    }
    BLOCK("group.zig:238-381 partition / subgroup / divide");
    {
        Shape s1 = Shape::sphere(), s2 = Shape::sphere(), s3 = Shape::sphere();
        s1.set_transform(Mat4::identity().translate(-2, 0, 0));
        s2.set_transform(Mat4::identity().translate(2, 0, 0));
        Shape g = Shape::group();
        g.add_child(s1);
        g.add_child(s2);
        g.add_child(s3);
        std::vector<Shape> l, r;
        g.partition_children(l, r);
        CHECK(g.children.size() == 1 && g.children[0].id == s3.id);
        CHECK(l.size() == 1 && l[0].id == s1.id && r.size() == 1 && r[0].id == s2.id);
        Shape h = Shape::group();
        std::vector<Shape> two = {Shape::sphere(), Shape::sphere()};
        h.make_subgroup(two);
        CHECK(h.children.size() == 1 && h.children[0].children.size() == 2);
        // divide(1)
        Shape a = Shape::sphere(), b = Shape::sphere(), c = Shape::sphere();
        a.set_transform(Mat4::identity().translate(-2, -2, 0));
        b.set_transform(Mat4::identity().translate(-2, 2, 0));
        c.set_transform(Mat4::identity().scale(4, 4, 4));
        Shape d = Shape::group();
        d.add_child(a);
        d.add_child(b);
        d.add_child(c);
        d.divide(1);
        CHECK(d.children.size() == 2 && d.children[0].id == c.id);
        const Shape& sub = d.children[1];
        CHECK(sub.children.size() == 2);
        CHECK(sub.children[0].children.size() == 1 && sub.children[0].children[0].id == a.id);
        CHECK(sub.children[1].children.size() == 1 && sub.children[1].children[0].id == b.id);
        // divide(3) with too few children at the top
        Shape q1 = Shape::sphere(), q2 = Shape::sphere(), q3 = Shape::sphere(), q4 = Shape::sphere();
        q1.set_transform(Mat4::identity().translate(-2, 0, 0));
        q2.set_transform(Mat4::identity().translate(2, 1, 0));
        q3.set_transform(Mat4::identity().translate(2, -1, 0));
        Shape sg = Shape::group();
        sg.add_child(q1);
        sg.add_child(q2);
        sg.add_child(q3);
        Shape top = Shape::group();
        top.add_child(sg);
        top.add_child(q4);
        top.divide(3);
        CHECK(top.children.size() == 2 && top.children[1].id == q4.id);
        CHECK(top.children[0].children.size() == 2);
        CHECK(top.children[0].children[0].children.size() == 1 && top.children[0].children[0].children[0].id == q1.id);
        CHECK(top.children[0].children[1].children.size() == 2 && top.children[0].children[1].children[0].id == q2.id &&
              top.children[0].children[1].children[1].id == q3.id);
    }
    BLOCK("csg.zig:143-258");
    {
        using rtc::CsgOp;
        const int U = 0, I = 1, D = 2;
        const int table[24][5] = {{U, 1, 1, 1, 0}, {U, 1, 1, 0, 1}, {U, 1, 0, 1, 0}, {U, 1, 0, 0, 1}, {U, 0, 1, 1, 0}, {U, 0, 1, 0, 0},
                                  {U, 0, 0, 1, 1}, {U, 0, 0, 0, 1}, {I, 1, 1, 1, 1}, {I, 1, 1, 0, 0}, {I, 1, 0, 1, 1}, {I, 1, 0, 0, 0},
                                  {I, 0, 1, 1, 1}, {I, 0, 1, 0, 1}, {I, 0, 0, 1, 0}, {I, 0, 0, 0, 0}, {D, 1, 1, 1, 0}, {D, 1, 1, 0, 1},
                                  {D, 1, 0, 1, 0}, {D, 1, 0, 0, 1}, {D, 0, 1, 1, 1}, {D, 0, 1, 0, 1}, {D, 0, 0, 1, 0}, {D, 0, 0, 0, 0}};
        const CsgOp ops[3] = {CsgOp::Union, CsgOp::Intersection, CsgOp::Difference};
        for (auto& k : table) CHECK(csg_intersection_allowed(ops[k[0]], k[1], k[2], k[3]) == (k[4] != 0));
        const int filt[3][3] = {{U, 0, 3}, {I, 1, 2}, {D, 0, 1}};
        for (auto& k : filt) {
            Shape c = Shape::csg(ops[k[0]], Shape::sphere(), Shape::cube());
            Intersections xs = {{1.0, c.left.get()}, {2.0, c.right.get()}, {3.0, c.left.get()}, {4.0, c.right.get()}};
            Intersections all = xs;
            csg_filter(c, xs, 0);
            CHECK(xs.size() == 2 && xs[0].t == all[(size_t)k[1]].t && xs[1].t == all[(size_t)k[2]].t);
        }
        Shape m = Shape::csg(CsgOp::Union, Shape::sphere(), Shape::cube());
        CHECK(isect(m, P(0, 2, -5), V(0, 0, 1)).empty());
        Shape s2 = Shape::sphere();
        s2.set_transform(Mat4::identity().translate(0, 0, 0.5));
        Shape hcsg = Shape::csg(CsgOp::Union, Shape::sphere(), s2);
        Intersections xs = isect(hcsg, P(0, 0, -5), V(0, 0, 1));
        CHECK(xs.size() == 2 && xs[0].t == 4.0 && xs[0].object == hcsg.left.get() && xs[1].t == 6.5 &&
              xs[1].object == hcsg.right.get());
    }
}

static rtc::PatternP solidp(double r, double g, double b) { return rtc::share(rtc::Pattern::solid({r, g, b})); }
static rtc::Pattern mk(rtc::Pattern::Kind k, rtc::PatternP a, rtc::PatternP b) { return rtc::Pattern::binary(k, a, b); }
static rtc::UvPattern align(rtc::PatternP c, rtc::PatternP ul, rtc::PatternP ur, rtc::PatternP bl, rtc::PatternP br) {
    rtc::UvPattern u;
    u.kind = rtc::UvPattern::AlignCheck;
    u.sub[0] = c, u.sub[1] = ul, u.sub[2] = ur, u.sub[3] = bl, u.sub[4] = br;
    return u;
}

static void shading_tests() {
    using rtc::Pattern;
    auto white = solidp(1, 1, 1), black = solidp(0, 0, 0);
    BLOCK("material.zig:78-163 Lighting");
    {
        rtc::Material m;
        Shape o = Shape::sphere();
        Tuple pos = P(0, 0, 0), n = V(0, 0, -1);
        Light l1{P(0, 0, -10), {1, 1, 1}}, l2{P(0, 10, -10), {1, 1, 1}}, l3{P(0, 0, 10), {1, 1, 1}};
        CHECK(capprox(lighting(m, l1, o, pos, V(0, 0, -1), n, false), 1.9, 1.9, 1.9));
        CHECK(capprox(lighting(m, l1, o, pos, V(0, 1 / S2, -1 / S2), n, false), 1.0, 1.0, 1.0));
        CHECK(capprox(lighting(m, l2, o, pos, V(0, 0, -1), n, false), 0.7364, 0.7364, 0.7364, 1e-4));
        CHECK(capprox(lighting(m, l2, o, pos, V(0, -1 / S2, -1 / S2), n, false), 1.63639, 1.63639, 1.63639, 1e-4));
        CHECK(capprox(lighting(m, l3, o, pos, V(0, 0, -1), n, false), 0.1, 0.1, 0.1));
        CHECK(capprox(lighting(m, l1, o, pos, V(0, 0, -1), n, true), 0.1, 0.1, 0.1));
        rtc::Material s;
        s.pattern = rtc::share(mk(Pattern::Stripes, white, black));
        s.ambient = 1.0;
        s.diffuse = 0.0;
        s.specular = 0.0;
        CHECK(capprox(lighting(s, l1, o, P(0.9, 0, 0), V(0, 0, -1), n, false), 1, 1, 1));
        CHECK(capprox(lighting(s, l1, o, P(1.1, 0, 0), V(0, 0, -1), n, false), 0, 0, 0));
    }
    BLOCK("patterns: stripes/rings/checkers/gradient/blend/test pattern");
    {
        Pattern st = mk(Pattern::Stripes, white, black);
        const double sx[6][2] = {{0, 1}, {0.9, 1}, {1.0, 0}, {-0.1, 0}, {-1.0, 0}, {-1.1, 1}};
        for (auto& k : sx) CHECK(pattern_at(st, P(k[0], 0, 0)).r == k[1]);
        CHECK(pattern_at(st, P(0, 1, 0)).r == 1 && pattern_at(st, P(0, 0, 2)).r == 1);
        Pattern rg = mk(Pattern::Rings, white, black);
        CHECK(pattern_at(rg, P(0, 0, 0)).r == 1 && pattern_at(rg, P(1, 0, 0)).r == 0 && pattern_at(rg, P(0, 0, 1)).r == 0 &&
              pattern_at(rg, P(0.708, 0, 0.708)).r == 0);
        Pattern ck = mk(Pattern::Checkers, white, black);
        CHECK(pattern_at(ck, P(0, 0, 0)).r == 1 && pattern_at(ck, P(0.99, 0, 0)).r == 1 && pattern_at(ck, P(1.01, 0, 0)).r == 0);
        CHECK(pattern_at(ck, P(0, 0.99, 0)).r == 1 && pattern_at(ck, P(0, 1.01, 0)).r == 0);
        CHECK(pattern_at(ck, P(0, 0, 0.99)).r == 1 && pattern_at(ck, P(0, 0, 1.01)).r == 0);
        Pattern gr = mk(Pattern::Gradient, white, black);
        CHECK(capprox(pattern_at(gr, P(0, 0, 0)), 1, 1, 1) && capprox(pattern_at(gr, P(0.25, 0, 0)), 0.75, 0.75, 0.75) &&
              capprox(pattern_at(gr, P(0.5, 0, 0)), 0.5, 0.5, 0.5) && capprox(pattern_at(gr, P(0.75, 0, 0)), 0.25, 0.25, 0.25));
        Pattern rd = mk(Pattern::RadialGradient, white, black);
        CHECK(capprox(pattern_at(rd, P(0.25, 0, 0)), 0.75, 0.75, 0.75) && capprox(pattern_at(rd, P(0, 0, 0.5)), 0.5, 0.5, 0.5) &&
              capprox(pattern_at(rd, P(0.353553, 0, 0.353553)), 0.5, 0.5, 0.5));
        Pattern rot = mk(Pattern::Stripes, white, black);
        rot.set_transform(Mat4::identity().rotate_y(PI / 2.0));
        Pattern bl = mk(Pattern::Blend, rtc::share(st), rtc::share(rot));
        CHECK(capprox(pattern_at(bl, P(0, 0, 0)), 1, 1, 1) && capprox(pattern_at(bl, P(0.5, 0, 0.5)), 0.5, 0.5, 0.5) &&
              capprox(pattern_at(bl, P(-0.5, 0, 0.5)), 0, 0, 0) && capprox(pattern_at(bl, P(-0.5, 0, -0.5)), 0.5, 0.5, 0.5));
        // pattern.zig:152-177
        Pattern tp;
        tp.kind = Pattern::Test;
        Shape s = Shape::sphere();
        s.set_transform(Mat4::identity().scale(2, 2, 2));
        CHECK(capprox(pattern_at_shape(tp, s, P(2, 3, 4)), 1, 1.5, 2));
        Pattern tp2 = tp;
        tp2.set_transform(Mat4::identity().scale(2, 2, 2));
        CHECK(capprox(pattern_at_shape(tp2, Shape::sphere(), P(2, 3, 4)), 1, 1.5, 2));
        Pattern tp3 = tp;
        tp3.set_transform(Mat4::identity().translate(0.5, 1, 1.5));
        CHECK(capprox(pattern_at_shape(tp3, s, P(2.5, 3, 3.5)), 0.75, 0.5, 0.25));
    }
    BLOCK("texture_map.zig:350-568");
    {
        rtc::UvPattern ck;
        ck.kind = rtc::UvPattern::Checkers;
        ck.width = 2;
        ck.height = 2;
        ck.sub[0] = black;
        ck.sub[1] = white;
        const double c2[5][3] = {{0, 0, 0}, {0.5, 0, 1}, {0, 0.5, 1}, {0.5, 0.5, 0}, {1, 1, 0}};
        for (auto& k : c2) CHECK(uv_pattern_at(ck, k[0], k[1], P(0, 0, 0)).r == k[2]);
        auto mapped = [&](Pattern::Mapping m, Tuple p, double u, double v) {
            Pattern t;
            t.kind = Pattern::TextureMap;
            t.mapping = m;
            t.uv[0].kind = rtc::UvPattern::Test;
            Color c = pattern_at(t, p);
            return near(c.r, u) && near(c.g, v);
        };
        const double sp[7][5] = {{0, 0, -1, 0, 0.5}, {1, 0, 0, 0.25, 0.5}, {0, 0, 1, 0.5, 0.5}, {-1, 0, 0, 0.75, 0.5},
                                 {0, 1, 0, 0.5, 1}, {0, -1, 0, 0.5, 0}, {1 / S2, 1 / S2, 0, 0.25, 0.75}};
        for (auto& k : sp) CHECK(mapped(Pattern::Spherical, P(k[0], k[1], k[2]), k[3], k[4]));
        const double pl[7][5] = {{0.25, 0, 0.5, 0.25, 0.5}, {0.25, 0, -0.25, 0.25, 0.75}, {0.25, 0.5, -0.25, 0.25, 0.75},
                                 {1.25, 0, 0.5, 0.25, 0.5}, {0.25, 0, -1.75, 0.25, 0.25}, {1, 0, -1, 0, 0}, {0, 0, 0, 0, 0}};
        for (auto& k : pl) CHECK(mapped(Pattern::Planar, P(k[0], k[1], k[2]), k[3], k[4]));
        const double cy[10][5] = {{0, 0, -1, 0, 0}, {0, 0.5, -1, 0, 0.5}, {0, 1, -1, 0, 0}, {0.70711, 0.5, -0.70711, 0.125, 0.5},
                                  {1, 0.5, 0, 0.25, 0.5}, {0.70711, 0.5, 0.70711, 0.375, 0.5}, {0, -0.25, 1, 0.5, 0.75},
                                  {-0.70711, 0.5, 0.70711, 0.625, 0.5}, {-1, 1.25, 0, 0.75, 0.25}, {-0.70711, 0.5, -0.70711, 0.875, 0.5}};
        for (auto& k : cy) CHECK(mapped(Pattern::Cylindrical, P(k[0], k[1], k[2]), k[3], k[4]));
        auto red = solidp(1, 0, 0), yellow = solidp(1, 1, 0), brown = solidp(1, 0.5, 0), green = solidp(0, 1, 0),
             cyan = solidp(0, 1, 1), blue = solidp(0, 0, 1), purple = solidp(1, 0, 1);
        rtc::UvPattern ac = align(white, red, yellow, green, cyan);
        CHECK(capprox(uv_pattern_at(ac, 0.5, 0.5, P(0, 0, 0)), 1, 1, 1) && capprox(uv_pattern_at(ac, 0.1, 0.9, P(0, 0, 0)), 1, 0, 0) &&
              capprox(uv_pattern_at(ac, 0.9, 0.9, P(0, 0, 0)), 1, 1, 0) && capprox(uv_pattern_at(ac, 0.1, 0.1, P(0, 0, 0)), 0, 1, 0) &&
              capprox(uv_pattern_at(ac, 0.9, 0.1, P(0, 0, 0)), 0, 1, 1));
        // mapped cube: face order front,back,left,right,up,down (texture_map.zig:216)
        Pattern cube;
        cube.kind = Pattern::TextureMap;
        cube.mapping = Pattern::Cubic;
        cube.uv[2] = align(yellow, cyan, red, blue, brown);      // left
        cube.uv[0] = align(cyan, red, yellow, brown, green);     // front
        cube.uv[3] = align(red, yellow, purple, green, white);   // right
        cube.uv[1] = align(green, purple, cyan, white, blue);    // back
        cube.uv[4] = align(brown, cyan, purple, red, yellow);    // up
        cube.uv[5] = align(purple, brown, green, blue, white);   // down
        struct { double x, y, z; rtc::PatternP c; } pts[] = {
            {-1, 0, 0, yellow}, {-1, 0.9, -0.9, cyan}, {-1, 0.9, 0.9, red}, {-1, -0.9, -0.9, blue}, {-1, -0.9, 0.9, brown},
            {0, 0, 1, cyan}, {-0.9, 0.9, 1, red}, {0.9, 0.9, 1, yellow}, {-0.9, -0.9, 1, brown}, {0.9, -0.9, 1, green},
            {1, 0, 0, red}, {1, 0.9, 0.9, yellow}, {1, 0.9, -0.9, purple}, {1, -0.9, 0.9, green}, {1, -0.9, -0.9, white},
            {0, 0, -1, green}, {0.9, 0.9, -1, purple}, {-0.9, 0.9, -1, cyan}, {0.9, -0.9, -1, white}, {-0.9, -0.9, -1, blue},
            {0, 1, 0, brown}, {-0.9, 1, -0.9, cyan}, {0.9, 1, -0.9, purple}, {-0.9, 1, 0.9, red}, {0.9, 1, 0.9, yellow},
            {0, -1, 0, purple}, {-0.9, -1, 0.9, brown}, {0.9, -1, 0.9, green}, {-0.9, -1, -0.9, blue}, {0.9, -1, -0.9, white}};
        for (auto& k : pts) {
            Color c = pattern_at(cube, P(k.x, k.y, k.z));
            CHECK(capprox(c, k.c->color.r, k.c->color.g, k.c->color.b));
        }
        // image lookup on a PPM canvas (texture_map.zig:540-568)
        std::string ppm = "P3\n10 10\n10\n";
        for (int y = 0; y < 10; y++) {
            for (int x = 0; x < 10; x++) {
                int v = (x + y) % 10;
                ppm += std::to_string(v) + " " + std::to_string(v) + " " + std::to_string(v) + "  ";
            }
            ppm += "\n";
        }
        rtc::UvPattern im;
        im.kind = rtc::UvPattern::ImageMap;
        im.image = rtc::decode_ppm(ppm);
        CHECK(capprox(uv_pattern_at(im, 0, 0, P(0, 0, 0)), 0.9, 0.9, 0.9) && capprox(uv_pattern_at(im, 0.3, 0, P(0, 0, 0)), 0.2, 0.2, 0.2) &&
              capprox(uv_pattern_at(im, 0.6, 0.3, P(0, 0, 0)), 0.1, 0.1, 0.1) && capprox(uv_pattern_at(im, 1, 1, P(0, 0, 0)), 0.9, 0.9, 0.9));
    }
    BLOCK("noise.zig:106-109 exact f64");
    {
        CHECK(noise(3.14, 42, 7) == 0.13691995878400012);
        CHECK(noise(-4.20, 10, 6) == 0.14208000000000043);
    }
}

static void host_tests() {
    BLOCK("camera.zig:128-187");
    {
        rtc::Camera c = rtc::Camera::make(200, 125, PI / 2.0);
        CHECK(near(c.pixel_size, 0.01));
        c = rtc::Camera::make(125, 200, PI / 2.0);
        CHECK(near(c.pixel_size, 0.01));
        c = rtc::Camera::make(201, 101, PI / 2.0);
        Ray r = ray_for_pixel(c, 100, 50);
        CHECK(tapprox(r.origin, P(0, 0, 0)) && tapprox(r.direction, V(0, 0, -1)));
        r = ray_for_pixel(c, 0, 0);
        CHECK(tapprox(r.origin, P(0, 0, 0)) && tapprox(r.direction, V(0.66519, 0.33259, -0.66851)));
        c.set_transform(Mat4::identity().translate(0, -2, 5).rotate_y(PI / 4.0));
        r = ray_for_pixel(c, 100, 50);
        CHECK(tapprox(r.origin, P(0, 2, -5)) && tapprox(r.direction, V(1 / S2, 0, -1 / S2)));
        World w = World::default_world();
        rtc::Camera k = rtc::Camera::make(11, 11, PI / 2.0);
        k.set_transform(Mat4::view_transform(P(0, 0, -5), P(0, 0, 0), V(0, 1, 0)));
        std::vector<double> img(11 * 11 * 3);
        RayCounts rc = render(k, w, img.data(), 2);
        const double* px = &img[(5 * 11 + 5) * 3];
        CHECK(near(px[0], 0.38066) && near(px[1], 0.47583) && near(px[2], 0.2855));
        CHECK(rc.primary == 121);
    }
    BLOCK("matrix.zig:330-684 (inverse, transforms, view)");
    {
        Mat4 a = Mat4::rows({{-5, 2, 6, -8}, {1, -5, 1, 8}, {7, 7, -6, -7}, {1, -3, 7, 4}});
        CHECK(near(a.det(), 532.0));
        Mat4 inv = a.inverse();
        Mat4 want = Mat4::rows({{0.21805, 0.45113, 0.24060, -0.04511},
                                {-0.80827, -1.45677, -0.44361, 0.52068},
                                {-0.07895, -0.22368, -0.05263, 0.19737},
                                {-0.52256, -0.81391, -0.30075, 0.30639}});
        CHECK(inv.approx(want));
        bool threw = false;
        try {
            Mat4::rows({{-4, 2, -2, -3}, {9, 6, 2, 6}, {0, -5, 1, -5}, {0, 0, 0, 0}}).inverse();
        } catch (const rtc::NotInvertible&) {
            threw = true;
        }
        CHECK(threw);
        CHECK(tapprox(Mat4::identity().translate(5, -3, 2).tuple_mul(P(-3, 4, 5)), P(2, 1, 7)));
        CHECK(tapprox(Mat4::identity().scale(2, 3, 4).tuple_mul(P(-4, 6, 8)), P(-8, 18, 32)));
        CHECK(tapprox(Mat4::identity().rotate_x(PI / 4).tuple_mul(P(0, 1, 0)), P(0, S2 / 2, S2 / 2)));
        CHECK(tapprox(Mat4::identity().rotate_y(PI / 4).tuple_mul(P(0, 0, 1)), P(S2 / 2, 0, S2 / 2)));
        CHECK(tapprox(Mat4::identity().rotate_z(PI / 4).tuple_mul(P(0, 1, 0)), P(-S2 / 2, S2 / 2, 0)));
        CHECK(tapprox(Mat4::identity().shear(0, 0, 0, 0, 0, 1).tuple_mul(P(2, 3, 4)), P(2, 3, 7)));
        // chained: rotate, scale, translate applied in that order
        CHECK(tapprox(Mat4::identity().rotate_x(PI / 2).scale(5, 5, 5).translate(10, 5, 7).tuple_mul(P(1, 0, 1)), P(15, 0, 7)));
        CHECK(Mat4::view_transform(P(0, 0, 0), P(0, 0, -1), V(0, 1, 0)).approx(Mat4::identity()));
        CHECK(Mat4::view_transform(P(0, 0, 0), P(0, 0, 1), V(0, 1, 0)).approx(Mat4::identity().scale(-1, 1, -1)));
        CHECK(Mat4::view_transform(P(0, 0, 8), P(0, 0, 0), V(0, 1, 0)).approx(Mat4::identity().translate(0, 0, -8)));
        Mat4 vt = Mat4::rows({{-0.50709, 0.50709, 0.67612, -2.36643},
                              {0.76772, 0.60609, 0.12122, -2.82843},
                              {-0.35857, 0.59761, -0.71714, 0.0},
                              {0, 0, 0, 1}});
        CHECK(Mat4::view_transform(P(1, 3, 2), P(4, -2, 8), V(1, 1, 0)).approx(vt));
    }
    BLOCK("color.zig:93-96 clamp / tuple.zig reflect");
    {
        CHECK(rtc::clamp_channel(0.1) == 26 && rtc::clamp_channel(0.5) == 128 && rtc::clamp_channel(-0.1) == 0 &&
              rtc::clamp_channel(21.0) == 255);
        CHECK(tapprox(V(1, -1, 0).reflect(V(0, 1, 0)), V(1, 1, 0)));
        CHECK(tapprox(V(0, -1, 0).reflect(V(S2 / 2, S2 / 2, 0)), V(1, 0, 0)));
        CHECK(tapprox(V(1, 2, 3).cross(V(2, 3, 4)), V(-1, 2, -1)));
    }
    BLOCK("obj.zig:288-544");
    {
        rtc::ObjParser g;
        g.load_obj("There was a young lady named Bright\nwho traveled much faster than light.\nShe set out one day\n"
                   "in a relative way,\nand came back the previous night.",
                   {}, false);
        CHECK(g.lines_ignored == 5);
        rtc::ObjParser v;
        v.load_obj("v -1 1 0\nv -1.0000 0.5000 0.0000\nv 1 0 0\nv 1 1 0", {}, false);
        CHECK(v.lines_ignored == 0 && v.vertices.size() == 4 && tapprox(v.vertices[1], P(-1, 0.5, 0)) && tapprox(v.vertices[3], P(1, 1, 0)));
        rtc::ObjParser f;
        f.load_obj("v -1 1 0\nv -1 0 0\nv 1 0 0\nv 1 1 0\n\nf 1 2 3\nf 1 3 4", {}, false);
        CHECK(f.lines_ignored == 0 && f.default_group.children.size() == 2);
        CHECK(tapprox(f.default_group.children[0].p1, f.vertices[0]) && tapprox(f.default_group.children[0].p3, f.vertices[2]));
        CHECK(tapprox(f.default_group.children[1].p2, f.vertices[2]) && tapprox(f.default_group.children[1].p3, f.vertices[3]));
        rtc::ObjParser poly;
        poly.load_obj("v -1 1 0\nv -1 0 0\nv 1 0 0\nv 1 1 0\nv 0 2 0\nf 1 2 3 4 5", {}, false);
        CHECK(poly.default_group.children.size() == 3);
        CHECK(tapprox(poly.default_group.children[2].p1, poly.vertices[0]) && tapprox(poly.default_group.children[2].p2, poly.vertices[3]) &&
              tapprox(poly.default_group.children[2].p3, poly.vertices[4]));
        rtc::ObjParser ng;
        ng.load_obj("v -1 1 0\nv -1 0 0\nv 1 0 0\nv 1 1 0\ng FirstGroup\nf 1 2 3\ng SecondGroup\nf 1 3 4", {}, false);
        CHECK(ng.lines_ignored == 0 && ng.named_group("FirstGroup") == &ng.default_group.children[0] &&
              ng.named_group("SecondGroup") == &ng.default_group.children[1]);
        CHECK(ng.default_group.children[0].children.size() == 1 && tapprox(ng.default_group.children[1].children[0].p3, ng.vertices[3]));
        rtc::ObjParser vn;
        vn.load_obj("vn 0 0 1\nvn 0.707 0 -0.707\nvn 1 2 3", {}, false);
        CHECK(vn.normals.size() == 3 && tapprox(vn.normals[1], V(0.707, 0, -0.707)) && tapprox(vn.normals[2], V(1, 2, 3)));
        rtc::ObjParser fn;
        fn.load_obj("v 0 1 0\nv -1 0 0\nv 1 0 0\nvn -1 0 0\nvn 1 0 0\nvn 0 1 0\nf 1//3 2//1 3//2\nf 1/0/3 2/102/1 3/14/2", {}, false);
        CHECK(fn.lines_ignored == 0 && fn.default_group.children.size() == 2);
        const Shape &t1 = fn.default_group.children[0], &t2 = fn.default_group.children[1];
        CHECK(t1.kind == rtc::ShapeKind::SmoothTriangle && tapprox(t1.n1, fn.normals[2]) && tapprox(t1.n2, fn.normals[0]) &&
              tapprox(t1.n3, fn.normals[1]));
        CHECK(tapprox(t2.p2, t1.p2) && tapprox(t2.n3, t1.n3));
    }
    BLOCK("scene.zig:664-774 parse round trip (own JSON, same schema)");
    {
        const char* js =
            "{\"camera\":{\"width\":100,\"height\":50,\"field-of-view\":0.785,\"from\":[-6,6,-10],\"to\":[6,0,6],\"up\":[-0.45,1,0]},"
            "\"lights\":[{\"point-light\":{\"position\":[50,100,-50],\"intensity\":[1,1,1]}}],"
            "\"shape-definitions\":[{\"name\":\"base\",\"value\":{\"type\":{\"sphere\":{}},\"material\":{\"ambient\":0.3},"
            "\"transform\":[{\"scale\":[2,2,2]}]}}],"
            "\"objects\":[{\"type\":{\"from-definition\":\"base\"},\"transform\":[{\"translate\":[1,0,0]}],\"material\":{\"diffuse\":0.5}},"
            "{\"type\":{\"group\":[{\"type\":{\"cube\":{}}},{\"type\":{\"cone\":{\"min\":-1,\"max\":0,\"closed\":true}},\"casts-shadow\":false}]},"
            "\"transform\":[{\"rotate-y\":0.5}],\"material\":{\"reflective\":0.25}}]}";
        rtc::SceneInfo si = rtc::parse_scene(js, rtc::directory_loader(""));
        CHECK(si.camera.hsize == 100 && si.camera.vsize == 50 && si.world.lights.size() == 1 && si.world.objects.size() == 2);
        const Shape& s = si.world.objects[0];
        CHECK(s.kind == rtc::ShapeKind::Sphere && s.material.ambient == 0.3 && s.material.diffuse == 0.5);
        CHECK(s.transform.approx(Mat4::identity().scale(2, 2, 2).translate(1, 0, 0)));
        const Shape& g = si.world.objects[1];
        CHECK(g.kind == rtc::ShapeKind::Group && g.children.size() == 2 && g.children[0].material.reflective == 0.25);
        CHECK(g.children[0].transform.approx(Mat4::identity().rotate_y(0.5)) && !g.children[1].casts_shadow &&
              g.children[1].closed && g.children[1].minimum == -1);
    }
}

int main() {
    world_tests();
    shape_tests();
    mesh_tests();
    group_csg_tests();
    shading_tests();
    host_tests();
    std::printf("kat_tests: %d checks, %d failures\n", g_checks, g_fail);
    return g_fail ? 1 : 0;
}
