// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.hpp).  World integrator and camera driver.
#include <atomic>
#include <thread>

#include "oracle.hpp"

namespace oracle {

// world.zig:71-83
void world_intersect(const World& w, const Ray& ray, Intersections& out) {
    out.clear();
    for (const Shape& o : w.objects) intersect(o, ray, out);
    sort_intersections(out, 0);
}

// PreComputations.new, world.zig:212-270
Comps prepare_computations(const Intersection& hit, const Ray& ray, const Intersections& xs) {
    const double epsilon = 1e-5;
    Comps c;
    c.intersection = hit;
    c.point = ray.position(hit.t);
    c.eyev = ray.direction.negate();
    c.normal = normal_at(*hit.object, c.point, hit);
    c.inside = false;
    if (c.normal.dot(c.eyev) < 0) {
        c.normal = c.normal.negate();
        c.inside = true;
    }
    c.over_point = c.point.add(c.normal.mul(epsilon));
    c.under_point = c.point.sub(c.normal.mul(epsilon));
    c.reflectv = ray.direction.reflect(c.normal);

    // containers walk, literally (including "no break when the list is empty after the hit")
    std::vector<const Shape*> containers;
    containers.reserve(xs.size());
    c.n1 = 1.0;
    c.n2 = 1.0;
    for (const Intersection& item : xs) {
        bool is_hit = item.t == hit.t && item.object->id == hit.object->id;
        if (is_hit && !containers.empty()) c.n1 = containers.back()->material.refractive_index;
        bool removed = false;
        for (size_t i = 0; i < containers.size(); i++) {
            if (containers[i]->id == item.object->id) {
                containers.erase(containers.begin() + (long)i);
                removed = true;
                break;
            }
        }
        if (!removed) containers.push_back(item.object);
        if (is_hit && !containers.empty()) {
            c.n2 = containers.back()->material.refractive_index;
            break;
        }
    }
    return c;
}

// world.zig:272-289
double schlick(const Comps& c) {
    double cosv = c.eyev.dot(c.normal);
    if (c.n1 > c.n2) {
        double n_ratio = c.n1 / c.n2;
        double sin2_t = n_ratio * n_ratio * (1.0 - cosv * cosv);
        if (sin2_t > 1.0) return 1.0;
        cosv = std::sqrt(1.0 - sin2_t);
    }
    double frac = (c.n1 - c.n2) / (c.n1 + c.n2);
    double r0 = frac * frac;
    return r0 + (1.0 - r0) * zig_pow(1 - cosv, 5);
}

// world.zig:126-154
bool is_shadowed(const World& w, const Tuple& point, const Light& light, RayCounts* rc) {
    Tuple direction = light.position.sub(point);
    double distance = direction.magnitude();
    Ray shadow_ray{point, direction.normalized()};
    if (rc) rc->shadow++;
    thread_local Intersections xs;  // shadow queries never nest
    world_intersect(w, shadow_ray, xs);
    bool shadowed = false;
    for (long i = hit_index(xs); i >= 0; i = hit_index(xs, (size_t)i + 1)) {
        if (xs[(size_t)i].t < distance && xs[(size_t)i].object->casts_shadow) {
            shadowed = true;
            break;
        }
    }
    return shadowed;
}

// world.zig:157-167
Color reflected_color(const World& w, const Comps& c, size_t remaining, RayCounts* rc) {
    const Material& m = c.intersection.object->material;
    if (remaining == 0 || m.reflective == 0.0) return {0, 0, 0};
    if (rc) rc->secondary++;
    Ray r{c.over_point, c.reflectv};
    return color_at(w, r, remaining - 1, rc).mul(m.reflective);
}

// world.zig:171-189
Color refracted_color(const World& w, const Comps& c, size_t remaining, RayCounts* rc) {
    const Material& m = c.intersection.object->material;
    double n_ratio = c.n1 / c.n2;
    double cos_i = c.eyev.dot(c.normal);
    double sin2_t = n_ratio * n_ratio * (1.0 - cos_i * cos_i);
    if (sin2_t > 1.0) return {0, 0, 0};
    if (remaining == 0 || m.transparency == 0.0) return {0, 0, 0};
    double cos_t = std::sqrt(1.0 - sin2_t);
    Tuple direction = c.normal.mul(n_ratio * cos_i - cos_t).sub(c.eyev.mul(n_ratio));
    if (rc) rc->secondary++;
    Ray r{c.under_point, direction};
    return color_at(w, r, remaining - 1, rc).mul(m.transparency);
}

// world.zig:86-108
Color shade_hit(const World& w, const Comps& c, size_t remaining, RayCounts* rc) {
    const Shape& obj = *c.intersection.object;
    Color surface{0, 0, 0};
    for (const Light& light : w.lights) {
        bool shadowed = is_shadowed(w, c.over_point, light, rc);
        surface = surface.add(lighting(obj.material, light, obj, c.over_point, c.eyev, c.normal, shadowed));
    }
    Color reflected = reflected_color(w, c, remaining, rc);
    Color refracted = refracted_color(w, c, remaining, rc);
    if (obj.material.reflective > 0.0 && obj.material.transparency > 0.0) {
        double reflectance = schlick(c);
        return surface.add(reflected.mul(reflectance)).add(refracted.mul(1.0 - reflectance));
    }
    return surface.add(reflected).add(refracted);
}

// world.zig:111-123
Color color_at(const World& w, const Ray& ray, size_t remaining, RayCounts* rc) {
    thread_local Intersections xs_pool[16];  // one list per recursion level, reused (arena stand-in)
    thread_local int depth = 0;
    Intersections& xs = xs_pool[depth & 15];
    depth++;
    world_intersect(w, ray, xs);
    Color out{0, 0, 0};
    long h = hit_index(xs);
    if (h >= 0) {
        Comps comps = prepare_computations(xs[(size_t)h], ray, xs);
        out = shade_hit(w, comps, remaining, rc);
    }
    depth--;
    return out;
}

// camera.zig:64-76
Ray ray_for_pixel(const Camera& c, size_t x, size_t y) {
    double xoffset = ((double)x + 0.5) * c.pixel_size;
    double yoffset = ((double)y + 0.5) * c.pixel_size;
    double world_x = c.half_width - xoffset;
    double world_y = c.half_height - yoffset;
    Tuple pixel = c.inverse.tuple_mul(Tuple::point(world_x, world_y, -1.0));
    Tuple origin = c.inverse.tuple_mul(Tuple::point(0, 0, 0));
    Tuple direction = pixel.sub(origin).normalized();
    return {origin, direction};
}

// camera.zig:80-125 : one task per scanline handed to a pool of `threads` workers.
RayCounts render(const Camera& c, const World& w, double* out, int threads, size_t y0, size_t y1,
                 size_t max_depth) {
    if (y1 > c.vsize) y1 = c.vsize;
    if (threads < 1) threads = 1;
    std::atomic<size_t> next{y0};
    std::vector<RayCounts> counts((size_t)threads);
    auto worker = [&](int tid) {
        RayCounts rc;
        for (;;) {
            size_t y = next.fetch_add(1);
            if (y >= y1) break;
            for (size_t x = 0; x < c.hsize; x++) {
                Ray ray = ray_for_pixel(c, x, y);
                rc.primary++;
                Color col = color_at(w, ray, max_depth, &rc);
                double* px = out + (y * c.hsize + x) * 3;
                px[0] = col.r;
                px[1] = col.g;
                px[2] = col.b;
            }
        }
        counts[(size_t)tid] = rc;
    };
    if (threads == 1) {
        worker(0);
    } else {
        std::vector<std::thread> pool;
        for (int t = 0; t < threads; t++) pool.emplace_back(worker, t);
        for (auto& t : pool) t.join();
    }
    RayCounts total;
    for (auto& r : counts) {
        total.primary += r.primary;
        total.shadow += r.shadow;
        total.secondary += r.secondary;
    }
    return total;
}

}  // namespace oracle
