// Wavefront form of the render path: Camera.render / World.colorAt (camera.zig:80-125, world.zig:86-189,
// 212-289) with the recursion turned into per-level queues.
//
// Why: in the per-pixel megakernel (render_kernel.cuh) one lane walks a pixel's whole tree of rays, a chain of
// up to ~40 dependent world queries; when the pixel queue runs dry the GPU idles while the longest chains
// finish (measured: 0.3-0.4 ms of dragons' 1.75 ms, 1.1 ms of cover's 4.6 ms, and the reason 8 GPUs gave only
// 2.7x).  Here the unit of work is ONE query:
//   level d launch = { closest-hit query of every ray of level d  ->  hit records + rays of level d+1 }
//                  + { shadow query of every (hit of level d-1, light) pair  ->  lighting() added to the pixel }
// so a frame is max_depth + 2 launches whose tails are one query long, warps always pull 32 tasks of the same
// kind (no mixed shading stages in a warp), and a thread carries no pixel state at all.
// shadeHit is linear in the colours its children return, so every contribution is weight x lighting() added
// to its pixel; the additions are integer (fixed point, scale chosen from the scene's light/material bounds),
// hence exact, order-independent and deterministic: any lane may add to any pixel with a plain RED.
// New rays / hit records are appended with one atomicAdd per warp (__ballot_sync + __popc ranking).
#pragma once
#include <cuda_runtime.h>

#include "render_params.h"
#include "rt_pattern.cuh"
#include "rt_trace.cuh"

namespace rtc {

#ifndef RTC_WF_BLOCKS_ANALYTIC
#define RTC_WF_BLOCKS_ANALYTIC 6
#endif
#ifndef RTC_WF_BLOCKS_MESH
#define RTC_WF_BLOCKS_MESH 4
#endif

// One full world query for the warp (live = this lane takes part): prims + CSG, then every mesh through the
// warp-converged while-while loop (rt_trace.cuh).
template <bool MESH>
__device__ __forceinline__ void wf_trace(Query& q, const DScene& S, int* stack, int* err, bool live) {
    MeshTrav trav;
    bool busy = trace_begin<MESH, false>(q, S, trav, err, live);
    if constexpr (MESH) {
        // the mailboxes (MB_SLOTS doubles per thread) precede the stacks in dynamic shared memory
        extern __shared__ double s_wf[];
        trace_meshes(q, S, trav, stack, s_wf + (threadIdx.x & ~31u), err, busy, live, threadIdx.x & 31u);
    }
    (void)trav;
}

// weight x colour added to a pixel: exact integer addition
__device__ __forceinline__ void wf_add(long long* acc, int pix, V3 c, double scale, int* err) {
    double v[3] = {c.x * scale, c.y * scale, c.z * scale};
#pragma unroll
    for (int k = 0; k < 3; k++) {
        if (!(fabs(v[k]) < 9.0e18)) {  // non-finite or beyond the scene's bound: the frame is flagged (ERR_ACC_RANGE)
            *err |= ERR_ACC_RANGE;
            continue;
        }
        long long i = __double2ll_rn(v[k]);
        if (i) atomicAdd(reinterpret_cast<unsigned long long*>(acc + 3ll * pix + k), (unsigned long long)i);
    }
}

// Appends one record per flagged lane to a queue: one atomicAdd per warp, lanes ranked with ballot + popc.
// Returns the slot (or ~0ull when the queue is full: the error bit is set and the frame is redone by the caller).
__device__ __forceinline__ unsigned long long wf_reserve(bool want, unsigned long long* counter, unsigned long long cap,
                                                        unsigned lane, unsigned lt_mask, int* err) {
    unsigned m = __ballot_sync(0xffffffffu, want);
    if (!m) return ~0ull;
    unsigned long long base = 0;
    int leader = __ffs(m) - 1;
    if ((int)lane == leader) base = atomicAdd(counter, (unsigned long long)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (!want) return ~0ull;
    unsigned long long slot = base + __popc(m & lt_mask);
    if (slot >= cap) {
        *err |= ERR_QUEUE;
        return ~0ull;
    }
    return slot;
}

__device__ __forceinline__ void wf_store_ray(WRay* dst, V3 o, V3 d, double w, int remaining, int pix) {
    double2* p = reinterpret_cast<double2*>(dst);
    p[0] = make_double2(o.x, o.y);
    p[1] = make_double2(o.z, d.x);
    p[2] = make_double2(d.y, d.z);
    p[3] = make_double2(w, __hiloint2double(pix, remaining));
}

// lighting() of one (hit, light) pair, shadow query included: World.isShadowed (world.zig:126-135) +
// Material.lighting (material.zig:40-74).  Warp-wide call (live = this lane has a pair).
template <bool MESH>
__device__ __forceinline__ V3 wf_light(const DScene& S, V3 over, V3 normal, V3 rd, V3 color, int mat, int li, bool live,
                                       int* stack, int* err, unsigned& n_culled) {
    const rtc_light& Lg = S.lights[li];
    V3 dir = v3(Lg.position[0], Lg.position[1], Lg.position[2]) - over;
    double dist = sqrt(dot(dir, dir));
    Query q;
    q.mode = Q_SHADOW;
    q.o = over;
    q.d = v3(dir.x / dist, dir.y / dist, dir.z / dist);
    q.t = dist;
    q.leaf = -1;
    q.a = q.b = 0.0;
    q.c_leaf = -1;
    q.flag = 0;
    // Light behind the surface: lighting() yields ambient whether or not the point is shadowed
    // (material.zig:55-63), so the ray need not be traced.  Counted as a shadow ray, reported as culled.
    const double ldn = dot(q.d, normal);
    const bool trace_it = live && (ldn >= 0.0);
    if (live && !trace_it) n_culled++;
    wf_trace<MESH>(q, S, stack, err, trace_it);
    V3 sum = v3(0, 0, 0);
    if (live) {
        const rtc_material& M = S.materials[mat];
        V3 li3 = v3(Lg.intensity[0], Lg.intensity[1], Lg.intensity[2]);
        V3 eff = v3(color.x * li3.x, color.y * li3.y, color.z * li3.z);
        sum = eff * M.ambient;
        if (trace_it && !q.flag) {
            V3 lightv = q.d;
            V3 diffuse = eff * (M.diffuse * ldn);
            V3 reflectv = lightv - normal * (2.0 * dot(lightv, normal));  // tuple.zig reflect
            double rde = dot(neg(reflectv), neg(rd));
            V3 spec = v3(0, 0, 0);
            if (rde > 0.0) spec = li3 * (M.specular * zig_pow(rde, M.shininess));
            sum = sum + diffuse + spec;
        }
    }
    return sum;
}

template <bool MESH>
__global__ void __launch_bounds__(RTC_BLOCK, MESH ? RTC_WF_BLOCKS_MESH : RTC_WF_BLOCKS_ANALYTIC)
wave_kernel(DScene S, WaveParams W) {
    extern __shared__ double s_wf[];  // MESH: MB_SLOTS x RTC_BLOCK doubles (mailboxes), then S.stack_cap x RTC_BLOCK ints, column per thread
    int* stack = reinterpret_cast<int*>(s_wf + MB_SLOTS * RTC_BLOCK) + threadIdx.x;
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    unsigned long long* C = W.counters;
    int err = 0;
    unsigned n_culled = 0;
    const int level = W.level;

    // ================= rays of this level: closest hit -> hit record + children =======================
    {
        const unsigned tiles_x = (unsigned)((W.cam.hsize + 7) / 8);
        unsigned long long n_rays;
        if (level == 0) n_rays = (unsigned long long)tiles_x * ((W.rows + 3) / 4) * 32ull;  // 8x4 pixel tiles
        else n_rays = C[WC_NRAYS + level] < W.cap ? C[WC_NRAYS + level] : W.cap;
        if (level > W.max_depth) n_rays = 0;
        for (;;) {
            unsigned long long base = 0;
            if (lane == 0) base = atomicAdd(C + WC_TAKE_RAY + level, 32ull);
            base = __shfl_sync(0xffffffffu, base, 0);
            if (base >= n_rays) break;
            const unsigned long long item = base + lane;
            bool live = item < n_rays;
            Query q;
            q.mode = Q_CLOSEST;
            q.o = v3(0, 0, 0);
            q.d = v3(0, 0, 1);
            double rw = 0.0;
            int remaining = 0, pix = 0;
            if (live) {
                if (level == 0) {
                    unsigned long long tile = item >> 5;
                    unsigned in = (unsigned)(item & 31u);
                    unsigned long long px = (tile % tiles_x) * 8ull + (in & 7u);
                    unsigned long long pr = (tile / tiles_x) * 4ull + (in >> 3);
                    if (px < W.cam.hsize && pr < W.rows) {
                        unsigned long long y = ((pr / W.tile_rows) * W.n_parts + W.part) * W.tile_rows + pr % W.tile_rows;
                        // Camera.rayForPixel (camera.zig:64-76)
                        double xoffset = ((double)px + 0.5) * W.cam.pixel_size;
                        double yoffset = ((double)y + 0.5) * W.cam.pixel_size;
                        double wx = W.cam.half_width - xoffset;
                        double wy = W.cam.half_height - yoffset;
                        const double* m = W.cam.inv;
                        V3 pixel = v3(m[0] * wx + m[1] * wy + m[2] * -1.0 + m[3], m[4] * wx + m[5] * wy + m[6] * -1.0 + m[7],
                                      m[8] * wx + m[9] * wy + m[10] * -1.0 + m[11]);
                        V3 origin = v3(m[3], m[7], m[11]);
                        q.o = origin;
                        q.d = normalized(pixel - origin);
                        rw = 1.0;
                        remaining = W.max_depth;
                        pix = (int)(pr * W.cam.hsize + px);
                    } else {
                        live = false;
                    }
                } else {
                    const double2* p = reinterpret_cast<const double2*>(W.rays_in + item);
                    double2 a = p[0], b = p[1], c = p[2], e = p[3];
                    q.o = v3(a.x, a.y, b.x);
                    q.d = v3(b.y, c.x, c.y);
                    rw = e.x;
                    remaining = __double2loint(e.y);
                    pix = __double2hiint(e.y);
                }
            }
            q.t = RTC_INF;
            q.leaf = -1;
            q.a = q.b = 0.0;
            q.c_leaf = -1;  // Q_CLOSEST: the leaf whose u,v are in q.a / q.b
            q.flag = 0;
            wf_trace<MESH>(q, S, stack, &err, live);

            const bool hit = live && q.leaf >= 0;  // colorAt: no hit -> black (world.zig:119-121)
            // ---- PreComputations.new (world.zig:212-227) ------------------------------------------
            V3 point = v3(0, 0, 0), normal = v3(0, 1, 0), color = v3(0, 0, 0);
            int mat = 0;
            const int hit_leaf = q.leaf;
            const double hit_t = q.t;
            double transparency = 0.0, reflective = 0.0;
            if (hit) {
                const rtc_leaf& L = S.leaves[hit_leaf];
                M34 inv = load_m34(S.transforms[L.transform].inverse);
                point = q.o + q.d * hit_t;
                int kind = L.kind;
                V3 lp = xf_point(inv, point);  // Shape.normalAt (shape.zig:338-350)
                double u = q.a, v = q.b;
                const double* nrm = nullptr;
                if (kind == RTC_TRIANGLE || kind == RTC_SMOOTH_TRIANGLE) {
                    nrm = S.normals[L.triangle].n;
                    if (kind == RTC_SMOOTH_TRIANGLE && q.c_leaf != hit_leaf) {
                        // u,v of a smooth triangle that is a CSG operand: same ray, same test, same values
                        const rtc_triangle& T = S.blob_tris[L.triangle];
                        double tt;
                        tri_intersect(v3(T.p1[0], T.p1[1], T.p1[2]), v3(T.e1[0], T.e1[1], T.e1[2]),
                                      v3(T.e2[0], T.e2[1], T.e2[2]), xf_point(inv, q.o), xf_vec(inv, q.d), tt, u, v);
                    }
                }
                V3 ln = leaf_local_normal(kind, L.minimum, L.maximum, lp, nrm, u, v);
                V3 n = xf_normal(inv, ln);
                if (dot(n, neg(q.d)) < 0.0) n = neg(n);
                normal = n;
                mat = L.material;
                const rtc_material& M = S.materials[mat];
                transparency = M.transparency;
                reflective = M.reflective;
                // Material.lighting evaluates the pattern at over_point for every light (material.zig:49);
                // it is the same colour each time, so once per hit.
                color = pattern_color(S, M.pattern, xf_point(inv, point + n * 1e-5), &err);
            }
            // ---- n1 / n2 (world.zig:229-255): only reach a result through refractedColor / schlick, both
            //      gated on transparency (world.zig:101-102, :178) -> second query for those lanes only
            double n1 = 1.0, n2 = 1.0;
            // (skipped when every material has refractive index 1.0: n1 = n2 = 1.0 whatever the list holds)
            const bool transp = hit && transparency != 0.0 && !S.ri_all_one;
            if (__any_sync(0xffffffffu, transp)) {
                Query c;
                c.mode = Q_CONTAINERS;
                c.o = q.o;
                c.d = q.d;
                c.t = hit_t;
                c.leaf = hit_leaf;
                c.a = -RTC_INF;
                c.b = -RTC_INF;
                c.c_leaf = -1;
                c.flag = 0;
                wf_trace<MESH>(c, S, stack, &err, transp);
                if (transp) {
                    double ri_hit = S.materials[mat].refractive_index;
                    double ri_c = c.c_leaf >= 0 ? S.materials[S.leaves[c.c_leaf].material].refractive_index : 1.0;
                    if (c.flag) {  // the hit leaf was a container itself
                        bool hit_is_last = c.c_leaf < 0 || c.b > c.a || (c.b == c.a && hit_leaf > c.c_leaf);
                        n1 = hit_is_last ? ri_hit : ri_c;
                        n2 = ri_c;
                    } else {
                        n1 = ri_c;
                        n2 = ri_hit;
                    }
                }
            }
            // ---- hit record for the shadow queries of the next launch (shadeHit, world.zig:89-96) ---
            if (S.n_lights > 0) {
                unsigned long long slot = wf_reserve(hit, C + WC_NHITS + level, W.cap, lane, lt_mask, &err);
                if (slot != ~0ull) {
                    V3 over = point + normal * 1e-5;
                    double2* p = reinterpret_cast<double2*>(W.hits_out + slot);
                    p[0] = make_double2(over.x, over.y);
                    p[1] = make_double2(over.z, normal.x);
                    p[2] = make_double2(normal.y, normal.z);
                    p[3] = make_double2(q.d.x, q.d.y);
                    p[4] = make_double2(q.d.z, color.x);
                    p[5] = make_double2(color.y, color.z);
                    p[6] = make_double2(rw, __hiloint2double(pix, mat));
                }
            }
            // ---- children (reflectedColor world.zig:157-167, refractedColor :171-189, schlick :272-289)
            bool spawn_refl = hit && remaining > 0 && reflective != 0.0;
            bool spawn_refr = false;
            double refl_w = spawn_refl ? reflective : 0.0, refr_w = 0.0;
            V3 refr_dir = v3(0, 0, 0);
            if (hit) {
                V3 eyev = neg(q.d);
                double n_ratio = n1 / n2;
                double cos_i = dot(eyev, normal);
                double sin2_t = n_ratio * n_ratio * (1.0 - cos_i * cos_i);
                spawn_refr = !(sin2_t > 1.0) && remaining > 0 && transparency != 0.0;
                if (spawn_refr) {
                    double cos_t = sqrt(1.0 - sin2_t);
                    refr_dir = normal * (n_ratio * cos_i - cos_t) - eyev * n_ratio;
                    refr_w = transparency;
                }
                if (reflective > 0.0 && transparency > 0.0) {
                    double cosv = cos_i;
                    double reflectance;
                    bool tir = false;
                    if (n1 > n2) {
                        double nr = n1 / n2;
                        double s2 = nr * nr * (1.0 - cosv * cosv);
                        if (s2 > 1.0) tir = true;
                        else cosv = sqrt(1.0 - s2);
                    }
                    if (tir) {
                        reflectance = 1.0;
                    } else {
                        double frac = (n1 - n2) / (n1 + n2);
                        double r0 = frac * frac;
                        reflectance = r0 + (1.0 - r0) * zig_pow(1.0 - cosv, 5.0);
                    }
                    refl_w *= reflectance;
                    refr_w *= (1.0 - reflectance);
                }
            }
            {
                unsigned long long slot = wf_reserve(spawn_refr, C + WC_NRAYS + level + 1, W.cap, lane, lt_mask, &err);
                if (slot != ~0ull) wf_store_ray(W.rays_out + slot, point - normal * 1e-5, refr_dir, rw * refr_w, remaining - 1, pix);
                slot = wf_reserve(spawn_refl, C + WC_NRAYS + level + 1, W.cap, lane, lt_mask, &err);
                if (slot != ~0ull) {
                    V3 reflectv = q.d - normal * (2.0 * dot(q.d, normal));
                    wf_store_ray(W.rays_out + slot, point + normal * 1e-5, reflectv, rw * refl_w, remaining - 1, pix);
                }
            }
        }
    }

    // ================= shadow queries of the previous level's hits: isShadowed + lighting ================
    if (level > 0 && S.n_lights > 0) {
        const unsigned long long L = (unsigned long long)S.n_lights;
        unsigned long long n_hits = C[WC_NHITS + level - 1] < W.cap ? C[WC_NHITS + level - 1] : W.cap;
        const unsigned long long n_tasks = n_hits * L;
        for (;;) {
            unsigned long long base = 0;
            if (lane == 0) base = atomicAdd(C + WC_TAKE_SHADOW + level - 1, 32ull);
            base = __shfl_sync(0xffffffffu, base, 0);
            if (base >= n_tasks) break;
            const unsigned long long task = base + lane;
            const bool live = task < n_tasks;
            const unsigned long long rec = live ? task / L : 0;
            const int li = live ? (int)(task % L) : 0;
            const double2* p = reinterpret_cast<const double2*>(W.hits_in + rec);
            double2 a = p[0], b = p[1], c = p[2], e = p[3], f = p[4], g = p[5], h = p[6];
            const V3 over = v3(a.x, a.y, b.x), normal = v3(b.y, c.x, c.y), rd = v3(e.x, e.y, f.x), color = v3(f.y, g.x, g.y);
            V3 sum = wf_light<MESH>(S, over, normal, rd, color, __double2loint(h.y), li, live, stack, &err, n_culled);
            if (live) wf_add(W.acc, __double2hiint(h.y), sum * h.x, W.scale, &err);
        }
    }

    // ---- counters: warp reduce, one atomic per warp ---------------------------------------------------------
    unsigned long long c_culled = n_culled;
    for (int s = 16; s > 0; s >>= 1) {
        c_culled += __shfl_down_sync(0xffffffffu, c_culled, s);
        err |= __shfl_down_sync(0xffffffffu, err, s);
    }
    if (lane == 0) {
        if (c_culled) atomicAdd(C + WC_CULLED, c_culled);
        if (err) atomicOr(C + WC_ERR, (unsigned long long)err);
    }
}

// Fixed-point accumulators -> Canvas.pixels (f64, packed rows or full frame) or clamp()-quantised RGBA8.
__device__ __forceinline__ unsigned char wf_quantise(double c) {  // color.zig:61-71
    double t = round(c * 255.0);
    if (!(t > 0.0)) return 0;
    if (t > 255.0) return 255;
    return (unsigned char)t;
}
static __global__ void resolve_kernel(ResolveParams R) {
    unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= R.n_pixels) return;
    double r = (double)R.acc[3 * i] * R.inv_scale, g = (double)R.acc[3 * i + 1] * R.inv_scale,
           b = (double)R.acc[3 * i + 2] * R.inv_scale;
    unsigned long long pr = i / R.hsize, px = i % R.hsize;
    unsigned long long o = i;
    if (R.full_frame) {
        unsigned long long y = ((pr / R.tile_rows) * R.n_parts + R.part) * R.tile_rows + pr % R.tile_rows;
        o = y * R.hsize + px;
    }
    if (R.out8) {
        reinterpret_cast<uchar4*>(R.out8)[o] = make_uchar4(wf_quantise(r), wf_quantise(g), wf_quantise(b), 255);
    } else {
        R.out[3 * o] = r;
        R.out[3 * o + 1] = g;
        R.out[3 * o + 2] = b;
    }
}

}  // namespace rtc
