// World.intersect (world.zig:71-83) re-expressed as three queries that never build the reference's
// sorted all-hit list:
//   Q_CLOSEST     hit() of the sorted list  = min over every intersection with t >= 0, ties broken by
//                 leaf order (what concatenation + stable sort gives the reference, shape.zig:64-80)
//   Q_SHADOW      isShadowed's scan (world.zig:141-151) = any intersection with 0 <= t < distance on a
//                 leaf that casts shadows
//   Q_CONTAINERS  the n1/n2 `containers` walk of PreComputations.new (world.zig:229-255), see Containers
// Meshes are traversed through an f32 outward-rounded BVH in object space; the triangle test itself
// is the reference's f64 Möller–Trumbore.  CSG objects run their compiled program (device_scene.h).
#pragma once
#include <cfloat>

#include "rt_math.cuh"

namespace rtc {

#include "render_params.h"
#define RTC_BLOCK RTC_BLOCK_THREADS
#define RTC_CSG_CAP 32
#define RTC_CSG_SEG 16

enum { Q_CLOSEST = 0, Q_SHADOW = 1, Q_CONTAINERS = 2 };
enum { ERR_TEXEL = 1, ERR_PAT_STACK = 2, ERR_CSG_CAP = 4, ERR_BVH_STACK = 8, ERR_QUEUE = 16, ERR_ACC_RANGE = 32, ERR_COPY = 64 };

// One world query.  The three modes never coexist, so they share their fields:
//                 t                 leaf        a       b            c_leaf      flag
//   Q_CLOSEST     best t so far     best leaf   u       v            leaf u,v are of   -
//   Q_SHADOW      light distance    -           -       -            -           shadowed
//   Q_CONTAINERS  t of the hit      hit leaf    c_t     hit_last_t   c_leaf      hit_odd
// In every mode nothing beyond t can change the result: t is the far end of the useful range.
struct Query {
    int mode;
    V3 o, d;  // world-space ray
    double t;
    int leaf;  // -1 = none
    double a, b;
    int c_leaf;
    int flag;
};

// The containers walk reduced to a streaming form.  Processing the sorted list up to the hit, an
// object is in `containers` iff it has an ODD number of intersections before the hit, and it sits
// at the position of its LAST such intersection (remove-or-append, world.zig:240-249).  So:
//   last container before the hit = the odd-parity leaf whose last pre-hit intersection has the
//   greatest (t, leaf order) key (kept in q.a / q.c_leaf); n1 is its refractive index (1.0 if none);
//   the hit toggles its own leaf: if it was inside (odd: q.flag, its last pre-hit t in q.b) it is
//   removed and n2 comes from the best OTHER leaf (or stays 1.0: the reference does not break on an
//   empty list), else it is appended last and n2 is its own index.
__device__ __forceinline__ void cont_add(Query& q, int leaf, int n, const double* ts) {
    int cnt = 0;
    double tl = -RTC_INF;
    for (int i = 0; i < n; i++) {
        bool before = ts[i] < q.t || (ts[i] == q.t && leaf < q.leaf);
        if (before) {
            cnt++;
            tl = fmax(tl, ts[i]);
        }
    }
    if (cnt & 1) {
        if (leaf == q.leaf) {
            q.flag = 1;
            q.b = tl;
        } else if (q.c_leaf < 0 || tl > q.a || (tl == q.a && leaf > q.c_leaf)) {
            q.a = tl;
            q.c_leaf = leaf;
        }
    }
}

// ---- conservative f32 slab tests ----------------------------------------------------------------------
// Used for every box that is only an OPTIMISATION (world pre-test boxes of prims and meshes, BVH
// nodes): the boxes are supersets rounded outward to f32, and the slab interval is widened by a
// rigorous bound on the f32 evaluation error, so a box the exact ray touches is never culled.
//   |error of (b - o) * (1/d) in f32|  <  2^-21 (|b| + 2|o|) |1/d|        (BVH)
// (rounding of o and d to f32, the f32 reciprocal, the rounded product o*(1/d), one fused multiply-add:
// together below 2^-22 (|b| + 2|o|) |1/d|; the bound used keeps a factor 2 in hand).
struct FRay {
    float ix, iy, iz;  // 1/d
    // -o * (1/d) -+ the error bound: the slab parameter of plane b is fma(b, 1/d, c); with the bound folded into
    // the constant (rounded outward), the near planes give a lower and the far planes an upper bound directly
    float nx, ny, nz, fx, fy, fz;
};
__device__ __forceinline__ float finv(double d) {
    // d == 0 (or below f32 range) gives +-inf: clamp, so that (b - o) == 0 yields 0 and not NaN
    return fminf(fmaxf(1.0f / (float)d, -FLT_MAX), FLT_MAX);
}
// babs: max |box coordinate| per axis over the boxes this ray will be tested against
__device__ __forceinline__ FRay make_fray(V3 o, V3 d, float bax, float bay, float baz) {
    FRay r;
    float ox = (float)o.x, oy = (float)o.y, oz = (float)o.z;
    r.ix = finv(d.x), r.iy = finv(d.y), r.iz = finv(d.z);
    // finite by construction when o is (|1/d| <= FLT_MAX); an overflowing product saturates towards its own
    // sign, which keeps the comparison against the (finite) query range right
    float cx = fminf(fmaxf(-ox * r.ix, -FLT_MAX), FLT_MAX);
    float cy = fminf(fmaxf(-oy * r.iy, -FLT_MAX), FLT_MAX);
    float cz = fminf(fmaxf(-oz * r.iz, -FLT_MAX), FLT_MAX);
    const float k = 4.76837158e-7f;  // 2^-21
    float ex = fminf(FLT_MAX, k * (bax + 2.0f * fabsf(ox)) * fabsf(r.ix));
    float ey = fminf(FLT_MAX, k * (bay + 2.0f * fabsf(oy)) * fabsf(r.iy));
    float ez = fminf(FLT_MAX, k * (baz + 2.0f * fabsf(oz)) * fabsf(r.iz));
    // (c - e) rounded down, (c + e) rounded up; a saturated e leaves the axis unconstrained (-inf / +inf, or a
    // NaN out of inf - inf in the fma below, which fminf / fmaxf drop)
    r.nx = __fsub_rd(cx, ex), r.ny = __fsub_rd(cy, ey), r.nz = __fsub_rd(cz, ez);
    r.fx = __fadd_ru(cx, ex), r.fy = __fadd_ru(cy, ey), r.fz = __fadd_ru(cz, ez);
    return r;
}
__device__ __forceinline__ bool fbox(const FRay& r, float lx, float ly, float lz, float hx, float hy, float hz,
                                     float t_lo, float t_hi, float& tnear) {
    // the plane the ray enters through is lo where it travels in +axis, hi where it travels in -axis
    const bool sx = r.ix < 0.0f, sy = r.iy < 0.0f, sz = r.iz < 0.0f;
    float ax = __fmaf_rn(sx ? hx : lx, r.ix, r.nx), bx = __fmaf_rn(sx ? lx : hx, r.ix, r.fx);
    float ay = __fmaf_rn(sy ? hy : ly, r.iy, r.ny), by = __fmaf_rn(sy ? ly : hy, r.iy, r.fy);
    float az = __fmaf_rn(sz ? hz : lz, r.iz, r.nz), bz = __fmaf_rn(sz ? lz : hz, r.iz, r.fz);
    float tn = fmaxf(fmaxf(ax, ay), fmaxf(az, t_lo));
    float tf = fminf(fminf(bx, by), fminf(bz, t_hi));
    tnear = tn;
    return tn <= tf;
}
// The t-range a query still cares about, rounded outward: nothing outside it can change the result.
__device__ __forceinline__ float query_t_lo(const Query& q) {
    return q.mode == Q_CONTAINERS ? -__int_as_float(0x7f800000) : 0.0f;
}
__device__ __forceinline__ float query_t_hi(const Query& q) {
    return __double2float_ru(q.t);
}

__device__ __forceinline__ void accept(Query& q, const DScene& S, int leaf, uint32_t flags, int n, const double* ts) {
    if (q.mode == Q_CLOSEST) {
        for (int i = 0; i < n; i++) {
            double t = ts[i];
            if (t >= 0.0 && (t < q.t || (t == q.t && leaf < q.leaf))) {
                q.t = t;
                q.leaf = leaf;
            }
        }
    } else if (q.mode == Q_SHADOW) {
        if (flags & RTC_CASTS_SHADOW)
            for (int i = 0; i < n; i++)
                if (ts[i] >= 0.0 && ts[i] < q.t) q.flag = 1;
    } else {
        cont_add(q, leaf, n, ts);
    }
}

// Top-level analytic leaves (and stand-alone triangles), two passes per chunk of 32:
//   prim_mask   conservative f32 world-box test of every prim -> bit mask (cheap, all lanes in step);
//   trace_prim  the reference's transform + closed-form intersection of ONE candidate.
// Each lane then walks its own candidate list, so lanes testing different prims still share the
// intersection code instead of idling while the warp steps through all prims in order.
__device__ __forceinline__ unsigned prim_mask(const Query& q, const DScene& S, const FRay& wr, int base) {
    unsigned mask = 0;
    const float t_lo = query_t_lo(q), t_hi = query_t_hi(q);
    const int n = min(32, S.n_prims - base);
#pragma unroll 1
    for (int i = 0; i < n; i++) {
        const float4* pp = reinterpret_cast<const float4*>(S.prims + base + i);
        float4 p0 = __ldg(pp), p1 = __ldg(pp + 1);
        uint32_t flags = __float_as_uint(p1.w);
        if (q.mode == Q_SHADOW && !(flags & RTC_CASTS_SHADOW)) continue;
        float tn;
        if (fbox(wr, p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, t_lo, t_hi, tn)) mask |= 1u << i;
    }
    return mask;
}

// LEAN (rtc_api.cu lean_scene): no guards and no stray triangles among the prims.
template <bool LEAN>
__device__ __forceinline__ void trace_prim(Query& q, const DScene& S, int i) {
    int leaf_id = S.prims[i].leaf;
    uint32_t flags = S.prims[i].flags;
    const rtc_leaf& L = S.leaves[leaf_id];
    if constexpr (!LEAN) {
        for (int g = L.guard; g >= 0;) {  // group.zig:46-50 for boxes that are not supersets
            const rtc_guard& G = S.guards[g];
            double a, b;
            if (!ref_slabs(q.o, q.d, G.bmin, G.bmax, a, b)) return;
            g = G.next;
        }
    }
    M34 inv = load_m34(S.transforms[L.transform].inverse);
    V3 oo = xf_point(inv, q.o), dd = xf_vec(inv, q.d);  // ray.zig:30-32
    double ts[4];
    int n;
    int kind = L.kind;
    if (!LEAN && (kind == RTC_TRIANGLE || kind == RTC_SMOOTH_TRIANGLE)) {
        const rtc_triangle& T = S.blob_tris[L.triangle];
        double u, v;
        n = tri_intersect(v3(T.p1[0], T.p1[1], T.p1[2]), v3(T.e1[0], T.e1[1], T.e1[2]), v3(T.e2[0], T.e2[1], T.e2[2]),
                          oo, dd, ts[0], u, v)
                ? 1
                : 0;
        const double t_before = q.t;
        const int leaf_before = q.leaf;
        accept(q, S, leaf_id, flags, n, ts);
        if (q.mode == Q_CLOSEST && (q.t != t_before || q.leaf != leaf_before)) {  // this triangle is the new closest hit
            q.a = u;
            q.b = v;
            q.c_leaf = leaf_id;  // whose u,v these are (the shading stage recomputes them for any other smooth triangle)
        }
        return;
    }
    n = leaf_intersect(kind, flags, L.minimum, L.maximum, oo, dd, ts);
    accept(q, S, leaf_id, flags, n, ts);
}

// ---- meshes ---------------------------------------------------------------------------------------
__device__ __forceinline__ void mesh_tris(Query& q, const DScene& S, int first, int count, V3 oo, V3 dd,
                                          const int32_t* __restrict__ lmap) {
    for (int i = 0; i < count; i++) {
        const double* p = reinterpret_cast<const double*>(S.tris + first + i);
        double2 a = ld2(p), b = ld2(p + 2), c = ld2(p + 4), e = ld2(p + 6), f = ld2(p + 8);
        double t, u, v;
        // (deciding u < 0, u > 1, v < 0, u + v > 1 on the undivided quantities first, with margins inside which the
        // reference's own operations decide, was measured on B200: dragons 0.871 ms against 0.849 with the plain test
        // -- the division is not what the time goes into; profiles/README.md)
        if (!tri_intersect(v3(a.x, a.y, b.x), v3(b.y, c.x, c.y), v3(e.x, e.y, f.x), oo, dd, t, u, v)) continue;
        const int rank = (int)(__double_as_longlong(f.y) & 0xffffffffLL);  // DTri.leaf: rank in the geometry
        if (q.mode == Q_CLOSEST) {
            if (t >= 0.0 && t <= q.t) {
                const int leaf = __ldg(lmap + rank);  // the instance's rtc_leaf: read only for a candidate
                if (t < q.t || leaf < q.leaf) {
                    q.t = t;
                    q.leaf = leaf;
                    q.a = u;
                    q.b = v;
                    q.c_leaf = leaf;  // whose u,v these are
                }
            }
        } else if (q.mode == Q_SHADOW) {
            if (t >= 0.0 && t < q.t) q.flag = 1;
        } else {
            cont_add(q, __ldg(lmap + rank), 1, &t);
        }
    }
}

// Work sharing inside the mesh traversal (trace_meshes, below).
// MEASURED on B200 (profiles/README.md, round 2): correct (frames and ray counts bit-identical, 70 GPU tests green) and
// it does raise the lanes active per instruction of this loop (10.5 -> 14.1), but the frame gets SLOWER -- dragons.json
// 1080p 0.87 -> 1.24 ms, teapot 0.66 -> 0.82, a 1/8 partition 0.27 -> 0.35 -- and stays slower when helping is made
// rare (only shadow queries, only with >= 28 idle lanes and >= 2 pending entries: 1.14 ms): the warp executes MORE
// instructions (578 M against 505 M: subtrees a closest-hit query would have culled, the bookkeeping at every step) and
// the loop no longer fits the instruction cache (stall_no_instruction 1.0 -> 2.5 per issue; 52 bytes of spills).  This
// kernel lives on a small, register-tight inner loop.  Off by default; -DRTC_STEAL=1 builds it.
// (the default, RTC_STEAL 0, is set in render_params.h: the host sizes shared memory by it)
#if RTC_STEAL
#define RTC_STACK_BASE(T) ((T).sb)
#else
#define RTC_STACK_BASE(T) 0
#endif

// Mesh traversal state of one lane.  It lives across trips of the render loop (render.cu): a lane
// whose query is finished leaves to be shaded and refilled while the others keep their place.
//   node >= 0        inner BvhNode index
//   node <  0        leaf: ~node = (first DTri << 3) | count       (count <= 4, first < 2^27)
//   node == MESH_END the traversal stack is empty: advance to the next candidate mesh
#define RTC_MESH_END ((int)0x80000000)
struct MeshTrav {
    int mbase;      // meshes [mbase, mbase + 32) are described by mask
    unsigned mask;  // candidate meshes not visited yet
    int node, sp;
#if RTC_STEAL
    int sb;         // the lane's stack is [sb, sp): entries below sb were handed to other lanes (trace_meshes)
#endif
    int lmap;       // DMesh.leaf_map of the mesh being traversed
    float t_hi;
    V3 oo, dd;  // the ray in the current mesh's object space
    FRay r;
};

// Candidate meshes among [mbase, mbase+32): world box reachable within the query's t-range.
__device__ __forceinline__ unsigned mesh_mask(const Query& q, const DScene& S, const FRay& wr, int mbase) {
    unsigned mask = 0;
    float t_lo = query_t_lo(q), t_hi = query_t_hi(q);
    int n = min(32, S.n_meshes - mbase);
#pragma unroll 1
    for (int i = 0; i < n; i++) {
        const float4* mp = reinterpret_cast<const float4*>(S.meshes + mbase + i);
        int4 hd = __ldg(reinterpret_cast<const int4*>(mp));
        if (q.mode == Q_SHADOW && !((uint32_t)hd.y & RTC_CASTS_SHADOW)) continue;
        float4 w0 = __ldg(mp + 4), w1 = __ldg(mp + 5);
        float tn;
        if (fbox(wr, w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, t_lo, t_hi, tn)) mask |= 1u << i;
    }
    return mask;
}

__device__ __forceinline__ void mesh_begin(const Query& q, const DScene& S, const FRay& wr, MeshTrav& T) {
    T.mbase = 0;
    T.mask = mesh_mask(q, S, wr, 0);
    T.node = RTC_MESH_END;
    T.sp = 0;
#if RTC_STEAL
    T.sb = 0;
#endif
    T.t_hi = query_t_hi(q);
}

// The "while-while" traversal, split at the point where the warp must reconverge (render.cu puts a
// full-mask ballot between the two halves):
//   mesh_inner  walks inner nodes -- opening the next candidate mesh when the stack runs dry -- until
//               the lane holds a leaf; returns false when the query has no mesh work left;
//   mesh_leaf   runs the reference's triangle test on the held leaf and pops the next node.
// Because the halves alternate warp-wide, the f64 Möller–Trumbore runs with every traversing lane
// active instead of the few that reach a leaf in the same step; and because all meshes go through
// this one loop, lanes inside different meshes and at different depths share the code.
// stack: this thread's column of the shared-memory traversal stack (stride RTC_BLOCK)
__device__ __forceinline__ bool mesh_inner(const Query& q, const DScene& S, MeshTrav& T, int* stack, int* err) {
    const float t_lo = query_t_lo(q);
    int node = T.node, sp = T.sp;
    for (;;) {
        if (node == RTC_MESH_END) {
            // ---- next candidate mesh: ray to object space (ray.zig:30-32, once per mesh) ----------
            if (T.mask == 0) {
                if (T.mbase + 32 >= S.n_meshes) return false;
                T.mbase += 32;
                T.mask = mesh_mask(q, S, make_fray(q.o, q.d, S.world_abs[0], S.world_abs[1], S.world_abs[2]), T.mbase);
                continue;
            }
            int b = __ffs(T.mask) - 1;
            T.mask &= T.mask - 1;
            const DMesh& M = S.meshes[T.mbase + b];
            M34 inv = load_m34(S.transforms[M.transform].inverse);
            T.oo = xf_point(inv, q.o), T.dd = xf_vec(inv, q.d);
            T.r = make_fray(T.oo, T.dd, M.bound_abs[0], M.bound_abs[1], M.bound_abs[2]);
            T.lmap = M.leaf_map;
            float tn;
            if (!fbox(T.r, M.lo[0], M.lo[1], M.lo[2], M.hi[0], M.hi[1], M.hi[2], t_lo, T.t_hi, tn)) continue;
            node = M.root;
        }
        const float t_hi = T.t_hi;
        while (node >= 0) {
            const float4* np = reinterpret_cast<const float4*>(S.bvh + node);
            float4 q0 = __ldg(np), q1 = __ldg(np + 1), q2 = __ldg(np + 2);
            int2 ci = __ldg(reinterpret_cast<const int2*>(np + 3));  // child references, leaves pre-encoded (scene_build.cpp)
            float n0, n1;
            bool h0 = fbox(T.r, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, t_lo, t_hi, n0);
            bool h1 = fbox(T.r, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, t_lo, t_hi, n1);
            const int c0 = ci.x, c1 = ci.y;
            if (h0 && h1) {
                bool swap = n1 < n0;  // nearer child first: it can shrink t_hi before the other is opened
                int nearc = swap ? c1 : c0, farc = swap ? c0 : c1;
                if (sp >= S.stack_cap) {
                    *err |= ERR_BVH_STACK;
                    return false;
                }
                stack[sp * RTC_BLOCK] = farc;
                sp++;
                node = nearc;
            } else if (h0) {
                node = c0;
            } else if (h1) {
                node = c1;
            } else if (sp > RTC_STACK_BASE(T)) {
                sp--;
                node = stack[sp * RTC_BLOCK];
            } else {
                node = RTC_MESH_END;
            }
        }
        if (node != RTC_MESH_END) break;  // holding a leaf
    }
    T.node = node;
    T.sp = sp;
    return true;
}

__device__ __forceinline__ bool mesh_leaf(Query& q, const DScene& S, MeshTrav& T, int* stack) {
    int enc = ~T.node;
    mesh_tris(q, S, enc >> 3, enc & 7, T.oo, T.dd, S.mesh_leaf + T.lmap);
    if (q.mode == Q_SHADOW && q.flag) return false;
    if (q.mode == Q_CLOSEST) T.t_hi = fminf(T.t_hi, __double2float_ru(q.t));  // (T.t_hi may hold a tighter bound found by a helper)
    if (T.sp > RTC_STACK_BASE(T)) {
        T.sp--;
        T.node = stack[T.sp * RTC_BLOCK];
    } else {
        T.node = RTC_MESH_END;
    }
    return true;
}

// ---- the mesh traversal of one warp trip, with work sharing between the lanes ---------------------------------------
// The lanes of a warp need very different numbers of steps for their queries (a ray that misses every mesh box needs
// none, one that grazes a dragon hundreds), and the trip lasts as long as the slowest: measured on dragons.json, 10 of
// 32 lanes were active per instruction of this loop.  So a lane that is through HELPS: at the point where the warp
// reconverges anyway, idle lanes are matched (ballot + popc ranks) with lanes that still have pending subtrees on their
// traversal stacks; the helper copies the donor's object-space ray by shuffles, takes the BOTTOM entry of the donor's
// stack (the largest pending subtree) and walks it as a query of its own.  Results meet in per-lane "mailboxes" in
// shared memory: closest hit = min over (t, leaf) -- associative, so the result is the same whoever finds what: exact;
// shadow = any hit.  Bounds travel too: a hit found by anyone working for a query shrinks t_hi of all of them (one
// shared-memory atomicMin per step), a shadow hit stops them.  Containers queries (parity of ALL hits) are not shared.
#ifndef RTC_STEAL_MIN_DEPTH
#define RTC_STEAL_MIN_DEPTH 1  // pending entries a lane must have on its stack before it gives the bottom one away
#endif
#ifndef RTC_STEAL_MIN_IDLE
#define RTC_STEAL_MIN_IDLE 1   // idle lanes a warp must have before any of them helps
#endif
#ifndef RTC_STEAL_MODES
#define RTC_STEAL_MODES 3      // bit 0: shadow queries are shared, bit 1: closest-hit queries are shared
#endif
enum { MB_T = 0, MB_LF = 1, MB_U = 2, MB_V = 3, MB_BOUND = 4, MB_SLOTS = RTC_MAILBOX_SLOTS };  // doubles per lane (stride RTC_BLOCK)
#define RTC_UV_NONE (-2)  // Query::c_leaf of a helper that has not found anything yet

// q, a finished Q_CLOSEST / Q_SHADOW task, merged into the mailbox of lane `tgt` (mbw = the warp's lane-0 column)
__device__ __forceinline__ void mb_merge(double* mbw, int tgt, const Query& q) {
    const double lf = mbw[MB_LF * RTC_BLOCK + tgt];  // hi word: flag, lo word: leaf
    if (q.mode == Q_CLOSEST) {
        if (q.leaf >= 0 && q.c_leaf != RTC_UV_NONE) {
            const double mt = mbw[MB_T * RTC_BLOCK + tgt];
            const int ml = __double2loint(lf);
            if (ml < 0 || q.t < mt || (q.t == mt && q.leaf < ml)) {
                mbw[MB_T * RTC_BLOCK + tgt] = q.t;
                mbw[MB_LF * RTC_BLOCK + tgt] = __hiloint2double(q.c_leaf, q.leaf);
                mbw[MB_U * RTC_BLOCK + tgt] = q.a;
                mbw[MB_V * RTC_BLOCK + tgt] = q.b;
            }
        }
    } else if (q.flag) {
        mbw[MB_LF * RTC_BLOCK + tgt] = __hiloint2double(1, __double2loint(lf));
    }
}

// Walks the meshes for the whole warp.  busy: this lane has mesh work set up in T (trace_begin).  own_live: q is a live
// query of this lane (its result matters).  stack: this thread's column; mbw: the warp's lane-0 mailbox column.
__device__ __forceinline__ void trace_meshes(Query& q, const DScene& S, MeshTrav& T, int* stack, double* mbw, int* err,
                                             bool busy, bool own_live, unsigned lane) {
    const unsigned FULL = 0xffffffffu;
#if RTC_STEAL
    const unsigned lt_mask = (1u << lane) - 1u;
    const int own_mode = q.mode;
    const bool can_help = !(own_live && own_mode == Q_CONTAINERS);  // its four result fields do not fit a mailbox
    bool shared = false;   // warp-uniform: mailboxes are in use in this trip
    bool own_in_q = true;  // q still is this lane's OWN query
    bool dirty = false;    // q holds a helper's task that is not in its owner's mailbox yet
    int owner = (int)lane; // whose query the lane works on
    // finished helpers put their results into their owners' mailboxes, one lane per mailbox at a time
    auto flush = [&](bool want, int tgt) {
        unsigned pending = __ballot_sync(FULL, want);
        while (pending) {
            const unsigned grp = __match_any_sync(FULL, want ? tgt : 32 + (int)lane);
            if (want && (int)lane == __ffs(grp) - 1) {
                mb_merge(mbw, tgt, q);
                want = false;
            }
            __syncwarp();
            pending = __ballot_sync(FULL, want);
        }
    };
#endif
#if !RTC_STEAL
    while (__any_sync(FULL, busy)) {
        if (busy) busy = mesh_inner(q, S, T, stack, err);
        if (!__any_sync(FULL, busy)) break;
        if (busy) busy = mesh_leaf(q, S, T, stack);
    }
    (void)mbw;
    (void)own_live;
    (void)lane;
#else
    while (__any_sync(FULL, busy)) {
        if (busy) busy = mesh_inner(q, S, T, stack, err);
        bool at_leaf = busy;  // holding a leaf for mesh_leaf (a lane that takes over a subtree below holds an inner node)
#if RTC_STEAL
        const unsigned idle = __ballot_sync(FULL, !busy && can_help);
        if (__popc(idle) >= RTC_STEAL_MIN_IDLE) {
            bool can_give = false;
            if (busy && (q.mode == Q_SHADOW ? (RTC_STEAL_MODES & 1) != 0 : q.mode == Q_CLOSEST && (RTC_STEAL_MODES & 2) != 0) &&
                T.sp - T.sb >= RTC_STEAL_MIN_DEPTH)
                can_give = stack[T.sb * RTC_BLOCK] >= 0;  // an inner node
            const unsigned donors = __ballot_sync(FULL, can_give);
            if (donors) {
                if (!shared) {
                    mbw[MB_T * RTC_BLOCK + lane] = RTC_INF;
                    mbw[MB_LF * RTC_BLOCK + lane] = __hiloint2double(0, -1);
                    reinterpret_cast<unsigned*>(mbw + MB_BOUND * RTC_BLOCK + lane)[0] = 0x7f800000u;  // +inf
                    shared = true;
                    __syncwarp();
                }
                const int n_d = __popc(donors), n_i = __popc(idle);
                const int my_rank = __popc(idle & lt_mask);
                const bool take = !busy && can_help && my_rank < n_d;
                const bool give = can_give && __popc(donors & lt_mask) < n_i;
                // what a taker holds goes to its mailbox first: its own finished query, or the task it helped with
                flush(take && (own_in_q ? own_live : dirty), owner);
                const int src = take ? (int)__fns(donors, 0, my_rank + 1) : (int)lane;
                const int d_sb = __shfl_sync(FULL, T.sb, src);
                const int d_owner = __shfl_sync(FULL, owner, src);
                const int d_mode = __shfl_sync(FULL, q.mode, src);
                const int d_leaf = __shfl_sync(FULL, q.leaf, src);
                const int d_lmap = __shfl_sync(FULL, T.lmap, src);
                const double d_t = __shfl_sync(FULL, q.t, src);
                const float d_thi = __shfl_sync(FULL, T.t_hi, src);
                const double ox = __shfl_sync(FULL, T.oo.x, src), oy = __shfl_sync(FULL, T.oo.y, src), oz = __shfl_sync(FULL, T.oo.z, src);
                const double dx = __shfl_sync(FULL, T.dd.x, src), dy = __shfl_sync(FULL, T.dd.y, src), dz = __shfl_sync(FULL, T.dd.z, src);
                FRay r;
                r.ix = __shfl_sync(FULL, T.r.ix, src), r.iy = __shfl_sync(FULL, T.r.iy, src), r.iz = __shfl_sync(FULL, T.r.iz, src);
                r.nx = __shfl_sync(FULL, T.r.nx, src), r.ny = __shfl_sync(FULL, T.r.ny, src), r.nz = __shfl_sync(FULL, T.r.nz, src);
                r.fx = __shfl_sync(FULL, T.r.fx, src), r.fy = __shfl_sync(FULL, T.r.fy, src), r.fz = __shfl_sync(FULL, T.r.fz, src);
                if (give) T.sb++;
                __syncwarp();
                if (take) {
                    T.node = (stack - lane + src)[d_sb * RTC_BLOCK];
                    T.sp = T.sb = 0;
                    T.mask = 0;
                    T.mbase = S.n_meshes;  // this subtree only: no further candidate meshes
                    T.lmap = d_lmap;
                    T.t_hi = d_thi;
                    T.oo = v3(ox, oy, oz);
                    T.dd = v3(dx, dy, dz);
                    T.r = r;
                    q.mode = d_mode;
                    q.t = d_t;        // the bound (and, with the leaf, the tie-break) the donor has reached
                    q.leaf = d_leaf;
                    q.c_leaf = RTC_UV_NONE;
                    q.flag = 0;
                    owner = d_owner;
                    own_in_q = false;
                    dirty = true;
                    busy = true;
                    at_leaf = false;
                }
            }
        }
        if (shared) {
            flush(!busy && dirty, owner);  // helpers that are through
            if (!busy) dirty = false;
            if (busy && q.mode == Q_CLOSEST) {  // the nearest hit anyone has found for this query bounds everyone's search
                unsigned* bp = reinterpret_cast<unsigned*>(mbw + MB_BOUND * RTC_BLOCK + owner);
                const unsigned mine = __float_as_uint(T.t_hi);  // t_hi >= 0: unsigned order is float order
                const unsigned old = atomicMin(bp, mine);
                T.t_hi = __uint_as_float(min(old, mine));
            } else if (busy && q.mode == Q_SHADOW) {
                if (__double2hiint(mbw[MB_LF * RTC_BLOCK + owner])) busy = false;  // someone found a hit: shadowed
            }
        }
#endif
        if (!__any_sync(FULL, busy)) break;
        if (busy && at_leaf) busy = mesh_leaf(q, S, T, stack);
    }
#if RTC_STEAL
    if (shared) {
        flush(dirty, owner);
        __syncwarp();
        if (own_live) {
            const double lf = mbw[MB_LF * RTC_BLOCK + lane];
            const int ml = __double2loint(lf), mf = __double2hiint(lf);
            if (own_mode == Q_CLOSEST) {
                const double mt = mbw[MB_T * RTC_BLOCK + lane];
                // own_in_q: the lane's own search result is still in q -- the better of the two; else the mailbox has all
                const bool from_mb = !own_in_q || (ml >= 0 && (q.leaf < 0 || mt < q.t || (mt == q.t && ml < q.leaf)));
                if (!own_in_q && ml < 0) {
                    q.t = RTC_INF;
                    q.leaf = -1;
                    q.a = q.b = 0.0;
                    q.c_leaf = 0;
                } else if (from_mb) {
                    q.t = mt;
                    q.leaf = ml;
                    q.c_leaf = mf;
                    q.a = mbw[MB_U * RTC_BLOCK + lane];
                    q.b = mbw[MB_V * RTC_BLOCK + lane];
                }
            } else if (own_mode == Q_SHADOW) {
                if (!own_in_q) q.flag = mf;
                else if (mf) q.flag = 1;
            }
            q.mode = own_mode;
        }
        __syncwarp();
    }
#endif
#endif
}

// ---- CSG ------------------------------------------------------------------------------------------
// csg.zig:115-121
__device__ __forceinline__ bool csg_allowed(int op, bool lhit, bool inl, bool inr) {
    if (op == RTC_CSG_UNION) return (lhit && !inr) || !(lhit || inl);
    if (op == RTC_CSG_INTERSECTION) return (lhit && inr) || (!lhit && inl);
    return (lhit && !inr) || (!lhit && inl);
}

// Runs the program of one top-level CSG object: Csg.localIntersect (csg.zig:75-96) for the whole
// subtree.  Leaves the filtered, sorted list in ts/ids and returns its length.
static __device__ __noinline__ int csg_eval(const DScene& S, int pc, V3 o, V3 d, double* ts, int* ids, int* err) {
    int seg[RTC_CSG_SEG];
    int nseg = 0, len = 0;
    for (;;) {
        int4 w = __ldg(S.csg_prog + pc);
        if (w.x == CSG_OP_END) break;
        if (w.x == CSG_OP_LEAF) {
            seg[nseg++] = len;
            const rtc_leaf& L = S.leaves[w.y];
            M34 inv = load_m34(S.transforms[L.transform].inverse);
            V3 oo = xf_point(inv, o), dd = xf_vec(inv, d);
            double lt[4];
            int n;
            if (L.kind == RTC_TRIANGLE || L.kind == RTC_SMOOTH_TRIANGLE) {
                const rtc_triangle& T = S.blob_tris[L.triangle];
                double u, v;
                n = tri_intersect(v3(T.p1[0], T.p1[1], T.p1[2]), v3(T.e1[0], T.e1[1], T.e1[2]),
                                  v3(T.e2[0], T.e2[1], T.e2[2]), oo, dd, lt[0], u, v)
                        ? 1
                        : 0;
            } else {
                n = leaf_intersect(L.kind, L.flags, L.minimum, L.maximum, oo, dd, lt);
            }
            if (len + n > RTC_CSG_CAP) {
                *err |= ERR_CSG_CAP;
                n = 0;
            }
            for (int i = 0; i < n; i++) {
                ts[len] = lt[i];
                ids[len] = w.y;
                len++;
            }
        } else if (w.x == CSG_OP_BBOX) {  // csg.zig:80-84 / group.zig:46-50
            const rtc_node& N = S.nodes[w.y];
            double a, b;
            if (!ref_slabs(o, d, N.bmin, N.bmax, a, b)) {
                seg[nseg++] = len;
                pc = w.z;
                continue;
            }
        } else if (w.x == CSG_OP_MERGE) {
            if (w.y == 0) seg[nseg++] = len;
            else nseg -= (w.y - 1);
        } else {  // CSG_OP_FILTER : sort (csg.zig:94) then filterIntersections (csg.zig:51-73)
            nseg -= 1;
            int start = seg[nseg - 1];
            for (int i = start + 1; i < len; i++) {  // stable insertion sort by t
                double kt = ts[i];
                int kid = ids[i];
                int j = i;
                while (j > start && kt < ts[j - 1]) {
                    ts[j] = ts[j - 1];
                    ids[j] = ids[j - 1];
                    j--;
                }
                ts[j] = kt;
                ids[j] = kid;
            }
            bool inl = false, inr = false;
            int wr = start;
            for (int i = start; i < len; i++) {
                int id = ids[i];
                bool lhit = id >= w.z && id < w.w;  // includes(left, leaf) (csg.zig:123-140)
                if (csg_allowed(w.y, lhit, inl, inr)) {
                    ts[wr] = ts[i];
                    ids[wr] = id;
                    wr++;
                }
                if (lhit) inl = !inl;
                else inr = !inr;
            }
            len = wr;
        }
        pc++;
    }
    return len;
}

__device__ __forceinline__ void trace_csg(Query& q, const DScene& S, const DCsgRoot& R, int* err) {
    double ts[RTC_CSG_CAP];
    int ids[RTC_CSG_CAP];
    int len = csg_eval(S, R.pc, q.o, q.d, ts, ids, err);
    if (q.mode == Q_CLOSEST) {
        for (int i = 0; i < len; i++) {  // sorted: the first t >= 0 is this object's candidate
            if (ts[i] >= 0.0) {
                if (ts[i] < q.t || (ts[i] == q.t && ids[i] < q.leaf)) {
                    q.t = ts[i];
                    q.leaf = ids[i];  // (a smooth triangle as CSG operand: q.c_leaf != q.leaf, u,v are recomputed by the caller)
                }
                break;
            }
        }
    } else if (q.mode == Q_SHADOW) {
        for (int i = 0; i < len; i++)
            if (ts[i] >= 0.0 && ts[i] < q.t && (S.leaves[ids[i]].flags & RTC_CASTS_SHADOW)) q.flag = 1;
    } else {
        // per-leaf parity over the filtered list: visit each leaf at its first entry
        for (int i = 0; i < len; i++) {
            int leaf = ids[i];
            bool first = true;
            for (int j = 0; j < i; j++)
                if (ids[j] == leaf) first = false;
            if (!first) continue;
            double lt[RTC_CSG_CAP];
            int n = 0;
            for (int j = i; j < len; j++)
                if (ids[j] == leaf) lt[n++] = ts[j];
            cont_add(q, leaf, n, lt);
        }
    }
}

// Starts one query.  Called by ALL 32 lanes of the warp (live = this lane has a query): analytic
// prims and CSG objects are finished here; the meshes are set up in T and walked by mesh_inner /
// mesh_leaf.  Returns true when mesh work remains.
// Prims: each lane's candidate mask is OR-reduced over the warp and the warp steps through the union
// in prim order -- prims no lane can reach cost nothing, and at every step the participating lanes
// run the SAME prim (same kind: no divergence inside the intersection switch).
template <bool MESH, bool LEAN>
__device__ __forceinline__ bool trace_begin(Query& q, const DScene& S, MeshTrav& T, int* err, bool live) {
    FRay wr = make_fray(q.o, q.d, S.world_abs[0], S.world_abs[1], S.world_abs[2]);
    for (int base = 0; base < S.n_prims; base += 32) {
        unsigned mine = live ? prim_mask(q, S, wr, base) : 0u;
        unsigned todo = __reduce_or_sync(0xffffffffu, mine);
        while (todo) {
            int i = __ffs(todo) - 1;
            todo &= todo - 1;
            if (((mine >> i) & 1u) && !(q.mode == Q_SHADOW && q.flag)) trace_prim<LEAN>(q, S, base + i);
        }
    }
    if (!live) return false;
    if constexpr (!LEAN) {
        for (int i = 0; i < S.n_csg; i++) {
            if (q.mode == Q_SHADOW && q.flag) return false;
            trace_csg(q, S, S.csg_roots[i], err);
        }
    }
    if constexpr (!MESH) return false;
    if (S.n_meshes == 0 || (q.mode == Q_SHADOW && q.flag)) return false;
    mesh_begin(q, S, wr, T);
    return T.mask != 0 || S.n_meshes > 32;
}

}  // namespace rtc
