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

#define RTC_BLOCK 128
#define RTC_BVH_STACK 64
#define RTC_CSG_CAP 32
#define RTC_CSG_SEG 16

enum { Q_CLOSEST = 0, Q_SHADOW = 1, Q_CONTAINERS = 2 };
enum { ERR_TEXEL = 1, ERR_PAT_STACK = 2, ERR_CSG_CAP = 4, ERR_BVH_STACK = 8 };

struct Hit {
    double t;
    int leaf;  // -1 = none
    double u, v;
};

// The containers walk reduced to a streaming form.  Processing the sorted list up to the hit, an
// object is in `containers` iff it has an ODD number of intersections before the hit, and it sits
// at the position of its LAST such intersection (remove-or-append, world.zig:240-249).  So:
//   last container before the hit = the odd-parity leaf whose last pre-hit intersection has the
//   greatest (t, leaf order) key; n1 is its refractive index (1.0 if none);
//   the hit toggles its own leaf: if it was inside (odd) it is removed and n2 comes from the best
//   OTHER leaf (or stays 1.0: the reference does not break on an empty list), else it is appended
//   last and n2 is its own index.
struct Containers {
    double hit_t;
    int hit_leaf;
    double c_t;   // best candidate among leaves other than the hit leaf
    int c_leaf;   // -1 = none
    int hit_odd;  // hit leaf has an odd number of intersections before the hit
    double hit_last_t;

    __device__ __forceinline__ void add(int leaf, int n, const double* ts) {
        int cnt = 0;
        double tl = -RTC_INF;
        for (int i = 0; i < n; i++) {
            bool before = ts[i] < hit_t || (ts[i] == hit_t && leaf < hit_leaf);
            if (before) {
                cnt++;
                tl = fmax(tl, ts[i]);
            }
        }
        if (cnt & 1) {
            if (leaf == hit_leaf) {
                hit_odd = 1;
                hit_last_t = tl;
            } else if (c_leaf < 0 || tl > c_t || (tl == c_t && leaf > c_leaf)) {
                c_t = tl;
                c_leaf = leaf;
            }
        }
    }
};

struct Query {
    int mode;
    V3 o, d;       // world-space ray
    V3 inv_d;      // 1/d, for the conservative box pre-tests
    double t_max;  // Q_SHADOW: distance to the light
    Hit best;      // Q_CLOSEST result
    bool shadowed; // Q_SHADOW result
    Containers cont;
};

// Conservative world-space box pre-test (an optimisation the reference does not have; the boxes are
// supersets, so it only skips work).  Returns the entry/exit parameters.
__device__ __forceinline__ bool box_range(const Query& q, const double mn[3], const double mx[3], double& tn,
                                          double& tf) {
    double ax = (mn[0] - q.o.x) * q.inv_d.x, bx = (mx[0] - q.o.x) * q.inv_d.x;
    double ay = (mn[1] - q.o.y) * q.inv_d.y, by = (mx[1] - q.o.y) * q.inv_d.y;
    double az = (mn[2] - q.o.z) * q.inv_d.z, bz = (mx[2] - q.o.z) * q.inv_d.z;
    tn = fmax(fmin(ax, bx), fmax(fmin(ay, by), fmin(az, bz)));
    tf = fmin(fmax(ax, bx), fmin(fmax(ay, by), fmax(az, bz)));
    double slack = 1e-9 * (fabs(tn) + fabs(tf));
    return !(tn > tf + slack);
}
__device__ __forceinline__ bool range_useful(const Query& q, double tn, double tf) {
    double slack = 1e-9 * (fabs(tn) + fabs(tf)) + 1e-12;
    if (q.mode == Q_CLOSEST) return !(tf + slack < 0.0) && !(tn - slack > q.best.t);
    if (q.mode == Q_SHADOW) return !(tf + slack < 0.0) && !(tn - slack > q.t_max);
    return !(tn - slack > q.cont.hit_t);
}

__device__ __forceinline__ void accept(Query& q, const DScene& S, int leaf, uint32_t flags, int n, const double* ts) {
    if (q.mode == Q_CLOSEST) {
        for (int i = 0; i < n; i++) {
            double t = ts[i];
            if (t >= 0.0 && (t < q.best.t || (t == q.best.t && leaf < q.best.leaf))) {
                q.best.t = t;
                q.best.leaf = leaf;
            }
        }
    } else if (q.mode == Q_SHADOW) {
        if (flags & RTC_CASTS_SHADOW)
            for (int i = 0; i < n; i++)
                if (ts[i] >= 0.0 && ts[i] < q.t_max) q.shadowed = true;
    } else {
        q.cont.add(leaf, n, ts);
    }
}

// One analytic leaf or stand-alone triangle at the top level.
__device__ __forceinline__ void trace_prim(Query& q, const DScene& S, int leaf_id) {
    const rtc_leaf& L = S.leaves[leaf_id];
    uint32_t flags = L.flags;
    if (q.mode == Q_SHADOW && !(flags & RTC_CASTS_SHADOW)) return;
    if (!(flags & RTC_UNBOUNDED)) {
        double tn, tf;
        if (!box_range(q, L.wmin, L.wmax, tn, tf) || !range_useful(q, tn, tf)) return;
    }
    for (int g = L.guard; g >= 0;) {  // group.zig:46-50 for boxes that are not supersets
        const rtc_guard& G = S.guards[g];
        double a, b;
        if (!ref_slabs(q.o, q.d, G.bmin, G.bmax, a, b)) return;
        g = G.next;
    }
    M34 inv = load_m34(S.transforms[L.transform].inverse);
    V3 oo = xf_point(inv, q.o), dd = xf_vec(inv, q.d);  // ray.zig:30-32
    double ts[4];
    int n;
    int kind = L.kind;
    if (kind == RTC_TRIANGLE || kind == RTC_SMOOTH_TRIANGLE) {
        const rtc_triangle& T = S.blob_tris[L.triangle];
        double u, v;
        n = tri_intersect(v3(T.p1[0], T.p1[1], T.p1[2]), v3(T.e1[0], T.e1[1], T.e1[2]), v3(T.e2[0], T.e2[1], T.e2[2]),
                          oo, dd, ts[0], u, v)
                ? 1
                : 0;
    } else {
        n = leaf_intersect(kind, flags, L.minimum, L.maximum, oo, dd, ts);
    }
    accept(q, S, leaf_id, flags, n, ts);
}

// ---- meshes ---------------------------------------------------------------------------------------
struct FRay {
    float ox, oy, oz, ix, iy, iz, ex, ey, ez;
};
__device__ __forceinline__ bool fbox(const FRay& r, float lx, float ly, float lz, float hx, float hy, float hz,
                                     float t_lo, float t_hi, float& tnear) {
    float ax = (lx - r.ox) * r.ix, bx = (hx - r.ox) * r.ix;
    float ay = (ly - r.oy) * r.iy, by = (hy - r.oy) * r.iy;
    float az = (lz - r.oz) * r.iz, bz = (hz - r.oz) * r.iz;
    float tn = fmaxf(fmaxf(fminf(ax, bx) - r.ex, fminf(ay, by) - r.ey), fmaxf(fminf(az, bz) - r.ez, t_lo));
    float tf = fminf(fminf(fmaxf(ax, bx) + r.ex, fmaxf(ay, by) + r.ey), fminf(fmaxf(az, bz) + r.ez, t_hi));
    tnear = tn;
    return tn <= tf;
}

__device__ __forceinline__ void mesh_tris(Query& q, const DScene& S, int first, int count, V3 oo, V3 dd,
                                          uint32_t flags) {
    for (int i = 0; i < count; i++) {
        const double* p = reinterpret_cast<const double*>(S.tris + first + i);
        double2 a = ld2(p), b = ld2(p + 2), c = ld2(p + 4), e = ld2(p + 6), f = ld2(p + 8);
        double t, u, v;
        if (!tri_intersect(v3(a.x, a.y, b.x), v3(b.y, c.x, c.y), v3(e.x, e.y, f.x), oo, dd, t, u, v)) continue;
        int leaf = (int)(__double_as_longlong(f.y) & 0xffffffffLL);
        if (q.mode == Q_CLOSEST) {
            if (t >= 0.0 && (t < q.best.t || (t == q.best.t && leaf < q.best.leaf))) {
                q.best.t = t;
                q.best.leaf = leaf;
                q.best.u = u;
                q.best.v = v;
            }
        } else if (q.mode == Q_SHADOW) {
            if (t >= 0.0 && t < q.t_max) q.shadowed = true;
        } else {
            q.cont.add(leaf, 1, &t);
        }
    }
}

// All meshes of the world for one query, as ONE loop shared by the whole warp ("while-while"
// traversal): every lane walks inner nodes until it holds a leaf, the warp reconverges, and only
// then are triangles tested -- so the f64 Möller–Trumbore runs with most lanes active instead of the
// few that happen to reach a leaf in the same step.  A lane that finishes one mesh moves on to its
// next candidate mesh inside the same loop, so lanes inside different dragons still share code.
//   node >= 0        inner BvhNode index
//   node <  0        leaf: ~node = (first DTri << 3) | count       (count <= 4, first < 2^27)
//   node == MESH_END this lane's traversal stack is empty: advance to the next mesh
// stack: this thread's column of the shared-memory traversal stack (stride RTC_BLOCK)
#define RTC_MESH_END ((int)0x80000000)
__device__ __forceinline__ void trace_meshes(Query& q, const DScene& S, int* stack, int* err) {
    const int mode = q.mode;
    const float t_lo = mode == Q_CONTAINERS ? -__int_as_float(0x7f800000) : 0.0f;
    float t_hi = __double2float_ru(mode == Q_CLOSEST ? q.best.t : (mode == Q_SHADOW ? q.t_max : q.cont.hit_t));
    int mi = 0, sp = 0, node = RTC_MESH_END;
    V3 oo = q.o, dd = q.d;
    FRay r = {};
    for (;;) {
        if (node == RTC_MESH_END) {
            // ---- next mesh whose world box the ray can reach (skips the ray transform otherwise) ----
            for (; mi < S.n_meshes; mi++) {
                const DMesh& M = S.meshes[mi];
                if (mode == Q_SHADOW && !(M.flags & RTC_CASTS_SHADOW)) continue;
                double tn, tf;
                if (!box_range(q, M.wmin, M.wmax, tn, tf) || !range_useful(q, tn, tf)) continue;
                M34 inv = load_m34(S.transforms[M.transform].inverse);
                oo = xf_point(inv, q.o), dd = xf_vec(inv, q.d);  // ray.zig:30-32, once per mesh
                r.ox = (float)oo.x, r.oy = (float)oo.y, r.oz = (float)oo.z;
                // f32 reciprocal of the f32-rounded direction: relative error <= 2^-23, inside the bound below
                r.ix = 1.0f / (float)dd.x, r.iy = 1.0f / (float)dd.y, r.iz = 1.0f / (float)dd.z;
                // |error| of (b - o) * (1/d) evaluated in f32 is below 2^-21 (|b| + 2|o|) |1/d|  (BVH)
                const float k = 4.76837158e-7f;
                r.ex = fminf(FLT_MAX, k * (M.bound_abs[0] + 2.0f * fabsf(r.ox)) * fabsf(r.ix));
                r.ey = fminf(FLT_MAX, k * (M.bound_abs[1] + 2.0f * fabsf(r.oy)) * fabsf(r.iy));
                r.ez = fminf(FLT_MAX, k * (M.bound_abs[2] + 2.0f * fabsf(r.oz)) * fabsf(r.iz));
                if (mode == Q_CLOSEST) t_hi = __double2float_ru(q.best.t);
                float tnf;
                if (!fbox(r, M.lo[0], M.lo[1], M.lo[2], M.hi[0], M.hi[1], M.hi[2], t_lo, t_hi, tnf)) continue;
                node = M.root;
                mi++;
                break;
            }
            if (node == RTC_MESH_END) return;
        }
        // ---- inner nodes, until this lane holds a leaf or runs out of stack ----------------------
        while (node >= 0) {
            const float4* np = reinterpret_cast<const float4*>(S.bvh + node);
            float4 q0 = __ldg(np), q1 = __ldg(np + 1), q2 = __ldg(np + 2);
            int4 ci = __ldg(reinterpret_cast<const int4*>(np + 3));
            float n0, n1;
            bool h0 = fbox(r, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, t_lo, t_hi, n0);
            bool h1 = fbox(r, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, t_lo, t_hi, n1);
            int c0 = ci.z > 0 ? ~((ci.x << 3) | ci.z) : ci.x;
            int c1 = ci.w > 0 ? ~((ci.y << 3) | ci.w) : ci.y;
            if (h0 && h1) {
                bool swap = n1 < n0;  // nearer child first: it can shrink t_hi before the other is opened
                int nearc = swap ? c1 : c0, farc = swap ? c0 : c1;
                if (sp >= RTC_BVH_STACK) {
                    *err |= ERR_BVH_STACK;
                    return;
                }
                stack[sp * RTC_BLOCK] = farc;
                sp++;
                node = nearc;
            } else if (h0) {
                node = c0;
            } else if (h1) {
                node = c1;
            } else if (sp > 0) {
                sp--;
                node = stack[sp * RTC_BLOCK];
            } else {
                node = RTC_MESH_END;
            }
        }
        // ---- leaf: the reference's triangle test on every record of the leaf --------------------
        if (node != RTC_MESH_END) {
            int enc = ~node;
            mesh_tris(q, S, enc >> 3, enc & 7, oo, dd, 0);
            if (mode == Q_SHADOW && q.shadowed) return;
            if (mode == Q_CLOSEST) t_hi = __double2float_ru(q.best.t);
            if (sp > 0) {
                sp--;
                node = stack[sp * RTC_BLOCK];
            } else {
                node = RTC_MESH_END;
            }
        }
    }
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
__device__ __noinline__ int csg_eval(const DScene& S, int pc, V3 o, V3 d, double* ts, int* ids, int* err) {
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
                if (ts[i] < q.best.t || (ts[i] == q.best.t && ids[i] < q.best.leaf)) {
                    q.best.t = ts[i];
                    q.best.leaf = ids[i];
                }
                break;
            }
        }
    } else if (q.mode == Q_SHADOW) {
        for (int i = 0; i < len; i++)
            if (ts[i] >= 0.0 && ts[i] < q.t_max && (S.leaves[ids[i]].flags & RTC_CASTS_SHADOW)) q.shadowed = true;
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
            q.cont.add(leaf, n, lt);
        }
    }
}

// The whole world for one query.
__device__ __forceinline__ void trace(Query& q, const DScene& S, int* stack, int* err) {
    q.inv_d = v3(1.0 / q.d.x, 1.0 / q.d.y, 1.0 / q.d.z);
    for (int i = 0; i < S.n_prims; i++) {
        trace_prim(q, S, __ldg(S.prims + i));
    }
    if (S.n_meshes > 0) {
        if (q.mode == Q_SHADOW && q.shadowed) return;
        trace_meshes(q, S, stack, err);
    }
    for (int i = 0; i < S.n_csg; i++) {
        if (q.mode == Q_SHADOW && q.shadowed) return;
        trace_csg(q, S, S.csg_roots[i], err);
    }
}

}  // namespace rtc
