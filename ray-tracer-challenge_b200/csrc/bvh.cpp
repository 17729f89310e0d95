// Binned-SAH BVH over the triangles of one mesh (see bvh.hpp).
#include "bvh.hpp"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <future>

namespace rtc {
namespace {

// Triangles per leaf: the traversal encodes a leaf as (first << 3) | count, so at most 7.  RTC_BVH_LEAF=1..7 overrides
// the default for experiments (profiles/README.md: leaf size against kernel time).
const int kLeafMax = [] {
    const char* e = std::getenv("RTC_BVH_LEAF");
    int v = e ? std::atoi(e) : 4;
    return v >= 1 && v <= 7 ? v : 4;
}();
constexpr int kBins = 16;
constexpr int kSahDepthLimit = 40;  // beyond this depth split by median count (bounds total depth)

struct Box {
    double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    void grow(const double p[3]) {
        for (int a = 0; a < 3; a++) {
            lo[a] = std::fmin(lo[a], p[a]);
            hi[a] = std::fmax(hi[a], p[a]);
        }
    }
    void grow(const Box& b) {
        grow(b.lo);
        grow(b.hi);
    }
    double area() const {
        double dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        if (!(dx >= 0)) return 0.0;
        return 2.0 * (dx * dy + dy * dz + dz * dx);
    }
};

struct Ref {
    Box box;
    double c[3];
    int tri;
};

inline float down(double v) {  // largest float <= v
    float f = (float)v;
    if ((double)f > v) f = std::nextafterf(f, -INFINITY);
    return f;
}
inline float up(double v) {
    float f = (float)v;
    if ((double)f < v) f = std::nextafterf(f, INFINITY);
    return f;
}

struct Ctx {
    std::vector<Ref> refs;
    const std::vector<DTri>* src;
    BvhBuild out;

    Box bounds(int b, int e) const {
        Box x;
        for (int i = b; i < e; i++) x.grow(refs[i].box);
        return x;
    }

    // chooses a split of refs[b,e) and partitions; returns mid
    int split(int b, int e, int depth) {
        int n = e - b;
        Box cb;
        for (int i = b; i < e; i++) cb.grow(refs[i].c);
        int best_axis = -1, best_bin = -1;
        double best_cost = INFINITY;
        if (depth < kSahDepthLimit) {
            for (int a = 0; a < 3; a++) {
                double ext = cb.hi[a] - cb.lo[a];
                if (!(ext > 0)) continue;
                Box bb[kBins];
                int cnt[kBins] = {0};
                double k = kBins / ext;
                for (int i = b; i < e; i++) {
                    int bi = std::min(kBins - 1, std::max(0, (int)((refs[i].c[a] - cb.lo[a]) * k)));
                    bb[bi].grow(refs[i].box);
                    cnt[bi]++;
                }
                double ra[kBins];
                int rc[kBins];
                Box acc;
                int c = 0;
                for (int i = kBins - 1; i > 0; i--) {
                    acc.grow(bb[i]);
                    c += cnt[i];
                    ra[i] = acc.area();
                    rc[i] = c;
                }
                Box l;
                int lc = 0;
                for (int i = 0; i < kBins - 1; i++) {
                    l.grow(bb[i]);
                    lc += cnt[i];
                    if (lc == 0 || rc[i + 1] == 0) continue;
                    double cost = l.area() * lc + ra[i + 1] * rc[i + 1];
                    if (cost < best_cost) best_cost = cost, best_axis = a, best_bin = i;
                }
            }
        }
        if (best_axis >= 0) {
            int a = best_axis;
            double ext = cb.hi[a] - cb.lo[a], k = kBins / ext;
            auto mid = std::partition(refs.begin() + b, refs.begin() + e, [&](const Ref& r) {
                int bi = std::min(kBins - 1, std::max(0, (int)((r.c[a] - cb.lo[a]) * k)));
                return bi <= best_bin;
            });
            int m = (int)(mid - refs.begin());
            if (m > b && m < e) return m;
        }
        // median by centroid along the widest axis (also the degenerate-centroid case)
        int a = 0;
        for (int k = 1; k < 3; k++)
            if (cb.hi[k] - cb.lo[k] > cb.hi[a] - cb.lo[a]) a = k;
        int m = b + n / 2;
        std::nth_element(refs.begin() + b, refs.begin() + m, refs.begin() + e,
                         [a](const Ref& x, const Ref& y) { return x.c[a] < y.c[a]; });
        return m;
    }

    // Subtrees handed to worker threads (build_bvh): refs[b,e) at `depth`, to be hung into child slot `slot` of node `node`.
    struct Job {
        int b, e, depth, node, slot;
    };
    std::vector<Job>* jobs = nullptr;  // non-null while the top of the tree is being split
    int job_depth = 0;

    // Fills child slot (box, c, n) for refs[b,e); recursion creates inner nodes.
    void child(int b, int e, int depth, float lo[3], float hi[3], int32_t& c, int32_t& n, int node, int slot) {
        Box bx = bounds(b, e);
        for (int a = 0; a < 3; a++) lo[a] = down(bx.lo[a]), hi[a] = up(bx.hi[a]);
        out.max_depth = std::max(out.max_depth, depth);
        if (e - b <= kLeafMax) {
            c = (int32_t)out.tris.size();
            n = e - b;
            for (int i = b; i < e; i++) out.tris.push_back((*src)[refs[i].tri]);
            return;
        }
        n = 0;
        if (jobs && depth >= job_depth && e - b >= 256) {  // a subtree of its own: built later, side by side with the others
            c = -1;
            jobs->push_back({b, e, depth, node, slot});
            return;
        }
        c = inner(b, e, depth);
    }

    int32_t inner(int b, int e, int depth) {
        int32_t id = (int32_t)out.nodes.size();
        out.nodes.emplace_back();
        int m = split(b, e, depth);
        BvhNode nd;
        child(b, m, depth + 1, nd.lo0, nd.hi0, nd.c0, nd.n0, id, 0);
        child(m, e, depth + 1, nd.lo1, nd.hi1, nd.c1, nd.n1, id, 1);
        out.nodes[id] = nd;
        return id;
    }
};

}  // namespace

BvhBuild build_bvh(const std::vector<DTri>& tris) {
    Ctx cx;
    cx.src = &tris;
    cx.refs.resize(tris.size());
    for (size_t i = 0; i < tris.size(); i++) {
        const DTri& t = tris[i];
        Ref& r = cx.refs[i];
        double p2[3], p3[3];
        // p2 = p1 + e1 is not bit-exact p2; widen each vertex box by a few ulps below (outward
        // rounding to f32 already adds ~1e-7 relative, which dwarfs the 1e-16 of this sum).
        for (int a = 0; a < 3; a++) p2[a] = t.p1[a] + t.e1[a], p3[a] = t.p1[a] + t.e2[a];
        r.box.grow(t.p1);
        r.box.grow(p2);
        r.box.grow(p3);
        for (int a = 0; a < 3; a++) {
            double pad = 4e-16 * std::fmax(std::fabs(r.box.lo[a]), std::fabs(r.box.hi[a]));
            r.box.lo[a] -= pad;
            r.box.hi[a] += pad;
            r.c[a] = 0.5 * (r.box.lo[a] + r.box.hi[a]);
        }
        r.tri = (int)i;
    }
    int n = (int)tris.size();
    Box root = cx.bounds(0, n);
    for (int a = 0; a < 3; a++) cx.out.lo[a] = down(root.lo[a]), cx.out.hi[a] = up(root.hi[a]);
    if (n <= kLeafMax) {
        BvhNode nd;
        for (int a = 0; a < 3; a++) {
            nd.lo0[a] = cx.out.lo[a], nd.hi0[a] = cx.out.hi[a];
            nd.lo1[a] = INFINITY, nd.hi1[a] = -INFINITY;
        }
        nd.c0 = 0, nd.n0 = n, nd.c1 = -1, nd.n1 = 0;
        cx.out.nodes.push_back(nd);
        cx.out.tris = tris;
        cx.out.max_depth = 1;
    } else if (n < 4096) {
        cx.inner(0, n, 0);
    } else {
        // A large mesh (nefertiti.obj: 99,938 triangles took 0.57 s on one thread): the top five levels are split here,
        // every subtree below is built by a task of its own over its slice of `refs` (split() only permutes inside
        // [b, e)), and the pieces are stitched together with their node / triangle indices moved.  Same tree as the
        // sequential build (the splits do not depend on who computes them), only its nodes are numbered differently.
        std::vector<Ctx::Job> jobs;
        cx.jobs = &jobs;
        cx.job_depth = n >= 65536 ? 5 : 4;
        cx.inner(0, n, 0);
        cx.jobs = nullptr;
        std::vector<std::future<BvhBuild>> parts;
        Ctx* top = &cx;
        for (const Ctx::Job& j : jobs)
            parts.push_back(std::async(std::launch::async, [top, j] {
                Ctx sub;
                sub.src = top->src;
                sub.refs.assign(top->refs.begin() + j.b, top->refs.begin() + j.e);
                sub.inner(0, j.e - j.b, j.depth);
                return std::move(sub.out);
            }));
        for (size_t k = 0; k < jobs.size(); k++) {
            BvhBuild part = parts[k].get();
            const int32_t node_base = (int32_t)cx.out.nodes.size(), tri_base = (int32_t)cx.out.tris.size();
            for (BvhNode nd : part.nodes) {
                if (nd.n0 > 0) nd.c0 += tri_base;
                else if (nd.c0 >= 0) nd.c0 += node_base;
                if (nd.n1 > 0) nd.c1 += tri_base;
                else if (nd.c1 >= 0) nd.c1 += node_base;
                cx.out.nodes.push_back(nd);
            }
            cx.out.tris.insert(cx.out.tris.end(), part.tris.begin(), part.tris.end());
            BvhNode& parent = cx.out.nodes[(size_t)jobs[k].node];
            (jobs[k].slot == 0 ? parent.c0 : parent.c1) = node_base;  // the subtree's root is its first node
            cx.out.max_depth = std::max(cx.out.max_depth, part.max_depth);
        }
    }
    return cx.out;
}

}  // namespace rtc
