// Blob -> host-side arrays of the device scene: validation, mesh grouping + BVH build, CSG program
// compilation.  Pure host code (no CUDA calls) so it can be unit-tested without a GPU.
#include "scene_build.hpp"

#include <chrono>
#include <cmath>
#include <cstdio>
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <map>
#include <future>
#include <stdexcept>

#include "bvh.hpp"

namespace rtc {
namespace {

template <class T>
const T* section(const uint8_t* base, size_t nbytes, uint64_t off, uint64_t count, const char* what) {
    if (off % 16 != 0 || off > nbytes || count > (nbytes - off) / sizeof(T))
        throw std::runtime_error(std::string("blob: section out of range: ") + what);
    return reinterpret_cast<const T*>(base + off);
}

void check_index(int64_t v, int64_t n, const char* what, bool allow_neg = false) {
    if (allow_neg && v == -1) return;
    if (v < 0 || v >= n) throw std::runtime_error(std::string("blob: index out of range: ") + what);
}

// World-space box of a mesh: the 8 corners of its object-space root box taken through the forward
// matrix (the affine inverse of the stored world->object rows), padded.  Only ever used to SKIP work.
inline float f32_down(double v) {  // largest float <= v
    float f = (float)v;
    if ((double)f > v) f = std::nextafterf(f, -INFINITY);
    return f;
}
inline float f32_up(double v) {  // smallest float >= v
    float f = (float)v;
    if ((double)f < v) f = std::nextafterf(f, INFINITY);
    return f;
}

void mesh_world_box(const rtc_transform& t, DMesh& m) {
    const double* a = t.inverse;  // rows r: a[4r..4r+2] linear part, a[4r+3] translation
    double A[3][3] = {{a[0], a[1], a[2]}, {a[4], a[5], a[6]}, {a[8], a[9], a[10]}};
    double det = A[0][0] * (A[1][1] * A[2][2] - A[1][2] * A[2][1]) - A[0][1] * (A[1][0] * A[2][2] - A[1][2] * A[2][0]) +
                 A[0][2] * (A[1][0] * A[2][1] - A[1][1] * A[2][0]);
    bool ok = std::isfinite(det) && det != 0.0;
    double F[3][3];
    if (ok) {
        F[0][0] = (A[1][1] * A[2][2] - A[1][2] * A[2][1]) / det;
        F[0][1] = (A[0][2] * A[2][1] - A[0][1] * A[2][2]) / det;
        F[0][2] = (A[0][1] * A[1][2] - A[0][2] * A[1][1]) / det;
        F[1][0] = (A[1][2] * A[2][0] - A[1][0] * A[2][2]) / det;
        F[1][1] = (A[0][0] * A[2][2] - A[0][2] * A[2][0]) / det;
        F[1][2] = (A[0][2] * A[1][0] - A[0][0] * A[1][2]) / det;
        F[2][0] = (A[1][0] * A[2][1] - A[1][1] * A[2][0]) / det;
        F[2][1] = (A[0][1] * A[2][0] - A[0][0] * A[2][1]) / det;
        F[2][2] = (A[0][0] * A[1][1] - A[0][1] * A[1][0]) / det;
    }
    double wmin[3], wmax[3];
    for (int k = 0; k < 3; k++) wmin[k] = INFINITY, wmax[k] = -INFINITY;
    for (int c = 0; c < 8 && ok; c++) {
        double p[3] = {(c & 1) ? m.hi[0] : m.lo[0], (c & 2) ? m.hi[1] : m.lo[1], (c & 4) ? m.hi[2] : m.lo[2]};
        double q[3] = {p[0] - a[3], p[1] - a[7], p[2] - a[11]};  // object = A*world + t  =>  world = F*(object - t)
        for (int k = 0; k < 3; k++) {
            double w = F[k][0] * q[0] + F[k][1] * q[1] + F[k][2] * q[2];
            if (!std::isfinite(w)) ok = false;
            wmin[k] = std::fmin(wmin[k], w);
            wmax[k] = std::fmax(wmax[k], w);
        }
    }
    for (int k = 0; k < 3; k++) {
        if (!ok) {
            m.wlo[k] = -INFINITY, m.whi[k] = INFINITY;
        } else {
            double pad = 1e-6 * (std::fabs(wmin[k]) + std::fabs(wmax[k]) + (wmax[k] - wmin[k])) + 1e-9;
            m.wlo[k] = f32_down(wmin[k] - pad);
            m.whi[k] = f32_up(wmax[k] + pad);
        }
    }
}

struct CsgCompiler {
    const std::vector<rtc_node>& nodes;
    const std::vector<int32_t>& children;
    std::vector<int4>& prog;
    int depth = 0, max_depth = 0;

    void push_seg() {
        depth++;
        if (depth > max_depth) max_depth = depth;
    }
    void compile(int32_t ni) {
        const rtc_node& n = nodes[(size_t)ni];
        size_t bbox_at = (size_t)-1;
        if (n.kind != RTC_NODE_LEAF && n.use_bbox) {
            bbox_at = prog.size();
            prog.push_back(make_int4(CSG_OP_BBOX, ni, 0, 0));
        }
        int d0 = depth;
        if (n.kind == RTC_NODE_LEAF) {
            prog.push_back(make_int4(CSG_OP_LEAF, n.a, 0, 0));
            push_seg();
        } else if (n.kind == RTC_NODE_GROUP) {
            if (n.b == 0) {
                prog.push_back(make_int4(CSG_OP_MERGE, 0, 0, 0));
                push_seg();
            }
            for (int32_t k = 0; k < n.b; k++) {
                compile(children[(size_t)(n.a + k)]);
                if (k > 0) {
                    prog.push_back(make_int4(CSG_OP_MERGE, 2, 0, 0));
                    depth--;
                }
            }
        } else {
            compile(n.a);
            compile(n.b);
            const rtc_node& l = nodes[(size_t)n.a];
            prog.push_back(make_int4(CSG_OP_FILTER, n.op, l.leaf_begin, l.leaf_end));
            depth--;
        }
        if (depth != d0 + 1) throw std::runtime_error("csg compile: segment accounting");
        if (bbox_at != (size_t)-1) prog[bbox_at].z = (int)prog.size();
    }
};

}  // namespace

namespace {
struct Lap {  // RTC_LOAD_TIMING=1: where the set-up time of a scene goes
    std::chrono::steady_clock::time_point t = std::chrono::steady_clock::now();
    bool on = std::getenv("RTC_LOAD_TIMING") != nullptr;
    void operator()(const char* what) {
        auto n = std::chrono::steady_clock::now();
        if (on) std::fprintf(stderr, "  scene build: %-28s %.3f s\n", what, std::chrono::duration<double>(n - t).count());
        t = n;
    }
};
}  // namespace

void build_host_scene(const void* blob, size_t nbytes, HostScene& hs) {
    Lap lap;
    if (!blob || nbytes < sizeof(rtc_blob_header)) throw std::runtime_error("blob: too small");
    const uint8_t* base = static_cast<const uint8_t*>(blob);
    rtc_blob_header h;
    std::memcpy(&h, base, sizeof h);
    if (h.magic != RTC_BLOB_MAGIC) throw std::runtime_error("blob: bad magic");
    if (h.version != RTC_BLOB_VERSION) throw std::runtime_error("blob: unsupported version");
    if (h.total_bytes != nbytes) throw std::runtime_error("blob: size mismatch");

    auto copy = [&](auto& vec, uint64_t off, uint64_t n, const char* what) {
        using T = typename std::remove_reference<decltype(vec)>::type::value_type;
        const T* p = section<T>(base, nbytes, off, n, what);
        vec.assign(p, p + n);
    };
    copy(hs.lights, h.off_lights, h.n_lights, "lights");
    copy(hs.materials, h.off_materials, h.n_materials, "materials");
    copy(hs.patterns, h.off_patterns, h.n_patterns, "patterns");
    copy(hs.uvs, h.off_uv_patterns, h.n_uv_patterns, "uv patterns");
    copy(hs.images, h.off_images, h.n_images, "images");
    copy(hs.transforms, h.off_transforms, h.n_transforms, "transforms");
    copy(hs.leaves, h.off_leaves, h.n_leaves, "leaves");
    copy(hs.blob_tris, h.off_triangles, h.n_triangles, "triangles");
    copy(hs.normals, h.off_tri_normals, h.n_triangles, "triangle normals");
    copy(hs.guards, h.off_guards, h.n_guards, "guards");
    copy(hs.nodes, h.off_nodes, h.n_nodes, "nodes");
    std::vector<int32_t> children, roots;
    copy(children, h.off_children, h.n_children, "children");
    copy(roots, h.off_csg_roots, h.n_csg_roots, "csg roots");
    copy(hs.texels, h.off_texels, h.texel_bytes, "texels");

    lap("copy sections");
    // ---- validate every index the kernels will follow --------------------------------------------
    for (auto& m : hs.materials) check_index(m.pattern, h.n_patterns, "material.pattern");
    for (auto& p : hs.patterns) {
        check_index(p.a, h.n_patterns, "pattern.a", true);
        check_index(p.b, h.n_patterns, "pattern.b", true);
        bool binary = p.kind == RTC_PAT_BLEND || p.kind == RTC_PAT_STRIPES || p.kind == RTC_PAT_GRADIENT ||
                      p.kind == RTC_PAT_RADIAL_GRADIENT || p.kind == RTC_PAT_RINGS || p.kind == RTC_PAT_CHECKERS;
        if (binary && (p.a < 0 || p.b < 0)) throw std::runtime_error("blob: pattern needs two children");
        if (p.kind == RTC_PAT_PERTURB && p.a < 0) throw std::runtime_error("blob: perturb needs a child");
        if (p.kind == RTC_PAT_TEXTURE_MAP) {
            int n = p.mapping == RTC_MAP_CUBIC ? 6 : 1;
            for (int i = 0; i < n; i++) check_index(p.uv[i], h.n_uv_patterns, "pattern.uv");
        }
        if (p.kind < 0 || p.kind > RTC_PAT_TEST) throw std::runtime_error("blob: pattern kind");
    }
    for (auto& u : hs.uvs) {
        if (u.kind == RTC_UV_IMAGE) check_index(u.image, h.n_images, "uv.image");
        int n = u.kind == RTC_UV_ALIGN_CHECK ? 5 : u.kind == RTC_UV_CHECKERS ? 2 : 0;
        for (int i = 0; i < n; i++) check_index(u.sub[i], h.n_patterns, "uv.sub");
    }
    for (auto& im : hs.images) {
        if (im.width <= 0 || im.height <= 0) throw std::runtime_error("blob: image size");
        uint64_t bytes = (uint64_t)im.width * im.height * (im.format == RTC_IMG_RGBA8 ? 4 : 24);
        if (im.offset % 16 || im.offset > h.texel_bytes || bytes > h.texel_bytes - im.offset)
            throw std::runtime_error("blob: image texels out of range");
    }
    for (auto& l : hs.leaves) {
        check_index(l.transform, h.n_transforms, "leaf.transform");
        check_index(l.material, h.n_materials, "leaf.material");
        check_index(l.guard, h.n_guards, "leaf.guard", true);
        bool tri = l.kind == RTC_TRIANGLE || l.kind == RTC_SMOOTH_TRIANGLE;
        if (tri) check_index(l.triangle, h.n_triangles, "leaf.triangle");
        if (l.kind < 0 || l.kind > RTC_TEST_SHAPE) throw std::runtime_error("blob: leaf kind");
    }
    for (auto& g : hs.guards) check_index(g.next, h.n_guards, "guard.next", true);
    for (size_t i = 0; i < hs.nodes.size(); i++) {
        const rtc_node& n = hs.nodes[i];
        if (n.kind == RTC_NODE_LEAF) check_index(n.a, h.n_leaves, "node.leaf");
        else if (n.kind == RTC_NODE_CSG) {
            // children precede parents (post-order emission) => no cycles
            if (n.a < 0 || n.b < 0 || (size_t)n.a >= i || (size_t)n.b >= i) throw std::runtime_error("blob: csg operands");
        } else if (n.kind == RTC_NODE_GROUP) {
            if (n.a < 0 || n.b < 0 || (uint64_t)n.a + (uint64_t)n.b > h.n_children) throw std::runtime_error("blob: group children");
            for (int32_t k = 0; k < n.b; k++)
                if (children[(size_t)(n.a + k)] < 0 || (size_t)children[(size_t)(n.a + k)] >= i)
                    throw std::runtime_error("blob: group child order");
        } else {
            throw std::runtime_error("blob: node kind");
        }
        if (n.leaf_begin < 0 || n.leaf_end < n.leaf_begin || (uint32_t)n.leaf_end > h.n_leaves)
            throw std::runtime_error("blob: node leaf range");
    }

    // ---- which leaves belong to CSG objects --------------------------------------------------------
    std::vector<char> in_csg(hs.leaves.size(), 0);
    for (int32_t r : roots) {
        check_index(r, h.n_nodes, "csg root");
        const rtc_node& n = hs.nodes[(size_t)r];
        for (int32_t i = n.leaf_begin; i < n.leaf_end; i++) in_csg[(size_t)i] = 1;
    }

    // ---- widen the pre-test boxes a little (they are an optimisation, never a decision) ------------
    for (auto& l : hs.leaves) {
        if (l.flags & RTC_UNBOUNDED) continue;
        for (int a = 0; a < 3; a++) {
            double pad = 1e-9 * (std::fabs(l.wmin[a]) + std::fabs(l.wmax[a])) + 1e-12;
            l.wmin[a] -= pad;
            l.wmax[a] += pad;
        }
    }

    lap("validate");
    // ---- top level: analytic prims, meshes ----------------------------------------------------------
    // A mesh INSTANCE = all non-CSG triangles sharing a world->object matrix and a casts_shadow flag.  Instances whose
    // object-space triangles (positions, edges, normals, flat/smooth) are the same multiset share ONE geometry: one
    // BVH, one DTri array (dragons.json: six instances of dragon.obj -> one 23,490-triangle geometry instead of six;
    // the reference re-parses the file per instance, scene.zig:455-488).  What differs per instance -- matrix,
    // material, and the ORDER of the leaves, which divide(8) derives from world-space boxes (group.zig:85-135) and
    // which decides ties at equal t -- stays per instance: DTri.leaf is the triangle's rank in the geometry's
    // canonical order and DMesh.leaf_map translates it into the instance's rtc_leaf index.
    struct Instance {
        int32_t transform;
        uint32_t flags;
        std::vector<int32_t> leaves;  // rtc_leaf indices, in leaf order
        std::vector<int32_t> canon;   // the same leaves in canonical (content) order
        uint64_t hash = 0;
        int geometry = -1;
    };
    std::map<std::pair<int32_t, uint32_t>, size_t> inst_of;
    std::vector<Instance> insts;
    for (size_t i = 0; i < hs.leaves.size(); i++) {
        if (in_csg[i]) continue;
        const rtc_leaf& l = hs.leaves[i];
        if (l.kind == RTC_TEST_SHAPE) continue;
        bool tri = l.kind == RTC_TRIANGLE || l.kind == RTC_SMOOTH_TRIANGLE;
        if (tri && l.guard < 0) {
            auto key = std::make_pair(l.transform, l.flags & (uint32_t)RTC_CASTS_SHADOW);
            auto it = inst_of.find(key);
            if (it == inst_of.end()) {
                it = inst_of.emplace(key, insts.size()).first;
                insts.emplace_back();
                insts.back().transform = key.first;
                insts.back().flags = key.second;
            }
            insts[it->second].leaves.push_back((int32_t)i);
        } else {
            DPrim p{};
            p.leaf = (int32_t)i;
            p.flags = l.flags;
            for (int a = 0; a < 3; a++) {
                bool open = (l.flags & RTC_UNBOUNDED) || !(l.wmin[a] <= l.wmax[a]);  // NaN corners: never cull
                p.lo[a] = open ? -INFINITY : f32_down(l.wmin[a]);
                p.hi[a] = open ? INFINITY : f32_up(l.wmax[a]);
            }
            hs.prims.push_back(p);
        }
    }
    // content of a triangle as far as the kernels read it through DTri / rtc_tri_normals / rtc_leaf.kind
    auto tri_cmp = [&](int32_t la, int32_t lb) {
        const rtc_leaf &A = hs.leaves[(size_t)la], &B = hs.leaves[(size_t)lb];
        int c = std::memcmp(hs.blob_tris[(size_t)A.triangle].p1, hs.blob_tris[(size_t)B.triangle].p1, 9 * sizeof(double));
        if (c) return c;
        c = std::memcmp(hs.normals[(size_t)A.triangle].n, hs.normals[(size_t)B.triangle].n, 9 * sizeof(double));
        if (c) return c;
        return A.kind < B.kind ? -1 : A.kind > B.kind ? 1 : 0;
    };
    static const bool no_instancing = [] {
        const char* e = std::getenv("RTC_INSTANCING");
        return e && e[0] == '0';
    }();
    // canonical order only where sharing is possible at all: another instance with the same triangle count
    std::map<size_t, int> by_count;
    for (const Instance& in : insts) by_count[in.leaves.size()]++;
    {
        std::vector<std::future<void>> jobs;
        for (Instance& in : insts) {
            in.canon = in.leaves;
            if (no_instancing || by_count[in.leaves.size()] < 2) continue;
            Instance* ip = &in;
            jobs.push_back(std::async(std::launch::async, [ip, &tri_cmp, &hs] {
                std::stable_sort(ip->canon.begin(), ip->canon.end(), [&](int32_t a, int32_t b) { return tri_cmp(a, b) < 0; });
                uint64_t h = 1469598103934665603ull;  // FNV-1a over the triangle contents in canonical order
                for (int32_t lf : ip->canon) {
                    const rtc_leaf& L = hs.leaves[(size_t)lf];
                    const unsigned char* p = reinterpret_cast<const unsigned char*>(hs.blob_tris[(size_t)L.triangle].p1);
                    for (size_t k = 0; k < 9 * sizeof(double); k += 8) {
                        uint64_t w;
                        std::memcpy(&w, p + k, 8);
                        h = (h ^ w) * 1099511628211ull;
                    }
                }
                ip->hash = h;
            }));
        }
        for (auto& j : jobs) j.get();
    }
    lap("instances, canonical order");
    struct Geometry {
        size_t first_inst;
        int32_t root = 0;  // BvhNode index
        float lo[3], hi[3];
    };
    std::vector<Geometry> geoms;
    for (size_t i = 0; i < insts.size(); i++) {
        Instance& in = insts[i];
        if (!no_instancing && by_count[in.leaves.size()] >= 2) {
            for (size_t g = 0; g < geoms.size() && in.geometry < 0; g++) {
                const Instance& o = insts[geoms[g].first_inst];
                if (o.hash != in.hash || o.canon.size() != in.canon.size()) continue;
                bool same = true;
                for (size_t k = 0; k < in.canon.size() && same; k++) same = tri_cmp(in.canon[k], o.canon[k]) == 0;
                if (same) in.geometry = (int)g;
            }
        }
        if (in.geometry < 0) {
            in.geometry = (int)geoms.size();
            geoms.push_back(Geometry{i});
        }
    }
    lap("match geometries");
    hs.bvh_max_depth = 0;
    // the geometries are independent: one builder thread each
    std::vector<std::vector<DTri>> gtris(geoms.size());
    std::vector<std::future<BvhBuild>> builds;
    for (size_t g = 0; g < geoms.size(); g++) {
        const Instance& in = insts[geoms[g].first_inst];
        std::vector<DTri>& v = gtris[g];
        v.resize(in.canon.size());
        for (size_t k = 0; k < in.canon.size(); k++) {
            const rtc_triangle& t = hs.blob_tris[(size_t)hs.leaves[(size_t)in.canon[k]].triangle];
            DTri& d = v[k];
            std::memcpy(d.p1, t.p1, sizeof d.p1);
            std::memcpy(d.e1, t.e1, sizeof d.e1);
            std::memcpy(d.e2, t.e2, sizeof d.e2);
            d.leaf = (int32_t)k;  // rank in the canonical order; DMesh.leaf_map makes it an rtc_leaf index
            d.pad = 0;
        }
        const std::vector<DTri>* tris = &v;
        // large meshes: LBVH on the device (bvh_gpu.cu), built when its turn comes, on the caller's thread and device
        if (bvh_build_on_gpu(v.size())) builds.push_back(std::async(std::launch::deferred, [tris] { return build_bvh_gpu(*tris); }));
        else builds.push_back(std::async(geoms.size() > 1 ? std::launch::async : std::launch::deferred, [tris] { return build_bvh(*tris); }));
    }
    lap("gather triangles");
    for (size_t g = 0; g < geoms.size(); g++) {
        BvhBuild b = builds[g].get();
        geoms[g].root = (int32_t)hs.bvh.size();
        for (int a = 0; a < 3; a++) geoms[g].lo[a] = b.lo[a], geoms[g].hi[a] = b.hi[a];
        int32_t node_base = (int32_t)hs.bvh.size(), tri_base = (int32_t)hs.tris.size();
        for (BvhNode n : b.nodes) {
            // child references as the traversal holds them: inner node index >= 0, or a leaf as
            // ~((first triangle << 3) | count) (rt_trace.cuh mesh_inner / mesh_leaf)
            if (n.n0 > 0) n.c0 = ~(((n.c0 + tri_base) << 3) | n.n0);
            else if (n.c0 >= 0) n.c0 += node_base;
            if (n.n1 > 0) n.c1 = ~(((n.c1 + tri_base) << 3) | n.n1);
            else if (n.c1 >= 0) n.c1 += node_base;
            hs.bvh.push_back(n);
        }
        hs.tris.insert(hs.tris.end(), b.tris.begin(), b.tris.end());
        if (b.max_depth > hs.bvh_max_depth) hs.bvh_max_depth = b.max_depth;
    }
    lap("geometries + BVH build");
    hs.n_geometries = (int)geoms.size();
    for (const Instance& in : insts) {
        const Geometry& G = geoms[(size_t)in.geometry];
        DMesh m{};
        m.transform = in.transform;
        m.flags = in.flags;
        m.root = G.root;
        m.n_tris = (int32_t)in.canon.size();
        m.leaf_map = (int32_t)hs.mesh_leaf.size();
        hs.mesh_leaf.insert(hs.mesh_leaf.end(), in.canon.begin(), in.canon.end());
        for (int a = 0; a < 3; a++) {
            m.lo[a] = G.lo[a];
            m.hi[a] = G.hi[a];
            m.bound_abs[a] = std::fmax(std::fabs(G.lo[a]), std::fabs(G.hi[a]));
        }
        mesh_world_box(hs.transforms[(size_t)m.transform], m);
        hs.meshes.push_back(m);
    }

    // scale of the f32 slab-test error bound for the world-space pre-tests (rt_trace.cuh make_fray)
    auto note_abs = [&](const float lo[3], const float hi[3]) {
        for (int a = 0; a < 3; a++) {
            if (std::isfinite(lo[a])) hs.world_abs[a] = std::fmax(hs.world_abs[a], std::fabs(lo[a]));
            if (std::isfinite(hi[a])) hs.world_abs[a] = std::fmax(hs.world_abs[a], std::fabs(hi[a]));
        }
    };
    for (const DPrim& p : hs.prims) note_abs(p.lo, p.hi);
    for (const DMesh& m : hs.meshes) note_abs(m.wlo, m.whi);

    // leaves are encoded on the traversal stack as ~((first << 3) | count)  (rt_trace.cuh trace_meshes)
    if (hs.tris.size() >= (size_t(1) << 27)) throw std::runtime_error("more than 2^27 mesh triangles");

    lap("instances -> meshes");
    // ---- CSG programs ------------------------------------------------------------------------------
    CsgCompiler cc{hs.nodes, children, hs.csg_prog};
    for (int32_t r : roots) {
        DCsgRoot root{};
        root.pc = (int32_t)hs.csg_prog.size();
        root.leaf_begin = hs.nodes[(size_t)r].leaf_begin;
        root.leaf_end = hs.nodes[(size_t)r].leaf_end;
        cc.depth = 0;
        cc.compile(r);
        hs.csg_prog.push_back(make_int4(CSG_OP_END, 0, 0, 0));
        hs.csg_roots.push_back(root);
    }
    hs.csg_max_segments = cc.max_depth;
}

}  // namespace rtc
