// BVH build ON THE DEVICE (what replaces divide(8), shape.zig:372-399 / group.zig:85-135, for
// traversal): a linear BVH -- Morton codes of the triangle centroids, one radix sort, Karras' parallel construction of
// the binary radix tree (every inner node finds its range and split by itself), boxes fitted bottom-up, subtrees of <= 4
// triangles collapsed into leaves -- emitted in the layout the host builder (bvh.cpp) produces, so everything after it
// is the same.  Boxes are f32 rounded OUTWARD from the f64 triangle corners exactly as bvh.cpp does (the same
// 4e-16 pad for p1 + e1 not being bit-exact p2): supersets, so the tree decides nothing -- which triangles a ray hits,
// and which of two hits at the same t wins, is decided by the f64 test and the leaf order (tests: frames bit-identical
// to the host-built tree's).  A sorted-by-Morton-code tree is built in a few hundred microseconds where the binned-SAH
// build takes tens to hundreds of milliseconds of host time, and is traversed slower (profiles/README.md has both).
#include <cuda_runtime.h>

#include <cfloat>
#include <cstdlib>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <stdexcept>
#include <string>

#include "bvh.hpp"

namespace rtc {
namespace {

constexpr int kLeafMax = 4;

struct Aabb {
    float lo[3], hi[3];
};

#define GPU_TRY(expr)                                                                                   \
    do {                                                                                                \
        cudaError_t e_ = (expr);                                                                        \
        if (e_ != cudaSuccess) throw std::runtime_error(std::string(#expr) + ": " + cudaGetErrorString(e_)); \
    } while (0)

// triangle box (outward f32) + centroid bounds of the mesh (atomics on order-preserving ints)
__device__ __forceinline__ int f2ord(float f) {
    int i = __float_as_int(f);
    return i >= 0 ? i : i ^ 0x7fffffff;
}
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

__global__ void tri_boxes_kernel(const DTri* __restrict__ tris, int n, Aabb* __restrict__ boxes, int* __restrict__ cmin, int* __restrict__ cmax) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const DTri& t = tris[i];
    Aabb b;
    for (int a = 0; a < 3; a++) {
        const double p1 = t.p1[a], p2 = t.p1[a] + t.e1[a], p3 = t.p1[a] + t.e2[a];
        double lo = fmin(p1, fmin(p2, p3)), hi = fmax(p1, fmax(p2, p3));
        const double pad = 4e-16 * fmax(fabs(lo), fabs(hi));  // bvh.cpp: p1 + e1 is not bit-exact p2
        lo -= pad;
        hi += pad;
        b.lo[a] = __double2float_rd(lo);
        b.hi[a] = __double2float_ru(hi);
        const float c = 0.5f * (b.lo[a] + b.hi[a]);
        atomicMin(cmin + a, f2ord(c));
        atomicMax(cmax + a, f2ord(c));
    }
    boxes[i] = b;
}

__device__ __forceinline__ unsigned expand10(unsigned v) {  // 10 bits -> every third bit
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

__global__ void morton_kernel(const Aabb* __restrict__ boxes, int n, const int* __restrict__ cmin, const int* __restrict__ cmax,
                              unsigned* __restrict__ keys, int* __restrict__ vals) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    unsigned q[3];
    for (int a = 0; a < 3; a++) {
        const float lo = ord2f(cmin[a]), hi = ord2f(cmax[a]);
        const float c = 0.5f * (boxes[i].lo[a] + boxes[i].hi[a]);
        const float ext = hi - lo;
        float u = ext > 0.0f ? (c - lo) / ext : 0.0f;
        u = fminf(fmaxf(u * 1024.0f, 0.0f), 1023.0f);
        q[a] = (unsigned)u;
    }
    keys[i] = (expand10(q[0]) << 2) | (expand10(q[1]) << 1) | expand10(q[2]);
    vals[i] = i;
}

// common prefix of the (code, position) keys of sorted triangles i and j; -1 outside the array (Karras 2012)
__device__ __forceinline__ int delta(const unsigned* __restrict__ codes, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const unsigned a = codes[i], b = codes[j];
    if (a != b) return __clz(a ^ b);
    return 32 + __clz((unsigned)i ^ (unsigned)j);
}

// inner node i of the binary radix tree over n sorted keys: its range, its two children (index >= 0: inner node,
// ~index: leaf = sorted triangle), and the children's parent links
__global__ void hierarchy_kernel(const unsigned* __restrict__ codes, int n, int2* __restrict__ child, int2* __restrict__ range,
                                 int* __restrict__ parent_inner, int* __restrict__ parent_leaf) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = delta(codes, n, i, i + 1) - delta(codes, n, i, i - 1) >= 0 ? 1 : -1;
    const int dmin = delta(codes, n, i, i - d);
    int lmax = 2;
    while (delta(codes, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(codes, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta(codes, n, i, j);
    int s = 0;
    for (int t = (l + 1) >> 1;; t = (t + 1) >> 1) {
        if (delta(codes, n, i, i + (s + t) * d) > dnode) s += t;
        if (t == 1) break;
    }
    const int gamma = i + s * d + min(d, 0);
    const int first = min(i, j), last = max(i, j);
    const int left = first == gamma ? ~gamma : gamma;
    const int right = last == gamma + 1 ? ~(gamma + 1) : gamma + 1;
    child[i] = make_int2(left, right);
    range[i] = make_int2(first, last);
    if (left >= 0) parent_inner[left] = i;
    else parent_leaf[~left] = i;
    if (right >= 0) parent_inner[right] = i;
    else parent_leaf[~right] = i;
    if (i == 0) parent_inner[0] = -1;
}

// (read at L2: another SM wrote the box, and a neighbouring box in the same line may sit in this SM's L1 from before)
__device__ __forceinline__ Aabb load_box_cg(const Aabb* p) {
    Aabb b;
    const float* f = reinterpret_cast<const float*>(p);
    for (int k = 0; k < 3; k++) b.lo[k] = __ldcg(f + k), b.hi[k] = __ldcg(f + 3 + k);
    return b;
}

// boxes of the inner nodes, bottom-up: the second thread to arrive at a node merges its children's boxes
__global__ void fit_kernel(const Aabb* __restrict__ tri_boxes, const int* __restrict__ sorted, int n, const int2* __restrict__ child,
                           const int* __restrict__ parent_inner, const int* __restrict__ parent_leaf, Aabb* __restrict__ inner_boxes,
                           Aabb* __restrict__ leaf_boxes, int* __restrict__ arrived) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    leaf_boxes[i] = tri_boxes[sorted[i]];
    __threadfence();
    int node = parent_leaf[i];
    while (node >= 0) {
        if (atomicAdd(arrived + node, 1) == 0) return;  // the sibling subtree is not done yet: its thread will come by
        __threadfence();
        const int2 c = child[node];
        const Aabb a = load_box_cg(c.x >= 0 ? inner_boxes + c.x : leaf_boxes + ~c.x);
        const Aabb b = load_box_cg(c.y >= 0 ? inner_boxes + c.y : leaf_boxes + ~c.y);
        Aabb m;
        for (int k = 0; k < 3; k++) m.lo[k] = fminf(a.lo[k], b.lo[k]), m.hi[k] = fmaxf(a.hi[k], b.hi[k]);
        inner_boxes[node] = m;
        __threadfence();
        node = parent_inner[node];
    }
}

__global__ void keep_kernel(const int2* __restrict__ range, int n_inner, int* __restrict__ keep) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_inner) keep[i] = (range[i].y - range[i].x + 1) > kLeafMax ? 1 : 0;
}

// kept inner nodes -> BvhNode records (children that span <= 4 triangles become leaves over their sorted range)
__global__ void emit_kernel(const int2* __restrict__ child, const int2* __restrict__ range, const int* __restrict__ keep,
                            const int* __restrict__ new_index, const Aabb* __restrict__ inner_boxes, const Aabb* __restrict__ leaf_boxes,
                            const int* __restrict__ parent_inner, int n_inner, BvhNode* __restrict__ out, int* __restrict__ max_depth) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_inner || !keep[i]) return;
    BvhNode nd;
    const int2 c = child[i];
    for (int side = 0; side < 2; side++) {
        const int ch = side ? c.y : c.x;
        const Aabb b = ch >= 0 ? inner_boxes[ch] : leaf_boxes[~ch];
        float* lo = side ? nd.lo1 : nd.lo0;
        float* hi = side ? nd.hi1 : nd.hi0;
        for (int k = 0; k < 3; k++) lo[k] = b.lo[k], hi[k] = b.hi[k];
        int32_t ref, cnt;
        if (ch >= 0 && keep[ch]) {
            ref = new_index[ch];
            cnt = 0;
        } else if (ch >= 0) {
            ref = range[ch].x;
            cnt = range[ch].y - range[ch].x + 1;
        } else {
            ref = ~ch;
            cnt = 1;
        }
        if (side) nd.c1 = ref, nd.n1 = cnt;
        else nd.c0 = ref, nd.n0 = cnt;
    }
    out[new_index[i]] = nd;
    int depth = 1;  // levels of kept nodes from here to the root, + the leaf level below
    for (int p = parent_inner[i]; p >= 0; p = parent_inner[p]) depth++;
    atomicMax(max_depth, depth + 1);
}

__global__ void gather_kernel(const DTri* __restrict__ tris, const int* __restrict__ sorted, int n, DTri* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = tris[sorted[i]];
}

struct DeviceBuffers {  // freed whatever happens
    void* p[16] = {};
    int n = 0;
    template <class T>
    T* get(size_t count) {
        void* q = nullptr;
        GPU_TRY(cudaMalloc(&q, count * sizeof(T) + 16));
        p[n++] = q;
        return static_cast<T*>(q);
    }
    ~DeviceBuffers() {
        for (int i = 0; i < n; i++) cudaFree(p[i]);
    }
};

}  // namespace

bool bvh_build_on_gpu(size_t n_tris) {
    const char* e = std::getenv("RTC_BVH");
    const std::string mode = e ? e : "auto";
    if (mode == "host") return false;
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev <= 0) {
        cudaGetLastError();
        return false;  // (inspecting a blob on a machine without a GPU: the host builder)
    }
    return mode == "gpu" || n_tris >= 65536;
}

BvhBuild build_bvh_gpu(const std::vector<DTri>& tris) {
    const int n = (int)tris.size();
    if (n <= 2 * kLeafMax) return build_bvh(tris);  // nothing to build in parallel
    DeviceBuffers buf;
    DTri* d_tris = buf.get<DTri>(n);
    DTri* d_sorted_tris = buf.get<DTri>(n);
    Aabb* d_tri_boxes = buf.get<Aabb>(n);
    Aabb* d_leaf_boxes = buf.get<Aabb>(n);
    Aabb* d_inner_boxes = buf.get<Aabb>(n);
    unsigned* d_keys = buf.get<unsigned>(2 * (size_t)n);
    int* d_vals = buf.get<int>(2 * (size_t)n);
    int2* d_child = buf.get<int2>(n);
    int2* d_range = buf.get<int2>(n);
    int* d_misc = buf.get<int>(5 * (size_t)n + 16);  // parent_inner, parent_leaf, arrived, keep, new_index, then 8 scalars
    int *d_parent_inner = d_misc, *d_parent_leaf = d_misc + n, *d_arrived = d_misc + 2 * n, *d_keep = d_misc + 3 * n,
        *d_new = d_misc + 4 * n, *d_scal = d_misc + 5 * n;
    BvhNode* d_nodes = buf.get<BvhNode>(n);
    GPU_TRY(cudaMemcpy(d_tris, tris.data(), (size_t)n * sizeof(DTri), cudaMemcpyHostToDevice));
    const int h_init[8] = {0x7fffffff, 0x7fffffff, 0x7fffffff, (int)0x80000000, (int)0x80000000, (int)0x80000000, 0, 0};
    GPU_TRY(cudaMemcpy(d_scal, h_init, sizeof h_init, cudaMemcpyHostToDevice));
    GPU_TRY(cudaMemset(d_arrived, 0, (size_t)n * sizeof(int)));
    const int T = 256, B = (n + T - 1) / T;
    tri_boxes_kernel<<<B, T>>>(d_tris, n, d_tri_boxes, d_scal, d_scal + 3);
    morton_kernel<<<B, T>>>(d_tri_boxes, n, d_scal, d_scal + 3, d_keys, d_vals);
    size_t tmp_bytes = 0;
    GPU_TRY(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, d_keys, d_keys + n, d_vals, d_vals + n, n, 0, 30));
    void* d_tmp = buf.get<char>(tmp_bytes);
    GPU_TRY(cub::DeviceRadixSort::SortPairs(d_tmp, tmp_bytes, d_keys, d_keys + n, d_vals, d_vals + n, n, 0, 30));
    const unsigned* codes = d_keys + n;
    const int* sorted = d_vals + n;
    hierarchy_kernel<<<B, T>>>(codes, n, d_child, d_range, d_parent_inner, d_parent_leaf);
    fit_kernel<<<B, T>>>(d_tri_boxes, sorted, n, d_child, d_parent_inner, d_parent_leaf, d_inner_boxes, d_leaf_boxes, d_arrived);
    keep_kernel<<<B, T>>>(d_range, n - 1, d_keep);
    size_t scan_bytes = 0;
    GPU_TRY(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, d_keep, d_new, n - 1));
    void* d_scan = buf.get<char>(scan_bytes);
    GPU_TRY(cub::DeviceScan::ExclusiveSum(d_scan, scan_bytes, d_keep, d_new, n - 1));
    emit_kernel<<<B, T>>>(d_child, d_range, d_keep, d_new, d_inner_boxes, d_leaf_boxes, d_parent_inner, n - 1, d_nodes, d_scal + 6);
    gather_kernel<<<B, T>>>(d_tris, sorted, n, d_sorted_tris);
    GPU_TRY(cudaGetLastError());
    int last_keep = 0, last_new = 0, h_scal[8];
    GPU_TRY(cudaMemcpy(&last_keep, d_keep + n - 2, sizeof(int), cudaMemcpyDeviceToHost));
    GPU_TRY(cudaMemcpy(&last_new, d_new + n - 2, sizeof(int), cudaMemcpyDeviceToHost));
    GPU_TRY(cudaMemcpy(h_scal, d_scal, sizeof h_scal, cudaMemcpyDeviceToHost));
    const int n_nodes = last_new + last_keep;
    BvhBuild out;
    out.nodes.resize((size_t)n_nodes);
    out.tris.resize((size_t)n);
    GPU_TRY(cudaMemcpy(out.nodes.data(), d_nodes, (size_t)n_nodes * sizeof(BvhNode), cudaMemcpyDeviceToHost));
    GPU_TRY(cudaMemcpy(out.tris.data(), d_sorted_tris, (size_t)n * sizeof(DTri), cudaMemcpyDeviceToHost));
    Aabb root;
    GPU_TRY(cudaMemcpy(&root, d_inner_boxes, sizeof root, cudaMemcpyDeviceToHost));
    for (int a = 0; a < 3; a++) out.lo[a] = root.lo[a], out.hi[a] = root.hi[a];
    out.max_depth = h_scal[6];
    return out;
}

}  // namespace rtc
