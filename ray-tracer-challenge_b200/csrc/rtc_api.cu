// extern "C" surface declared in include/rtc_b200.h: scene upload, render calls, host loader.
#include <cuda_runtime.h>

#include <chrono>
#include <cmath>
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "host/flatten.hpp"
#include "host/loader.hpp"
#include "render_params.h"
#include "scene_build.hpp"

using namespace rtc;
struct rtc_scene;

#include <atomic>

static thread_local std::string g_error;
// Process-wide (not per thread: the lib.zig twin creates its scene on a worker thread, rtc_renderer.cpp).
static std::atomic<int> g_device{0};     // rtc_set_device: first device of scenes created afterwards
static std::atomic<int> g_n_devices{1};  // rtc_set_devices: how many devices a scene is replicated on

static int fail(int code, const std::string& msg) {
    g_error = msg;
    return code;
}
#define CUDA_TRY(expr)                                                                             \
    do {                                                                                           \
        cudaError_t e_ = (expr);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return fail(RTC_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_));         \
    } while (0)

struct rtc_scene {
    int device = 0;
    void* slab = nullptr;  // one allocation holding every array of the scene
    size_t slab_bytes = 0;
    DScene d{};
    unsigned long long* d_counters = nullptr;  // 8 x u64
    double* d_frame = nullptr;                 // scratch framebuffer for the host-output entry points
    size_t d_frame_bytes = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;  // wavefront path (synchronous)
    // Frames in flight.  A render call only ENQUEUES work on its stream (counters memset, render kernel, the two tile
    // order kernels, a 64-byte copy of the counters into pinned memory); the statistics of a frame are made out of
    // its ring slot when somebody asks for them -- the synchronous entry points at once, rtc_stats_fetch later.
    static constexpr int kRing = 16;
    struct Frame {
        int state = 0;  // 0 free, 1 on the stream, 2 statistics ready
        int best_slot = 0, lpp = 1, rc = 0;
        bool lpt = false, order_made = false, ordered = false;
        uint64_t pixels = 0;
        rtc_stats stats{};
    };
    Frame ring[kRing];
    cudaEvent_t rev0[kRing] = {}, rev1[kRing] = {}, rev2[kRing] = {};  // kernel start / end, counters landed
    unsigned long long* h_counters = nullptr;  // pinned, kRing x 8 u64
    uint64_t seq_next = 0, seq_pop = 0;        // frames enqueued / handed out
    cudaStream_t own_stream = nullptr;         // replicas (rtc_set_devices) render on a stream of their own
    std::vector<rtc_scene*> replicas;          // the same scene on the other devices (primary only)
    int grid = 0;
    size_t smem = 0;  // dynamic shared memory per CTA: parked state + the BVH traversal stacks
    // wavefront path (render_wavefront.cuh): per-level queues, fixed-point accumulators, counters
    int wf_grid = 0;
    size_t wf_smem = 0;
    double wf_scale = 1.0;
    unsigned long long wf_cap = 0;   // entries per queue
    size_t wf_pixels = 0;            // pixels the accumulators hold
    WRay* wf_rays[2] = {nullptr, nullptr};
    WHit* wf_hits[2] = {nullptr, nullptr};
    long long* wf_acc = nullptr;
    unsigned long long* wf_counters = nullptr;
    bool has_secondary = false;  // some material is reflective or transparent: rays beyond level 0 can exist
    int feat = 3;                // RTC_FEAT_* the scene needs: render_kernel<MESH, LPP, FEAT>
    // Which of the two kernels renders this scene faster is MEASURED (the answer depends on the scene: the
    // wavefront wins where pixels have long or branching ray trees, the megakernel where they do not and on
    // small frames): frames 1-4 alternate between the heuristic choice and the other kernel, later frames use the faster one.
    int frames = 0;
    // [0] megakernel, [1] wavefront, [2..4] megakernel with four lanes per pixel on the tiles that cost >= 8x / 4x / 2x the
    // mean in the frame the order was made from (small partitions only, see mixed_candidate); per pixel-count of the last frame
    float best_ms[2 + RTC_ORDER_BUCKETS] = {-1.0f, -1.0f, -1.0f, -1.0f, -1.0f, -1.0f, -1.0f, -1.0f, -1.0f, -1.0f};  // ([2 ..]: one per threshold of the order's bucket set)
    size_t tuned_pixels = 0;
    // Tile order of the megakernel: frame k adds the cost of every tile to cost buffer k & 1; frame k + 1 first makes
    // the permutation out of it (tile_order.cu), clears it for frame k + 2, then renders in that order.  Two buffers,
    // because renderers sharing a frame (share.n > 1) write into EACH OTHER's buffers: the one a peer may already be
    // filling for the next frame is never the one being read or cleared here (rtc_scene_share_tiles).
    unsigned* d_cost_own[2] = {nullptr, nullptr};  // own allocations (unused when the caller provided shared buffers)
    unsigned* d_scratch = nullptr;                 // bucket counts of tile_order.cu + the heavy-tile count
    unsigned* d_tile_order = nullptr;
    unsigned* d_row_work = nullptr;  // tile_order.cu launch_row_order: keys, sorted keys, tile numbers; then cub's temporary storage
    size_t row_work_cap = 0, row_temp_bytes = 0;
    size_t tile_cap = 0;        // tiles the own cost buffers hold
    size_t order_cap = 0;       // tiles d_scratch / d_tile_order hold
    // A frame records costs only when the order at hand was not made from this very view (camera, depth, partition): a
    // re-render of the same view keeps its order and pays nothing for it -- no cost atomics, no memset, no order kernels;
    // a camera that moves records every frame and re-orders every frame.
    struct View {
        rtc_camera cam;
        int depth;
        unsigned tile_rows, part, n_parts, deal_rank, deal_n;
        int last_bucket;  // tile_order.cu: how many cost buckets the order distinguishes; < 0: bands of -last_bucket rows (launch_row_order)
        int by_work;      // ... cheapest band first (1) or top to bottom (0)
        int bucket_set;   // tile_order.cu: 0 = buckets at 8x / 4x / 2x the mean, 1 = at 12x / 8x / 7x / 6x / 4x / 2x (shared frames)
    };
    size_t pending = 0;         // tiles the last frame recorded costs for (0: none), in cost buffer (parity ^ 1)
    View pending_view{};        // ... and the view it rendered
    size_t order_tiles = 0;     // tiles d_tile_order is a permutation of (0: no order)
    View order_view{};          // the view whose costs it was made from
    int parity = 0;             // cost buffer the NEXT recording frame writes
    unsigned* d_done = nullptr;  // RenderParams::done_list + copy_ctl (decoupled host canvases)
    size_t done_cap = 0;
    bool last_copied = false;    // the last frame's copy warps delivered it to the host canvas
    // ... or the HOST moves them: the kernel announces every finished row of tiles in h_done (pinned, mapped) and the
    // calling thread hands runs of finished rows to the copy engine on copy_stream while the kernel renders the rest
    unsigned* h_done = nullptr;      // host address
    unsigned* h_done_dev = nullptr;  // the same memory as the device addresses it
    size_t h_done_cap = 0;
    static constexpr int kCopyStreams = 4;
    cudaStream_t copy_stream[kCopyStreams] = {};  // runs of rows go to them in turn: the set-up of one copy hides behind the bytes of another
    bool last_announced = false;  // the last frame announced its rows in h_done
    struct Share {
        unsigned rank = 0, n = 1;
        unsigned* cost[2][RTC_MAX_SHARE] = {};  // [buffer][renderer], as THIS device addresses them
        size_t words = 0;                       // capacity of each buffer
        bool external = false;                  // buffers belong to the caller (rtc_scene_share_tiles)
    } share;
};

// words of one cost buffer for n_tiles tiles: the costs, padded to an even count, then the 64-bit sum
static size_t cost_words(size_t n_tiles) { return ((n_tiles + 1) & ~size_t(1)) + 2; }
static size_t frame_tiles(uint64_t hsize, uint64_t rows) {
    return (size_t)((hsize + RTC_TILE_W - 1) / RTC_TILE_W) * (size_t)((rows + RTC_TILE_H - 1) / RTC_TILE_H);
}

// The replicas of a scene (rtc_set_devices) share their frames tile by tile: cost buffers on every device, each
// addressable from all others (peer access, all pairs).  Grown when a frame needs more tiles.
static int ensure_replica_share(rtc_scene* s, size_t n_tiles) {
    const unsigned n = 1u + (unsigned)s->replicas.size();
    if (s->share.n == n && s->tile_cap >= n_tiles) return RTC_OK;
    if (n > RTC_MAX_SHARE) return fail(RTC_ERR_INVALID, "more devices than renderers that can share a frame");
    std::vector<rtc_scene*> all{s};
    all.insert(all.end(), s->replicas.begin(), s->replicas.end());
    const size_t words = cost_words(n_tiles);
    for (rtc_scene* r : all) {
        CUDA_TRY(cudaSetDevice(r->device));
        CUDA_TRY(cudaDeviceSynchronize());
        for (rtc_scene* o : all) {
            if (o == r) continue;
            cudaError_t e = cudaDeviceEnablePeerAccess(o->device, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
            else if (e != cudaSuccess) return fail(RTC_ERR_CUDA, std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e));
        }
        for (int i = 0; i < 2; i++) {
            if (r->d_cost_own[i]) cudaFree(r->d_cost_own[i]);
            r->d_cost_own[i] = nullptr;
            CUDA_TRY(cudaMalloc(&r->d_cost_own[i], words * sizeof(unsigned)));
            CUDA_TRY(cudaMemset(r->d_cost_own[i], 0, words * sizeof(unsigned)));
        }
    }
    for (unsigned k = 0; k < n; k++) {
        rtc_scene* r = all[k];
        r->share.rank = k;
        r->share.n = n;
        r->share.words = words;
        r->share.external = false;
        for (int i = 0; i < 2; i++)
            for (unsigned j = 0; j < n; j++) r->share.cost[i][j] = all[j]->d_cost_own[i];
        r->tile_cap = n_tiles;
        r->pending = 0;
        r->order_tiles = 0;
        r->parity = 0;
    }
    CUDA_TRY(cudaSetDevice(s->device));
    return RTC_OK;
}

static int launch_mega(const rtc_scene* s, const RenderParams& P, int lpp, cudaStream_t stream) {
    return launch_render(s->d.n_meshes > 0, s->feat, s->d, P, lpp, s->grid, s->smem, stream);
}

// RTC_KERNEL=mega|wave pins the kernel; default (auto) measures both on the first four frames of a scene.
static int forced_kernel() {  // read at every render: tests switch it between frames
    const char* e = std::getenv("RTC_KERNEL");
    if (!e) return -1;
    if (std::string(e) == "mega") return 0;
    if (std::string(e) == "wave") return 1;
    return -1;
}

// Lanes per pixel of the megakernel (render_kernel.cuh LPP).  RTC_LPP=1|2|4 pins it; -1 = not pinned.
static int forced_lpp(const rtc_scene* s) {
    const char* e = std::getenv("RTC_LPP");  // read at every render: tests switch it between frames
    if (!e) return -1;
    int lpp = std::atoi(e);
    if (lpp != 0 && lpp != 1 && lpp != 2 && lpp != 4) return -1;
    if (lpp == 0) return s->d.n_lights >= 2 && s->d.n_lights <= RTC_HEAVY_LANES - (RTC_HEAVY_LANES == 8 ? 1 : 0) ? 0 : 1;  // mixed: several lanes per pixel on last frame's expensive tiles
    while (lpp > 1 && s->d.n_lights < lpp) lpp >>= 1;  // lanes without a light would only repeat the closest-hit query
    return lpp;
}
// The mixed kernel (render_kernel<MESH, 0>: four lanes per pixel on the tiles that cost >= 8x the mean in the last
// frame, one elsewhere) is worth MEASURING (third candidate of the first-frames measurement below) where the frame
// time is the chain of trips behind the most expensive tile rather than the machine's throughput: partitions that
// give a resident warp fewer than 8 tiles (a 1/4 or 1/8 partition of a 1080p frame on a B200), in scenes with >= 2
// lights.  Measured on dragons.json 1080p, L2 flushed (kernel ms, one lane / two lanes / mixed): 1/4 partition
// 0.46 / 0.45 / 0.36, 1/8 partition 0.48 / 0.39 / 0.33; 1/2 partition 0.58 / - / 0.63, whole frame 1.04 / 1.60 / 1.19
// (profiles/README.md).
// Frames shared tile by tile (dealt): up to 16 tiles per warp -- a 1/2 share, whose time is its heaviest one-lane tile
// (507 us of a 0.49 ms kernel) -- and six thresholds to choose from (tile_order.cu, set 1).
static bool mixed_candidate(const rtc_scene* s, size_t n_px, bool dealt) {
    const size_t warps = (size_t)s->grid * (RTC_BLOCK_THREADS / 32);
    return s->d.n_lights >= 2 && s->d.n_lights <= RTC_HEAVY_LANES - (RTC_HEAVY_LANES == 8 ? 1 : 0) && n_px / 32 < (dealt ? 16 : 8) * warps;
}

extern "C" {

const char* rtc_last_error(void) { return g_error.c_str(); }

int rtc_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

int rtc_set_device(int device) {
    if (device < 0 || device >= rtc_device_count()) return fail(RTC_ERR_CUDA, "rtc_set_device: no such CUDA device");
    g_device = device;
    return RTC_OK;
}

int rtc_set_devices(int n) {
    const int have = rtc_device_count();
    if (n < 1 || g_device.load() + n > have) {
        char buf[128];
        std::snprintf(buf, sizeof buf, "rtc_set_devices(%d): devices %d..%d wanted, %d visible", n, g_device.load(), g_device.load() + n - 1, have);
        return fail(RTC_ERR_CUDA, buf);
    }
    g_n_devices = n;
    return RTC_OK;
}

uint64_t rtc_partition_rows(uint64_t vsize, const rtc_partition* part) {
    if (!part || part->n_parts <= 1) return vsize;
    uint64_t tr = part->tile_rows ? part->tile_rows : 1, rows = 0;
    uint64_t n_tiles = (vsize + tr - 1) / tr;
    for (uint64_t t = part->part; t < n_tiles; t += part->n_parts) {
        uint64_t y0 = t * tr, y1 = y0 + tr < vsize ? y0 + tr : vsize;
        rows += y1 - y0;
    }
    return rows;
}

void rtc_scene_destroy(rtc_scene* s) {
    if (!s) return;
    for (rtc_scene* r : s->replicas) rtc_scene_destroy(r);
    s->replicas.clear();
    cudaSetDevice(s->device);
    cudaDeviceSynchronize();  // frames may still be on a stream
    if (s->slab) cudaFree(s->slab);
    if (s->d_counters) cudaFree(s->d_counters);
    for (int i = 0; i < 2; i++)
        if (s->d_cost_own[i]) cudaFree(s->d_cost_own[i]);
    if (s->d_scratch) cudaFree(s->d_scratch);
    if (s->d_done) cudaFree(s->d_done);
    if (s->d_row_work) cudaFree(s->d_row_work);
    if (s->h_done) cudaFreeHost(s->h_done);
    for (cudaStream_t cs : s->copy_stream)
        if (cs) cudaStreamDestroy(cs);
    if (s->d_tile_order) cudaFree(s->d_tile_order);
    for (int i = 0; i < 2; i++) {
        if (s->wf_rays[i]) cudaFree(s->wf_rays[i]);
        if (s->wf_hits[i]) cudaFree(s->wf_hits[i]);
    }
    if (s->wf_acc) cudaFree(s->wf_acc);
    if (s->wf_counters) cudaFree(s->wf_counters);
    if (s->d_frame) cudaFree(s->d_frame);
    if (s->h_counters) cudaFreeHost(s->h_counters);
    if (s->ev0) cudaEventDestroy(s->ev0);
    if (s->ev1) cudaEventDestroy(s->ev1);
    for (int i = 0; i < rtc_scene::kRing; i++) {
        if (s->rev0[i]) cudaEventDestroy(s->rev0[i]);
        if (s->rev1[i]) cudaEventDestroy(s->rev1[i]);
        if (s->rev2[i]) cudaEventDestroy(s->rev2[i]);
    }
    if (s->own_stream) cudaStreamDestroy(s->own_stream);
    delete s;
}

// Host arrays of a scene -> one staging buffer with 256-byte aligned sections (uploaded with ONE copy); the offsets
// of the sections are what DScene's pointers are made of.
struct SlabLayout {
    size_t lights, materials, patterns, uvs, images, texels, transforms, leaves, normals, blob_tris, guards, nodes, prims, meshes,
        bvh, tris, csg_prog, csg_roots, mesh_leaf;
};
static SlabLayout stage_slab(const HostScene& hs, std::vector<uint8_t>& stage) {
    auto put = [&](const void* p, size_t bytes) -> size_t {
        size_t off = (stage.size() + 255) & ~size_t(255);
        stage.resize(off + bytes);
        if (bytes) std::memcpy(&stage[off], p, bytes);
        return off;
    };
#define PUT(vec) put((vec).data(), (vec).size() * sizeof((vec)[0]))
    SlabLayout o;
    o.lights = PUT(hs.lights), o.materials = PUT(hs.materials), o.patterns = PUT(hs.patterns), o.uvs = PUT(hs.uvs);
    o.images = PUT(hs.images), o.texels = PUT(hs.texels), o.transforms = PUT(hs.transforms), o.leaves = PUT(hs.leaves);
    o.normals = PUT(hs.normals), o.blob_tris = PUT(hs.blob_tris), o.guards = PUT(hs.guards), o.nodes = PUT(hs.nodes);
    o.prims = PUT(hs.prims), o.meshes = PUT(hs.meshes), o.bvh = PUT(hs.bvh), o.tris = PUT(hs.tris);
    o.csg_prog = PUT(hs.csg_prog), o.csg_roots = PUT(hs.csg_roots), o.mesh_leaf = PUT(hs.mesh_leaf);
#undef PUT
    return o;
}

int rtc_blob_inspect(const void* blob, size_t nbytes, rtc_scene_info* info) {
    if (!info) return fail(RTC_ERR_INVALID, "rtc_blob_inspect: info_out is NULL");
    std::memset(info, 0, sizeof *info);
    HostScene hs;
    auto t0 = std::chrono::steady_clock::now();
    try {
        build_host_scene(blob, nbytes, hs);
    } catch (const std::exception& e) {
        return fail(RTC_ERR_INVALID, e.what());
    }
    std::vector<uint8_t> stage;
    stage_slab(hs, stage);
    info->build_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    info->n_prims = (uint32_t)hs.prims.size();
    info->n_meshes = (uint32_t)hs.meshes.size();
    info->n_geometries = (uint32_t)hs.n_geometries;
    info->n_csg = (uint32_t)hs.csg_roots.size();
    info->n_bvh_nodes = hs.bvh.size();
    info->n_mesh_triangles = hs.tris.size();
    info->device_bytes = stage.size() + 256;
    info->bvh_max_depth = (uint32_t)hs.bvh_max_depth;
    info->n_leaves = (uint32_t)hs.leaves.size();
    return RTC_OK;
}

// The scene on ONE device.  src_slab != NULL: the slab is copied from that device's memory (peer copy over NVLink)
// instead of from the host staging buffer.
static int create_on_device(const HostScene& hs, const std::vector<uint8_t>& stage, const SlabLayout& o, int device,
                            const void* src_slab, int src_device, rtc_scene** out) {
    rtc_scene* s = new rtc_scene();
    s->device = device;
    cudaError_t e = cudaSetDevice(s->device);
    if (e != cudaSuccess) {
        delete s;
        return fail(RTC_ERR_CUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(e));
    }
    s->slab_bytes = stage.size() + 256;
    auto bail = [&](cudaError_t err, const char* what) {
        std::string msg = std::string(what) + ": " + cudaGetErrorString(err);
        rtc_scene_destroy(s);
        return fail(RTC_ERR_CUDA, msg);
    };
    if ((e = cudaMalloc(&s->slab, s->slab_bytes)) != cudaSuccess) return bail(e, "cudaMalloc(scene)");
    // ---- one slab, 256-byte aligned sections, one copy -------------------------------------------
    if (src_slab) e = cudaMemcpyPeer(s->slab, device, src_slab, src_device, stage.size());
    else e = cudaMemcpy(s->slab, stage.data(), stage.size(), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) return bail(e, "cudaMemcpy(scene)");
    if ((e = cudaMalloc(&s->d_counters, 8 * sizeof(unsigned long long))) != cudaSuccess) return bail(e, "cudaMalloc(counters)");
    if ((e = cudaHostAlloc(&s->h_counters, rtc_scene::kRing * 8 * sizeof(unsigned long long), cudaHostAllocDefault)) != cudaSuccess)
        return bail(e, "cudaHostAlloc(counters)");
    if ((e = cudaEventCreate(&s->ev0)) != cudaSuccess) return bail(e, "cudaEventCreate");
    if ((e = cudaEventCreate(&s->ev1)) != cudaSuccess) return bail(e, "cudaEventCreate");
    for (int i = 0; i < rtc_scene::kRing; i++) {
        if ((e = cudaEventCreate(&s->rev0[i])) != cudaSuccess) return bail(e, "cudaEventCreate");
        if ((e = cudaEventCreate(&s->rev1[i])) != cudaSuccess) return bail(e, "cudaEventCreate");
        if ((e = cudaEventCreateWithFlags(&s->rev2[i], cudaEventDisableTiming)) != cudaSuccess) return bail(e, "cudaEventCreate");
    }
    const uint8_t* b = static_cast<const uint8_t*>(s->slab);
    DScene& d = s->d;
    d.lights = (const rtc_light*)(b + o.lights);
    d.materials = (const rtc_material*)(b + o.materials);
    d.patterns = (const rtc_pattern*)(b + o.patterns);
    d.uvs = (const rtc_uv_pattern*)(b + o.uvs);
    d.images = (const rtc_image*)(b + o.images);
    d.texels = b + o.texels;
    d.transforms = (const rtc_transform*)(b + o.transforms);
    d.leaves = (const rtc_leaf*)(b + o.leaves);
    d.normals = (const rtc_tri_normals*)(b + o.normals);
    d.blob_tris = (const rtc_triangle*)(b + o.blob_tris);
    d.guards = (const rtc_guard*)(b + o.guards);
    d.nodes = (const rtc_node*)(b + o.nodes);
    d.prims = (const DPrim*)(b + o.prims);
    d.meshes = (const DMesh*)(b + o.meshes);
    d.bvh = (const BvhNode*)(b + o.bvh);
    d.tris = (const DTri*)(b + o.tris);
    d.mesh_leaf = (const int32_t*)(b + o.mesh_leaf);
    d.csg_prog = (const int4*)(b + o.csg_prog);
    d.csg_roots = (const DCsgRoot*)(b + o.csg_roots);
    d.n_lights = (int32_t)hs.lights.size();
    d.n_prims = (int32_t)hs.prims.size();
    d.n_meshes = (int32_t)hs.meshes.size();
    d.n_csg = (int32_t)hs.csg_roots.size();
    d.n_leaves = (int32_t)hs.leaves.size();
    d.bvh_max_depth = hs.bvh_max_depth;
    for (int a = 0; a < 3; a++) d.world_abs[a] = hs.world_abs[a];

    // persistent grid: every SM filled to the kernel's occupancy
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, s->device)) != cudaSuccess) return bail(e, "cudaGetDeviceProperties");
    int per_sm = 0;
    d.stack_cap = d.n_meshes > 0 ? hs.bvh_max_depth + 2 : 0;
    // parked per-lane state (render_kernel.cuh PK_SLOTS doubles per thread) + the BVH traversal stacks
    s->smem = (size_t)RTC_PARK_SLOTS * RTC_BLOCK_THREADS * sizeof(double) +
              (size_t)d.stack_cap * RTC_BLOCK_THREADS * sizeof(int);
    {
        // which of the rarer machinery the scene needs; RTC_LEAN=0 keeps the general instantiation (tests compare them)
        const char* le = std::getenv("RTC_LEAN");
        s->feat = 0;
        if (!hs.csg_roots.empty() || !hs.guards.empty()) s->feat |= RTC_FEAT_CSG;
        for (const rtc_material& m : hs.materials) {  // rt_pattern.cuh simple_pattern_color covers exactly these
            const rtc_pattern& pt = hs.patterns[(size_t)m.pattern];
            const bool two_solids = pt.a >= 0 && pt.b >= 0 && hs.patterns[(size_t)pt.a].kind == RTC_PAT_SOLID &&
                                    hs.patterns[(size_t)pt.b].kind == RTC_PAT_SOLID;
            const bool simple = pt.kind == RTC_PAT_SOLID ||
                                (two_solids && (pt.kind == RTC_PAT_STRIPES || pt.kind == RTC_PAT_RINGS || pt.kind == RTC_PAT_CHECKERS ||
                                                pt.kind == RTC_PAT_GRADIENT || pt.kind == RTC_PAT_RADIAL_GRADIENT || pt.kind == RTC_PAT_BLEND));
            if (!simple) s->feat |= RTC_FEAT_PATTERNS;
        }
        // (CSG with simple patterns only -- csg.json -- measured no faster than the general instantiation: 1.43 against
        // 1.38 ms at 3840x2160; that combination is not built)
        if ((s->feat & RTC_FEAT_CSG) || (le && le[0] == '0')) s->feat = RTC_FEAT_CSG | RTC_FEAT_PATTERNS;
    }
    e = (cudaError_t)occupancy_render(d.n_meshes > 0, s->feat, &per_sm, s->smem);
    if (e != cudaSuccess) return bail(e, "occupancy query");
    if (per_sm < 1) per_sm = 1;
    s->grid = prop.multiProcessorCount * per_sm;
    // wavefront kernels: no parked state, only the mailboxes and the traversal stacks
    s->wf_smem = d.n_meshes > 0 ? (size_t)RTC_MAILBOX_SLOTS * RTC_BLOCK_THREADS * sizeof(double) + (size_t)d.stack_cap * RTC_BLOCK_THREADS * sizeof(int) : 0;
    d.ri_all_one = 1;
    double branch = 2.0;  // sum of the weights of a hit's children relative to its own: reflective + transparency
    for (const rtc_material& m : hs.materials) {
        if (m.reflective != 0.0 || m.transparency != 0.0) s->has_secondary = true;
        if (m.refractive_index != 1.0) d.ri_all_one = 0;
        branch = std::fmax(branch, std::fabs(m.reflective) + std::fabs(m.transparency));
    }
    int wf_per_sm = 0;
    e = (cudaError_t)(d.n_meshes > 0 ? occupancy_wave_mesh(&wf_per_sm, s->wf_smem) : occupancy_wave_analytic(&wf_per_sm, s->wf_smem));
    if (e != cudaSuccess) return bail(e, "occupancy query (wavefront)");
    if (wf_per_sm < 1) wf_per_sm = 1;
    s->wf_grid = prop.multiProcessorCount * wf_per_sm;
    // Fixed-point unit of the pixel accumulators: a pixel is the sum over the nodes of its ray tree of weight x lighting();
    // the weights of level d add up to at most branch^d (branch = max over materials of reflective + transparency, at
    // least 2: the JSON allows values above 1), levels 0..7; |lighting()| <= sum over lights |I| x (ambient + diffuse +
    // specular) x max |colour|.  A term beyond the range (NaN included) sets ERR_ACC_RANGE instead of wrapping.
    {
        double li = 0.0, coef = 1.0, col = 1.0, levels = 0.0, w = 1.0;
        for (int lv = 0; lv <= 7; lv++, w *= branch) levels += w;
        for (const rtc_light& l : hs.lights)
            li += std::fmax(std::fabs(l.intensity[0]), std::fmax(std::fabs(l.intensity[1]), std::fabs(l.intensity[2])));
        for (const rtc_material& m : hs.materials)
            coef = std::fmax(coef, std::fabs(m.ambient) + std::fabs(m.diffuse) + std::fabs(m.specular));
        for (const rtc_pattern& pt : hs.patterns)
            for (int k = 0; k < 3; k++) col = std::fmax(col, std::fabs(pt.color[k]));
        double bound = levels * std::fmax(li, 1.0) * coef * col;
        int ex = 0;
        if (!(bound < 1e300)) bound = 1e300;
        std::frexp(bound, &ex);          // bound < 2^ex
        int sh = 62 - ex;
        if (sh > 52) sh = 52;            // never finer than a double's own resolution at 1.0
        if (sh < -900) sh = -900;
        s->wf_scale = std::ldexp(1.0, sh);
    }
    *out = s;
    return RTC_OK;
}

int rtc_scene_create(const void* blob, size_t nbytes, rtc_scene** out) {
    if (!out) return fail(RTC_ERR_INVALID, "rtc_scene_create: out is NULL");
    *out = nullptr;
    if (rtc_device_count() <= 0) return fail(RTC_ERR_CUDA, "rtc_scene_create: no CUDA device (there is no CPU fallback)");
    const int first = g_device.load(), n_dev = g_n_devices.load();
    CUDA_TRY(cudaSetDevice(first));  // (large meshes build their BVH on it)
    HostScene hs;
    try {
        build_host_scene(blob, nbytes, hs);
    } catch (const std::exception& e) {
        return fail(RTC_ERR_INVALID, e.what());
    }
    if (hs.bvh_max_depth > 190) return fail(RTC_ERR_CAPACITY, "BVH deeper than the traversal stack can be made");
    if (hs.csg_max_segments > 16) return fail(RTC_ERR_CAPACITY, "CSG nesting deeper than the segment stack");
    std::vector<uint8_t> stage;
    const SlabLayout o = stage_slab(hs, stage);
    if (first + n_dev > rtc_device_count()) return fail(RTC_ERR_CUDA, "rtc_scene_create: fewer CUDA devices than rtc_set_devices asked for");
    rtc_scene* s = nullptr;
    int rc = create_on_device(hs, stage, o, first, nullptr, 0, &s);
    if (rc) return rc;
    // rtc_set_devices(n): the scene is parsed and built ONCE and its slab replicated device to device; every replica
    // may store into device 0's memory (frames assembled there) -- camera.zig:88-97 hands one row to each of n_threads
    // workers of ONE process, this hands interleaved row tiles to the n devices of one process.
    for (int k = 1; k < n_dev; k++) {
        rtc_scene* r = nullptr;
        rc = create_on_device(hs, stage, o, first + k, s->slab, first, &r);
        if (rc) {
            rtc_scene_destroy(s);
            return rc;
        }
        s->replicas.push_back(r);
        cudaError_t e = cudaStreamCreateWithFlags(&r->own_stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) {
            e = cudaDeviceEnablePeerAccess(first, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) {
                cudaGetLastError();
                e = cudaSuccess;
            }
        }
        if (e != cudaSuccess) {
            std::string msg = std::string("replica on device ") + std::to_string(first + k) + ": " + cudaGetErrorString(e);
            rtc_scene_destroy(s);
            return fail(RTC_ERR_CUDA, msg);
        }
    }
    cudaSetDevice(first);
    *out = s;
    return RTC_OK;
}

// Grows (never shrinks) the wavefront buffers of a scene for a partition of n_pixels pixels.
static int ensure_wavefront(rtc_scene* s, size_t n_pixels) {
    if (!s->wf_counters) CUDA_TRY(cudaMalloc(&s->wf_counters, WC_COUNT * sizeof(unsigned long long)));
    if (s->wf_pixels >= n_pixels) return RTC_OK;
    for (int i = 0; i < 2; i++) {
        if (s->wf_rays[i]) cudaFree(s->wf_rays[i]);
        if (s->wf_hits[i]) cudaFree(s->wf_hits[i]);
        s->wf_rays[i] = nullptr;
        s->wf_hits[i] = nullptr;
    }
    if (s->wf_acc) cudaFree(s->wf_acc);
    s->wf_acc = nullptr;
    s->wf_pixels = 0;
    // a level holds at most min(2^d, ...) x pixels rays; 3 x pixels covers every shipped scene with room to
    // spare, and a frame that overflows is redone by the megakernel (never truncated)
    unsigned long long cap = 3ull * n_pixels + (1ull << 16);
    for (int i = 0; i < 2; i++) {
        CUDA_TRY(cudaMalloc(&s->wf_rays[i], cap * sizeof(WRay)));
        CUDA_TRY(cudaMalloc(&s->wf_hits[i], cap * sizeof(WHit)));
    }
    CUDA_TRY(cudaMalloc(&s->wf_acc, n_pixels * 3 * sizeof(long long)));
    s->wf_cap = cap;
    s->wf_pixels = n_pixels;
    return RTC_OK;
}

// The frame as max_depth + 2 wavefront launches + one resolve (render_wavefront.cuh).
static int render_wavefront(rtc_scene* s, const RenderParams& P, cudaStream_t stream, rtc_stats* stats) {
    const size_t n_pixels = (size_t)(P.rows * P.cam.hsize);
    if (n_pixels >= (1ull << 31)) return fail(RTC_ERR_INVALID, "render: partition larger than 2^31 pixels");
    int rc = ensure_wavefront(s, n_pixels);
    if (rc) return rc;
    WaveParams W{};
    W.cam = P.cam;
    W.max_depth = P.max_depth;
    W.tile_rows = P.tile_rows;
    W.part = P.part;
    W.n_parts = P.n_parts;
    W.rows = P.rows;
    W.n_pixels = n_pixels;
    W.cap = s->wf_cap;
    W.acc = s->wf_acc;
    W.scale = s->wf_scale;
    W.counters = s->wf_counters;
    CUDA_TRY(cudaMemsetAsync(s->wf_counters, 0, WC_COUNT * sizeof(unsigned long long), stream));
    CUDA_TRY(cudaEventRecord(s->ev0, stream));
    CUDA_TRY(cudaMemsetAsync(s->wf_acc, 0, n_pixels * 3 * sizeof(long long), stream));
    // without reflective / transparent materials no ray beyond level 0 can exist: level 0 + its shadows
    const int depth = s->has_secondary ? P.max_depth : 0;
    const int n_launch = depth + 2;  // levels 0..depth, + the shadow queries of the last level's hits
    for (int level = 0; level < n_launch; level++) {
        W.level = level;
        W.rays_in = s->wf_rays[level & 1];
        W.rays_out = s->wf_rays[(level + 1) & 1];
        W.hits_in = s->wf_hits[(level + 1) & 1];
        W.hits_out = s->wf_hits[level & 1];
        CUDA_TRY((cudaError_t)(s->d.n_meshes > 0 ? launch_wave_mesh(s->d, W, s->wf_grid, s->wf_smem, stream)
                                                 : launch_wave_analytic(s->d, W, s->wf_grid, s->wf_smem, stream)));
    }
    ResolveParams R{};
    R.hsize = P.cam.hsize;
    R.rows = P.rows;
    R.n_pixels = n_pixels;
    R.tile_rows = P.tile_rows;
    R.part = P.part;
    R.n_parts = P.n_parts;
    R.full_frame = P.full_frame;
    R.acc = s->wf_acc;
    R.inv_scale = 1.0 / s->wf_scale;
    R.out = P.out;
    R.out8 = P.out8;
    CUDA_TRY((cudaError_t)launch_resolve(R, stream));
    CUDA_TRY(cudaEventRecord(s->ev1, stream));
    unsigned long long c[WC_COUNT];
    CUDA_TRY(cudaMemcpyAsync(c, s->wf_counters, sizeof c, cudaMemcpyDeviceToHost, stream));
    CUDA_TRY(cudaStreamSynchronize(stream));
    float ms = 0;
    CUDA_TRY(cudaEventElapsedTime(&ms, s->ev0, s->ev1));
    if (c[WC_ERR] & 16ull) return RTC_ERR_CAPACITY + 1000;  // ERR_QUEUE: the caller redoes the frame with the megakernel
    if (stats) {
        unsigned long long hits = 0, secondary = 0;
        for (int l = 0; l <= P.max_depth; l++) hits += c[WC_NHITS + l];
        for (int l = 1; l <= P.max_depth; l++) secondary += c[WC_NRAYS + l];
        stats->primary_rays = n_pixels;
        stats->shadow_rays = hits * (unsigned long long)s->d.n_lights;
        stats->secondary_rays = secondary;
        stats->pixels = n_pixels;
        stats->kernel_ms = ms;
        stats->launches = (unsigned long long)n_launch + 1;  // + resolve
        stats->kernel = 1u;
        stats->shadow_rays_culled = c[WC_CULLED];
    }
    if (c[WC_ERR]) {
        char buf[224];
        std::snprintf(buf, sizeof buf,
                      "render: per-ray capacity exceeded (bits %llu: 1=texture lookup out of range, 2=pattern stack, "
                      "4=CSG list, 8=BVH stack, 32=pixel sum out of the accumulator's range)",
                      c[WC_ERR]);
        return fail(RTC_ERR_CAPACITY, buf);
    }
    return RTC_OK;
}

// ---- frames --------------------------------------------------------------------------------------------------------
struct Target {              // where a frame goes
    double* rgb = nullptr;   // device-addressable f64 canvas (device memory, peer memory or mapped host memory), or
    unsigned char* rgba = nullptr;  // RGBA8
    bool full_frame = false; // rows of the partition at their frame positions (else packed)
    bool spread_rows = false;  // canvas in mapped host memory: strided row order, no cost order (see enqueue_frame)
    double* copy_out = nullptr;  // mapped host canvas the frame in `rgb` (device memory) is moved to BY THE KERNEL's copy warps
    bool announce = false;       // the frame in `rgb` / `rgba` (device memory) is moved by the HOST: finished rows of tiles are announced in h_done
};

// Makes the statistics of the frame in ring slot `slot` (waits for it if it is still on the stream).
static int resolve_frame(rtc_scene* s, int slot) {
    rtc_scene::Frame& f = s->ring[slot];
    if (f.state != 1) return f.state == 2 ? f.rc : RTC_OK;
    CUDA_TRY(cudaSetDevice(s->device));
    CUDA_TRY(cudaEventSynchronize(s->rev2[slot]));
    float ms = 0;
    CUDA_TRY(cudaEventElapsedTime(&ms, s->rev0[slot], s->rev1[slot]));
    const unsigned long long* c = s->h_counters + 8 * slot;
#ifdef RTC_TRIP_PROBE
    std::fprintf(stderr, "trip probe: max trip %.1f us, trips %llu, > 50 us: %llu, > 20 us: %llu (kernel %.3f ms)\n", c[6] / 1965.0,
                 c[7] & 0xffffffffull, (c[7] >> 32) & 0xffffull, c[7] >> 48, ms);
#endif
    f.stats.primary_rays = c[1];
    f.stats.shadow_rays = c[2];
    f.stats.secondary_rays = c[3];
    f.stats.pixels = f.pixels;
    f.stats.kernel_ms = ms;  // render kernel + the two tile order kernels behind it
    f.stats.launches = f.order_made ? 3 : 1;  // render kernel (+ the two kernels of tile_order.cu when this frame made a new order)
    f.stats.kernel = f.lpp != 1 ? 2u : 0u;
    f.stats.shadow_rays_culled = c[5];
    if (s->best_ms[f.best_slot] < 0.0f || ms < s->best_ms[f.best_slot]) s->best_ms[f.best_slot] = ms;
    static const bool trace = std::getenv("RTC_TRACE_FRAMES") != nullptr;
    if (trace)
        std::fprintf(stderr, "frame: dev %d px %llu candidate %d lanes/px %d ordered %d recorded %d order made %d  %.4f ms\n", s->device,
                     (unsigned long long)f.pixels, f.best_slot, f.lpp, (int)f.ordered, (int)f.lpt, (int)f.order_made, ms);
    f.state = 2;
    f.rc = RTC_OK;
    if (c[4]) {
        char buf[224];
        std::snprintf(buf, sizeof buf,
                      "render: per-ray capacity exceeded (bits %llu: 1=texture lookup out of range, 2=pattern stack, "
                      "4=CSG list, 8=BVH stack, 32=pixel sum out of the accumulator's range, 64=host canvas copy)",
                      c[4]);
        f.rc = fail(RTC_ERR_CAPACITY, buf);
    }
    return f.rc;
}

// Puts one frame (one partition on one device) on `stream`.  Returns the ring slot in *slot_out; with wait = true the
// frame's statistics are ready on return (the wavefront path and the first frames of a scene, whose times pick the
// kernel, always are).
static int enqueue_frame(rtc_scene* s, const rtc_camera* cam, int max_depth, const rtc_partition* part, const Target& tg,
                         cudaStream_t stream, bool wait, int* slot_out, bool dealt = false) {
    if (!s || !cam) return fail(RTC_ERR_INVALID, "render: NULL scene or camera");
    if (max_depth < 0 || max_depth > 7) return fail(RTC_ERR_INVALID, "render: max_depth must be in [0,7]");
    if (cam->hsize == 0 || cam->vsize == 0) return fail(RTC_ERR_INVALID, "render: empty frame");
    rtc_partition whole = {1, 0, 1};
    if (!part) part = &whole;
    if (part->n_parts == 0 || part->part >= part->n_parts || part->tile_rows == 0)
        return fail(RTC_ERR_INVALID, "render: bad partition");
    CUDA_TRY(cudaSetDevice(s->device));
    s->last_copied = s->last_announced = false;
    // a free ring slot: the oldest frame nobody fetched is resolved (its time still counts for the kernel choice) and dropped
    if (s->seq_next - s->seq_pop == (uint64_t)rtc_scene::kRing) {
        const int old = (int)(s->seq_pop % rtc_scene::kRing);
        resolve_frame(s, old);
        s->ring[old].state = 0;
        s->seq_pop++;
    }
    const int slot = (int)(s->seq_next % rtc_scene::kRing);
    rtc_scene::Frame& F = s->ring[slot];
    F = rtc_scene::Frame();
    // The frames of a scene share its counters, tile costs and tile order: a frame put on ANOTHER stream than the one
    // before it must still run behind it (and behind the order kernels that read its costs).
    if (s->seq_next > 0) CUDA_TRY(cudaStreamWaitEvent(stream, s->rev2[(int)((s->seq_next - 1) % rtc_scene::kRing)], 0));

    RenderParams P{};
    P.cam.hsize = cam->hsize;
    P.cam.vsize = cam->vsize;
    P.cam.half_width = cam->half_width;
    P.cam.half_height = cam->half_height;
    P.cam.pixel_size = cam->pixel_size;
    std::memcpy(P.cam.inv, cam->inverse, sizeof P.cam.inv);
    P.max_depth = max_depth;
    P.tile_rows = part->tile_rows;
    P.part = part->part;
    P.n_parts = part->n_parts;
    P.rows = rtc_partition_rows(cam->vsize, part);
    P.tiles_x = (unsigned)((cam->hsize + RTC_TILE_W - 1) / RTC_TILE_W);
    const unsigned n_tile_rows = (unsigned)((P.rows + RTC_TILE_H - 1) / RTC_TILE_H);
    P.n_items = (unsigned long long)P.tiles_x * n_tile_rows * 32ull;
    // dealt: this renderer and share.n - 1 others render ONE whole frame between them, tile by tile (RenderParams::deal_n)
    P.deal_n = dealt ? s->share.n : 1u;
    P.deal_rank = dealt ? s->share.rank : 0u;
    // runs of 8 consecutive entries of the order per renderer: a renderer's tiles stay neighbours (shared BVH nodes and
    // texels in L1) while every cost bucket is still spread evenly (measured on one B200 rendering 1/8 shares: dealt one by
    // one cover.json 0.73 ms, teapot 0.133 against 0.49 / 0.114 for row partitions; RTC_DEAL_SHIFT overrides)
    static const unsigned deal_shift = [] {
        const char* e = std::getenv("RTC_DEAL_SHIFT");
        int v = e ? std::atoi(e) : 3;
        return (unsigned)(v >= 0 && v <= 8 ? v : 3);
    }();
    P.deal_shift = dealt ? deal_shift : 0u;
    const size_t n_tiles = (size_t)(P.n_items / 32);  // of the frame (dealt) or of the partition
    if (dealt) {
        if (P.n_parts != 1 || !tg.full_frame || tg.spread_rows) return fail(RTC_ERR_INVALID, "render: a shared frame is a whole frame in device memory");
        const size_t run = size_t(1) << P.deal_shift, round = run * P.deal_n, rem = n_tiles % round, lo = (size_t)P.deal_rank * run;
        const size_t mine = n_tiles / round * run + (rem > lo ? std::min(rem - lo, run) : 0);
        P.n_items = (unsigned long long)mine * 32ull;
    }
    {
        // Order of the tile rows.  Frame in device memory: top to bottom (neighbouring rows share BVH nodes and
        // texels).  Frame in mapped host memory: groups of rows in golden-ratio stride order, so that pixels are
        // produced at an even rate whatever the picture shows (RenderParams::row_stride).
        static const long forced = [] {
            const char* e = std::getenv("RTC_ROW_STRIDE");
            return e ? std::atol(e) : -1L;
        }();
        static const long group = [] {
            const char* e = std::getenv("RTC_ROW_GROUP");
            return e ? std::atol(e) : 4L;
        }();
        P.row_group = (unsigned)(group > 0 ? group : 1);
        P.n_row_groups = n_tile_rows / P.row_group;
        unsigned st = 1;
        static const bool copy_spread = [] {
            const char* e = std::getenv("RTC_COPY_SPREAD");
            return !(e && e[0] == '0');
        }();
        static const bool dma_spread = [] {  // rows announced to the host: measured both ways (profiles/README.md)
            const char* e = std::getenv("RTC_DMA_SPREAD");
            return e && e[0] == '1';
        }();
        if (forced >= 0 ? forced != 1 : (tg.spread_rows || (tg.copy_out && copy_spread) || (tg.announce && dma_spread))) {
            st = forced > 1 ? (unsigned)forced : (unsigned)(0.381966 * P.n_row_groups);
            auto gcd = [](unsigned a, unsigned b) {
                while (b) {
                    unsigned t = a % b;
                    a = b;
                    b = t;
                }
                return a;
            };
            while (st > 1 && gcd(st, P.n_row_groups) != 1) st--;
            if (st == 0) st = 1;
        }
        P.row_stride = P.n_row_groups > 1 ? st % P.n_row_groups : 1;
        if (P.row_stride == 0 || P.n_row_groups == 0) {
            P.row_stride = 1;
            P.n_row_groups = 0;
        }
    }
    P.full_frame = tg.full_frame ? 1 : 0;
    P.scale = s->wf_scale;
    P.out = tg.rgb;
    P.out8 = tg.rgba;
    P.counters = s->d_counters;
    const size_t n_px = dealt ? (size_t)(P.n_items) : (size_t)(P.rows * cam->hsize);  // (dealt: pixel slots of its tiles)
    F.pixels = n_px;
    if (P.rows == 0 || P.n_items == 0) {  // an empty partition: nothing to do, statistics all zero
        F.state = 2;
        s->seq_next++;
        *slot_out = slot;
        return RTC_OK;
    }
    if (s->tuned_pixels != n_px) {  // another frame size (or partition): measure again
        s->tuned_pixels = n_px;
        s->frames = 0;
        for (float& b : s->best_ms) b = -1.0f;
    }
    int which = forced_kernel();
    if (dealt && which == 1) which = 0;  // the wavefront path renders row partitions only
    bool measuring = false;
    if (which < 0) {
        const int guess = 0;
        const bool three = forced_lpp(s) < 0 && mixed_candidate(s, n_px, dealt);
        const int n_thr = dealt ? 6 : 3;  // thresholds of the order's bucket set (tile_order.cu)
        // Frames 0-2: megakernel (the first is cold, the second makes the tile order, the third is what a steady frame
        // costs), frames 3-4: wavefront (the first loads its kernels), then -- small partitions -- two frames each with
        // the tiles that cost >= 8x, >= 4x, >= 2x the mean four-fold.  The best time of each candidate counts.
        (void)guess;
        if (s->frames < 3) {
            which = 0;
            measuring = true;
        } else if (s->frames < 5 && !dealt) {
            which = 1;
            measuring = true;
        } else if (three && s->frames < (dealt ? 3 : 5) + 2 * n_thr) {  // (a shared frame has no wavefront frames)
            which = 2 + (s->frames - (dealt ? 3 : 5)) / 2;
            measuring = true;
        } else {
            which = (!dealt && s->best_ms[1] >= 0.0f && (s->best_ms[0] < 0.0f || s->best_ms[1] < s->best_ms[0])) ? 1 : 0;
            if (three)
                for (int k = 2; k < 2 + n_thr; k++)
                    if (s->best_ms[k] >= 0.0f && s->best_ms[k] < s->best_ms[which]) which = k;
            // A frame that is moved to a host canvas while it renders (rows announced, or copy warps): only the megakernel
            // with one lane per pixel finishes rows before the end.  A wavefront frame, or one with several lanes per pixel,
            // is copied afterwards, kernel + link, where the megakernel's costs about max(kernel, link) (csg.json 2160p:
            // wavefront 1.59 + 3.6 ms against megakernel 1.6 beside 3.6).
            if (which != 0 && (tg.announce || tg.copy_out) && s->best_ms[0] > 0.0f) {
                const double link_ms = (double)P.rows * cam->hsize * (tg.rgba ? 4.0 : 24.0) / 55e6;
                if (s->best_ms[which] + link_ms > 1.05 * std::max((double)s->best_ms[0], link_ms)) which = 0;
            }
        }
    }
    int lpp = forced_lpp(s);
    if (lpp < 0) lpp = which >= 2 ? 0 : 1;
    F.best_slot = which >= 2 ? which : 0;  // where a megakernel frame's time is recorded
    s->frames++;
    if (which == 1) {
        rtc_stats local{};
        int rc = render_wavefront(s, P, stream, &local);  // synchronous: its queue counters decide about a redo
        if (rc != RTC_ERR_CAPACITY + 1000) {
            if (rc == RTC_OK && (s->best_ms[1] < 0.0f || local.kernel_ms < s->best_ms[1])) s->best_ms[1] = local.kernel_ms;
            if (std::getenv("RTC_TRACE_FRAMES")) std::fprintf(stderr, "frame: dev %d px %llu wavefront  %.4f ms\n", s->device, (unsigned long long)n_px, local.kernel_ms);
            CUDA_TRY(cudaEventRecord(s->rev2[slot], stream));
            F.stats = local;
            F.state = 2;
            F.rc = rc;
            s->seq_next++;
            *slot_out = slot;
            return rc;
        }
        s->best_ms[1] = 1e30f;  // queue overflow: this scene stays with the megakernel; redo the frame below
    }
    // Tiles in the order of their cost in the previous frame of this size (tile_order.cu); RTC_LPT=0 keeps the plain order.
    // Not for canvases in mapped host memory: there the time a warp spends on a tile is mostly the back-pressure of its
    // stores on the PCIe link, and the even byte rate of the strided order matters more (measured: e2e 1.49 -> 1.91 ms;
    // with the cost counted in trips of the render loop instead of clock ticks 1.80: it is the burst of cheap tiles at the
    // end that costs, not the noise of the measurement).
    static const int lpt_env = [] {
        const char* e = std::getenv("RTC_LPT");
        return e ? std::atoi(e) : 1;
    }();
    const bool lpt = dealt || lpt_env == 2 || (lpt_env == 1 && !tg.spread_rows);
    F.lpt = lpt;
    // A whole frame for a mapped host canvas, one lane per pixel: rendered into device memory, moved by copy warps.
    static const unsigned copy_ctas = [] {
        const char* e = std::getenv("RTC_COPY_CTAS");
        int v = e ? std::atoi(e) : 4;  // measured on B200, dragons 1080p e2e ms per frame: 4 CTAs 1.27, 8: 1.28, 16: 1.29 (direct stores: 1.54)
        return (unsigned)(v >= 0 && v <= 64 ? v : 4);
    }();
    const bool by_warps = tg.copy_out && tg.rgb && copy_ctas > 0 && (int)copy_ctas * 2 <= s->grid;
    const bool by_host = tg.announce && (tg.rgb || tg.rgba);
    if ((by_warps || by_host) && tg.full_frame && !dealt && (P.n_parts == 1 || P.tile_rows % RTC_TILE_H == 0) && lpp == 1) {
        // per row of tiles: a counter of finished tiles; the list of finished rows; the two cursors -- all set to 0xffffffff
        const size_t n_rows = n_tile_rows, words = 2 * n_rows + 2;
        if (s->done_cap < words) {
            CUDA_TRY(cudaStreamSynchronize(stream));
            if (s->d_done) cudaFree(s->d_done);
            s->d_done = nullptr;
            s->done_cap = 0;
            CUDA_TRY(cudaMalloc(&s->d_done, words * sizeof(unsigned)));
            s->done_cap = words;
        }
        CUDA_TRY(cudaMemsetAsync(s->d_done, 0xff, words * sizeof(unsigned), stream));
        P.done_list = s->d_done;
        P.copy_ctl = s->d_done + n_rows;
        P.row_count = s->d_done + n_rows + 2;
        P.copy_rows = (unsigned)n_rows;
        if (by_host) {
            if (s->h_done_cap < n_rows) {
                CUDA_TRY(cudaStreamSynchronize(stream));
                if (s->h_done) cudaFreeHost(s->h_done);
                s->h_done = s->h_done_dev = nullptr;
                s->h_done_cap = 0;
                CUDA_TRY(cudaHostAlloc(&s->h_done, n_rows * sizeof(unsigned), cudaHostAllocMapped));
                CUDA_TRY(cudaHostGetDevicePointer(&s->h_done_dev, s->h_done, 0));
                s->h_done_cap = n_rows;
            }
            for (cudaStream_t& cs : s->copy_stream)
                if (!cs) CUDA_TRY(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
            // (the frame before this one was collected by the caller before it came back: nobody reads the list any more)
            std::memset(s->h_done, 0xff, n_rows * sizeof(unsigned));
            P.copy_out = tg.rgb ? tg.rgb : reinterpret_cast<double*>(tg.rgba);  // (here it only says: count the finished rows)
            P.done_list = s->h_done_dev;
            P.copy_ctas = 0;  // no copy warps
            s->last_announced = true;
        } else {
            P.copy_out = tg.copy_out;
            P.copy_ctas = copy_ctas;
            s->last_copied = true;
        }
    }
    CUDA_TRY(cudaMemsetAsync(s->d_counters, 0, 8 * sizeof(unsigned long long), stream));
    CUDA_TRY(cudaEventRecord(s->rev0[slot], stream));
    if (lpt) {
        const size_t words = cost_words(n_tiles);
        if (s->share.external) {
            if (words > s->share.words) return fail(RTC_ERR_INVALID, "render: the shared tile buffers are too small for this frame");
        } else if (s->tile_cap < n_tiles) {
            if (s->share.n > 1) return fail(RTC_ERR_INVALID, "render: frame larger than the replicas' tile buffers");
            CUDA_TRY(cudaStreamSynchronize(stream));  // frames of the old size may still use the arrays
            for (int i = 0; i < 2; i++) {
                if (s->d_cost_own[i]) cudaFree(s->d_cost_own[i]);
                s->d_cost_own[i] = nullptr;
                CUDA_TRY(cudaMalloc(&s->d_cost_own[i], words * sizeof(unsigned)));
                CUDA_TRY(cudaMemsetAsync(s->d_cost_own[i], 0, words * sizeof(unsigned), stream));
                s->share.cost[i][0] = s->d_cost_own[i];
            }
            s->share.words = words;
            s->tile_cap = n_tiles;
            s->pending = 0;
        }
        if (s->order_cap < n_tiles) {  // the arrays that are never shared: bucket counts, the permutation
            CUDA_TRY(cudaStreamSynchronize(stream));
            if (s->d_scratch) cudaFree(s->d_scratch);
            if (s->d_tile_order) cudaFree(s->d_tile_order);
            s->d_scratch = s->d_tile_order = nullptr;
            s->order_cap = 0;
            CUDA_TRY(cudaMalloc(&s->d_scratch, RTC_ORDER_BUCKETS * (n_tiles / 1024 + 2) * sizeof(unsigned)));
            CUDA_TRY(cudaMalloc(&s->d_tile_order, n_tiles * sizeof(unsigned)));
            s->order_cap = n_tiles;
            s->order_tiles = 0;
        }
        const int b = s->parity;
        unsigned* mine_prev = s->share.cost[b ^ 1][s->share.rank];
        rtc_scene::View view{};
        view.cam = *cam;
        view.depth = max_depth;
        view.tile_rows = P.tile_rows, view.part = P.part, view.n_parts = P.n_parts, view.deal_rank = P.deal_rank, view.deal_n = P.deal_n;
        // A frame whose finished tiles are streamed to a host canvas keeps the bulk of its tiles in memory order (only the
        // tiles that cost >= 8x the mean come first).  Measured on B200, dragons.json 1080p, e2e ms per frame: four
        // buckets 1.62 (the cheap 85 % of the bytes all come at the end), plain order 1.25, >= 8x first 1.21; the expensive
        // buckets interleaved evenly with the cheap tiles 1.21 -- and skybox.json 4K 9.6 instead of 2.3: writes that
        // jump about the host canvas are slow, writes that walk through it are not (profiles/README.md).
        // < 0: bands of -last_bucket rows of tiles, cheapest band first, every tile started ahead of its row by what it
        // costs (tile_order.cu launch_row_order).  A band is what one copy moves: ~1.5 MB (a copy is set up in 3.5 us), or the
        // block of rows a partition holds together in the frame.  RTC_COPY_BUCKETS >= 0 keeps the bucket orders.
        static const int copy_buckets = [] {
            const char* e = std::getenv("RTC_COPY_BUCKETS");
            int v = e ? std::atoi(e) : -1;
            return v >= -1 && v <= 3 ? v : -1;
        }();
        view.last_bucket = dealt ? 6 : 3;
        view.bucket_set = dealt ? 1 : 0;
        if (tg.copy_out || tg.announce) {
            view.last_bucket = copy_buckets;
            if (copy_buckets < 0) {
                static const long band_kb = [] {
                    const char* e = std::getenv("RTC_BAND_KB");
                    long v = e ? std::atol(e) : 1536;
                    return v > 0 ? v : 1536;
                }();
                const size_t tile_row_bytes = (size_t)cam->hsize * RTC_TILE_H * (tg.rgba ? 4 : 24);
                size_t band = std::max<size_t>(1, ((size_t)band_kb << 10) / tile_row_bytes);
                if (P.n_parts > 1) band = std::max<size_t>(1, P.tile_rows / RTC_TILE_H);
                view.last_bucket = -(int)std::min<size_t>(band, 4096);
                // cheapest band first where the kernel takes about as long as the link (the link must never wait); top to
                // bottom where the link takes much longer anyway and long runs of neighbouring rows serve it best (measured:
                // teapot 1080p 0.99 against 1.03 ms, skybox 2160p 3.62 against 3.83; dragons 1.24 against 1.10)
                static const int forced_seq = [] {
                    const char* e = std::getenv("RTC_BAND_BY_WORK");
                    return e ? std::atoi(e) : -1;
                }();
                const double link_ms = (double)P.rows * cam->hsize * (tg.rgba ? 4.0 : 24.0) / 55e6;
                view.by_work = forced_seq >= 0 ? (forced_seq != 0) : !(s->best_ms[0] > 0.0f && link_ms > 1.4 * s->best_ms[0]);
            }
        }
        static const bool fine_dealt = [] {  // shared frames: the order by cost in eighths of the mean (tile_order.cu launch_fine_order)
            const char* e = std::getenv("RTC_FINE_ORDER");
            return !(e && e[0] == '0');
        }();
        const bool sorted_order = s->pending == n_tiles && (s->pending_view.last_bucket < 0 || (s->pending_view.bucket_set == 1 && fine_dealt));
        if (sorted_order && s->row_work_cap < n_tiles) {  // keys, sorted keys, tile numbers, row tables; then cub's temporary storage
            CUDA_TRY(cudaStreamSynchronize(stream));
            if (s->d_row_work) cudaFree(s->d_row_work);
            s->d_row_work = nullptr;
            s->row_work_cap = 0;
            s->row_temp_bytes = row_order_temp_bytes((unsigned)n_tiles);
            CUDA_TRY(cudaMalloc(&s->d_row_work, (3 * n_tiles + 3 * (n_tiles + 8) + 8) * sizeof(unsigned) + 256 + s->row_temp_bytes));  // (rows <= tiles)
            s->row_work_cap = n_tiles;
        }
        void* const sort_temp = s->d_row_work ? reinterpret_cast<char*>(s->d_row_work) + (((3 * s->row_work_cap + 3 * (s->row_work_cap + 8) + 8) * sizeof(unsigned) + 255) & ~size_t(255)) : nullptr;
        if (sorted_order && s->pending_view.last_bucket >= 0) {
            CUDA_TRY((cudaError_t)launch_fine_order(mine_prev, reinterpret_cast<const unsigned long long*>(mine_prev + cost_words(n_tiles) - 2), s->d_row_work,
                                                    s->d_tile_order, s->d_scratch + RTC_ORDER_BUCKETS * ((n_tiles + 1023) / 1024), (unsigned)n_tiles,
                                                    sort_temp, s->row_temp_bytes, stream));
            s->order_tiles = n_tiles;
            s->order_view = s->pending_view;
            F.order_made = true;
        } else if (sorted_order) {
            static const unsigned lead_pct = [] {  // how far ahead, in per cent of the estimate (warps of the grid working side by side)
                const char* e = std::getenv("RTC_ROW_LEAD");
                int v = e ? std::atoi(e) : 100;
                return (unsigned)(v >= 0 && v <= 1000 ? v : 100);
            }();
            const unsigned lead = (unsigned)((unsigned long long)s->grid * (RTC_BLOCK_THREADS / 32) * lead_pct / 100u);
            void* temp = sort_temp;
            CUDA_TRY((cudaError_t)launch_row_order(mine_prev, reinterpret_cast<const unsigned long long*>(mine_prev + cost_words(n_tiles) - 2),
                                                   s->d_row_work, s->d_tile_order, (unsigned)n_tiles, P.tiles_x, (unsigned)-s->pending_view.last_bucket, s->pending_view.by_work, lead,
                                                   temp, s->row_temp_bytes, stream));
            s->order_tiles = n_tiles;
            s->order_view = s->pending_view;
            F.order_made = true;
        } else if (s->pending == n_tiles) {
            // the previous frame's costs (complete in EVERY renderer's copy: a shared frame ends with a barrier) -> the order
            CUDA_TRY((cudaError_t)launch_tile_order(mine_prev, reinterpret_cast<const unsigned long long*>(mine_prev + cost_words(n_tiles) - 2),
                                                    s->d_scratch, s->d_tile_order, (unsigned)n_tiles, s->pending_view.last_bucket,
                                                    s->pending_view.bucket_set, stream));
            s->order_tiles = n_tiles;
            s->order_view = s->pending_view;
            F.order_made = true;
        }
        if (s->pending) CUDA_TRY(cudaMemsetAsync(mine_prev, 0, s->share.words * sizeof(unsigned), stream));  // clean for its next turn
        s->pending = 0;
        P.tile_order = s->order_tiles == n_tiles ? s->d_tile_order : nullptr;
        static const int forced_idx = [] {  // with RTC_LPP=0: which threshold of the set (experiments)
            const char* e = std::getenv("RTC_HEAVY_IDX");
            int v = e ? std::atoi(e) : 0;
            return v >= 0 && v < RTC_ORDER_BUCKETS - 1 ? v : 0;
        }();
        P.heavy_tiles = s->d_scratch + RTC_ORDER_BUCKETS * ((n_tiles + 1023) / 1024) + (which >= 2 ? which - 2 : forced_idx);
        static const bool always = [] {
            const char* e = std::getenv("RTC_LPT_ALWAYS");
            return e && e[0] == '1';
        }();
        const bool record = always || !(P.tile_order && std::memcmp(&s->order_view, &view, sizeof view) == 0);
        F.ordered = P.tile_order != nullptr;
        F.lpt = record;
        if (record) {
            for (unsigned r = 0; r < P.deal_n; r++) P.cost_peer[r] = dealt ? s->share.cost[b][r] : s->share.cost[b][s->share.rank];
            P.cost_tiles = (unsigned)(cost_words(n_tiles) - 2);
            s->parity ^= 1;
            s->pending = n_tiles;
            s->pending_view = view;
        }
    }
    // The mixed kernel needs the order of a previous frame.  Without one (first frame of a size; canvases in mapped host
    // memory, which keep the strided order) the candidate is two lanes per pixel everywhere: measured on the 1/8 partition
    // of dragons.json 1080p without an order, 0.49 -> 0.39 ms.
    if (lpp == 0 && !P.tile_order) lpp = 2;
    F.lpp = lpp;
    P.n_items *= (unsigned long long)(lpp ? lpp : 1);
    CUDA_TRY((cudaError_t)launch_mega(s, P, lpp, stream));
    CUDA_TRY(cudaEventRecord(s->rev1[slot], stream));
    CUDA_TRY(cudaMemcpyAsync(s->h_counters + 8 * slot, s->d_counters, 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, stream));
    CUDA_TRY(cudaEventRecord(s->rev2[slot], stream));
    F.state = 1;
    s->seq_next++;
    *slot_out = slot;
    if (wait || measuring) return resolve_frame(s, slot);
    return RTC_OK;
}

// Hands out the statistics of the OLDEST frame not handed out yet and frees its slot.
static int pop_frame(rtc_scene* s, rtc_stats* stats) {
    if (s->seq_pop == s->seq_next) return fail(RTC_ERR_INVALID, "no frame in flight");
    const int slot = (int)(s->seq_pop % rtc_scene::kRing);
    int rc = resolve_frame(s, slot);
    if (stats) *stats = s->ring[slot].stats;
    s->ring[slot].state = 0;
    s->seq_pop++;
    return rc;
}

static void add_stats(rtc_stats* sum, const rtc_stats& a) {
    sum->primary_rays += a.primary_rays;
    sum->shadow_rays += a.shadow_rays;
    sum->secondary_rays += a.secondary_rays;
    sum->pixels += a.pixels;
    sum->shadow_rays_culled += a.shadow_rays_culled;
    sum->launches += a.launches;
    sum->kernel_ms = std::fmax(sum->kernel_ms, a.kernel_ms);  // the devices work side by side
    sum->kernel = a.kernel > sum->kernel ? a.kernel : sum->kernel;
}

// One frame, all devices of the scene.  A scene on one device: `part` as given.  A replicated scene (rtc_set_devices)
// rendering a WHOLE frame: device k of n takes the interleaved 8-row tiles k, k + n, ... and stores them at their frame
// positions into the same target (device 0's memory through peer access, or mapped host memory each over its own PCIe
// link); the caller's stream continues when all of them are through.
static int render_frame(rtc_scene* s, const rtc_camera* cam, int max_depth, const rtc_partition* part, Target tg,
                        cudaStream_t stream, bool wait, rtc_stats* stats) {
    if (stats) std::memset(stats, 0, sizeof *stats);
    if (!s) return fail(RTC_ERR_INVALID, "render: NULL scene");
    const bool whole = !part || part->n_parts <= 1;
    if (s->replicas.empty() || !whole) {
        int slot = -1;
        // a scene whose tiles are shared with other processes (rtc_scene_share_tiles) renders ITS share of a whole frame
        const bool dealt = whole && s->share.external && s->share.n > 1 && tg.full_frame && !tg.spread_rows;
        int rc = enqueue_frame(s, cam, max_depth, part, tg, stream, wait, &slot, dealt);
        if (!wait || slot < 0) return rc;
        while (s->seq_next - s->seq_pop > 1) pop_frame(s, nullptr);  // frames put on the stream earlier and never fetched
        const int rp = pop_frame(s, stats);
        return rc ? rc : rp;
    }
    const unsigned n = 1u + (unsigned)s->replicas.size();
    tg.full_frame = true;
    // Frame in device memory: the tiles are DEALT by last frame's cost (every device gets its share of the expensive
    // ones).  Canvas in mapped host memory: interleaved 8-row tiles in strided row order (even byte rate on every link).
    const bool dealt = !tg.spread_rows;
    if (dealt) {
        int rs = ensure_replica_share(s, frame_tiles(cam->hsize, cam->vsize));
        if (rs) return rs;
    }
    CUDA_TRY(cudaSetDevice(s->device));
    // the replicas start when the caller's stream has reached this point (its earlier work may still read the target)
    CUDA_TRY(cudaEventRecord(s->ev0, stream));
    int rc = RTC_OK;
    std::vector<int> slots(n, -1);
    std::vector<cudaEvent_t> before(n, nullptr);  // end of the frame each device rendered last
    for (unsigned j = 0; j < n; j++) {
        rtc_scene* o = j == 0 ? s : s->replicas[j - 1];
        if (o->seq_next > 0) before[j] = o->rev2[(int)((o->seq_next - 1) % rtc_scene::kRing)];
    }
    for (unsigned k = 0; k < n && rc == RTC_OK; k++) {
        rtc_scene* r = k == 0 ? s : s->replicas[k - 1];
        rtc_partition p = {8u, k, n};
        cudaStream_t st = k == 0 ? stream : r->own_stream;
        CUDA_TRY(cudaSetDevice(r->device));
        if (k > 0) CUDA_TRY(cudaStreamWaitEvent(st, s->ev0, 0));
        if (dealt) {
            // a frame starts when the frame before it is through on EVERY device (it wrote costs into all of them)
            for (unsigned j = 0; j < n; j++)
                if (j != k && before[j]) CUDA_TRY(cudaStreamWaitEvent(st, before[j], 0));
        }
        rc = enqueue_frame(r, cam, max_depth, dealt ? nullptr : &p, tg, st, false, &slots[k], dealt);
    }
    CUDA_TRY(cudaSetDevice(s->device));
    for (unsigned k = 1; k < n; k++) {
        rtc_scene* r = s->replicas[k - 1];
        if (slots[k] >= 0 && r->ring[slots[k]].state != 0) CUDA_TRY(cudaStreamWaitEvent(stream, r->rev2[slots[k]], 0));
    }
    if (!wait) return rc;
    for (unsigned k = 0; k < n; k++) {
        rtc_scene* r = k == 0 ? s : s->replicas[k - 1];
        if (slots[k] < 0) continue;
        while (r->seq_next - r->seq_pop > 1) pop_frame(r, nullptr);
        rtc_stats one{};
        int rk = pop_frame(r, &one);
        if (rk && !rc) rc = rk;
        if (stats) add_stats(stats, one);
    }
    CUDA_TRY(cudaSetDevice(s->device));
    return rc;
}

int rtc_render_device(rtc_scene* s, const rtc_camera* cam, int max_depth, const rtc_partition* part, double* d_rgb,
                      void* stream, rtc_stats* stats) {
    if (!d_rgb) return fail(RTC_ERR_INVALID, "rtc_render_device: d_rgb is NULL");
    Target tg;
    tg.rgb = d_rgb;
    return render_frame(s, cam, max_depth, part, tg, (cudaStream_t)stream, true, stats);
}

int rtc_render_device_frame(rtc_scene* s, const rtc_camera* cam, int max_depth, const rtc_partition* part,
                            double* d_frame, void* stream, rtc_stats* stats) {
    if (!d_frame) return fail(RTC_ERR_INVALID, "rtc_render_device_frame: d_frame is NULL");
    Target tg;
    tg.rgb = d_frame;
    tg.full_frame = true;
    return render_frame(s, cam, max_depth, part, tg, (cudaStream_t)stream, true, stats);
}

int rtc_render_device_async(rtc_scene* s, const rtc_camera* cam, int max_depth, const rtc_partition* part, double* d_frame,
                            void* stream) {
    if (!d_frame) return fail(RTC_ERR_INVALID, "rtc_render_device_async: d_frame is NULL");
    Target tg;
    tg.rgb = d_frame;
    tg.full_frame = true;
    return render_frame(s, cam, max_depth, part, tg, (cudaStream_t)stream, false, nullptr);
}

uint64_t rtc_tile_share_words(uint64_t hsize, uint64_t vsize) { return cost_words(frame_tiles(hsize, vsize)); }

int rtc_scene_share_tiles(rtc_scene* s, const rtc_tile_share* sh) {
    if (!s) return fail(RTC_ERR_INVALID, "rtc_scene_share_tiles: NULL scene");
    if (!s->replicas.empty()) return fail(RTC_ERR_INVALID, "rtc_scene_share_tiles: the scene is replicated inside this process already");
    CUDA_TRY(cudaSetDevice(s->device));
    CUDA_TRY(cudaDeviceSynchronize());
    s->pending = 0;
    s->order_tiles = 0;
    s->parity = 0;
    s->tuned_pixels = 0;
    if (!sh || sh->n_ranks <= 1) {  // back to a renderer by itself
        s->share = rtc_scene::Share();
        for (int i = 0; i < 2; i++) s->share.cost[i][0] = s->d_cost_own[i];
        s->share.words = s->tile_cap ? cost_words(s->tile_cap) : 0;
        for (int i = 0; i < 2; i++)
            if (s->d_cost_own[i]) CUDA_TRY(cudaMemset(s->d_cost_own[i], 0, s->share.words * sizeof(unsigned)));
        return RTC_OK;
    }
    if (sh->n_ranks > RTC_MAX_SHARE || sh->rank >= sh->n_ranks) return fail(RTC_ERR_INVALID, "rtc_scene_share_tiles: bad rank / n_ranks (at most 8 renderers)");
    for (int i = 0; i < 2; i++)
        for (unsigned r = 0; r < sh->n_ranks; r++)
            if (!sh->cost[i][r]) return fail(RTC_ERR_INVALID, "rtc_scene_share_tiles: NULL buffer");
    s->share.rank = sh->rank;
    s->share.n = sh->n_ranks;
    s->share.words = (size_t)sh->words;
    s->share.external = true;
    for (int i = 0; i < 2; i++)
        for (unsigned r = 0; r < sh->n_ranks; r++) s->share.cost[i][r] = static_cast<unsigned*>(sh->cost[i][r]);
    // this renderer's own two buffers start clean (the caller puts a barrier behind the call: see rtc_b200.h)
    for (int i = 0; i < 2; i++) CUDA_TRY(cudaMemset(s->share.cost[i][sh->rank], 0, s->share.words * sizeof(unsigned)));
    return RTC_OK;
}

int rtc_stats_fetch(rtc_scene* s, rtc_stats* stats_out, uint32_t max_frames, uint32_t* n_out) {
    if (n_out) *n_out = 0;
    if (!s) return fail(RTC_ERR_INVALID, "rtc_stats_fetch: NULL scene");
    int rc = RTC_OK;
    uint32_t n = 0;
    while (s->seq_pop != s->seq_next && (n < max_frames || !stats_out)) {
        rtc_stats sum{};
        int rk = pop_frame(s, &sum);
        if (rk && !rc) rc = rk;
        for (rtc_scene* r : s->replicas) {  // a replicated frame has one entry per device
            if (r->seq_pop == r->seq_next) continue;
            rtc_stats one{};
            rk = pop_frame(r, &one);
            if (rk && !rc) rc = rk;
            add_stats(&sum, one);
        }
        if (stats_out) stats_out[n] = sum;
        n++;
    }
    if (n_out) *n_out = n;
    cudaSetDevice(s->device);
    return rc;
}

static int ensure_frame(rtc_scene* s, size_t bytes) {
    if (s->d_frame_bytes >= bytes) return RTC_OK;
    if (s->d_frame) cudaFree(s->d_frame);
    s->d_frame = nullptr;
    s->d_frame_bytes = 0;
    CUDA_TRY(cudaMalloc(&s->d_frame, bytes));
    s->d_frame_bytes = bytes;
    return RTC_OK;
}

// Copies the packed partition rows into their places in the caller's full frame: the tiles of a partition are
// tile_rows rows each, n_parts tiles apart -- one strided 2-D copy for all full tiles, one more for a ragged last tile.
static int scatter_rows(const rtc_camera* cam, const rtc_partition* part, const void* d_src, void* h_dst, size_t px_bytes) {
    const size_t row_bytes = (size_t)cam->hsize * px_bytes;
    if (!part || part->n_parts <= 1) {
        CUDA_TRY(cudaMemcpy(h_dst, d_src, row_bytes * cam->vsize, cudaMemcpyDeviceToHost));
        return RTC_OK;
    }
    const uint64_t tr = part->tile_rows, n_tiles = (cam->vsize + tr - 1) / tr;
    if (part->part >= n_tiles) return RTC_OK;
    const uint64_t mine = (n_tiles - part->part + part->n_parts - 1) / part->n_parts;  // tiles of this partition
    const uint64_t last = part->part + (mine - 1) * part->n_parts;                       // index of its last tile
    const bool ragged = (last + 1) * tr > cam->vsize;
    const uint64_t full = ragged ? mine - 1 : mine;
    const size_t tile_bytes = (size_t)tr * row_bytes;
    if (full)
        CUDA_TRY(cudaMemcpy2D((char*)h_dst + (size_t)part->part * tile_bytes, (size_t)part->n_parts * tile_bytes, d_src, tile_bytes,
                              tile_bytes, full, cudaMemcpyDeviceToHost));
    if (ragged)
        CUDA_TRY(cudaMemcpy((char*)h_dst + (size_t)last * tile_bytes, (const char*)d_src + (size_t)full * tile_bytes,
                            (size_t)(cam->vsize - last * tr) * row_bytes, cudaMemcpyDeviceToHost));
    return RTC_OK;
}

// The rows of a partition from a frame-sized device buffer to the same places of a frame-sized host buffer.
static int copy_partition_in_place(const rtc_camera* cam, const rtc_partition* part, const void* d_src, void* h_dst, size_t px_bytes) {
    const size_t row_bytes = (size_t)cam->hsize * px_bytes;
    const uint64_t tr = part->tile_rows, n_tiles = (cam->vsize + tr - 1) / tr;
    if (part->part >= n_tiles) return RTC_OK;
    const uint64_t mine = (n_tiles - part->part + part->n_parts - 1) / part->n_parts;
    const uint64_t last = part->part + (mine - 1) * part->n_parts;
    const bool ragged = (last + 1) * tr > cam->vsize;
    const uint64_t full = ragged ? mine - 1 : mine;
    const size_t tile_bytes = (size_t)tr * row_bytes, first = (size_t)part->part * tile_bytes, pitch = (size_t)part->n_parts * tile_bytes;
    if (full) CUDA_TRY(cudaMemcpy2D((char*)h_dst + first, pitch, (const char*)d_src + first, pitch, tile_bytes, full, cudaMemcpyDeviceToHost));
    if (ragged)
        CUDA_TRY(cudaMemcpy((char*)h_dst + (size_t)last * tile_bytes, (const char*)d_src + (size_t)last * tile_bytes,
                            (size_t)(cam->vsize - last * tr) * row_bytes, cudaMemcpyDeviceToHost));
    return RTC_OK;
}

// A frame (or row partition) for a canvas in PINNED host memory, moved by the copy engine while it renders: the kernel
// renders into the scene's frame-sized device buffer at full speed (cost order, no store crosses PCIe) and announces every
// finished row of tiles -- RTC_TILE_H scanlines, contiguous in the frame -- in a list in mapped host memory; this thread,
// which would only wait for the frame anyway, polls the list and hands runs of neighbouring finished rows to
// cudaMemcpyAsync on a stream of its own.  The link works at DMA speed (57 GB/s measured against 50-52 for SM-written host
// memory) from the first finished rows on and no CTA is given up for copying.  px_bytes: 24 (f64 RGB) or 4 (RGBA8).
static int render_to_pinned(rtc_scene* s, const rtc_camera* cam, int max_depth, const rtc_partition* part, void* h_dst, size_t px_bytes,
                            rtc_stats* stats) {
    if (stats) std::memset(stats, 0, sizeof *stats);
    const bool whole = !part || part->n_parts <= 1;
    const size_t row_bytes = (size_t)cam->hsize * px_bytes, frame_bytes = row_bytes * cam->vsize;
    int rc = ensure_frame(s, frame_bytes + 16);
    if (rc) return rc;
    Target tg;
    if (px_bytes == 4) tg.rgba = reinterpret_cast<unsigned char*>(s->d_frame);
    else tg.rgb = s->d_frame;
    tg.full_frame = true;
    tg.announce = true;
    int slot = -1;
    rc = enqueue_frame(s, cam, max_depth, part, tg, 0, false, &slot, false);
    if (rc && rc != RTC_ERR_CAPACITY) return rc;
    auto collect = [&](int rc0) {
        while (s->seq_next - s->seq_pop > 1) pop_frame(s, nullptr);
        const int rp = pop_frame(s, stats);
        return rc0 ? rc0 : rp;
    };
    if (!s->last_announced) {  // rendered another way (wavefront, several lanes per pixel, an empty partition): copied afterwards
        rc = collect(rc);
        if (rc) return rc;
        if (whole) {
            CUDA_TRY(cudaMemcpy(h_dst, s->d_frame, frame_bytes, cudaMemcpyDeviceToHost));
            return RTC_OK;
        }
        return copy_partition_in_place(cam, part, s->d_frame, h_dst, px_bytes);
    }
    const rtc_partition whole_part = {1, 0, 1};
    const rtc_partition& pt = part ? *part : whole_part;
    const uint64_t rows = rtc_partition_rows(cam->vsize, &pt);
    const unsigned n_rows = (unsigned)((rows + RTC_TILE_H - 1) / RTC_TILE_H);
    // scanlines [y0, y1) of the frame that row r of tiles of the partition covers (render_kernel.cuh copy_finished_rows)
    auto span = [&](unsigned r, uint64_t& y0, uint64_t& y1) {
        const uint64_t pr0 = (uint64_t)r * RTC_TILE_H, pr1 = std::min<uint64_t>(pr0 + RTC_TILE_H, rows);
        y0 = ((pr0 / pt.tile_rows) * pt.n_parts + pt.part) * pt.tile_rows + pr0 % pt.tile_rows;
        y1 = y0 + (pr1 - pr0);
    };
    // A copy is set up in ~3.5 us whatever its size (measured: 49.8 MB in 1 MB pieces 47 GB/s, in 4 MB pieces 54, in one
    // piece 56.4; two streams change nothing), so pieces should be as large as the link allows without waiting for them:
    // a stream takes a new run of finished rows only when its last copy is through, and then the whole run that has
    // gathered meanwhile.  While the link keeps up the pieces stay small and early; when it falls behind they grow.
    static const size_t min_run = [] {  // bytes below which a run that can still grow is not worth a copy of its own
        const char* e = std::getenv("RTC_DMA_MIN_KB");
        long v = e ? std::atol(e) : 1024;
        return (size_t)(v > 0 ? v : 1) << 10;
    }();
    static const int n_streams = [] {
        const char* e = std::getenv("RTC_DMA_STREAMS");
        int v = e ? std::atoi(e) : 2;
        return v >= 1 && v <= rtc_scene::kCopyStreams ? v : 2;
    }();
    static const bool trace = std::getenv("RTC_DMA_TRACE") != nullptr;  // timeline of one frame on stderr
    const auto t_begin = std::chrono::steady_clock::now();
    auto now_us = [&] { return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t_begin).count(); };
    std::vector<double> t_row;                      // when row k of the list was seen
    std::vector<std::pair<double, double>> t_copy;  // when a copy was handed over, its megabytes
    bool busy[rtc_scene::kCopyStreams] = {};
    std::vector<unsigned char> st(n_rows, 0);  // 0 rendering, 1 finished, 2 handed to the copy engine
    volatile unsigned* list = s->h_done;
    unsigned taken = 0, copied = 0, first_open = 0;  // first_open: no row below it is still waiting for a copy
    bool frame_over = false;
    int cerr = RTC_OK;
    unsigned idle = 0;
    while (copied < n_rows && cerr == RTC_OK) {
        bool news = false;
        while (taken < n_rows) {
            const unsigned r = list[taken];
            if (r == 0xffffffffu) break;
            if (r < n_rows) st[r] = 1;
            taken++;
            news = true;
            if (trace) t_row.push_back(now_us());
        }
        if (news) std::atomic_thread_fence(std::memory_order_acquire);
        if (!news && !frame_over && taken < n_rows && (++idle & 63u) == 0) {  // now and then: is the frame over, or broken?
            cudaError_t q = cudaEventQuery(s->rev2[slot]);
            if (q == cudaSuccess) {
                frame_over = true;
                continue;  // (read the list once more: everything the kernel announced is there now)
            }
            if (q != cudaErrorNotReady) cerr = fail(RTC_ERR_CUDA, std::string("render: ") + cudaGetErrorString(q));
        }
        if (frame_over && !news && taken < n_rows) {  // the kernel ended without announcing every row: its error bits say why
            for (unsigned k = 0; k < n_rows; k++)
                if (st[k] == 0) st[k] = 1;  // whatever is there goes out; the frame's status below is what counts
            taken = n_rows;
        }
        const bool all_in = taken == n_rows;
        for (int k = 0; k < n_streams && copied < n_rows && cerr == RTC_OK; k++) {
            if (busy[k]) {
                cudaError_t q = cudaStreamQuery(s->copy_stream[k]);
                if (q == cudaErrorNotReady) continue;
                if (q != cudaSuccess) {
                    cerr = fail(RTC_ERR_CUDA, std::string("copy stream: ") + cudaGetErrorString(q));
                    break;
                }
                busy[k] = false;
            }
            while (first_open < n_rows && st[first_open] == 2) first_open++;
            for (unsigned a = first_open; a < n_rows;) {
                if (st[a] != 1) {
                    a++;
                    continue;
                }
                // the run of finished rows starting here that is also one run of bytes in the frame
                uint64_t y0, y1;
                span(a, y0, y1);
                unsigned b = a + 1;
                for (; b < n_rows && st[b] == 1; b++) {
                    uint64_t z0, z1;
                    span(b, z0, z1);
                    if (z0 != y1) break;
                    y1 = z1;
                }
                // taken when large enough, when it cannot grow any more, or when the frame is through
                const bool closed = (a == 0 || st[a - 1] == 2) && (b == n_rows || st[b] == 2);
                if ((y1 - y0) * row_bytes >= min_run || closed || all_in || !whole) {
                    cudaError_t e = cudaMemcpyAsync((char*)h_dst + y0 * row_bytes, (const char*)s->d_frame + y0 * row_bytes, (y1 - y0) * row_bytes,
                                                    cudaMemcpyDeviceToHost, s->copy_stream[k]);
                    if (e != cudaSuccess) {
                        cerr = fail(RTC_ERR_CUDA, std::string("cudaMemcpyAsync: ") + cudaGetErrorString(e));
                        break;
                    }
                    for (unsigned j = a; j < b; j++) st[j] = 2;
                    copied += b - a;
                    busy[k] = true;
                    if (trace) t_copy.push_back({now_us(), (double)((y1 - y0) * row_bytes) / 1e6});
                    break;
                }
                a = b;
            }
        }
    }
    cudaError_t ce = cudaSuccess;
    for (int k = 0; k < n_streams; k++) {
        cudaError_t e = cudaStreamSynchronize(s->copy_stream[k]);
        if (e != cudaSuccess) ce = e;
    }
    if (trace && !t_row.empty()) {
        const size_t n = t_row.size();
        std::fprintf(stderr, "dma: rows seen at us: first %.0f, 10%% %.0f, 25%% %.0f, 50%% %.0f, 75%% %.0f, 90%% %.0f, last %.0f; all copied at %.0f; %zu copies:", t_row[0],
                     t_row[n / 10], t_row[n / 4], t_row[n / 2], t_row[3 * n / 4], t_row[9 * n / 10], t_row[n - 1], now_us(), t_copy.size());
        for (auto& c : t_copy) std::fprintf(stderr, " %.0f:%.1fMB", c.first, c.second);
        std::fprintf(stderr, "\n");
    }
    rc = collect(rc);
    if (cerr) return cerr;
    if (ce != cudaSuccess) return fail(RTC_ERR_CUDA, std::string("cudaStreamSynchronize: ") + cudaGetErrorString(ce));
    return rc;
}

// Pinned, device-mapped host memory for canvases (rtc_b200.h).
void* rtc_host_alloc(size_t nbytes) {
    void* p = nullptr;
    if (cudaHostAlloc(&p, nbytes ? nbytes : 1, cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) {
        g_error = std::string("rtc_host_alloc: ") + cudaGetErrorString(cudaGetLastError());
        return nullptr;
    }
    return p;
}
void rtc_host_free(void* p) {
    if (p) cudaFreeHost(p);
}

// How a frame reaches a canvas in pinned host memory.  RTC_HOST_COPY=dma (default): render_to_pinned; warps: the
// kernel's copy warps where the link time matters, else direct; direct: the render warps store into the mapped canvas.
static int host_copy_mode() {
    static const int mode = [] {
        const char* e = std::getenv("RTC_HOST_COPY");
        const std::string v = e ? e : "dma";
        return v == "warps" ? 1 : v == "direct" ? 2 : 0;  // (dma_all: as dma, row partitions included)
    }();
    return mode;
}

// The device-side address of a pinned + mapped host buffer, or NULL for pageable memory.
static void* mapped_device_pointer(const void* host) {
    static const bool off = [] {
        const char* e = std::getenv("RTC_NO_ZEROCOPY");
        return e && e[0] == '1';
    }();
    if (off) return nullptr;
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, host) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    if (at.type != cudaMemoryTypeHost || !at.devicePointer) return nullptr;
    return at.devicePointer;
}

int rtc_render(rtc_scene* s, const rtc_camera* cam, int max_depth, const rtc_partition* part, double* rgb_out,
               rtc_stats* stats) {
    if (!s || !cam || !rgb_out) return fail(RTC_ERR_INVALID, "rtc_render: NULL argument");
    CUDA_TRY(cudaSetDevice(s->device));
    // Canvas in pinned memory (rtc_host_alloc): the kernel stores every finished pixel straight into it
    // across PCIe, overlapped with the rendering of the others -- no staging frame, no copy afterwards.
    const bool whole = !part || part->n_parts <= 1;
    if (void* mapped = mapped_device_pointer(rgb_out)) {
        // whole frames: the copy engine (render_to_pinned).  Row partitions (N processes filling one shared canvas) lie in
        // blocks of tile_rows scanlines, too small for a copy of their own (3.5 us of set-up each): the kernel's copy warps
        // move those (measured at 2 GPUs, e2e: 0.79 ms against 0.99; RTC_HOST_COPY=dma_all sends them to the copy engine too)
        const bool tiles_ok = whole || part->tile_rows % RTC_TILE_H == 0;  // a row of tiles lies together in the frame
        static const bool dma_parts = [] {
            const char* e = std::getenv("RTC_HOST_COPY");
            return e && std::string(e) == "dma_all";
        }();
        if (host_copy_mode() == 0 && (whole || dma_parts) && tiles_ok && s->replicas.empty())
            return render_to_pinned(s, cam, max_depth, part, rgb_out, 3 * sizeof(double), stats);
        static const bool decouple = [] {
            const char* e = std::getenv("RTC_COPY_CTAS");
            return !(e && std::atoi(e) == 0) && host_copy_mode() != 2;
        }();
        // (Until the per-tile fence was found to drop the L1, this path was used only where the link time was at least half
        // the kernel time.  Since then it wins everywhere -- the direct stores also cost the frame its cost order: a 1/4
        // partition of dragons 1080p 0.80 ms with direct stores, its kernel in cost order 0.42 -- so it is unconditional.)
        const bool worth = true;
        const bool rows_ok = whole || part->tile_rows % RTC_TILE_H == 0;  // a row of tiles lies together in the frame
        if (decouple && worth && rows_ok && s->replicas.empty()) {
            // rendered into device memory (a frame-sized staging buffer, pixels at their frame positions) in cost order at
            // full speed while copy warps move every finished row of tiles to the canvas (enqueue_frame); a frame the
            // library renders another way (wavefront, several lanes per pixel) is copied afterwards.  A partition of the
            // frame works the same way: N processes, each with its own link, fill one shared pinned canvas.
            const size_t frame_bytes = (size_t)cam->vsize * cam->hsize * 3 * sizeof(double);
            int rc = ensure_frame(s, frame_bytes + 16);
            if (rc) return rc;
            Target tg;
            tg.rgb = s->d_frame;
            tg.full_frame = true;
            tg.copy_out = static_cast<double*>(mapped);
            rc = render_frame(s, cam, max_depth, part, tg, 0, true, stats);
            if (rc || s->last_copied) return rc;
            if (whole) {
                CUDA_TRY(cudaMemcpy(rgb_out, s->d_frame, frame_bytes, cudaMemcpyDeviceToHost));
                return RTC_OK;
            }
            return copy_partition_in_place(cam, part, s->d_frame, rgb_out, 3 * sizeof(double));
        }
        Target tg;
        tg.rgb = static_cast<double*>(mapped);
        tg.full_frame = tg.spread_rows = true;
        return render_frame(s, cam, max_depth, part, tg, 0, true, stats);
    }
    uint64_t rows = rtc_partition_rows(cam->vsize, part);
    int rc = ensure_frame(s, (size_t)rows * cam->hsize * 3 * sizeof(double) + 16);
    if (rc) return rc;
    Target tg;
    tg.rgb = s->d_frame;
    tg.full_frame = whole;
    rc = render_frame(s, cam, max_depth, part, tg, 0, true, stats);
    if (rc) return rc;
    return scatter_rows(cam, part, s->d_frame, rgb_out, 3 * sizeof(double));
}

int rtc_render_rgba8(rtc_scene* s, const rtc_camera* cam, int max_depth, const rtc_partition* part, uint8_t* rgba_out,
                     rtc_stats* stats) {
    if (!s || !cam || !rgba_out) return fail(RTC_ERR_INVALID, "rtc_render_rgba8: NULL argument");
    CUDA_TRY(cudaSetDevice(s->device));
    // as rtc_render: a pinned + mapped buffer (rtc_host_alloc; the lib.zig twin's pixel buffer is one) is written by
    // the kernel itself, 4 bytes per pixel over PCIe while the frame renders
    if (void* mapped = mapped_device_pointer(rgba_out)) {
        const bool whole8 = !part || part->n_parts <= 1;
        if (host_copy_mode() == 0 && whole8 && s->replicas.empty()) return render_to_pinned(s, cam, max_depth, part, rgba_out, 4, stats);
        Target tg;
        tg.rgba = static_cast<unsigned char*>(mapped);
        tg.full_frame = tg.spread_rows = true;
        return render_frame(s, cam, max_depth, part, tg, 0, true, stats);
    }
    const bool whole = !part || part->n_parts <= 1;
    uint64_t rows = rtc_partition_rows(cam->vsize, part);
    int rc = ensure_frame(s, (size_t)rows * cam->hsize * 4 + 16);
    if (rc) return rc;
    Target tg;
    tg.rgba = reinterpret_cast<unsigned char*>(s->d_frame);
    tg.full_frame = whole;
    rc = render_frame(s, cam, max_depth, part, tg, 0, true, stats);
    if (rc) return rc;
    return scatter_rows(cam, part, s->d_frame, rgba_out, 4);
}

// ---- host loader -------------------------------------------------------------------------------------
int rtc_load_scene_json(const char* json_text, const char* asset_dir, long width, long height, void** blob_out,
                        size_t* nbytes_out, rtc_camera* cam_out) {
    if (!json_text || !blob_out || !nbytes_out || !cam_out) return fail(RTC_ERR_INVALID, "rtc_load_scene_json: NULL argument");
    try {
        auto T0 = std::chrono::steady_clock::now();
        SceneInfo si = parse_scene(json_text, directory_loader(asset_dir ? asset_dir : ""), width, height);
        auto T1 = std::chrono::steady_clock::now();
        std::vector<uint8_t> blob = flatten_world(si.world);
        auto T2 = std::chrono::steady_clock::now();
        if (std::getenv("RTC_LOAD_TIMING"))
            std::fprintf(stderr, "load: parse %.3f flatten %.3f s\n", std::chrono::duration<double>(T1 - T0).count(),
                         std::chrono::duration<double>(T2 - T1).count());
        void* p = std::malloc(blob.size());
        if (!p) return fail(RTC_ERR_INVALID, "out of memory");
        std::memcpy(p, blob.data(), blob.size());
        *blob_out = p;
        *nbytes_out = blob.size();
        *cam_out = flatten_camera(si.camera);
        return RTC_OK;
    } catch (const std::exception& e) {
        return fail(RTC_ERR_PARSE, e.what());
    }
}

void rtc_blob_free(void* blob) { std::free(blob); }

int rtc_camera_look_at(rtc_camera* cam, const double from[3], const double to[3], const double up[3]) {
    if (!cam || !from || !to || !up) return fail(RTC_ERR_INVALID, "rtc_camera_look_at: NULL argument");
    try {
        Mat4 view = Mat4::view_transform(Tuple::point(from[0], from[1], from[2]), Tuple::point(to[0], to[1], to[2]),
                                         Tuple::vec3(up[0], up[1], up[2]));
        Mat4 inv = view.inverse();
        for (int r = 0; r < 4; r++)
            for (int c = 0; c < 4; c++) cam->inverse[r * 4 + c] = inv.d[r][c];
        return RTC_OK;
    } catch (const std::exception& e) {
        return fail(RTC_ERR_INVALID, e.what());
    }
}

}  // extern "C"

// ---- a render call that allocates nothing ---------------------------------------------------------------------------
// (No counterpart in the reference; added for callers that cannot call cudaMalloc:
// code under CUDA graph capture, tight serving loops, harnesses that forbid it.)  The megakernel's scratch -- the ray
// counters, the two per-tile cost buffers, the tile order made from them and the bucket counts of tile_order.cu, which
// the other entry points allocate per scene and grow on demand -- lives in ONE file-scope device array of fixed size;
// every GPU that loads the library holds it (3.1 MB) until the process exits.  In host code the array's name is only a
// stand-in for the symbol: its device address is asked for once per device (cudaGetSymbolAddress with that device
// current) and that pointer is what the kernels are given.
#define RTC_STATIC_MAX_TILES (1u << 18)  // 8x4-pixel tiles: 8,388,608 pixels; 3840x2160 (259,200 tiles) fits
namespace {
struct StaticScratch {
    unsigned long long counters[8];                              // RenderParams::counters
    unsigned cost[2][RTC_STATIC_MAX_TILES + 2];                  // tile costs + the 64-bit sum, two buffers taking turns
    unsigned order[RTC_STATIC_MAX_TILES];                        // RenderParams::tile_order
    unsigned buckets[RTC_ORDER_BUCKETS * (RTC_STATIC_MAX_TILES / 1024 + 2)];  // tile_order.cu: bucket counts + heavy-tile counts
};
}  // namespace
static __device__ __align__(256) unsigned char g_static_scratch[sizeof(StaticScratch)];

namespace {
constexpr int kStaticDevices = 64;
// What the array of a device holds from the calls before (it is shared by every call on that device, whatever the scene).
struct StaticState {
    StaticScratch* ptr = nullptr;  // device address of g_static_scratch on this device
    const rtc_scene* scene = nullptr;
    rtc_scene::View pending_view{}, order_view{};
    size_t pending = 0, order_tiles = 0;  // as rtc_scene::pending / order_tiles
    int parity = 0;
};
StaticState g_static[kStaticDevices];
}  // namespace

// Set-up, once per device (the device must be current).  No allocation: the array came with the module.
static int static_scratch_setup(int device, StaticState** out) {
    if (device < 0 || device >= kStaticDevices) return fail(RTC_ERR_INVALID, "rtc_render_device_static: device index out of range");
    StaticState& st = g_static[device];
    if (!st.ptr) {
        void* p = nullptr;
        CUDA_TRY(cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p || (reinterpret_cast<uintptr_t>(p) & 255u)) return fail(RTC_ERR_CUDA, "rtc_render_device_static: bad address of the static scratch array");
        CUDA_TRY(cudaMemset(p, 0, sizeof(StaticScratch)));  // (zero at module load already; the cost buffers MUST start at zero)
        st.ptr = static_cast<StaticScratch*>(p);
    }
    *out = &st;
    return RTC_OK;
}

extern "C" {

uint64_t rtc_static_max_tiles(void) { return RTC_STATIC_MAX_TILES; }

int rtc_render_device_static(rtc_scene* s, const rtc_camera* cam, int max_depth, double* d_rgb, void* stream_, rtc_stats* stats) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!s || !cam || !d_rgb) return fail(RTC_ERR_INVALID, "rtc_render_device_static: NULL scene, camera or frame");
    if (max_depth < 0 || max_depth > 7) return fail(RTC_ERR_INVALID, "render: max_depth must be in [0,7]");
    if (cam->hsize == 0 || cam->vsize == 0) return fail(RTC_ERR_INVALID, "render: empty frame");
    const size_t n_tiles = frame_tiles(cam->hsize, cam->vsize);
    if (cam->hsize > (1ull << 24) || cam->vsize > (1ull << 24) || n_tiles > RTC_STATIC_MAX_TILES) {
        char buf[200];
        std::snprintf(buf, sizeof buf,
                      "rtc_render_device_static: a %llux%llu frame is %zu tiles of %ux%u pixels, the static scratch holds %u "
                      "(use rtc_render_device)",
                      (unsigned long long)cam->hsize, (unsigned long long)cam->vsize, n_tiles, RTC_TILE_W, RTC_TILE_H, RTC_STATIC_MAX_TILES);
        return fail(RTC_ERR_INVALID, buf);
    }
    CUDA_TRY(cudaSetDevice(s->device));
    StaticState* st = nullptr;
    if (int rc = static_scratch_setup(s->device, &st)) return rc;
    StaticScratch* sc = st->ptr;
    const bool same_scene = st->scene == s;
    if (!same_scene) {  // another scene: the costs and the order at hand say nothing about this one
        st->scene = s;
        st->order_tiles = 0;
    }

    RenderParams P{};
    P.cam.hsize = cam->hsize;
    P.cam.vsize = cam->vsize;
    P.cam.half_width = cam->half_width;
    P.cam.half_height = cam->half_height;
    P.cam.pixel_size = cam->pixel_size;
    std::memcpy(P.cam.inv, cam->inverse, sizeof P.cam.inv);
    P.max_depth = max_depth;
    P.tile_rows = 1, P.part = 0, P.n_parts = 1;  // the whole frame
    P.rows = cam->vsize;
    P.tiles_x = (unsigned)((cam->hsize + RTC_TILE_W - 1) / RTC_TILE_W);
    const unsigned n_tile_rows = (unsigned)((P.rows + RTC_TILE_H - 1) / RTC_TILE_H);
    P.n_items = (unsigned long long)n_tiles * 32ull;
    P.deal_n = 1;
    P.row_group = 4;  // tile rows top to bottom (only read while there is no order yet)
    P.n_row_groups = n_tile_rows / P.row_group;
    P.row_stride = 1;
    P.scale = s->wf_scale;
    P.out = d_rgb;
    P.counters = sc->counters;

    // the same protocol as enqueue_frame: last call's costs -> this call's order; a view rendered in the order made
    // from its own costs records nothing
    rtc_scene::View view{};
    view.cam = *cam;
    view.depth = max_depth;
    view.tile_rows = 1, view.part = 0, view.n_parts = 1, view.deal_rank = 0, view.deal_n = 1;
    view.last_bucket = 3;
    const size_t words = cost_words(n_tiles);
    unsigned* prev = sc->cost[st->parity ^ 1];
    bool order_made = false;
    CUDA_TRY(cudaMemsetAsync(sc->counters, 0, sizeof sc->counters, stream));
    const bool timed = stats != nullptr;
    if (timed) CUDA_TRY(cudaEventRecord(s->ev0, stream));
    if (same_scene && st->pending == n_tiles) {
        CUDA_TRY((cudaError_t)launch_tile_order(prev, reinterpret_cast<const unsigned long long*>(prev + words - 2), sc->buckets, sc->order,
                                                (unsigned)n_tiles, 3, 0, stream));
        st->order_tiles = n_tiles;
        st->order_view = st->pending_view;
        order_made = true;
    }
    if (st->pending) CUDA_TRY(cudaMemsetAsync(prev, 0, cost_words(st->pending) * sizeof(unsigned), stream));  // clean for its next turn
    st->pending = 0;
    P.tile_order = st->order_tiles == n_tiles ? sc->order : nullptr;
    P.heavy_tiles = sc->buckets + RTC_ORDER_BUCKETS * ((n_tiles + 1023) / 1024);
    if (!(P.tile_order && std::memcmp(&st->order_view, &view, sizeof view) == 0)) {
        P.cost_peer[0] = sc->cost[st->parity];
        P.cost_tiles = (unsigned)(words - 2);
        st->parity ^= 1;
        st->pending = n_tiles;
        st->pending_view = view;
    }
    CUDA_TRY((cudaError_t)launch_mega(s, P, 1, stream));
    if (!timed) return RTC_OK;  // enqueued only: nothing is synchronised, capacity errors are not reported

    CUDA_TRY(cudaEventRecord(s->ev1, stream));
    unsigned long long c[8];
    CUDA_TRY(cudaMemcpyAsync(c, sc->counters, sizeof c, cudaMemcpyDeviceToHost, stream));
    CUDA_TRY(cudaStreamSynchronize(stream));
    float ms = 0;
    CUDA_TRY(cudaEventElapsedTime(&ms, s->ev0, s->ev1));
    *stats = rtc_stats{};
    stats->primary_rays = c[1];
    stats->shadow_rays = c[2];
    stats->secondary_rays = c[3];
    stats->shadow_rays_culled = c[5];
    stats->pixels = cam->hsize * cam->vsize;
    stats->kernel_ms = ms;
    stats->launches = order_made ? 3 : 1;
    if (c[4]) {
        char buf[160];
        std::snprintf(buf, sizeof buf, "render: per-ray capacity exceeded (bits %llu, see rtc_render)", c[4]);
        return fail(RTC_ERR_CAPACITY, buf);
    }
    return RTC_OK;
}

}  // extern "C"
