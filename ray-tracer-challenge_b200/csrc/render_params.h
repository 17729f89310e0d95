// Launch parameters of render_kernel, shared by render.cu (device) and rtc_api.cu (host).
#pragma once
#include <cuda_runtime.h>

#include "device_scene.h"

namespace rtc {

#ifndef RTC_BLOCK_THREADS
#define RTC_BLOCK_THREADS 128  // threads per CTA of the render kernels
#endif

// The pixel tile one warp works on at a time: RTC_TILE_W x RTC_TILE_H = 32 pixels (megakernel only; the wavefront's
// level-0 generation keeps 8x4).  log2 values.
#ifndef RTC_TILE_W_LOG2
#define RTC_TILE_W_LOG2 3
#endif
#define RTC_TILE_H_LOG2 (5 - RTC_TILE_W_LOG2)
#define RTC_TILE_W (1u << RTC_TILE_W_LOG2)
#define RTC_TILE_H (1u << RTC_TILE_H_LOG2)

// lanes per pixel on the heavy tiles of the mixed kernel (render_kernel.cuh): 4, or 8 = pipelined (lane 0 the chain of
// closest hits, lanes 1 .. n_lights the shadow rays of the hit before).  MEASURED on B200, dragons.json 1080p, one GPU
// rendering a dealt 1/8 share (profiles/README.md): 8 lanes pipelined 0.272 ms against 0.242 with 4 (1/4 share: 0.38 / 0.30):
// the chain of the heaviest tile does get shorter (197 -> 150 us of warp time), but a trip now lasts as long as the
// slower of a closest-hit and a shadow traversal and the tile holds twice the warps.  4 stays.
#ifndef RTC_HEAVY_LANES
#define RTC_HEAVY_LANES 4
#endif
// groups of RTC_HEAVY_LANES lanes per pixel on the heavy tiles, which split the pixel's ray tree (render_kernel.cuh SPLIT)
#ifndef RTC_HEAVY_GROUPS
#define RTC_HEAVY_GROUPS 1
#endif
#define RTC_FEAT_CSG 1       // render_kernel<MESH, LPP, FEAT>: see render_kernel.cuh
#define RTC_FEAT_PATTERNS 2
#define RTC_MAX_SHARE 8    // renderers (GPUs) that can share one frame tile by tile (RenderParams::cost_peer)
// Work sharing inside the mesh traversal (rt_trace.cuh trace_meshes): built, measured slower on B200, off by default.
#ifndef RTC_STEAL
#define RTC_STEAL 0
#endif
#if RTC_STEAL
#define RTC_MAILBOX_SLOTS 5  // rt_trace.cuh MB_SLOTS: per-lane mailbox in shared memory
#else
#define RTC_MAILBOX_SLOTS 0
#endif
// doubles of per-lane state parked in shared memory (render_kernel.cuh PK_*), mailbox included.  Shared memory is taken
// out of the L1 cache the BVH nodes live in: every kilobyte counts (27 slots instead of 22 cost teapot.json 5 %).
#define RTC_PARK_SLOTS (22 + RTC_MAILBOX_SLOTS)

struct DCamera {  // Camera(T) fields read by rayForPixel (camera.zig:22-30)
    unsigned long long hsize, vsize;
    double half_width, half_height, pixel_size;
    double inv[16];
};

struct RenderParams {
    DCamera cam;
    int max_depth;
    unsigned int tile_rows, part, n_parts;
    unsigned long long rows;       // rows of this partition
    unsigned long long n_items;    // tiles_x * tiles_y * 32 pixels (some fall outside the frame) x lanes per pixel
    unsigned int tiles_x;
    // Groups of tile rows are handed out in the order (k * row_stride) % n_row_groups, k = 0, 1, ... (row_stride coprime
    // to n_row_groups, about 0.38 of it; the rows past the last full group keep their place): consecutive work is spread over the frame, so cheap and expensive regions of
    // the picture are rendered side by side and pixels reach the canvas at an even rate (a canvas in host memory is
    // written over PCIe while the kernel runs; a burst of cheap rows at the end would have to wait for the link).
    unsigned int row_stride, n_row_groups, row_group;  // in groups of row_group tile rows (neighbouring rows share BVH nodes)
    double scale;                  // fixed-point units per 1.0 of the pixel accumulators (a power of two)
    int full_frame;                // 0: out holds the packed rows of the partition; 1: the whole frame
    double* out;                   // 3 doubles per pixel (may be peer-mapped memory of another GPU)
    unsigned char* out8;           // or RGBA8 (NULL when `out` is used)
    // Tiles are handed out in the order tile_order[0], tile_order[1], ... (NULL: 0, 1, ...): a permutation of the tiles
    // of this partition, the most expensive ones of the previous frame first (rtc_api.cu).  tile_cost[tile] receives
    // the time the warps spent on the tile in this frame (NULL: not recorded).
    const unsigned int* tile_order;
    // Cost recording: cost_peer[0] = this renderer's array (NULL: not recorded); entry [cost_tiles] (8-byte aligned, two
    // words) receives the sum of all costs.  Several renderers sharing one frame (one per GPU; deal_n of them, this one
    // is deal_rank) take the entries deal_rank, deal_rank + deal_n, ... of the SAME order -- the tiles of every cost
    // bucket are dealt round -- and add the cost of each tile they render to the arrays of ALL of them (cost_peer[r] =
    // renderer r's array as this device addresses it: peer memory over NVLink), so that each can make the same next
    // order from its own copy.  deal_n == 1: a renderer by itself.
    unsigned int* cost_peer[RTC_MAX_SHARE];
    unsigned int cost_tiles;
    unsigned int deal_n, deal_rank;
    unsigned int deal_shift;  // the order is dealt in runs of 1 << deal_shift consecutive entries (neighbouring tiles stay with one renderer)
    // Frame for a canvas in mapped HOST memory, decoupled from the link: the render warps store their pixels into `out`
    // in device memory and count the finished tiles of every ROW of tiles; a row that is complete is announced in
    // done_list[], and the warps of the first copy_ctas CTAs do nothing but wait for announcements and move finished
    // rows -- RTC_TILE_H full scanlines, contiguous -- to copy_out over PCIe.  The link is busy from the first finished
    // row on, no render warp ever waits for it, and the host canvas is written front to back in long runs
    // (render_kernel.cuh copy_finished_rows).
    double* copy_out;          // NULL: off
    unsigned int* done_list;   // [copy_rows] rows of tiles in the order they were finished (0xffffffff: not yet)
    unsigned int* copy_ctl;    // [0] rows announced, [1] rows taken by copy warps (both start at 0xffffffff)
    unsigned int* row_count;   // [copy_rows] finished tiles of the row (starts at 0xffffffff)
    unsigned int copy_ctas, copy_rows;
    const unsigned int* heavy_tiles;  // device word: how many leading entries of tile_order cost >= 8x the mean (LPP == 0)
    unsigned long long* counters;  // [0] next item, [1] primary, [2] shadow, [3] secondary, [4] error bits, [5] culled shadow rays
};

// ---- wavefront path (render_wavefront.cuh) ------------------------------------------------------------
// One ray of recursion level d waiting for its closest-hit query (colorAt, world.zig:111-123).  64 bytes.
struct WRay {
    double ox, oy, oz, dx, dy, dz;
    double w;  // product of the reflective / transparency / Schlick factors on the path from the eye
    int remaining;  // remaining_recursions (world.zig:160, :178)
    int pix;        // pixel of the partition (packed row-major index)
};
static_assert(sizeof(WRay) == 64, "WRay must be 64 bytes");
// One shaded hit waiting for its shadow queries: everything lighting() needs (material.zig:40-74).  128 bytes.
struct WHit {
    double over[3];    // over_point
    double normal[3];  // normalv (flipped towards the eye)
    double rd[3];      // ray direction (eyev = -rd)
    double color[3];   // pattern colour at over_point
    double w;
    int mat, pix;
    double pad[2];
};
static_assert(sizeof(WHit) == 128, "WHit must be 128 bytes");

#define RTC_WF_LEVELS 9  // recursion levels 0..max_depth (<= 7) + 1
// counters (u64), all zeroed before a frame
enum {
    WC_ERR = 0,        // error bits
    WC_SHADOW = 1,     // shadow rays (incl. culled)
    WC_CULLED = 2,
    WC_NRAYS = 8,      // [level] rays queued for that level (level 0: unused, primaries are generated)
    WC_NHITS = 24,     // [level] hit records of that level
    WC_TAKE_RAY = 40,  // [level] consumer cursor over the rays of that level
    WC_TAKE_SHADOW = 56,  // [level] consumer cursor over (hit record, light) pairs of that level
    WC_COUNT = 72
};

struct WaveParams {
    DCamera cam;
    int max_depth, level;
    unsigned int tile_rows, part, n_parts;
    unsigned long long rows, n_pixels;  // of this partition
    const WRay* rays_in;   // rays of `level` (level 0: NULL, generated from the camera)
    WRay* rays_out;        // rays of level + 1
    const WHit* hits_in;   // hits of level - 1: their shadow queries run in this launch
    WHit* hits_out;        // hits of `level`
    unsigned long long cap;  // capacity (entries) of each queue
    long long* acc;          // fixed-point RGB accumulators, 3 per pixel of the partition
    double scale;            // accumulator units per 1.0
    unsigned long long* counters;
};

struct ResolveParams {
    unsigned long long hsize, rows, n_pixels;
    unsigned int tile_rows, part, n_parts;
    int full_frame;
    const long long* acc;
    double inv_scale;
    double* out;
    unsigned char* out8;
};

int launch_wave_mesh(const DScene& S, const WaveParams& W, int grid, size_t smem, cudaStream_t stream);
int launch_wave_analytic(const DScene& S, const WaveParams& W, int grid, size_t smem, cudaStream_t stream);
int occupancy_wave_mesh(int* blocks_per_sm, size_t smem);
int occupancy_wave_analytic(int* blocks_per_sm, size_t smem);
int launch_resolve(const ResolveParams& R, cudaStream_t stream);
// tile_order.cu: order[] = the tiles 0..n_tiles-1, stably sorted into four cost buckets (most expensive first)
#define RTC_ORDER_BUCKETS 8  // buckets at most; scratch = RTC_ORDER_BUCKETS x (blocks of 1024 tiles + 1) words
int launch_tile_order(const unsigned* cost, const unsigned long long* sum, unsigned* scratch, unsigned* order, unsigned n_tiles,
                      int last_bucket, int set, cudaStream_t stream);
// ... or for frames streamed to a host canvas: bands of rows, cheapest band first, every tile started ahead of its row by what it costs
// ... or for frames shared tile by tile: by cost in steps of 1/8 of the mean (heavy: six counters, cost >= 12x .. 2x the mean)
int launch_fine_order(const unsigned* cost, const unsigned long long* sum, unsigned* work, unsigned* order, unsigned* heavy, unsigned n_tiles,
                      void* temp, size_t temp_bytes, cudaStream_t stream);
size_t row_order_temp_bytes(unsigned n_tiles);
size_t row_order_work_words(unsigned n_tiles, unsigned tiles_x);
int launch_row_order(const unsigned* cost, const unsigned long long* sum, unsigned* work, unsigned* order, unsigned n_tiles, unsigned tiles_x,
                     unsigned band_rows, int by_work, unsigned lead, void* temp, size_t temp_bytes, cudaStream_t stream);

// render_inst_*.cu: the instantiations of render_kernel<MESH, LPP, FEAT>, one translation unit per (MESH, FEAT).
// smem = dynamic shared memory per CTA (parked state + the BVH traversal stacks).
// lpp = lanes per pixel (1, 2 or 4; 0 = four on last frame's expensive tiles, one elsewhere);
// P.n_items counts lanes, i.e. pixels x lpp (lpp 0: pixels, the kernel adds the extra lanes of the expensive tiles).
int launch_render(bool mesh, int feat, const DScene& S, const RenderParams& P, int lpp, int grid, size_t smem, cudaStream_t stream);
int occupancy_render(bool mesh, int feat, int* blocks_per_sm, size_t smem);
#define RTC_DECLARE_INST(M, F)                                                                                          \
    int launch_render_##M##_##F(const DScene& S, const RenderParams& P, int lpp, int grid, size_t smem, cudaStream_t stream); \
    int occupancy_render_##M##_##F(int* blocks_per_sm, size_t smem);
RTC_DECLARE_INST(0, 0) RTC_DECLARE_INST(0, 2) RTC_DECLARE_INST(0, 3)
RTC_DECLARE_INST(1, 0) RTC_DECLARE_INST(1, 2) RTC_DECLARE_INST(1, 3)

}  // namespace rtc
