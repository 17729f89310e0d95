// The render kernel: Camera.render / renderWorker (camera.zig:80-125) + World.colorAt and everything
// below it (world.zig:86-189, 212-289) for every pixel of a partition of the frame.
//
// Execution model:
//   * persistent threads: the grid is sized to the machine (SMs x resident CTAs); warps pull pixels
//     from a global atomic counter, 32-pixel 8x4 tiles at a time when a whole warp is idle, single
//     pixels (ballot + popc ranked) when only some lanes are;
//   * the recursion colorAt -> shadeHit -> reflectedColor/refractedColor -> colorAt is unrolled
//     into a per-lane state machine.  shadeHit is linear in the colours its children return, so a
//     ray carries the product of the factors on its path (reflective, transparency, Schlick) and
//     adds weight x surface to the pixel; pending children sit on a per-lane stack of <= 6 rays;
//   * every trip through the loop each live lane runs exactly ONE world query (closest hit, shadow
//     any-hit or the containers scan), so lanes in different shading stages still share the
//     traversal code and the warp stays converged where the time goes.
#pragma once
#include <cuda_runtime.h>

#include "render_params.h"
#include "rt_pattern.cuh"
#include "rt_trace.cuh"

namespace rtc {

struct RayTask {
    double ox, oy, oz, dx, dy, dz, w;
    int remaining, pad;
};
#define RTC_RAY_STACK 8

enum { ST_IDLE = 0, ST_CLOSEST = 1, ST_CONTAINERS = 2, ST_SHADOW = 3, ST_WAIT = 4 };  // ST_WAIT: a lane of a pipelined pixel between two jobs

// color.zig:61-71 : round half away from zero, clamped to [0,255]
__device__ __forceinline__ unsigned char quantise(double c) {
    double t = round(c * 255.0);
    if (!(t > 0.0)) return 0;
    if (t > 255.0) return 255;
    return (unsigned char)t;
}

// Per-lane state that is NOT needed while a query is being traced lives in shared memory, one
// column of doubles per thread (slot k of thread t at park[k * RTC_BLOCK + t]: conflict-free).  The
// kernel is latency-bound and wants many resident warps, i.e. few registers per thread; with this
// state in registers ptxas spilled traversal-loop variables to local memory instead (the per-thread
// frames of a full grid exceed L2: ncu showed 650 MB of DRAM traffic per frame from spills alone).
enum {
    PK_RO = 0,       // current ray origin
    PK_RD = 3,       // current ray direction
    PK_PCOL = 6,     // colour accumulated for the current pixel: three int64 in fixed point (bit-cast into the slots)
    PK_NORMAL = 9,   // hit normal (flipped towards the eye)
    PK_COLOR = 12,   // pattern colour at the hit
    PK_SPARE = 15,   // (three free slots)
    PK_RW = 18,      // product of the reflective / transparency / Schlick factors along the path
    PK_N1 = 19,
    PK_N2 = 20,
    PK_T = 21,       // t of the current hit
    PK_MB = 22,      // MB_SLOTS slots: the lane's mailbox for the work sharing of the mesh traversal (rt_trace.cuh trace_meshes)
    PK_SLOTS = RTC_PARK_SLOTS
};
static_assert(PK_MB + MB_SLOTS == PK_SLOTS, "RTC_PARK_SLOTS (render_params.h) out of date");
__device__ __forceinline__ V3 pk_ld(const double* pk, int k) {
    return v3(pk[k * RTC_BLOCK], pk[(k + 1) * RTC_BLOCK], pk[(k + 2) * RTC_BLOCK]);
}
__device__ __forceinline__ void pk_st(double* pk, int k, V3 v) {
    pk[k * RTC_BLOCK] = v.x;
    pk[(k + 1) * RTC_BLOCK] = v.y;
    pk[(k + 2) * RTC_BLOCK] = v.z;
}
#define PK(k) pk[(k) * RTC_BLOCK]

// The copy warps of a frame that goes to mapped host memory (RenderParams::copy_out): take the finished rows of tiles in
// the order they are announced and move each -- RTC_TILE_H scanlines, one contiguous run of hsize x 3 x RTC_TILE_H doubles
// -- to the host canvas, 32 lanes x 8 bytes side by side, several loads in flight per lane.  (Tile by tile, 192 bytes at
// a time, the same warps moved a 1080p frame at 43 GB/s and a 2160p frame at 21 GB/s: the host side wants long runs.)
__device__ __forceinline__ void copy_finished_rows(const RenderParams& P) {
    const unsigned lane = threadIdx.x & 31u;
    const unsigned long long row_doubles = P.cam.hsize * 3ull;
    for (;;) {
        unsigned idx = 0, row = 0xffffffffu;
        if (lane == 0) idx = atomicAdd(P.copy_ctl + 1, 1u) + 1u;  // (the cursors start at 0xffffffff: one memset sets everything up)
        idx = __shfl_sync(0xffffffffu, idx, 0);
        if (idx >= P.copy_rows) return;
        if (lane == 0) {
            volatile unsigned* slot = P.done_list + idx;
            unsigned spins = 0;
            while ((row = *slot) == 0xffffffffu && ++spins < (1u << 22)) __nanosleep(500);  // (~2 s: a frame takes milliseconds)
            if (row == 0xffffffffu) atomicOr(P.counters + 4, (unsigned long long)ERR_COPY);  // never announced: flag the frame, do not hang
            __threadfence();  // the row's pixels were written before it was announced
        }
        row = __shfl_sync(0xffffffffu, row, 0);
        if (row == 0xffffffffu) return;
        // `row` counts the rows of tiles of this PARTITION (packed rows pr0 ...); its scanlines lie together in the frame
        // as long as tile_rows is a multiple of RTC_TILE_H (the host checks): y as in the refill above
        const unsigned long long pr0 = (unsigned long long)row * RTC_TILE_H;
        const unsigned long long pr1 = pr0 + RTC_TILE_H < P.rows ? pr0 + RTC_TILE_H : P.rows;
        const unsigned long long y0 = ((pr0 / P.tile_rows) * P.n_parts + P.part) * P.tile_rows + pr0 % P.tile_rows;
        const unsigned long long first = y0 * row_doubles, last = (y0 + (pr1 - pr0)) * row_doubles;
        const double* __restrict__ src = P.out;
        double* __restrict__ dst = P.copy_out;
        unsigned long long i = first + lane;
        for (; i + 7ull * 32ull < last; i += 8ull * 32ull) {
            double v[8];
#pragma unroll
            for (int k = 0; k < 8; k++) v[k] = __ldcg(src + i + 32ull * k);
#pragma unroll
            for (int k = 0; k < 8; k++) dst[i + 32ull * k] = v[k];
        }
        for (; i < last; i += 32ull) dst[i] = __ldcg(src + i);
    }
}

// MESH: the scene has triangle meshes (BVH traversal compiled in; the traversal stacks follow the
// parked state in dynamic shared memory).  Scenes without meshes run the instantiation without it.
// Resident CTAs per SM the register allocation aims for, measured on B200 (the kernel is latency-bound:
// occupancy against spills).  Analytic scenes: 6 (80 registers); mesh scenes keep the traversal state of
// mesh_inner / mesh_leaf in registers: 4 (128 registers).
// Idle lanes a warp needs before it fetches new pixels.  32 = a warp takes a whole 8x4 tile when all its lanes are
// through with the previous one: the lanes then stay in step (closest hit, shadow ray to light 0, to light 1, ...),
// the rays of one trip through the world are alike and the traversal loop runs with far fewer idle lanes.  Measured
// on B200 (kernel ms at 1 / 8 / 16 / 32): dragons 1.52 / 1.40 / 1.37 / 1.23, cover 4.8 / 4.2 / 3.7 / 3.3, fresnel
// 0.96 / 0.92 / 0.89 / 0.83, teapot 0.60 throughout.  (1 = every lane for itself, the first design.)
// (The lanes-per-pixel instantiations below rely on whole-warp refills: since they exist, only 32 compiles; the
// measurements above were taken with -DRTC_REFILL_MIN=... variant builds of the one-lane kernel.)
#ifndef RTC_REFILL_MIN
#define RTC_REFILL_MIN 32
#endif
#ifndef RTC_MIN_BLOCKS_ANALYTIC
#define RTC_MIN_BLOCKS_ANALYTIC 6
#endif
// ... analytic scenes that need CSG or the general pattern evaluator (FEAT != 0): 5 (96 registers).  Measured on B200, kernel
// ms at 6 / 5 / 4 CTAs per SM: csg.json 2160p 1.472 / 1.398 / 1.406, skybox.json 2160p 1.257 / 1.228 / 1.280; the lean
// instantiation: fresnel 600^2 0.530 / 0.505 / 0.517 but cover.json 1080p 2.044 / 2.201 / 2.347 -- it stays at 6.
#ifndef RTC_MIN_BLOCKS_ANALYTIC_GENERAL
#define RTC_MIN_BLOCKS_ANALYTIC_GENERAL 5
#endif
#ifndef RTC_MIN_BLOCKS_MESH
#define RTC_MIN_BLOCKS_MESH 4
#endif
// LPP = lanes per pixel (1, 2 or 4; rtc_api.cu picks it per partition size).  With LPP > 1 the LPP neighbouring lanes
// of a warp take the SAME pixel: they trace the same closest-hit (and containers) queries side by side -- same ray, same
// nodes, same result, same stack of child rays -- and share the LIGHTS of every hit, lane `sub` tracing the shadow rays
// to lights sub, sub + LPP, ...  Each lane adds its lights' terms to its own fixed-point accumulator; the LPP partial
// sums are added (integers: exact, the frame stays bit-identical) when the warp's lanes are all through.  The chain of
// dependent trips behind a pixel shrinks from 1 + n_lights per hit to 1 + ceil(n_lights / LPP), at the price of LPP
// times the closest-hit work: a loss while there are enough pixels to fill the machine, a gain for the small
// partitions of a multi-GPU frame, whose time is the chain of their most expensive tile (profiles/README.md).
// FEAT = what of the rarer machinery the scene needs (rtc_api.cu decides per scene):
//   RTC_FEAT_CSG       CSG objects, guard boxes, stray triangles among the prims (csg.json);
//   RTC_FEAT_PATTERNS  patterns beyond a solid colour or ONE simple pattern over two solid colours: texture maps, perturb,
//                      nested patterns (skybox / earth / align_check / nefertiti .json).
// An instantiation carries only what its scene needs: fewer instructions to fetch (the render loop is larger than the
// instruction cache: stall_no_instruction leads the ncu captures), no calls in the loop, fewer live registers.  FEAT = 0
// ("lean") serves 9 of the 14 shipped scenes, dragons / teapot / cover / fresnel among them.
#ifdef RTC_MAXNREG_MESH  // experiment: an explicit register limit for the mesh instantiations instead of a CTA count
#define RTC_KERNEL_BOUNDS __launch_bounds__(RTC_BLOCK) __maxnreg__(MESH ? RTC_MAXNREG_MESH : FEAT ? 96 : 80)
#else
#define RTC_KERNEL_BOUNDS \
    __launch_bounds__(RTC_BLOCK, MESH ? RTC_MIN_BLOCKS_MESH : FEAT ? RTC_MIN_BLOCKS_ANALYTIC_GENERAL : RTC_MIN_BLOCKS_ANALYTIC)
#endif
template <bool MESH, int LPP, int FEAT>
__global__ void RTC_KERNEL_BOUNDS render_kernel(DScene S, RenderParams P) {
    constexpr bool LEAN = (FEAT & RTC_FEAT_CSG) == 0;            // no CSG / guards / stray triangles
    constexpr bool SIMPLE_PAT = (FEAT & RTC_FEAT_PATTERNS) == 0;  // rt_pattern.cuh simple_pattern_color suffices
    static_assert(LPP == 1 || RTC_REFILL_MIN == 32, "lanes sharing a pixel rely on whole-warp refills");
    // LPP == 0 ("mixed"): four lanes per pixel on the tiles the last frame found expensive (the first *P.heavy_tiles
    // entries of tile_order: cost >= 8x the mean), one lane per pixel on all others.  The few per cent of the tiles
    // whose chains of dependent trips set the time of a small partition get the short chain; the rest of the frame
    // is not traced twice.
    constexpr bool MIXED = LPP == 0;
    if constexpr (LPP == 1) {  // (a decoupled host canvas is a whole frame with one lane per pixel)
        if (P.copy_out && blockIdx.x < P.copy_ctas) {
            copy_finished_rows(P);
            return;
        }
    }
    // Lanes per pixel of a heavy tile (mixed kernel).  4: as LPP = 4, all four lanes trace a hit's closest-hit query and
    // split its lights.  8 (PIPE): the pixel's lanes work as a PIPELINE -- lane 0 only walks the chain of closest-hit
    // queries (hit -> children -> next ray -> hit ...), and every hit it finds is handed, through the parked state in
    // shared memory, to lanes 1 .. n_lights, which trace its shadow rays in the NEXT trip while lane 0 is already on the
    // next ray.  A level of the ray tree then costs one trip instead of two (1 + n_lights / lanes): the chain behind the
    // deepest pixels of a small partition -- glass boxes, 6 levels -- is what its frame time is made of.
    constexpr int HL = RTC_HEAVY_LANES;
    constexpr bool PIPE = MIXED && HL == 8;
    // RTC_HEAVY_GROUPS = G > 1: G such groups of four lanes per pixel, HT = 4 G lanes in all, which split the pixel's RAY
    // TREE between them.  Every group starts at the primary ray; at the first hit that spawns BOTH a reflected and a
    // refracted ray the groups whose next path bit is 0 follow the reflection and the others the refraction, and so on
    // until the log2(G) bits are used up -- from there a group walks its subtree with its own ray stack as before.  A hit
    // on the shared upper part of the tree is traced by several groups but SHADED (shadow rays, lighting, ray counts) by
    // exactly one, the group whose remaining path bits are all 0; the partial sums of all HT lanes are added as integers
    // at the end, so the pixel is bit-identical.  The chain of dependent trips behind a glass pixel (63 hits at depth 5)
    // becomes log2(G) + 63 / G hits.
    constexpr int HG = MIXED ? RTC_HEAVY_GROUPS : 1;
    constexpr int HT = HL * HG;
    constexpr bool SPLIT = HG > 1;
    static_assert(HG == 1 || HG == 2 || HG == 4 || HG == 8, "RTC_HEAVY_GROUPS: 1, 2, 4 or 8");
    static_assert(!SPLIT || HL == 4, "groups split the tree of a pixel only in the four-lane mode");
    constexpr int HG_BITS = HG == 8 ? 3 : HG == 4 ? 2 : HG == 2 ? 1 : 0;
    // (of the heavy tiles at the head of the order this renderer is dealt its runs, like of all others)
    unsigned long long heavy_items = 0ull;
    if (MIXED && P.tile_order) {
        const unsigned h = __ldg(P.heavy_tiles), run = 1u << P.deal_shift, round = run * P.deal_n;
        const unsigned rem = h % round, lo = P.deal_rank * run;
        heavy_items = (unsigned long long)(h / round * run + (rem > lo ? min(rem - lo, run) : 0u)) * (32ull * HT);
    }
    const unsigned long long n_items = P.n_items + heavy_items / HT * (HT - 1);  // a heavy tile is 32 x HT lane items instead of 32
    int lpp = MIXED ? 1 : LPP;  // lanes per pixel of the tile this warp is working on
    extern __shared__ double s_mem[];  // PK_SLOTS x RTC_BLOCK doubles, then (MESH) S.stack_cap x RTC_BLOCK ints
    double* pk = s_mem + threadIdx.x;
    int* stack = reinterpret_cast<int*>(s_mem + PK_SLOTS * RTC_BLOCK) + threadIdx.x;
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    unsigned sub = MIXED ? 0u : lane & (unsigned)(LPP - 1);  // which of the pixel's lanes this is
    bool wpend = false;                                // LPP > 1: the pixel is through, its sum not yet written
    unsigned gid = 0, grp = 0;  // SPLIT: the lane's group among the pixel's; the bits of it not used for a branch yet
    int nb = 0;                 // ... and how many of them are left
    (void)gid, (void)grp, (void)nb;

    int state = ST_IDLE;
    bool done = false;
    PK(PK_SPARE) = -1.0;  // lane 0: tile in flight (none)
    int err = 0;
    unsigned n_primary = 0, n_shadow = 0, n_secondary = 0, n_culled = 0;  // per thread: far below 2^32

    int remaining = 0;       // recursion budget of the current ray
    long long pix = -1;      // output index of the current pixel
    int hit_leaf = -1;       // current hit (valid in ST_CONTAINERS / ST_SHADOW)
    int h_mat = 0, h_light = 0;

    RayTask rays[RTC_RAY_STACK];
    int rsp = 0;

    Query q;
    q.mode = Q_CLOSEST;
    q.o = v3(0, 0, 0);
    q.d = v3(0, 0, 1);
    q.t = 0.0;
    q.leaf = -1;
    q.a = q.b = 0.0;
    q.c_leaf = -1;
    q.flag = 0;
    for (;;) {
        if constexpr (LPP != 1) {
            // every lane of the warp is through with its pixel: add up the partial sums of the lanes that shared one
            if (__all_sync(0xffffffffu, state == ST_IDLE) && __any_sync(0xffffffffu, lpp > 1)) {
                long long a0 = __double_as_longlong(PK(PK_PCOL)), a1 = __double_as_longlong(PK(PK_PCOL + 1)),
                          a2 = __double_as_longlong(PK(PK_PCOL + 2));
#pragma unroll
                for (int sft = 1; sft < (MIXED ? HT : LPP); sft <<= 1) {
                    a0 += __shfl_xor_sync(0xffffffffu, a0, sft);
                    a1 += __shfl_xor_sync(0xffffffffu, a1, sft);
                    a2 += __shfl_xor_sync(0xffffffffu, a2, sft);
                }
                if (wpend && sub == 0 && gid == 0) {
                    const double inv_scale = 1.0 / P.scale;  // a power of two: exact
                    V3 pcol = v3((double)a0 * inv_scale, (double)a1 * inv_scale, (double)a2 * inv_scale);
                    if (P.out8) {
                        reinterpret_cast<uchar4*>(P.out8)[pix] = make_uchar4(quantise(pcol.x), quantise(pcol.y), quantise(pcol.z), 255);
                    } else {
                        double* o = P.out + pix * 3;
                        o[0] = pcol.x;
                        o[1] = pcol.y;
                        o[2] = pcol.z;
                    }
                }
                wpend = false;
            }
        }
        // ---- refill idle lanes ------------------------------------------------------------------
        unsigned need = __ballot_sync(0xffffffffu, state == ST_IDLE && !done);
#if RTC_REFILL_MIN > 1
        // refill in groups (see RTC_REFILL_MIN): idle lanes wait until enough of them have gathered
        if (__popc(need) < RTC_REFILL_MIN && __any_sync(0xffffffffu, state != ST_IDLE)) need = 0;
#endif
        if (need) {
            // cost of the tile this warp is through with (clock ticks / 64), for the order of the NEXT frame
            // (added to the cost arrays of all renderers sharing the frame, so that all copies stay equal)
            if (LPP == 1 && P.copy_out) __syncwarp();  // every lane's pixel stores before lane 0 announces the tile
            if ((P.cost_peer[0] || (LPP == 1 && P.copy_out)) && lane == 0 && PK(PK_SPARE) >= 0.0) {
                const unsigned done_tile = (unsigned)PK(PK_SPARE);
                if (P.cost_peer[0]) {
                    const long long t0 = __double_as_longlong(PK(PK_SPARE + 1));
                    const unsigned cost = (unsigned)((clock64() - t0) >> 6);
                    for (unsigned r = 0; r < P.deal_n; r++) {
                        unsigned* cp = P.cost_peer[r];
                        atomicAdd(cp + done_tile, cost);
                        atomicAdd(reinterpret_cast<unsigned long long*>(cp + P.cost_tiles), (unsigned long long)cost);  // the mean of the frame (tile_order.cu)
                    }
                }
                if (LPP == 1 && P.copy_out) {  // for the copy warps (copy_finished_rows): one more tile of its row is done
                    const unsigned row = (unsigned)PK(PK_SPARE + 2) / P.tiles_x;
                    // the tile's pixels before the count: a RELEASE on the counter, not __threadfence() -- that one compiles to
                    // MEMBAR.SC + CCTL.IVALL, and dropping the SM's whole L1 (the BVH nodes) after every tile cost 15-20 % of the
                    // kernel (measured; profiles/README.md).  The one thread per row that goes on to announce it fences fully.
                    unsigned before;
                    asm volatile("atom.release.gpu.global.add.u32 %0, [%1], 1;" : "=r"(before) : "l"(P.row_count + row) : "memory");
                    if (before + 2u == P.tiles_x) {  // (counts start at 0xffffffff) the row's last tile
                        __threadfence_system();  // (the list may be in host memory, read by the host: rtc_api.cu render_to_pinned)
                        P.done_list[atomicAdd(P.copy_ctl, 1u) + 1u] = row;
                    }
                }
                PK(PK_SPARE) = -1.0;
            }
            unsigned long long base = 0;
            const int leader = __ffs(need) - 1;
            if ((int)lane == leader) base = atomicAdd(P.counters, (unsigned long long)__popc(need));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (state == ST_IDLE && !done) {
                unsigned long long item = base + __popc(need & lt_mask);
                if (item >= n_items) {
                    done = true;
                } else {
                    // 8x4 pixel tiles, row-major tiles over the packed rows of this partition
                    unsigned long long pitem;  // the lanes of a pixel hold consecutive items
                    if constexpr (MIXED) {
                        const bool heavy = item < heavy_items;  // the same for the whole warp: heavy_items % 32 == 0
                        lpp = heavy ? HL : 1;
                        sub = heavy ? (unsigned)item & (unsigned)(HL - 1) : 0u;
                        pitem = heavy ? item / (unsigned)HT : item - heavy_items / HT * (HT - 1);
                        if constexpr (SPLIT) {
                            gid = grp = heavy ? ((unsigned)item / (unsigned)HL) & (unsigned)(HG - 1) : 0u;
                            nb = heavy ? HG_BITS : 0;
                        }
                    } else {
                        pitem = item / (unsigned)LPP;
                    }
                    unsigned long long tile = ((((unsigned)(pitem >> 5) >> P.deal_shift) * P.deal_n + P.deal_rank) << P.deal_shift) |
                                              ((unsigned)(pitem >> 5) & ((1u << P.deal_shift) - 1u));  // the entry of the order this renderer is dealt
                    unsigned in = (unsigned)(pitem & 31u);
                    if (P.tile_order) tile = __ldg(P.tile_order + tile);  // most expensive tiles of the last frame first
                    if ((P.cost_peer[0] || (LPP == 1 && P.copy_out)) && lane == 0) {
                        PK(PK_SPARE) = (double)tile;
                        PK(PK_SPARE + 1) = __longlong_as_double(clock64());
                    }
                    unsigned long long px = (tile % P.tiles_x) * RTC_TILE_W + (in & (RTC_TILE_W - 1u));
                    unsigned long long trow = tile / P.tiles_x;
                    if (P.row_stride != 1) {  // strided order of the row groups (render_params.h), 32-bit arithmetic
                        const unsigned t = (unsigned)trow, g = t / P.row_group;
                        if (g < P.n_row_groups)
                            trow = (unsigned)(((unsigned long long)g * P.row_stride) % P.n_row_groups) * P.row_group + t % P.row_group;
                    }
                    if (LPP == 1 && P.copy_out && lane == 0) PK(PK_SPARE + 2) = (double)(unsigned)(trow * P.tiles_x + tile % P.tiles_x);
                    unsigned long long pr = trow * RTC_TILE_H + (in >> RTC_TILE_W_LOG2);
                    if (px < P.cam.hsize && pr < P.rows) {
                        unsigned long long y = ((pr / P.tile_rows) * P.n_parts + P.part) * P.tile_rows + pr % P.tile_rows;
                        // Camera.rayForPixel (camera.zig:64-76)
                        double xoffset = ((double)px + 0.5) * P.cam.pixel_size;
                        double yoffset = ((double)y + 0.5) * P.cam.pixel_size;
                        double wx = P.cam.half_width - xoffset;
                        double wy = P.cam.half_height - yoffset;
                        const double* m = P.cam.inv;
                        V3 pixel = v3(m[0] * wx + m[1] * wy + m[2] * -1.0 + m[3], m[4] * wx + m[5] * wy + m[6] * -1.0 + m[7],
                                      m[8] * wx + m[9] * wy + m[10] * -1.0 + m[11]);
                        V3 origin = v3(m[3], m[7], m[11]);
                        pk_st(pk, PK_RO, origin);
                        pk_st(pk, PK_RD, normalized(pixel - origin));
                        PK(PK_RW) = 1.0;
                        remaining = P.max_depth;
                        pix = (long long)((P.full_frame ? y : pr) * P.cam.hsize + px);
                        pk_st(pk, PK_PCOL, v3(0, 0, 0));  // +0.0 is also the bit pattern of int64 0
                        rsp = 0;
                        state = ST_CLOSEST;
                        if (PIPE && sub != 0) state = ST_WAIT;  // (lpp == HL: only there sub != 0) a shadow lane: waits for lane 0's hits
                        if (sub == 0 && gid == 0) n_primary++;
                    }
                    // items outside the frame are simply dropped: the lane asks again next trip
                }
            }
        }
        if (__all_sync(0xffffffffu, done && state == ST_IDLE)) break;

        // ---- one world query per live lane --------------------------------------------------------
        MeshTrav trav;
        bool busy = false, skip = false;
        (void)trav;
        const bool live = state != ST_IDLE && !(PIPE && state == ST_WAIT);
        bool publish = false;  // PIPE, lane 0 of a pixel: found a hit in this trip
        (void)publish;
        if (live) {
            q.o = pk_ld(pk, PK_RO);
            q.d = pk_ld(pk, PK_RD);
            if (state == ST_CLOSEST) {
                q.mode = Q_CLOSEST;
                q.t = RTC_INF;
                q.leaf = -1;
                q.a = q.b = 0.0;
                q.c_leaf = -1;
            } else if (state == ST_SHADOW) {  // World.isShadowed (world.zig:126-135)
                // over_point = point + normal * EPSILON, recomputed from the parked ray and hit
                // (same operations as PreComputations.new, world.zig:219-224: same bits)
                V3 n = pk_ld(pk, PK_NORMAL);
                V3 over = (q.o + q.d * PK(PK_T)) + n * 1e-5;
                const rtc_light& Lg = S.lights[h_light];
                V3 dir = v3(Lg.position[0], Lg.position[1], Lg.position[2]) - over;
                double dist = sqrt(dot(dir, dir));
                q.mode = Q_SHADOW;
                q.o = over;
                q.d = v3(dir.x / dist, dir.y / dist, dir.z / dist);
                q.t = dist;
                q.flag = 0;
                n_shadow++;
                // Light behind the surface: lighting() yields ambient whether or not the point is
                // shadowed (material.zig:55-63: in_shadow -> ambient; light_dot_normal < 0 -> ambient
                // + black + black), so the ray need not be traced.  Counted, and reported as culled.
                if (!(dot(q.d, n) >= 0.0)) {
                    skip = true;
                    q.flag = 1;
                    n_culled++;
                }
            } else {
                q.mode = Q_CONTAINERS;
                q.t = PK(PK_T);
                q.leaf = hit_leaf;
                q.a = -RTC_INF;
                q.c_leaf = -1;
                q.flag = 0;
                q.b = -RTC_INF;
            }
        }
        // analytic prims and CSG objects are finished here (warp-wide call); meshes are set up in trav
#ifdef RTC_TRIP_PROBE
        const long long probe_t0 = clock64();
#endif
        busy = trace_begin<MESH, LEAN>(q, S, trav, &err, live && !skip);
        // every mesh through ONE warp-wide traversal loop; the full-mask votes are the points where
        // the warp reconverges, so inner-node steps and triangle tests each run with all traversing lanes
        // (and lanes that are through help those that are not: trace_meshes)
        if constexpr (MESH) trace_meshes(q, S, trav, stack, s_mem + PK_MB * RTC_BLOCK + (threadIdx.x & ~31u), &err, busy, live, lane);

#ifdef RTC_TRIP_PROBE
        {   // how long does one warp trip through the world take? [6] max cycles, [7] trips | long trips << 32
            const long long dt = clock64() - probe_t0;
            if (lane == 0) {
                atomicMax(P.counters + 6, (unsigned long long)dt);
                atomicAdd(P.counters + 7, 1ull + (dt > 98000 ? (1ull << 32) : 0ull) + (dt > 39000 ? (1ull << 48) : 0ull));
            }
        }
#endif
        // ---- consume the result -----------------------------------------------------------------------
        bool pop_next = false;
        if (state == ST_CLOSEST) {
            if (q.leaf < 0) {
                pop_next = true;  // colorAt: no hit -> black (world.zig:119-121)
            } else {
                // PreComputations.new (world.zig:212-227); q.o / q.d still hold the ray
                hit_leaf = q.leaf;
                PK(PK_T) = q.t;
                const rtc_leaf& L = S.leaves[hit_leaf];
                M34 inv = load_m34(S.transforms[L.transform].inverse);
                V3 point = q.o + q.d * q.t;
                int kind = L.kind;
                V3 lp = xf_point(inv, point);  // Shape.normalAt (shape.zig:338-350)
                double u = q.a, v = q.b;
                const double* nrm = nullptr;
                if (kind == RTC_TRIANGLE || kind == RTC_SMOOTH_TRIANGLE) {
                    nrm = S.normals[L.triangle].n;
                    if (!LEAN && kind == RTC_SMOOTH_TRIANGLE && q.c_leaf != hit_leaf) {
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
                pk_st(pk, PK_NORMAL, n);
                V3 over = point + n * 1e-5;
                h_mat = L.material;
                const rtc_material& M = S.materials[h_mat];
                // Material.lighting evaluates the pattern at over_point for every light (material.zig:49);
                // it is the same colour each time, so once per hit.
                if constexpr (SIMPLE_PAT) {  // every pattern is a solid colour or one simple pattern over two of them
                    pk_st(pk, PK_COLOR, simple_pattern_color(S, M.pattern, xf_point(inv, over)));
                } else {
                    pk_st(pk, PK_COLOR, pattern_color(S, M.pattern, xf_point(inv, over), &err));
                }
                h_light = (int)sub;
                PK(PK_N1) = 1.0;
                PK(PK_N2) = 1.0;
                // n1/n2 only reach a result through refractedColor / schlick, both gated on
                // transparency (world.zig:101-102, :178)
                // ... and when every material of the scene has refractive index 1.0 (dragons.json's glass boxes),
                // n1 = n2 = 1.0 whatever the containers list holds (world.zig:236-254): the walk is skipped.
                // (SPLIT: a group that is not the one shading this hit goes straight on to the children)
                if (M.transparency != 0.0 && !S.ri_all_one) state = ST_CONTAINERS;
                else state = S.n_lights > (int)sub && !(SPLIT && grp != 0) ? ST_SHADOW : -1;  // no lights (left for this lane): straight to the children below
                if (PIPE && lpp == HL) {  // lane 0 of a pipelined pixel: its shadow rays are the other lanes' business
                    publish = S.n_lights > 0;
                    if (state == ST_SHADOW) state = -1;
                }
            }
        } else if (state == ST_CONTAINERS) {
            double ri_hit = S.materials[S.leaves[hit_leaf].material].refractive_index;
            double ri_c = q.c_leaf >= 0 ? S.materials[S.leaves[q.c_leaf].material].refractive_index : 1.0;
            if (q.flag) {  // the hit leaf was a container itself
                bool hit_is_last = q.c_leaf < 0 || q.b > q.a || (q.b == q.a && hit_leaf > q.c_leaf);
                PK(PK_N1) = hit_is_last ? ri_hit : ri_c;
                PK(PK_N2) = ri_c;
            } else {
                PK(PK_N1) = ri_c;
                PK(PK_N2) = ri_hit;
            }
            state = S.n_lights > (int)sub && !(SPLIT && grp != 0) ? ST_SHADOW : -1;
            if (PIPE && lpp == HL) state = -1;
        } else if (state == ST_SHADOW) {
            // Material.lighting (material.zig:40-74)
            const rtc_light& Lg = S.lights[h_light];
            const rtc_material& M = S.materials[h_mat];
            V3 li = v3(Lg.intensity[0], Lg.intensity[1], Lg.intensity[2]);
            V3 color = pk_ld(pk, PK_COLOR);
            V3 eff = v3(color.x * li.x, color.y * li.y, color.z * li.z);
            V3 sum = eff * M.ambient;
            if (!q.flag) {
                V3 lightv = q.d;  // == normalize(light.position - over_point)
                V3 normal = pk_ld(pk, PK_NORMAL);
                double ldn = dot(lightv, normal);
                if (ldn >= 0.0) {
                    V3 diffuse = eff * (M.diffuse * ldn);
                    V3 reflectv = lightv - normal * (2.0 * dot(lightv, normal));  // tuple.zig reflect
                    double rde = dot(neg(reflectv), neg(pk_ld(pk, PK_RD)));
                    V3 spec = v3(0, 0, 0);
                    if (rde > 0.0) spec = li * (M.specular * zig_pow(rde, M.shininess));
                    sum = sum + diffuse + spec;
                }
            }
            // shadeHit is linear: weight x lighting() of THIS light goes straight into the pixel, as an exact
            // integer addition (fixed point, P.scale units per 1.0).  Exact means order-independent: the frame is
            // bit-identical whatever the partition, the lane assignment or the kernel (the wavefront path adds
            // the very same terms, render_wavefront.cuh wf_add).
            {
                V3 c = sum * PK(PK_RW);
                double v[3] = {c.x * P.scale, c.y * P.scale, c.z * P.scale};
#pragma unroll
                for (int k = 0; k < 3; k++) {
                    // outside the range the scene's bound guarantees (rtc_api.cu accumulator_scale: e.g. a NaN out of
                    // a degenerate normal): the frame is flagged, never silently wrong
                    if (!(fabs(v[k]) < 9.0e18)) {
                        err |= ERR_ACC_RANGE;
                        continue;
                    }
                    long long a = __double_as_longlong(PK(PK_PCOL + k)) + __double2ll_rn(v[k]);
                    PK(PK_PCOL + k) = __longlong_as_double(a);
                }
            }
            h_light += lpp;
            if (h_light >= S.n_lights) state = -1;
            if (PIPE && lpp == HL) state = ST_WAIT;  // a shadow lane: one light per hit, then the next hit
        }

        if constexpr (PIPE) {
            if (lpp == HL) {  // (the same for the whole warp)
                // lane 0 of every pixel hands the hit it found in this trip to the pixel's shadow lanes: they copy what
                // lighting() needs out of its parked state (ray, t, normal, colour, weight) before it moves on
                const int src = (int)(lane & ~(unsigned)(HL - 1));
                const bool pub = __shfl_sync(0xffffffffu, (int)publish, src) != 0;
                const int pmat = __shfl_sync(0xffffffffu, h_mat, src);
                const bool pdone = __shfl_sync(0xffffffffu, (int)(state == ST_IDLE), src) != 0;  // its pixel has no rays left
                __syncwarp();
                if (sub != 0) {
                    if (pub && (int)sub <= S.n_lights) {
                        const double* ppk = pk - sub;
#pragma unroll
                        for (int k = 0; k < 3; k++) {
                            PK(PK_RO + k) = ppk[(PK_RO + k) * RTC_BLOCK];
                            PK(PK_RD + k) = ppk[(PK_RD + k) * RTC_BLOCK];
                            PK(PK_NORMAL + k) = ppk[(PK_NORMAL + k) * RTC_BLOCK];
                            PK(PK_COLOR + k) = ppk[(PK_COLOR + k) * RTC_BLOCK];
                        }
                        PK(PK_T) = ppk[PK_T * RTC_BLOCK];
                        PK(PK_RW) = ppk[PK_RW * RTC_BLOCK];
                        h_mat = pmat;
                        h_light = (int)sub - 1;
                        state = ST_SHADOW;
                    } else if (state == ST_WAIT && pdone) {
                        state = ST_IDLE;
                    }
                }
                __syncwarp();  // before lane 0 overwrites its slots with the next ray
            }
        }

        if (state == -1) {
            // shadeHit tail (world.zig:97-107) with the children deferred onto the ray stack
            const rtc_material& M = S.materials[h_mat];
            const double rw = PK(PK_RW), h_n1 = PK(PK_N1), h_n2 = PK(PK_N2);
            const V3 ro = pk_ld(pk, PK_RO), rd = pk_ld(pk, PK_RD), h_normal = pk_ld(pk, PK_NORMAL);
            V3 eyev = neg(rd);
            double refl_w = 0.0, refr_w = 0.0;
            V3 refr_dir = v3(0, 0, 0);
            bool spawn_refl = remaining > 0 && M.reflective != 0.0;  // reflectedColor (world.zig:157-167)
            if (spawn_refl) refl_w = M.reflective;
            // refractedColor (world.zig:171-189)
            double n_ratio = h_n1 / h_n2;
            double cos_i = dot(eyev, h_normal);
            double sin2_t = n_ratio * n_ratio * (1.0 - cos_i * cos_i);
            bool spawn_refr = !(sin2_t > 1.0) && remaining > 0 && M.transparency != 0.0;
            if (spawn_refr) {
                double cos_t = sqrt(1.0 - sin2_t);
                refr_dir = h_normal * (n_ratio * cos_i - cos_t) - eyev * n_ratio;
                refr_w = M.transparency;
            }
            if (M.reflective > 0.0 && M.transparency > 0.0) {  // schlick (world.zig:272-289)
                double cosv = cos_i;
                double reflectance;
                bool tir = false;
                if (h_n1 > h_n2) {
                    double nr = h_n1 / h_n2;
                    double s2 = nr * nr * (1.0 - cosv * cosv);
                    if (s2 > 1.0) tir = true;
                    else cosv = sqrt(1.0 - s2);
                }
                if (tir) {
                    reflectance = 1.0;
                } else {
                    double frac = (h_n1 - h_n2) / (h_n1 + h_n2);
                    double r0 = frac * frac;
                    reflectance = r0 + (1.0 - r0) * zig_pow(1.0 - cosv, 5.0);
                }
                refl_w *= reflectance;
                refr_w *= (1.0 - reflectance);
            }
            bool take_refr = spawn_refr, take_refl = spawn_refl, count = sub == 0;
            if constexpr (SPLIT) {
                count = count && grp == 0;  // the group that shades this hit counts its children
                if (nb > 0 && spawn_refr && spawn_refl) {  // the tree branches: the groups on this hit part here
                    if (grp & 1u) take_refl = false;
                    else take_refr = false;
                    grp >>= 1;
                    nb--;
                }
            }
            if (spawn_refr || spawn_refl) {
                V3 h_point = ro + rd * PK(PK_T);  // world.zig:219
                if (spawn_refr) {
                    V3 under = h_point - h_normal * 1e-5;
                    if (take_refr && rsp < RTC_RAY_STACK) {
                        rays[rsp++] = {under.x, under.y, under.z, refr_dir.x, refr_dir.y, refr_dir.z, rw * refr_w, remaining - 1, 0};
                    }
                    if (count) n_secondary++;
                }
                if (spawn_refl) {
                    V3 h_over = h_point + h_normal * 1e-5;
                    V3 reflectv = rd - h_normal * (2.0 * dot(rd, h_normal));
                    if (take_refl && rsp < RTC_RAY_STACK) {
                        rays[rsp++] = {h_over.x, h_over.y, h_over.z, reflectv.x, reflectv.y, reflectv.z, rw * refl_w, remaining - 1, 0};
                    }
                    if (count) n_secondary++;
                }
            }
            pop_next = true;
        }

        if (pop_next) {
            if (rsp > 0) {
                RayTask tk = rays[--rsp];
                pk_st(pk, PK_RO, v3(tk.ox, tk.oy, tk.oz));
                pk_st(pk, PK_RD, v3(tk.dx, tk.dy, tk.dz));
                PK(PK_RW) = tk.w;
                remaining = tk.remaining;
                state = ST_CLOSEST;
            } else if (LPP != 1 && lpp > 1) {
                wpend = true;  // written at the top of the loop, once the lanes sharing this pixel are all through
                state = ST_IDLE;
            } else {
                const double inv_scale = 1.0 / P.scale;  // a power of two: exact
                V3 pcol = v3((double)__double_as_longlong(PK(PK_PCOL)) * inv_scale,
                             (double)__double_as_longlong(PK(PK_PCOL + 1)) * inv_scale,
                             (double)__double_as_longlong(PK(PK_PCOL + 2)) * inv_scale);
                if (P.out8) {
                    uchar4 c = make_uchar4(quantise(pcol.x), quantise(pcol.y), quantise(pcol.z), 255);
                    reinterpret_cast<uchar4*>(P.out8)[pix] = c;
                } else {
                    double* o = P.out + pix * 3;
                    o[0] = pcol.x;
                    o[1] = pcol.y;
                    o[2] = pcol.z;
                }
                state = ST_IDLE;
            }
        }
    }

    // ---- ray counters: warp reduce, one atomic per warp -------------------------------------------
    unsigned long long c_primary = n_primary, c_shadow = n_shadow, c_secondary = n_secondary, c_culled = n_culled;
    for (int s = 16; s > 0; s >>= 1) {
        c_primary += __shfl_down_sync(0xffffffffu, c_primary, s);
        c_shadow += __shfl_down_sync(0xffffffffu, c_shadow, s);
        c_secondary += __shfl_down_sync(0xffffffffu, c_secondary, s);
        c_culled += __shfl_down_sync(0xffffffffu, c_culled, s);
        err |= __shfl_down_sync(0xffffffffu, err, s);
    }
    if (lane == 0) {
        atomicAdd(P.counters + 1, c_primary);
        atomicAdd(P.counters + 2, c_shadow);
        atomicAdd(P.counters + 3, c_secondary);
        atomicAdd(P.counters + 5, c_culled);
        if (err) atomicOr(P.counters + 4, (unsigned long long)err);
    }
}

}  // namespace rtc
