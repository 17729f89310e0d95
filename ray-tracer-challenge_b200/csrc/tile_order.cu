// Order in which the megakernel hands out the tiles of the NEXT frame, made on the device from the time its warps
// spent on every tile of THIS frame (RenderParams::tile_cost / tile_order, render_kernel.cuh).
//
// Why: the tiles of a frame differ in cost by more than an order of magnitude (dragons.json 1920x1080: mean 38 us of
// warp time per 8x4 tile, 1 % of the tiles above 300 us, the longest 600 us -- pixels whose ray trees run through the
// glass boxes), and a persistent grid that meets such a tile late finishes late: the tail behind an empty work queue
// was a quarter of the frame.  Longest-processing-time-first scheduling needs the costs in advance; the previous frame
// of the same scene and frame size is a good predictor (re-renders of the bench, orbiting cameras of lib.zig:166-190).
// Nothing is cached or skipped: every pixel is traced in every frame, only the ORDER of the work changes, and the
// frame does not depend on it (each pixel is accumulated by itself in exact integers).
//
// The order is a STABLE sort into four buckets -- cost >= 8x, 4x, 2x the mean, the rest -- so that tiles of like cost
// keep the order they had (top to bottom, or the strided order of a mapped canvas): neighbouring tiles stay neighbours
// in time, which a full sort by cost loses (measured: teapot.json 0.61 -> 0.66 ms with a full sort).
//
// Two small launches behind the render kernel (a single CTA doing all of it took 0.14 ms for the 64,800 tiles of a
// 1080p frame, latency-bound): bucket counts per block of 1024 consecutive tiles, then every block scans the counts of
// the blocks before it (a few hundred numbers) and scatters its tiles, ranked by ballot so that the order inside a
// bucket is the tile order.  The mean comes from the render kernel (counters[6] = sum of all tile costs).
#include <cuda_runtime.h>

#include <cub/device/device_radix_sort.cuh>

#include "render_params.h"

namespace rtc {

namespace {
constexpr int T = 1024;

// last: buckets beyond it are merged into it (last = 3: all four; last = 1: the tiles that cost >= 8x the mean, then ALL
// others in their plain order -- frames whose finished tiles are streamed to a host canvas want the bulk of the tiles
// in memory order, rtc_api.cu)
// set 0: the four buckets above.  set 1 (frames shared tile by tile between renderers, rtc_api.cu): seven buckets, cost >=
// 12x, 8x, 7x, 6x, 4x, 2x the mean, the rest -- the counts of the leading buckets are the candidates for "how many tiles
// get four lanes per pixel" (render_kernel.cuh MIXED), and which threshold is best depends on the size of the share:
// tools/order_sim.py replays the recorded tile costs of dragons.json and predicts 12x for a 1/2 share, 6-8x for 1/4, 6x
// for 1/8 (249 -> 199 us).
constexpr int NB = 8;  // buckets at most
__device__ __forceinline__ int bucket_of(unsigned cost, unsigned long long n, unsigned long long sum, int last, int set) {
    if (sum == 0) return last;  // nothing measured (every tile below one tick): keep the plain order
    const unsigned long long c = (unsigned long long)cost * n;  // cost / mean = c / sum; cost < 2^32, n < 2^24
    int b;
    if (set == 0) b = c >= 8 * sum ? 0 : c >= 4 * sum ? 1 : c >= 2 * sum ? 2 : 3;
    else b = c >= 12 * sum ? 0 : c >= 8 * sum ? 1 : c >= 7 * sum ? 2 : c >= 6 * sum ? 3 : c >= 4 * sum ? 4 : c >= 2 * sum ? 5 : 6;
    return b < last ? b : last;
}

// counts[b * gridDim.x + block] = tiles of bucket b among the block's 1024
__global__ void __launch_bounds__(T) tile_count_kernel(const unsigned* __restrict__ cost, const unsigned long long* __restrict__ sum_p,
                                                       unsigned* __restrict__ counts, unsigned n, int last, int set) {
    __shared__ unsigned s_cnt[NB];
    if (threadIdx.x < NB) s_cnt[threadIdx.x] = 0;
    __syncthreads();
    const unsigned i = blockIdx.x * T + threadIdx.x;
    const int b = i < n ? bucket_of(cost[i], n, *sum_p, last, set) : -1;
#pragma unroll
    for (int k = 0; k < NB; k++) {
        const unsigned m = __ballot_sync(0xffffffffu, b == k);
        if ((threadIdx.x & 31u) == 0 && m) atomicAdd(&s_cnt[k], __popc(m));
    }
    __syncthreads();
    if (threadIdx.x < NB) counts[threadIdx.x * gridDim.x + blockIdx.x] = s_cnt[threadIdx.x];
}

__global__ void __launch_bounds__(T) tile_scatter_kernel(const unsigned* __restrict__ cost, const unsigned long long* __restrict__ sum_p,
                                                         const unsigned* __restrict__ counts, unsigned* __restrict__ order,
                                                         unsigned* __restrict__ heavy, unsigned n, int last, int set) {
    __shared__ unsigned s_base[NB];           // where this block's tiles of bucket b start
    __shared__ unsigned s_start[NB + 1];      // where bucket b starts (s_start[NB] = n)
    __shared__ unsigned s_warp[NB][T / 32];   // tiles of bucket b in each warp, then their exclusive prefix
    __shared__ unsigned s_red[2][T / 32];
    const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    // flattened (bucket, block) index of this block's bucket b is b * gridDim.x + blockIdx.x: sum everything before it
    for (int b = 0; b < NB; b++) {
        const unsigned first = b * gridDim.x, end = first + blockIdx.x;
        unsigned before = 0, mine = 0;  // tiles of the buckets before b (all blocks); tiles of bucket b in the blocks before this one
        for (unsigned j = tid; j < end; j += T) {
            const unsigned c = counts[j];
            if (j < first) before += c;
            else mine += c;
        }
        for (int s = 16; s > 0; s >>= 1) {
            before += __shfl_down_sync(0xffffffffu, before, s);
            mine += __shfl_down_sync(0xffffffffu, mine, s);
        }
        if (lane == 0) s_red[0][warp] = before, s_red[1][warp] = mine;
        __syncthreads();
        if (tid == 0) {
            unsigned t0 = 0, t1 = 0;
            for (int w = 0; w < T / 32; w++) t0 += s_red[0][w], t1 += s_red[1][w];
            s_start[b] = t0;
            s_base[b] = t0 + t1;
        }
        __syncthreads();
    }
    if (tid == 0) s_start[NB] = n;
    if (blockIdx.x == 0 && tid == 0)
        for (int k = 0; k + 1 < NB; k++) heavy[k] = s_start[k + 1];  // tiles that cost >= the k-th threshold (set 0: 8x, 4x, 2x the mean)
    const unsigned i = blockIdx.x * T + tid;
    const int b = i < n ? bucket_of(cost[i], n, *sum_p, last, set) : -1;
    unsigned rank = 0;  // among the tiles of my bucket in my warp
#pragma unroll
    for (int k = 0; k < NB; k++) {
        const unsigned m = __ballot_sync(0xffffffffu, b == k);
        if (b == k) rank = __popc(m & ((1u << lane) - 1u));
        if (lane == 0) s_warp[k][warp] = __popc(m);
    }
    __syncthreads();
    if (warp < NB) {  // warp k: exclusive prefix over the 32 warp counts of bucket k
        const unsigned c = s_warp[warp][lane];
        unsigned incl = c;
        for (int s = 1; s < 32; s <<= 1) {
            const unsigned up = __shfl_up_sync(0xffffffffu, incl, s);
            if ((int)lane >= s) incl += up;
        }
        s_warp[warp][lane] = incl - c;
    }
    __syncthreads();
    if (b < 0) return;
    const unsigned pos = s_base[b] + s_warp[b][warp] + rank;
    order[pos] = i;
}
// Order for frames whose finished ROWS of tiles are moved to a host canvas while the rest renders (rtc_api.cu
// render_to_pinned, copy warps).  What such a frame needs is a link that is never idle: finished rows at least as fast as
// the link takes them, from the first microseconds to the end of the kernel.  Two things stand in the way (measured on
// dragons.json 1080p, tools/order_sim.py replays them from the recorded tile costs):
//   * a row is complete only when its most expensive tile is, and in memory order that tile holds its row back by its
//     own duration (tiles of 130-500 us in a 0.9 ms frame) -> every tile is started AHEAD of its row by what it costs;
//   * the work is not spread evenly over the rows (dragons: the top 50 and bottom 50 rows of 270 hold 6 % of the work,
//     the middle 110 rows 80 %): top to bottom, the cheap rows at the bottom arrive last, together with the expensive
//     ones before them, and the link has a quarter of the frame left when the kernel ends -> the rows are taken in
//     BANDS of a few rows (one copy each), cheapest band first: cheap bands keep the link busy while the expensive
//     tiles, started at once because they are due so far ahead, work through their chains.
// key = position of the tile's row in that sequence - cost x (rows the frame advances per unit of cost); a stable sort
// by key.  The frame advances R rows in sum / W units of cost (W warps side by side; `lead` = W x a tuning factor).
__global__ void band_work_kernel(const unsigned* __restrict__ cost, unsigned n, unsigned per_band, unsigned long long* __restrict__ band_work) {
    __shared__ unsigned long long s_sum[32];
    const unsigned first = blockIdx.x * per_band, last = min(first + per_band, n);
    unsigned long long acc = 0;
    for (unsigned i = first + threadIdx.x; i < last; i += blockDim.x) acc += cost[i];
    for (int s = 16; s > 0; s >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, s);
    if ((threadIdx.x & 31u) == 0) s_sum[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t = 0;
        for (unsigned w = 0; w < blockDim.x / 32; w++) t += s_sum[w];
        band_work[blockIdx.x] = t;
    }
}
// seq[row] = place of the row in the sequence: bands by ascending work (ties: top first), the rows of a band together
// (by_work == 0: top to bottom -- frames whose link time is well above their kernel time: there the link is busy anyway
// and wants long runs of neighbouring rows, not the cheapest ones)
__global__ void band_rank_kernel(const unsigned long long* __restrict__ band_work, unsigned n_bands, unsigned band_rows, unsigned n_rows,
                                 int by_work, unsigned* __restrict__ seq) {
    if (!by_work) {
        for (unsigned r = threadIdx.x; r < n_rows; r += blockDim.x) seq[r] = r;
        return;
    }
    for (unsigned b = threadIdx.x; b < n_bands; b += blockDim.x) {
        const unsigned long long w = band_work[b];
        unsigned rows_before = 0;
        for (unsigned j = 0; j < n_bands; j++) {
            const unsigned long long wj = band_work[j];
            if (wj < w || (wj == w && j < b)) rows_before += min(band_rows, n_rows - j * band_rows);
        }
        for (unsigned r = b * band_rows; r < min((b + 1) * band_rows, n_rows); r++) seq[r] = rows_before + (r - b * band_rows);
    }
}
__global__ void row_key_kernel(const unsigned* __restrict__ cost, const unsigned long long* __restrict__ sum_p, unsigned n, unsigned tiles_x,
                               unsigned n_rows, unsigned lead, const unsigned* __restrict__ seq, unsigned* __restrict__ keys,
                               unsigned* __restrict__ vals) {
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned long long sum = *sum_p;
    const unsigned row = seq[i / tiles_x];
    // cost < 2^32, n_rows x lead < 2^31
    const unsigned long long ahead = sum ? ((unsigned long long)cost[i] * ((unsigned long long)n_rows * lead)) / sum : 0ull;
    keys[i] = ahead < row ? row - (unsigned)ahead : 0u;
    vals[i] = i;
}
// Order of a frame shared tile by tile (rtc_api.cu, dealt): by cost in steps of 1/8 of the mean, most expensive first, the
// tiles below the mean in one last bucket; memory order among equals (stable sort).  The seven buckets of set 1 leave
// a tail where a share is a few tiles per warp: tiles of 1-2x the mean handed out in memory order at the end
// (tools/order_sim.py, 1/4 share of dragons: 301 us with the buckets, 255 with this order, 249 with a full sort by cost).
// heavy[k] = tiles that cost >= 12x, 8x, 7x, 6x, 4x, 2x the mean: whole steps, so exactly the leading entries.
constexpr unsigned kFineLevels = 1023;  // key = kFineLevels - min(cost / (mean / 8), kFineLevels - 1); below the mean: kFineLevels
__global__ void fine_key_kernel(const unsigned* __restrict__ cost, const unsigned long long* __restrict__ sum_p, unsigned n,
                                unsigned* __restrict__ keys, unsigned* __restrict__ vals, unsigned* __restrict__ heavy) {
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned long long sum = *sum_p;
    unsigned long long lv = 0;  // cost in eighths of the mean (cost < 2^32, 8 n < 2^27)
    if (i < n && sum) lv = ((unsigned long long)cost[i] * (8ull * n)) / sum;
    if (i < n) {
        keys[i] = lv >= 8 ? kFineLevels - (unsigned)(lv < kFineLevels - 1 ? lv : kFineLevels - 1) : kFineLevels;
        vals[i] = i;
    }
    const unsigned thr[6] = {96, 64, 56, 48, 32, 16};
#pragma unroll
    for (int k = 0; k < 6; k++) {
        const unsigned m = __ballot_sync(0xffffffffu, i < n && lv >= thr[k]);
        if ((threadIdx.x & 31u) == 0 && m) atomicAdd(heavy + k, __popc(m));
    }
}
}  // namespace

// work / temp as for launch_row_order; heavy: six counters (zeroed here)
int launch_fine_order(const unsigned* cost, const unsigned long long* sum, unsigned* work, unsigned* order, unsigned* heavy, unsigned n_tiles,
                      void* temp, size_t temp_bytes, cudaStream_t stream) {
    if (!n_tiles) return 0;
    unsigned *keys = work, *keys_out = work + n_tiles, *vals = work + 2 * (size_t)n_tiles;
    cudaError_t e = cudaMemsetAsync(heavy, 0, 6 * sizeof(unsigned), stream);
    if (e != cudaSuccess) return (int)e;
    fine_key_kernel<<<(n_tiles + 255) / 256, 256, 0, stream>>>(cost, sum, n_tiles, keys, vals, heavy);
    e = cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys, keys_out, vals, order, (int)n_tiles, 0, 10, stream);
    if (e != cudaSuccess) return (int)e;
    return (int)cudaGetLastError();
}

size_t row_order_temp_bytes(unsigned n_tiles) {
    size_t bytes = 0;
    unsigned* p = nullptr;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, p, p, p, p, (int)n_tiles, 0, 32);
    return bytes;
}
// words of `work` for a frame of n_tiles tiles in rows of tiles_x
size_t row_order_work_words(unsigned n_tiles, unsigned tiles_x) {
    const size_t n_rows = (n_tiles + tiles_x - 1) / tiles_x;
    return 3 * (size_t)n_tiles + 3 * n_rows + 8;  // keys, sorted keys, tile numbers; seq[n_rows]; band_work[<= n_rows] (64-bit)
}

// lead = warps working side by side (x a tuning factor); band_rows = rows of tiles per band
int launch_row_order(const unsigned* cost, const unsigned long long* sum, unsigned* work, unsigned* order, unsigned n_tiles, unsigned tiles_x,
                     unsigned band_rows, int by_work, unsigned lead, void* temp, size_t temp_bytes, cudaStream_t stream) {
    if (!n_tiles) return 0;
    const unsigned n_rows = (n_tiles + tiles_x - 1) / tiles_x;
    if (band_rows < 1) band_rows = 1;
    const unsigned n_bands = (n_rows + band_rows - 1) / band_rows;
    unsigned *keys = work, *keys_out = work + n_tiles, *vals = work + 2 * (size_t)n_tiles, *seq = work + 3 * (size_t)n_tiles;
    unsigned long long* band_work = reinterpret_cast<unsigned long long*>(work + ((3 * (size_t)n_tiles + n_rows + 1) & ~size_t(1)));
    if (by_work) band_work_kernel<<<n_bands, 256, 0, stream>>>(cost, n_tiles, band_rows * tiles_x, band_work);
    band_rank_kernel<<<1, 256, 0, stream>>>(band_work, n_bands, band_rows, n_rows, by_work, seq);
    row_key_kernel<<<(n_tiles + 255) / 256, 256, 0, stream>>>(cost, sum, n_tiles, tiles_x, n_rows, lead, seq, keys, vals);
    int bits = 1;
    while ((1u << bits) < n_rows) bits++;
    cudaError_t e = cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys, keys_out, vals, order, (int)n_tiles, 0, bits, stream);
    if (e != cudaSuccess) return (int)e;
    return (int)cudaGetLastError();
}

// scratch: RTC_ORDER_BUCKETS * ceil(n_tiles / 1024) + RTC_ORDER_BUCKETS unsigned (the last ones receive the number of tiles
// that cost >= the set's thresholds: 8x, 4x, 2x the mean / 12x, 8x, 7x, 6x, 4x, 2x); sum: the render kernel's sum of all
// tile costs (device memory); last_bucket: buckets beyond it are merged into it (set 0: <= 3, set 1: <= 6)
int launch_tile_order(const unsigned* cost, const unsigned long long* sum, unsigned* scratch, unsigned* order, unsigned n_tiles,
                      int last_bucket, int set, cudaStream_t stream) {
    static_assert(NB == RTC_ORDER_BUCKETS, "render_params.h RTC_ORDER_BUCKETS out of date");
    if (!n_tiles) return 0;
    const unsigned blocks = (n_tiles + T - 1) / T;
    tile_count_kernel<<<blocks, T, 0, stream>>>(cost, sum, scratch, n_tiles, last_bucket, set);
    tile_scatter_kernel<<<blocks, T, 0, stream>>>(cost, sum, scratch, order, scratch + NB * blocks, n_tiles, last_bucket, set);
    return (int)cudaGetLastError();
}

}  // namespace rtc
