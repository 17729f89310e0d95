// Launch parameters of render_kernel, shared by render.cu (device) and rtc_api.cu (host).
#pragma once
#include <cuda_runtime.h>

#include "device_scene.h"

namespace rtc {

#ifndef RTC_BLOCK_THREADS
#define RTC_BLOCK_THREADS 128  // threads per CTA of the render kernels
#endif

#define RTC_PARK_SLOTS 22  // doubles of per-lane state parked in shared memory (render_kernel.cuh PK_*)

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
    unsigned long long n_items;    // tiles_x * tiles_y * 32 work items (some fall outside the frame)
    unsigned int tiles_x;
    int full_frame;                // 0: out holds the packed rows of the partition; 1: the whole frame
    double* out;                   // 3 doubles per pixel (may be peer-mapped memory of another GPU)
    unsigned char* out8;           // or RGBA8 (NULL when `out` is used)
    unsigned long long* counters;  // [0] next item, [1] primary, [2] shadow, [3] secondary, [4] error bits, [5] culled shadow rays
};

// render_mesh.cu / render_analytic.cu: the two instantiations of render_kernel<MESH>.
// smem = dynamic shared memory per CTA (the BVH traversal stacks; 0 for the analytic kernel).
int launch_render_mesh(const DScene& S, const RenderParams& P, int grid, size_t smem, cudaStream_t stream);
int launch_render_analytic(const DScene& S, const RenderParams& P, int grid, size_t smem, cudaStream_t stream);
int occupancy_render_mesh(int* blocks_per_sm, size_t smem);
int occupancy_render_analytic(int* blocks_per_sm, size_t smem);

}  // namespace rtc
