// Launch parameters of render_kernel, shared by render.cu (device) and rtc_api.cu (host).
#pragma once
#include <cuda_runtime.h>

#include "device_scene.h"

namespace rtc {

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
    double* out;                   // packed rows of the partition, 3 doubles per pixel
    unsigned char* out8;           // or RGBA8 (NULL when `out` is used)
    unsigned long long* counters;  // [0] next item, [1] primary, [2] shadow, [3] secondary, [4] error bits
};

int launch_render(const DScene& S, const RenderParams& P, int grid, cudaStream_t stream);
int render_kernel_occupancy(int* blocks_per_sm);

}  // namespace rtc
