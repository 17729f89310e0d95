// Launch / occupancy plumbing shared by the per-instantiation translation units (render_mesh.cu, render_mesh_lean.cu,
// render_analytic.cu, render_analytic_lean.cu: one TU per <MESH, FEAT> pair so that `make -j` compiles them side by side).
#pragma once
#include "render_kernel.cuh"

namespace rtc {

template <bool MESH, int FEAT>
int launch_render_impl(const DScene& S, const RenderParams& P, int lpp, int grid, size_t smem, cudaStream_t stream) {
    if (lpp == 0) render_kernel<MESH, 0, FEAT><<<grid, RTC_BLOCK, smem, stream>>>(S, P);
    else if (lpp == 4) render_kernel<MESH, 4, FEAT><<<grid, RTC_BLOCK, smem, stream>>>(S, P);
    else if (lpp == 2) render_kernel<MESH, 2, FEAT><<<grid, RTC_BLOCK, smem, stream>>>(S, P);
    else render_kernel<MESH, 1, FEAT><<<grid, RTC_BLOCK, smem, stream>>>(S, P);
    return (int)cudaGetLastError();
}

// The dynamic shared memory limit is a property of the KERNEL on the current device, not of a scene: it is raised to
// the device's opt-in maximum (a scene with a deep BVH needs more than the 48 KB default for its traversal stacks) and
// never lowered, so scenes created later cannot take it away from scenes created earlier.
template <bool MESH, int FEAT>
int occupancy_render_impl(int* blocks_per_sm, size_t smem) {
    int dev = 0, optin = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    if (e == cudaSuccess && smem > (size_t)optin) return (int)cudaErrorInvalidValue;
    if (e == cudaSuccess) e = cudaFuncSetAttribute(render_kernel<MESH, 1, FEAT>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(render_kernel<MESH, 2, FEAT>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(render_kernel<MESH, 4, FEAT>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(render_kernel<MESH, 0, FEAT>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin);
    if (e != cudaSuccess) return (int)e;
    // the instantiations share their launch bounds; the grid is sized by the one-lane-per-pixel kernel
    return (int)cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, render_kernel<MESH, 1, FEAT>, RTC_BLOCK, smem);
}

#define RTC_DEFINE_INST(M, F)                                                                                            \
    int launch_render_##M##_##F(const DScene& S, const RenderParams& P, int lpp, int grid, size_t smem, cudaStream_t stream) { \
        return launch_render_impl<(M) != 0, F>(S, P, lpp, grid, smem, stream);                                              \
    }                                                                                                                      \
    int occupancy_render_##M##_##F(int* blocks_per_sm, size_t smem) { return occupancy_render_impl<(M) != 0, F>(blocks_per_sm, smem); }

}  // namespace rtc
