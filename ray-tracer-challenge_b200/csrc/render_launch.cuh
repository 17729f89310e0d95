// Launch / occupancy plumbing shared by the per-instantiation translation units (render_mesh.cu, render_mesh_lean.cu,
// render_analytic.cu, render_analytic_lean.cu: one TU per <MESH, LEAN> pair so that `make -j` compiles them side by side).
#pragma once
#include "render_kernel.cuh"

namespace rtc {

template <bool MESH, bool LEAN>
int launch_render_impl(const DScene& S, const RenderParams& P, int lpp, int grid, size_t smem, cudaStream_t stream) {
    if (lpp == 0) render_kernel<MESH, 0, LEAN><<<grid, RTC_BLOCK, smem, stream>>>(S, P);
    else if (lpp == 4) render_kernel<MESH, 4, LEAN><<<grid, RTC_BLOCK, smem, stream>>>(S, P);
    else if (lpp == 2) render_kernel<MESH, 2, LEAN><<<grid, RTC_BLOCK, smem, stream>>>(S, P);
    else render_kernel<MESH, 1, LEAN><<<grid, RTC_BLOCK, smem, stream>>>(S, P);
    return (int)cudaGetLastError();
}

// The dynamic shared memory limit is a property of the KERNEL on the current device, not of a scene: it is raised to
// the device's opt-in maximum (a scene with a deep BVH needs more than the 48 KB default for its traversal stacks) and
// never lowered, so scenes created later cannot take it away from scenes created earlier.
template <bool MESH, bool LEAN>
int occupancy_render_impl(int* blocks_per_sm, size_t smem) {
    int dev = 0, optin = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    if (e == cudaSuccess && smem > (size_t)optin) return (int)cudaErrorInvalidValue;
    if (e == cudaSuccess) e = cudaFuncSetAttribute(render_kernel<MESH, 1, LEAN>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(render_kernel<MESH, 2, LEAN>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(render_kernel<MESH, 4, LEAN>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(render_kernel<MESH, 0, LEAN>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin);
    if (e != cudaSuccess) return (int)e;
    // the instantiations share their launch bounds; the grid is sized by the one-lane-per-pixel kernel
    return (int)cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, render_kernel<MESH, 1, LEAN>, RTC_BLOCK, smem);
}

}  // namespace rtc
