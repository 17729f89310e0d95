// Instantiations of the render kernels for scenes WITH triangle meshes (see render_kernel.cuh): the general one.
#include "render_launch.cuh"
#include "render_wavefront.cuh"

namespace rtc {

int launch_render_mesh(const DScene& S, const RenderParams& P, int lpp, int grid, size_t smem, cudaStream_t stream) {
    return launch_render_impl<true, false>(S, P, lpp, grid, smem, stream);
}
int occupancy_render_mesh(int* blocks_per_sm, size_t smem) { return occupancy_render_impl<true, false>(blocks_per_sm, smem); }

int launch_wave_mesh(const DScene& S, const WaveParams& W, int grid, size_t smem, cudaStream_t stream) {
    wave_kernel<true><<<grid, RTC_BLOCK, smem, stream>>>(S, W);
    return (int)cudaGetLastError();
}

int occupancy_wave_mesh(int* blocks_per_sm, size_t smem) {
    int dev = 0, optin = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(wave_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin);
    if (e != cudaSuccess) return (int)e;
    return (int)cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, wave_kernel<true>, RTC_BLOCK, smem);
}

}  // namespace rtc
