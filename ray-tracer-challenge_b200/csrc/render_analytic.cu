// Instantiations of the render kernels for scenes WITHOUT triangle meshes (see render_kernel.cuh): the general one.
#include "render_launch.cuh"
#include "render_wavefront.cuh"

namespace rtc {

int launch_render_analytic(const DScene& S, const RenderParams& P, int lpp, int grid, size_t smem, cudaStream_t stream) {
    return launch_render_impl<false, false>(S, P, lpp, grid, smem, stream);
}
int occupancy_render_analytic(int* blocks_per_sm, size_t smem) { return occupancy_render_impl<false, false>(blocks_per_sm, smem); }

int launch_wave_analytic(const DScene& S, const WaveParams& W, int grid, size_t smem, cudaStream_t stream) {
    wave_kernel<false><<<grid, RTC_BLOCK, smem, stream>>>(S, W);
    return (int)cudaGetLastError();
}

int occupancy_wave_analytic(int* blocks_per_sm, size_t smem) {
    int dev = 0, optin = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(wave_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, optin);
    if (e != cudaSuccess) return (int)e;
    return (int)cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, wave_kernel<false>, RTC_BLOCK, smem);
}

int launch_resolve(const ResolveParams& R, cudaStream_t stream) {
    unsigned long long blocks = (R.n_pixels + 255) / 256;
    if (blocks) resolve_kernel<<<(unsigned)blocks, 256, 0, stream>>>(R);
    return (int)cudaGetLastError();
}

}  // namespace rtc
