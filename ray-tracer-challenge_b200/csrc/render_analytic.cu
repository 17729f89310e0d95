// Instantiation of the render kernel for scenes WITHOUT triangle meshes (see render_kernel.cuh).
#include "render_kernel.cuh"
#include "render_wavefront.cuh"

namespace rtc {

int launch_render_analytic(const DScene& S, const RenderParams& P, int lpp, int grid, size_t smem, cudaStream_t stream) {
    if (lpp == 0) render_kernel<false, 0><<<grid, RTC_BLOCK, smem, stream>>>(S, P);
    else if (lpp == 4) render_kernel<false, 4><<<grid, RTC_BLOCK, smem, stream>>>(S, P);
    else if (lpp == 2) render_kernel<false, 2><<<grid, RTC_BLOCK, smem, stream>>>(S, P);
    else render_kernel<false, 1><<<grid, RTC_BLOCK, smem, stream>>>(S, P);
    return (int)cudaGetLastError();
}

int occupancy_render_analytic(int* blocks_per_sm, size_t smem) {
    cudaError_t e = cudaFuncSetAttribute(render_kernel<false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(render_kernel<false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(render_kernel<false, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(render_kernel<false, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    // the three instantiations share their launch bounds; the grid is sized by the one-lane-per-pixel kernel
    return (int)cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, render_kernel<false, 1>, RTC_BLOCK, smem);
}

int launch_wave_analytic(const DScene& S, const WaveParams& W, int grid, size_t smem, cudaStream_t stream) {
    wave_kernel<false><<<grid, RTC_BLOCK, smem, stream>>>(S, W);
    return (int)cudaGetLastError();
}

int occupancy_wave_analytic(int* blocks_per_sm, size_t smem) {
    cudaError_t e = cudaFuncSetAttribute(wave_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    return (int)cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, wave_kernel<false>, RTC_BLOCK, smem);
}

int launch_resolve(const ResolveParams& R, cudaStream_t stream) {
    unsigned long long blocks = (R.n_pixels + 255) / 256;
    if (blocks) resolve_kernel<<<(unsigned)blocks, 256, 0, stream>>>(R);
    return (int)cudaGetLastError();
}

}  // namespace rtc
