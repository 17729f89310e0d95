// Instantiation of the render kernel for scenes WITHOUT triangle meshes (see render_kernel.cuh).
#include "render_kernel.cuh"

namespace rtc {

int launch_render_analytic(const DScene& S, const RenderParams& P, int grid, size_t smem, cudaStream_t stream) {
    render_kernel<false><<<grid, RTC_BLOCK, smem, stream>>>(S, P);
    return (int)cudaGetLastError();
}

int occupancy_render_analytic(int* blocks_per_sm, size_t smem) {
    cudaError_t e = cudaFuncSetAttribute(render_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    return (int)cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, render_kernel<false>, RTC_BLOCK, smem);
}

}  // namespace rtc
