// The wavefront kernel for scenes WITH triangle meshes (render_wavefront.cuh) and the dispatch over the
// render_kernel<MESH, LPP, FEAT> instantiations (render_inst_*.cu).
#include "render_wavefront.cuh"

namespace rtc {

int launch_render(bool mesh, int feat, const DScene& S, const RenderParams& P, int lpp, int grid, size_t smem, cudaStream_t stream) {
    switch ((mesh ? 4 : 0) | (feat & 3)) {
        case 0: return launch_render_0_0(S, P, lpp, grid, smem, stream);
        case 2: return launch_render_0_2(S, P, lpp, grid, smem, stream);
        case 3: return launch_render_0_3(S, P, lpp, grid, smem, stream);
        case 4: return launch_render_1_0(S, P, lpp, grid, smem, stream);
        case 6: return launch_render_1_2(S, P, lpp, grid, smem, stream);
        default: return launch_render_1_3(S, P, lpp, grid, smem, stream);
    }
}

int occupancy_render(bool mesh, int feat, int* blocks_per_sm, size_t smem) {
    switch ((mesh ? 4 : 0) | (feat & 3)) {
        case 0: return occupancy_render_0_0(blocks_per_sm, smem);
        case 2: return occupancy_render_0_2(blocks_per_sm, smem);
        case 3: return occupancy_render_0_3(blocks_per_sm, smem);
        case 4: return occupancy_render_1_0(blocks_per_sm, smem);
        case 6: return occupancy_render_1_2(blocks_per_sm, smem);
        default: return occupancy_render_1_3(blocks_per_sm, smem);
    }
}

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
