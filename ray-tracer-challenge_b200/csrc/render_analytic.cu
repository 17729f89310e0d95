// The wavefront kernel for scenes WITHOUT triangle meshes and the resolve kernel (render_wavefront.cuh).
#include "render_wavefront.cuh"

namespace rtc {

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
