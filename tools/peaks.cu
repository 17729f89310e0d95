// Measured denominators for the ALU / L2 rooflines of SURVEY.md 8(d) that MEASURED_PEAKS.json does not hold:
// FP64 and FP32 FMA issue peak, L2-resident copy bandwidth, and an HBM copy for cross-checking.
// build: make tools/peaks      run (on the B200): tools/peaks > profiles/r1_alu_l2_peaks.json
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>

#define CHECK(x)                                                                         \
    do {                                                                                 \
        cudaError_t e_ = (x);                                                            \
        if (e_ != cudaSuccess) {                                                         \
            std::fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_));                \
            std::exit(1);                                                                \
        }                                                                                \
    } while (0)

// 8 independent FMA chains per thread: enough ILP to fill the pipe at full occupancy
template <typename T>
__global__ void fma_kernel(T* out, int iters, T a, T b) {
    T x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; i++) {
        x0 = x0 * a + b;
        x1 = x1 * a + b;
        x2 = x2 * a + b;
        x3 = x3 * a + b;
        x4 = x4 * a + b;
        x5 = x5 * a + b;
        x6 = x6 * a + b;
        x7 = x7 * a + b;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}

__global__ void copy_kernel(const float4* __restrict__ src, float4* __restrict__ dst, size_t n) {
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) dst[i] = src[i];
}

template <typename F>
static float best_ms(F launch, int reps) {
    cudaEvent_t a, b;
    CHECK(cudaEventCreate(&a));
    CHECK(cudaEventCreate(&b));
    float best = 1e30f;
    for (int r = 0; r < reps; r++) {
        CHECK(cudaEventRecord(a));
        launch();
        CHECK(cudaEventRecord(b));
        CHECK(cudaEventSynchronize(b));
        float ms;
        CHECK(cudaEventElapsedTime(&ms, a, b));
        if (ms < best) best = ms;
    }
    return best;
}

int main() {
    cudaDeviceProp p;
    CHECK(cudaGetDeviceProperties(&p, 0));
    const int sms = p.multiProcessorCount, threads = 256, blocks = sms * 8, iters = 1 << 14;
    void* out;
    CHECK(cudaMalloc(&out, (size_t)blocks * threads * sizeof(double)));
    fma_kernel<double><<<blocks, threads>>>((double*)out, 16, 1.0000001, 1e-9);  // warm-up
    fma_kernel<float><<<blocks, threads>>>((float*)out, 16, 1.0000001f, 1e-9f);
    CHECK(cudaDeviceSynchronize());
    float ms64 = best_ms([&] { fma_kernel<double><<<blocks, threads>>>((double*)out, iters, 1.0000001, 1e-9); }, 5);
    float ms32 = best_ms([&] { fma_kernel<float><<<blocks, threads>>>((float*)out, iters, 1.0000001f, 1e-9f); }, 5);
    const double fmas = (double)blocks * threads * iters * 8.0;
    const double tf64 = 2.0 * fmas / (ms64 * 1e-3) / 1e12, tf32 = 2.0 * fmas / (ms32 * 1e-3) / 1e12;

    // L2-resident copy: 2 x 24 MB (B200 L2: 126 MB); HBM copy: 2 x 2 GB
    auto copy_gbs = [&](size_t bytes, int inner) {
        float4 *s, *d;
        CHECK(cudaMalloc(&s, bytes));
        CHECK(cudaMalloc(&d, bytes));
        CHECK(cudaMemset(s, 1, bytes));
        const size_t n = bytes / sizeof(float4);
        for (int i = 0; i < 3; i++) copy_kernel<<<sms * 16, 256>>>(s, d, n);
        CHECK(cudaDeviceSynchronize());
        float ms = best_ms(
            [&] {
                for (int i = 0; i < inner; i++) copy_kernel<<<sms * 16, 256>>>(s, d, n);
            },
            5);
        CHECK(cudaFree(s));
        CHECK(cudaFree(d));
        return 2.0 * (double)bytes * inner / (ms * 1e-3) / 1e9;  // read + write
    };
    const double l2 = copy_gbs((size_t)24 << 20, 50), hbm = copy_gbs((size_t)2 << 30, 2);
    std::printf(
        "{\"gpu\": \"%s\", \"sms\": %d, \"sm_clock_mhz\": %d, \"fp64_fma_tflops\": %.2f, \"fp32_fma_tflops\": %.2f, "
        "\"l2_copy_gbs\": %.0f, \"hbm_copy_gbs\": %.0f, "
        "\"how\": \"8 independent FMA chains per thread, %d CTAs x %d threads x %d iterations, best of 5 (2 flops per FMA); "
        "copy kernel with float4 loads/stores, read + write bytes, 2 x 24 MB resident in L2 (50 back-to-back passes) and 2 x 2 GB through HBM\"}\n",
        p.name, sms, p.clockRate / 1000, tf64, tf32, l2, hbm, blocks, threads, iters);
    return 0;
}
