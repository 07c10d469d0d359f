// ffma_peak.cu — measures the FP32 FMA roofline denominator on the box (MEASURED_PEAKS.json has
// no FP32 figure).  Every thread runs 16 independent dependent-FFMA chains with register operands;
// the result is stored so nothing is eliminated.  Timed with CUDA events, best of `reps`.
#include <cuda_runtime.h>
#include <stdint.h>

__global__ void __launch_bounds__(256) ffma_chain_kernel(float* out, float a, float b, int iters) {
    float x[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = (float)(threadIdx.x + i) * 1e-3f;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int i = 0; i < 16; ++i) x[i] = fmaf(x[i], a, b);
        }
    }
    float s = 0.0f;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

extern "C" __attribute__((visibility("default")))
int ffma_peak(int device, int reps, double* tflops_best, double* tflops_median_unused, int* sm_count_out) {
    (void)tflops_median_unused;
    if (cudaSetDevice(device) != cudaSuccess) return -1;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return -1;
    const int sms = prop.multiProcessorCount;
    if (sm_count_out) *sm_count_out = sms;
    const int blocks = sms * 8, threads = 256, iters = 4096;
    float* out = nullptr;
    if (cudaMalloc(&out, (size_t)blocks * threads * sizeof(float)) != cudaSuccess) return -2;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    double best = 0.0;
    for (int r = 0; r < reps + 2; ++r) {
        cudaEventRecord(e0);
        ffma_chain_kernel<<<blocks, threads>>>(out, 0.999f, 1e-3f, iters);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(out); return -3; }
        float ms = 0.0f;
        cudaEventElapsedTime(&ms, e0, e1);
        const double flops = 2.0 * 16 * 8 * (double)iters * blocks * threads;
        const double tf = flops / (ms * 1e-3) / 1e12;
        if (r >= 2 && tf > best) best = tf;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(out);
    if (tflops_best) *tflops_best = best;
    return 0;
}
