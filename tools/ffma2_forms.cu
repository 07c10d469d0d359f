// ffma2_forms.cu — developer microbenchmark: FFMA2 throughput by operand pattern (sm_100a).
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>

// MODE 0: acc = ffma2(acc, c1, c2)                       (two loop-invariant pair operands)
// MODE 1: acc[i] = ffma2(bcast(s[j]), u[i], acc[i])      (scalar-broadcast form, product kernel's shape)
// MODE 2: acc[i] = ffma2(u[i], v[i], acc[i])             (three distinct pair operands)
// MODE 3: like 1 but chained depth 3 per (i, j) with fresh accumulators (h = s.x*ux + (s.y*uy + (s.z*uz + s.w)))
template <int MODE, int N>
__global__ void __launch_bounds__(256) k(float* out, const float4* __restrict__ sp, int iters) {
    float2 acc[N], u[N], v[N], w[N];
#pragma unroll
    for (int i = 0; i < N; ++i) {
        acc[i] = make_float2(threadIdx.x * 1e-3f + i, threadIdx.x * 2e-3f - i);
        u[i] = make_float2(1.0f + i * 1e-3f + threadIdx.x * 1e-6f, 1.0f - i * 1e-3f);
        v[i] = make_float2(0.5f + i * 1e-3f, 0.25f - i * 1e-3f + threadIdx.x * 1e-6f);
        w[i] = make_float2(0.75f + i * 1e-3f, 0.125f - i * 1e-3f + threadIdx.x * 1e-6f);
    }
    __shared__ float4 s_sp[64];
    if (threadIdx.x < 64) s_sp[threadIdx.x] = sp[threadIdx.x];
    __syncthreads();
    int sink = -1;
    if (MODE < 4) for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float4 s = s_sp[(it * 8 + j) & 63];
#pragma unroll
            for (int i = 0; i < N; ++i) {
                if (MODE == 0) acc[i] = __ffma2_rn(acc[i], u[0], v[0]);
                if (MODE == 1) {
                    acc[i] = __ffma2_rn(make_float2(s.x, s.x), u[i], acc[i]);
                    acc[i] = __ffma2_rn(make_float2(s.y, s.y), v[i], acc[i]);
                    acc[i] = __ffma2_rn(make_float2(s.z, s.z), w[i], acc[i]);
                }
                if (MODE == 2) {
                    acc[i] = __ffma2_rn(u[i], v[i], acc[i]);
                    acc[i] = __ffma2_rn(v[i], w[i], acc[i]);
                    acc[i] = __ffma2_rn(w[i], u[i], acc[i]);
                }
                if (MODE == 3) {
                    float2 h = __ffma2_rn(make_float2(s.x, s.x), u[i], __ffma2_rn(make_float2(s.y, s.y), v[i],
                               __ffma2_rn(make_float2(s.z, s.z), w[i], make_float2(s.w, s.w))));
                    sink &= __float_as_int(h.x) & __float_as_int(h.y);
                }
            }
        }
    }
    if (MODE == 4 || MODE == 5) {
        // MODE 4: ray pair fixed (64-bit operand reusable), spheres vary: h[i][j] chains, 4 spheres in registers
        // MODE 5: same arithmetic, sphere fixed, ray pairs vary (scalar operand reusable)
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int g = 0; g < 2; ++g) {
                float4 s[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) s[j] = s_sp[(it * 8 + g * 4 + j) & 63];
                float2 h[N][4];
                if (MODE == 4) {
#pragma unroll
                    for (int i = 0; i < N; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) h[i][j] = __ffma2_rn(make_float2(s[j].z, s[j].z), w[i], make_float2(s[j].w, s[j].w));
#pragma unroll
                    for (int i = 0; i < N; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) h[i][j] = __ffma2_rn(make_float2(s[j].y, s[j].y), v[i], h[i][j]);
#pragma unroll
                    for (int i = 0; i < N; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) h[i][j] = __ffma2_rn(make_float2(s[j].x, s[j].x), u[i], h[i][j]);
                } else {
#pragma unroll
                    for (int j = 0; j < 4; ++j)
#pragma unroll
                        for (int i = 0; i < N; ++i) h[i][j] = __ffma2_rn(make_float2(s[j].z, s[j].z), w[i], make_float2(s[j].w, s[j].w));
#pragma unroll
                    for (int j = 0; j < 4; ++j)
#pragma unroll
                        for (int i = 0; i < N; ++i) h[i][j] = __ffma2_rn(make_float2(s[j].y, s[j].y), v[i], h[i][j]);
#pragma unroll
                    for (int j = 0; j < 4; ++j)
#pragma unroll
                        for (int i = 0; i < N; ++i) h[i][j] = __ffma2_rn(make_float2(s[j].x, s[j].x), u[i], h[i][j]);
                }
#pragma unroll
                for (int i = 0; i < N; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) sink &= __float_as_int(h[i][j].x) & __float_as_int(h[i][j].y);
            }
        }
    }
    float r = (float)sink;
#pragma unroll
    for (int i = 0; i < N; ++i) r += acc[i].x + acc[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <int MODE, int N>
static void run(const char* name, float* out, const float4* sp, int sms, int bps) {
    const int iters = 2048, blocks = sms * bps;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    double best = 1e30;
    for (int r = 0; r < 5; ++r) {
        cudaEventRecord(e0);
        k<MODE, N><<<blocks, 256>>>(out, sp, iters);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (r >= 1 && ms < best) best = ms;
    }
    const double per = (MODE == 0) ? 1 : 3;
    const double flops = (double)blocks * 256 * iters * 8 * N * per * 4;
    cudaFuncAttributes fa; cudaFuncGetAttributes(&fa, k<MODE, N>);
    printf("%-40s N=%2d bps=%d regs=%3d : %6.2f TFLOP/s (%s)\n", name, N, bps, fa.numRegs, flops / (best * 1e-3) / 1e12,
           cudaGetErrorString(cudaGetLastError()));
}

int main() {
    cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
    float* out; float4* sp;
    cudaMalloc(&out, (size_t)prop.multiProcessorCount * 8 * 256 * 4); cudaMalloc(&sp, 64 * 16);
    float4 h[64]; for (int i = 0; i < 64; ++i) h[i] = make_float4(1e-3f * i, -1e-3f * i, 2e-3f, -1.0f);
    cudaMemcpy(sp, h, sizeof(h), cudaMemcpyHostToDevice);
    const int sms = prop.multiProcessorCount;
    for (int bps = 2; bps <= 4; bps += 2) {
        run<0, 8>("invariant pair operands", out, sp, sms, bps);
        run<0, 16>("invariant pair operands", out, sp, sms, bps);
        run<1, 4>("scalar-broadcast, accumulate", out, sp, sms, bps);
        run<1, 8>("scalar-broadcast, accumulate", out, sp, sms, bps);
        run<2, 4>("three distinct pairs", out, sp, sms, bps);
        run<2, 8>("three distinct pairs", out, sp, sms, bps);
        run<3, 2>("broadcast chain depth 3 + LOP3", out, sp, sms, bps);
        run<3, 4>("broadcast chain depth 3 + LOP3", out, sp, sms, bps);
        run<3, 8>("broadcast chain depth 3 + LOP3", out, sp, sms, bps);
        run<4, 2>("chain, ray-pair operand reusable", out, sp, sms, bps);
        run<4, 4>("chain, ray-pair operand reusable", out, sp, sms, bps);
        run<5, 2>("chain, sphere scalar reusable", out, sp, sms, bps);
        run<5, 4>("chain, sphere scalar reusable", out, sp, sms, bps);
    }
    return 0;
}
