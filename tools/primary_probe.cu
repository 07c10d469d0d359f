// primary_probe.cu — developer microbenchmark for the common-origin broad phase.
// Times only the inner scan: R rays per thread against 1024 shared-memory sphere records, in the
// candidate-test formulations considered for the product kernel.  Not part of the product.
//
//   form A: g = fma(h, h, -c), candidate <=> sign(g) clear; sign bits AND-ed with LOP3      (4 FMA / test)
//   form B: candidate <=> !(h < q); one FSETP per test                                        (3 FMA / test)
//   form C: candidate <=> !(max_r h_r < q); FMNMX3 tree over the thread's rays, FSETP/sphere  (3 FMA / test)
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>

__device__ __forceinline__ float max3(float a, float b, float c) {
    float d;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}
__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}

template <int PAIRS>
__device__ __forceinline__ void dot_pairs(const float4 sp, const float2 (&ux)[PAIRS], const float2 (&uy)[PAIRS],
                                          const float2 (&uz)[PAIRS], float2 (&h)[PAIRS]) {
    const float2 sx = make_float2(sp.x, sp.x), sy = make_float2(sp.y, sp.y), sz = make_float2(sp.z, sp.z);
#pragma unroll
    for (int q = 0; q < PAIRS; ++q) h[q] = __ffma2_rn(sx, ux[q], __ffma2_rn(sy, uy[q], __fmul2_rn(sz, uz[q])));
}

__constant__ float4 c_geom[1024 + 16];

// form 7: records read from the constant bank (uniform datapath), FFMA2 + LOP3
template <int PAIRS, int UNROLL>
__global__ void __launch_bounds__(256) scan_const_kernel(const float* __restrict__ dirs, unsigned* __restrict__ out, int n, int reps) {
    unsigned hits = 0;
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    for (int rep = 0; rep < reps; ++rep) {
        float2 ux[PAIRS], uy[PAIRS], uz[PAIRS];
#pragma unroll
        for (int q = 0; q < PAIRS; ++q) {
            const float* d = dirs + ((size_t)(tid + rep * 7) * PAIRS + q) % 4096 * 6;
            ux[q] = make_float2(d[0], d[3]); uy[q] = make_float2(d[1], d[4]); uz[q] = make_float2(d[2], d[5]);
        }
#pragma unroll 1
        for (int s = 0; s < n; s += UNROLL) {
            int acc = -1;
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                const float4 sp = c_geom[s + u];
                const float2 sx = make_float2(sp.x, sp.x), sy = make_float2(sp.y, sp.y), sz = make_float2(sp.z, sp.z),
                             sw = make_float2(sp.w, sp.w);
#pragma unroll
                for (int q = 0; q < PAIRS; ++q) {
                    const float2 h = __ffma2_rn(sx, ux[q], __ffma2_rn(sy, uy[q], __ffma2_rn(sz, uz[q], sw)));
                    acc &= __float_as_int(h.x) & __float_as_int(h.y);
                }
            }
            if (acc >= 0) hits += 1;
        }
    }
    out[tid] = hits;
}

template <int FORM, int PAIRS, int UNROLL>
__global__ void __launch_bounds__(256) scan_kernel(const float4* __restrict__ geom, const float* __restrict__ dirs,
                                                   unsigned* __restrict__ out, int n, int reps) {
    extern __shared__ float4 s_geom[];
    for (int i = threadIdx.x; i < n + UNROLL; i += blockDim.x) s_geom[i] = geom[i];
    __syncthreads();
    unsigned hits = 0;
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    for (int rep = 0; rep < reps; ++rep) {
        float2 ux[PAIRS], uy[PAIRS], uz[PAIRS];
#pragma unroll
        for (int q = 0; q < PAIRS; ++q) {
            const float* d = dirs + ((size_t)(tid + rep * 7) * PAIRS + q) % 4096 * 6;
            ux[q] = make_float2(d[0], d[3]); uy[q] = make_float2(d[1], d[4]); uz[q] = make_float2(d[2], d[5]);
        }
        uint32_t addr = (uint32_t)__cvta_generic_to_shared(s_geom);
        float4 cur[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) cur[u] = lds128(addr + 16u * u);
#pragma unroll 1
        for (int s = 0; s < n; s += UNROLL) {
            addr += 16u * UNROLL;
            float4 nxt[UNROLL];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) nxt[u] = lds128(addr + 16u * u);
            bool any = false;
            if (FORM == 0) {
                int acc = -1;
#pragma unroll
                for (int u = 0; u < UNROLL; ++u) {
                    float2 h[PAIRS];
                    dot_pairs<PAIRS>(cur[u], ux, uy, uz, h);
                    const float2 mc = make_float2(-cur[u].w, -cur[u].w);
#pragma unroll
                    for (int q = 0; q < PAIRS; ++q) {
                        const float2 g = __ffma2_rn(h[q], h[q], mc);
                        acc &= __float_as_int(g.x) & __float_as_int(g.y);
                    }
                }
                any = acc >= 0;
            } else if (FORM == 1) {
#pragma unroll
                for (int u = 0; u < UNROLL; ++u) {
                    float2 h[PAIRS];
                    dot_pairs<PAIRS>(cur[u], ux, uy, uz, h);
#pragma unroll
                    for (int q = 0; q < PAIRS; ++q) {
                        any |= !(h[q].x < cur[u].w);
                        any |= !(h[q].y < cur[u].w);
                    }
                }
            } else if (FORM == 3 || FORM == 4) {
                // FFMA2 only: h' = C.u - q, candidate <=> h' >= 0.  3: sign bits AND-ed (LOP3); 4: running FMNMX3
                int acc = -1;
                float m = -1.0f;
#pragma unroll
                for (int u = 0; u < UNROLL; ++u) {
                    const float4 sp = cur[u];
                    const float2 sx = make_float2(sp.x, sp.x), sy = make_float2(sp.y, sp.y), sz = make_float2(sp.z, sp.z),
                                 sw = make_float2(sp.w, sp.w);
#pragma unroll
                    for (int q = 0; q < PAIRS; ++q) {
                        const float2 h = __ffma2_rn(sx, ux[q], __ffma2_rn(sy, uy[q], __ffma2_rn(sz, uz[q], sw)));
                        if (FORM == 3) acc &= __float_as_int(h.x) & __float_as_int(h.y);
                        else m = max3(m, h.x, h.y);
                    }
                }
                any = (FORM == 3) ? (acc >= 0) : !(m < 0.0f);
            } else if (FORM == 9) {
                // sphere-pair layout: cur[2k] = (x0, x1, y0, y1), cur[2k+1] = (z0, z1, w0, w1) of spheres (2k, 2k+1);
                // the ray components are the broadcast scalars, the 64-bit sphere operands repeat across rays
                int acc = -1;
#pragma unroll
                for (int k = 0; k < UNROLL / 2; ++k) {
                    const float2 sx = make_float2(cur[2 * k].x, cur[2 * k].y), sy = make_float2(cur[2 * k].z, cur[2 * k].w);
                    const float2 sz = make_float2(cur[2 * k + 1].x, cur[2 * k + 1].y), sw = make_float2(cur[2 * k + 1].z, cur[2 * k + 1].w);
#pragma unroll
                    for (int q = 0; q < PAIRS; ++q) {
                        const float2 h0 = __ffma2_rn(sx, make_float2(ux[q].x, ux[q].x),
                                          __ffma2_rn(sy, make_float2(uy[q].x, uy[q].x), __ffma2_rn(sz, make_float2(uz[q].x, uz[q].x), sw)));
                        const float2 h1 = __ffma2_rn(sx, make_float2(ux[q].y, ux[q].y),
                                          __ffma2_rn(sy, make_float2(uy[q].y, uy[q].y), __ffma2_rn(sz, make_float2(uz[q].y, uz[q].y), sw)));
                        acc &= __float_as_int(h0.x) & __float_as_int(h0.y);
                        acc &= __float_as_int(h1.x) & __float_as_int(h1.y);
                    }
                }
                any = acc >= 0;
            } else if (FORM == 8) {
                // like 3, but ray-pair major: the 64-bit ray operand repeats across consecutive FFMA2s
                int acc = -1;
#pragma unroll
                for (int q = 0; q < PAIRS; ++q) {
                    float2 h[UNROLL];
#pragma unroll
                    for (int u = 0; u < UNROLL; ++u) h[u] = __ffma2_rn(make_float2(cur[u].z, cur[u].z), uz[q], make_float2(cur[u].w, cur[u].w));
#pragma unroll
                    for (int u = 0; u < UNROLL; ++u) h[u] = __ffma2_rn(make_float2(cur[u].y, cur[u].y), uy[q], h[u]);
#pragma unroll
                    for (int u = 0; u < UNROLL; ++u) h[u] = __ffma2_rn(make_float2(cur[u].x, cur[u].x), ux[q], h[u]);
#pragma unroll
                    for (int u = 0; u < UNROLL; ++u) acc &= __float_as_int(h[u].x) & __float_as_int(h[u].y);
                }
                any = acc >= 0;
            } else if (FORM == 5 || FORM == 6) {
                // 5: scalar FFMA for every ray; 6: pairs [0, PAIRS/2) packed, the rest scalar.  Sign bits AND-ed.
                int acc = -1;
                constexpr int kPacked = (FORM == 6) ? PAIRS / 2 : 0;
#pragma unroll
                for (int u = 0; u < UNROLL; ++u) {
                    const float4 sp = cur[u];
                    const float2 sx = make_float2(sp.x, sp.x), sy = make_float2(sp.y, sp.y), sz = make_float2(sp.z, sp.z),
                                 sw = make_float2(sp.w, sp.w);
#pragma unroll
                    for (int q = 0; q < PAIRS; ++q) {
                        if (q < kPacked) {
                            const float2 h = __ffma2_rn(sx, ux[q], __ffma2_rn(sy, uy[q], __ffma2_rn(sz, uz[q], sw)));
                            acc &= __float_as_int(h.x) & __float_as_int(h.y);
                        } else {
                            const float h0 = fmaf(sp.x, ux[q].x, fmaf(sp.y, uy[q].x, fmaf(sp.z, uz[q].x, sp.w)));
                            const float h1 = fmaf(sp.x, ux[q].y, fmaf(sp.y, uy[q].y, fmaf(sp.z, uz[q].y, sp.w)));
                            acc &= __float_as_int(h0) & __float_as_int(h1);
                        }
                    }
                }
                any = acc >= 0;
            } else {
#pragma unroll
                for (int u = 0; u < UNROLL; ++u) {
                    float2 h[PAIRS];
                    dot_pairs<PAIRS>(cur[u], ux, uy, uz, h);
                    float m = h[0].x;
                    if (PAIRS == 1) m = fmaxf(m, h[0].y);
                    else {
                        m = max3(m, h[0].y, h[1].x);
#pragma unroll
                        for (int q = 1; q + 1 < PAIRS; ++q) m = max3(m, h[q].y, h[q + 1].x);
                        m = fmaxf(m, h[PAIRS - 1].y);
                    }
                    any |= !(m < cur[u].w);
                }
            }
            if (any) hits += 1;
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) cur[u] = nxt[u];
        }
    }
    out[tid] = hits;
}

template <int FORM, int PAIRS, int UNROLL>
static void run(const char* name, const float4* geom, const float* dirs, unsigned* out, int n, int sms, int bps) {
    auto kern = scan_kernel<FORM, PAIRS, UNROLL>;
    const int smem = (n + UNROLL) * 16;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, kern);
    int occ = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 256, smem);
    if (bps > occ) bps = occ;
    const int blocks = sms * bps, reps = 48 / PAIRS;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    double best = 1e30;
    for (int r = 0; r < 6; ++r) {
        cudaEventRecord(e0);
        kern<<<blocks, 256, smem>>>(geom, dirs, out, n, reps);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (r >= 2 && ms < best) best = ms;
    }
    const double tests = (double)blocks * 256 * reps * PAIRS * 2 * n;
    const double gt = tests / (best * 1e-3) / 1e9;
    printf("%-26s regs %3d occ %d blocks/SM %d : %8.1f Gtests/s  = %7.1f Mrays/s @1024  (%s)\n", name, fa.numRegs, occ, bps, gt,
           gt * 1e3 / 1024.0, cudaGetErrorString(cudaGetLastError()));
}

template <int PAIRS, int UNROLL>
static void run_const(const char* name, const float* dirs, unsigned* out, int n, int sms, int bps) {
    auto kern = scan_const_kernel<PAIRS, UNROLL>;
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, kern);
    const int blocks = sms * bps, reps = 48 / PAIRS;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    double best = 1e30;
    for (int r = 0; r < 6; ++r) {
        cudaEventRecord(e0);
        kern<<<blocks, 256>>>(dirs, out, n, reps);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (r >= 2 && ms < best) best = ms;
    }
    const double tests = (double)blocks * 256 * reps * PAIRS * 2 * n;
    const double gt = tests / (best * 1e-3) / 1e9;
    printf("%-26s regs %3d blocks/SM %d : %8.1f Gtests/s  = %7.1f Mrays/s @1024  (%s)\n", name, fa.numRegs, bps, gt,
           gt * 1e3 / 1024.0, cudaGetErrorString(cudaGetLastError()));
}

int main() {
    cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
    const int n = 1024, sms = prop.multiProcessorCount;
    float4* geom; float* dirs; unsigned* out;
    cudaMalloc(&geom, (n + 16) * 16); cudaMalloc(&dirs, 4096 * 6 * 4); cudaMalloc(&out, (size_t)sms * 8 * 256 * 4);
    float4* hg = (float4*)malloc((n + 16) * 16); float* hd = (float*)malloc(4096 * 6 * 4);
    srand(1);
    for (int i = 0; i < n + 16; ++i) hg[i] = make_float4(rand() / (float)RAND_MAX * 20 - 10, rand() / (float)RAND_MAX, rand() / (float)RAND_MAX * 20 - 10, -1e30f);
    for (int i = 0; i < 4096 * 6; ++i) hd[i] = rand() / (float)RAND_MAX - 0.5f;
    cudaMemcpy(geom, hg, (n + 16) * 16, cudaMemcpyHostToDevice); cudaMemcpy(dirs, hd, 4096 * 6 * 4, cudaMemcpyHostToDevice);
#define RUN(F, P, U, B) run<F, P, U>("form " #F " pairs " #P " unroll " #U, geom, dirs, out, n, sms, B)
    cudaMemcpyToSymbol(c_geom, hg, (n + 16) * 16);
    for (int bps = 2; bps <= 3; ++bps) {
        RUN(3, 4, 4, bps); RUN(3, 4, 8, bps);
        RUN(9, 2, 4, bps); RUN(9, 2, 8, bps); RUN(9, 4, 4, bps); RUN(9, 4, 8, bps); RUN(9, 8, 4, bps);
    }
    return 0;
}
