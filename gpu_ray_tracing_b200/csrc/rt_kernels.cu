// rt_kernels.cu — CUDA kernels (sm_100a) for the `update` pass of compute_shader.wgsl:333-364.
//
// Build with -fmad=false: every a*b+c below is two roundings, exactly as the oracle evaluates it.
// Fused multiply-adds appear only as explicit fmaf() in the BROAD PHASE, whose result is never
// used for anything but "does this (ray, sphere) pair need the exact quadratic".
//
// Kernel structure:
//   broad phase  — per (ray, sphere): 7 FFMA + 1 FSETP.  With unit direction u = d/|d|,
//                  disc/|d|^2 = (C.u - O.u)^2 - (C.(-2O) + (|C|^2 - r^2)) - |O|^2, so with the
//                  per-sphere float4 (C, k = |C|^2 - r^2 - slack) broadcast from shared memory and
//                  8 per-ray registers (u, -O.u, -2O, |O|^2 - slack) the sign test is two 3-FFMA
//                  dot chains, one FFMA and one compare.  The slack makes it conservative: it can
//                  pass a sphere the exact test rejects, never the reverse.
//   narrow phase — for the rare candidates only, the reference's quadratic verbatim
//                  (rtc::sphere_root), in sphere-index order with the shrinking t_max of
//                  sphere_list_hit (compute_shader.wgsl:164-180).  Results are therefore
//                  bit-identical to evaluating every sphere exactly.
#include <cooperative_groups.h>
#include <type_traits>

#include "rt_kernels.cuh"
#include "rt_core.cuh"
#include "rt_tree_layout.h"

namespace rtk {

using rtc::f3;

namespace {

// Absolute image row of a window row.  A contiguous window starts at row_begin; an interleaved one (rt_create_bands)
// owns every band_stride / band_rows-th band of band_rows rows (a multiple of the 8-row tile, so a tile never
// straddles two bands).  Pixel seeds use absolute coordinates (compute_shader.wgsl:309-310).
__device__ __forceinline__ int abs_row(const UpdateParams& P, int row) {
    if (P.band_rows == 0u) return (int)P.row_begin + row;
    const uint32_t b = (uint32_t)row / P.band_rows;
    return (int)(P.row_begin + b * P.band_stride + ((uint32_t)row - b * P.band_rows));
}

constexpr float kTMin = 0.001f;    // compute_shader.wgsl:266
constexpr float kTMax = 3.4e35f;
constexpr unsigned kFull = 0xFFFFFFFFu;
constexpr int kUnroll = 4;         // spheres per broad-phase iteration
constexpr int kGeomPad = 8;        // geometry arrays are padded to a multiple of this ...
constexpr int kGeomTail = 8;       // ... plus a tail the software prefetch may read (never computed)

// Per-thread state of kPix rays.  Every index is a compile-time constant after unrolling, so the
// arrays live in registers.
constexpr int kPairs = kPix / 2;
static_assert(kPix % 2 == 0, "the broad phase works on pixel pairs (FFMA2)");

struct Lanes {
    float dx[kPix], dy[kPix], dz[kPix];     // ray direction, as the reference holds it
    float ox[kPix], oy[kPix], oz[kPix];     // ray origin
    float a[kPix];                          // dot(d, d) with the reference's rounding
    // broad phase, two rays per 64-bit register pair (.x = ray 2q, .y = ray 2q+1)
    float2 ux[kPairs], uy[kPairs], uz[kPairs];     // unit direction
    float2 h0[kPairs];                             // -dot(O, u)
    float2 mox[kPairs], moy[kPairs], moz[kPairs];  // -2 O
    float thr[kPix];                        // |O|^2 (1 - kappa); +inf for idle rays
    float best_t[kPix];
    int best_i[kPix];
    bool alive[kPix];
};

__device__ __forceinline__ float& half(float2& v, int p) { return (p & 1) ? v.y : v.x; }
__device__ __forceinline__ float half(const float2& v, int p) { return (p & 1) ? v.y : v.x; }

// Broad phase for the ray pair q against one sphere record: 7 FFMA2 (fma.rn.f32x2, sm_100).
// The sphere scalars enter as broadcast operands; each half is the IEEE fused operation, so the
// values equal rtc::broad_g() per ray.
__device__ __forceinline__ float2 broad_g2(const float4 sp, const Lanes& L, int q) {
    const float2 sx = make_float2(sp.x, sp.x), sy = make_float2(sp.y, sp.y), sz = make_float2(sp.z, sp.z);
    float2 h = __ffma2_rn(sx, L.ux[q], __ffma2_rn(sy, L.uy[q], __ffma2_rn(sz, L.uz[q], L.h0[q])));
    float2 c = __ffma2_rn(sx, L.mox[q], __ffma2_rn(sy, L.moy[q], __ffma2_rn(sz, L.moz[q], make_float2(sp.w, sp.w))));
    return __ffma2_rn(h, h, make_float2(-c.x, -c.y));
}

__device__ __forceinline__ float broad_g(const float4 sp, const Lanes& L, int p) {
    const int q = p >> 1;
    return rtc::broad_g(sp.x, sp.y, sp.z, sp.w, half(L.ux[q], p), half(L.uy[q], p), half(L.uz[q], p), half(L.h0[q], p),
                        half(L.mox[q], p), half(L.moy[q], p), half(L.moz[q], p));
}

template <int F = 0>
__device__ __forceinline__ void arm_ray(Lanes& L, int p, f3 o, f3 d) {
    const int q = p >> 1;
    L.dx[p] = d.x; L.dy[p] = d.y; L.dz[p] = d.z;
    L.ox[p] = o.x; L.oy[p] = o.y; L.oz[p] = o.z;
    const float a = rtc::dot(d, d);
    L.a[p] = a;
    const rtc::BroadRay b = rtc::broad_arm<F>(o, d, a);
    half(L.mox[q], p) = b.mox; half(L.moy[q], p) = b.moy; half(L.moz[q], p) = b.moz;
    half(L.ux[q], p) = b.ux; half(L.uy[q], p) = b.uy; half(L.uz[q], p) = b.uz;
    half(L.h0[q], p) = b.h0;
    L.thr[p] = b.thr;
    L.best_t[p] = kTMax;
    L.best_i[p] = -1;
    L.alive[p] = true;
}

// A ray whose first traversal has already been done by the primary pass: only what scatter() and the
// sky colour need.  It is armed later if it turns out to be untraced (adopt_first_hits).
__device__ __forceinline__ void park_ray(Lanes& L, int p, f3 o, f3 d) {
    L.dx[p] = d.x; L.dy[p] = d.y; L.dz[p] = d.z;
    L.ox[p] = o.x; L.oy[p] = o.y; L.oz[p] = o.z;
    const int q = p >> 1;
    L.a[p] = 1.0f;
    half(L.mox[q], p) = 0.0f; half(L.moy[q], p) = 0.0f; half(L.moz[q], p) = 0.0f;
    half(L.ux[q], p) = 0.0f; half(L.uy[q], p) = 0.0f; half(L.uz[q], p) = 0.0f;
    half(L.h0[q], p) = 0.0f;
    L.thr[p] = __int_as_float(0x7f800000);
    L.best_t[p] = kTMax;
    L.best_i[p] = -1;
    L.alive[p] = true;
}

__device__ __forceinline__ void idle_ray(Lanes& L, int p) {
    const int q = p >> 1;
    L.dx[p] = L.dy[p] = L.dz[p] = 0.0f;
    L.ox[p] = L.oy[p] = L.oz[p] = 0.0f;
    L.a[p] = 1.0f;
    half(L.mox[q], p) = 0.0f; half(L.moy[q], p) = 0.0f; half(L.moz[q], p) = 0.0f;
    half(L.ux[q], p) = 0.0f; half(L.uy[q], p) = 0.0f; half(L.uz[q], p) = 0.0f;
    half(L.h0[q], p) = 0.0f;
    L.thr[p] = __int_as_float(0x7f800000);
    L.best_t[p] = kTMax;
    L.best_i[p] = -1;
    L.alive[p] = false;
}

__device__ __forceinline__ f3 origin_of(const Lanes& L, int p) { return rtc::mk(L.ox[p], L.oy[p], L.oz[p]); }

// Narrow phase over `count` consecutive spheres starting at shared-memory slot `slot`
// (global index `index0`): the reference's sphere_hit root for every candidate ray.
__device__ __forceinline__ void narrow_phase(const float4* __restrict__ s_geom, int slot, int count, int index0,
                                          const GpuSphere* __restrict__ spheres, uint32_t n_spheres, Lanes& L,
                                          uint32_t& exact_tests) {
#pragma unroll 1
    for (int u = 0; u < count; ++u) {
        const int idx = index0 + u;
        if ((uint32_t)idx >= n_spheres) break;   // padding
        const float4 sp = s_geom[slot + u];
        const float radius = __ldg(&spheres[idx].radius);
        const f3 center = rtc::mk(sp.x, sp.y, sp.z);
#pragma unroll
        for (int p = 0; p < kPix; ++p) {
            if (L.alive[p] && !(broad_g(sp, L, p) < L.thr[p])) {
                ++exact_tests;
                float root;
                if (rtc::sphere_root(origin_of(L, p), rtc::mk(L.dx[p], L.dy[p], L.dz[p]), L.a[p], center, radius,
                                     kTMin, L.best_t[p], &root)) {
                    L.best_t[p] = root;
                    L.best_i[p] = idx;
                }
            }
        }
    }
}

__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}

// Broad phase over `count` (multiple of kUnroll) spheres resident at s_geom[0..count); the
// kUnroll records after the end must be readable (prefetch).  The next group's records are
// loaded before the current group's arithmetic so the LDS latency is off the critical path.
__device__ __forceinline__ void scan_spheres(const float4* __restrict__ s_geom, int count, int index0,
                                             const GpuSphere* __restrict__ spheres, uint32_t n_spheres, Lanes& L,
                                             uint32_t& exact_tests) {
    uint32_t addr = (uint32_t)__cvta_generic_to_shared(s_geom);
    float4 cur[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) cur[u] = lds128(addr + 16u * u);
#pragma unroll 2
    for (int s = 0; s < count; s += kUnroll) {
        addr += 16u * kUnroll;
        float4 nxt[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) nxt[u] = lds128(addr + 16u * u);
        bool any = false;
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
#pragma unroll
            for (int q = 0; q < kPairs; ++q) {
                const float2 g = broad_g2(cur[u], L, q);
                any |= !(g.x < L.thr[2 * q]);
                any |= !(g.y < L.thr[2 * q + 1]);
            }
        }
        if (any) narrow_phase(s_geom, s, kUnroll, index0 + s, spheres, n_spheres, L, exact_tests);
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) cur[u] = nxt[u];
    }
}

__device__ __forceinline__ GpuSphere load_sphere(const GpuSphere* __restrict__ spheres, int idx) {
    const float4* q = reinterpret_cast<const float4*>(spheres + idx);
    float4 g = __ldg(q), m = __ldg(q + 1);
    GpuSphere s;
    s.position.x = g.x; s.position.y = g.y; s.position.z = g.z; s.radius = g.w;
    s.color[0] = m.x; s.color[1] = m.y; s.color[2] = m.z; s.color[3] = m.w;
    return s;
}

// Per-pixel bookkeeping of `update` around the traversal.
struct PixelJob {
    f3 factor[kPix];
    uint32_t bseed[kPix];   // ray_color's `seed` argument (compute_shader.wgsl:355)
    uint32_t n[kPix];       // samples_until_now after the reset rule
    bool sampling[kPix];    // n < samples_per_pixel and inside the image
    bool black[kPix];       // sample absorbed
    bool inb[kPix];
};

}  // namespace

namespace {

// mean += (color - mean) / f32(n + 1).  The first sample divides by 1.0, which is exact for every operand: skipped
// (same bits).  Tolerance mode multiplies by the MUFU reciprocal.
template <int F>
__device__ __forceinline__ f3 running_mean(f3 mean, f3 color, uint32_t n) {
    const f3 delta = rtc::sub(color, mean);
    if (n == 0u) return rtc::add(mean, delta);
    if constexpr (F) return rtc::axpy_t<1>(rtc::fastm::rcp((float)(n + 1u)), delta, mean);
    else return rtc::add(mean, rtc::divs(delta, (float)(n + 1u)));
}
// Load the texels of one warp tile, apply the reset / sample-count rule of `update`
// (compute_shader.wgsl:339-353) and arm the primary rays (get_ray, :305-325).
__device__ __forceinline__ void begin_tile(const UpdateParams& P, uint32_t tile, int lane, uint32_t seed_bits,
                                           uint32_t spp, bool reset, Lanes& L, PixelJob& J, int& row, int (&px)[kPix]) {
    const uint32_t tx = tile % P.tiles_x, ty = tile / P.tiles_x;
    row = (int)(ty * kWarpTileH) + (lane >> 3);
    const int y_abs = abs_row(P, row);
#pragma unroll
    for (int p = 0; p < kPix; ++p) {
        px[p] = (int)(tx * kWarpTileW) + (lane & 7) + 8 * p;
        J.inb[p] = (uint32_t)px[p] < P.width && (uint32_t)row < P.row_count;
        J.black[p] = false;
        J.factor[p] = rtc::mk(1.0f, 1.0f, 1.0f);
        uint32_t n = 0;
        if (J.inb[p] && !reset) n = rtc::f2u(P.accum[(size_t)row * P.width + px[p]].w);
        J.n[p] = n;
        J.sampling[p] = J.inb[p] && n < spp;
        if (J.sampling[p]) {
            const uint32_t seed = 1u + n + seed_bits;   // compute_shader.wgsl:353
            rtc::Ray r = rtc::get_ray(P.cam, px[p], y_abs, seed, seed_bits);
            J.bseed[p] = seed + 1u;
            if (P.first_hit != nullptr) park_ray(L, p, r.o, r.d);
            else arm_ray(L, p, r.o, r.d);
        } else {
            J.bseed[p] = 0u;
            idle_ray(L, p);
        }
    }
}

// After a traversal: miss -> the ray leaves for the sky; hit -> scatter (compute_shader.wgsl:267-290).
__device__ __forceinline__ void after_traversal(const UpdateParams& P, uint32_t bounce, bool last, Lanes& L, PixelJob& J) {
#pragma unroll
    for (int p = 0; p < kPix; ++p) {
        if (!L.alive[p]) continue;
        if (L.best_i[p] < 0) {
            // keep direction for the sky colour, stop tracing
            L.alive[p] = false;
            L.thr[p] = __int_as_float(0x7f800000);
            continue;
        }
        const GpuSphere sp = load_sphere(P.spheres, L.best_i[p]);
        rtc::Ray r;
        r.o = origin_of(L, p);
        r.d = rtc::mk(L.dx[p], L.dy[p], L.dz[p]);
        const uint32_t s = rtc::hash(J.bseed[p] + bounce * 1000u);   // compute_shader.wgsl:268
        const bool ok = rtc::scatter(&r, &J.factor[p], L.best_t[p], sp, s);
        if (!ok) {
            J.black[p] = true;
            L.alive[p] = false;
            L.thr[p] = __int_as_float(0x7f800000);
            L.dx[p] = r.d.x; L.dy[p] = r.d.y; L.dz[p] = r.d.z;
        } else if (last) {
            // no further traversal: keep the scattered direction for the sky colour (compute_shader.wgsl:293-296)
            L.alive[p] = false;
            L.dx[p] = r.d.x; L.dy[p] = r.d.y; L.dz[p] = r.d.z;
        } else {
            arm_ray(L, p, r.o, r.d);
        }
    }
}

// Running mean and texel store (compute_shader.wgsl:356-363).
__device__ __forceinline__ void end_tile(const UpdateParams& P, const Lanes& L, const PixelJob& J, int row,
                                         const int (&px)[kPix], bool reset) {
#pragma unroll
    for (int p = 0; p < kPix; ++p) {
        if (!J.inb[p]) continue;
        float4* texel = &P.accum[(size_t)row * P.width + px[p]];
        f3 mean = rtc::mk(0.0f, 0.0f, 0.0f);
        if (!reset) {
            float4 t = *texel;
            mean = rtc::mk(t.x, t.y, t.z);
        }
        uint32_t n = J.n[p];
        if (J.sampling[p]) {
            f3 color = J.black[p] ? rtc::mk(0.0f, 0.0f, 0.0f)
                                  : rtc::sky_times(J.factor[p], rtc::mk(L.dx[p], L.dy[p], L.dz[p]));
            mean = running_mean<0>(mean, color, n);
            n += 1u;
        }
        *texel = make_float4(mean.x, mean.y, mean.z, (float)n);
    }
}

// Bounce 0 when the primary pass already traced the ray: take (t, index) from first_hit and park the
// ray (not alive) for the duration of the scan that the untraced rays (index code -2) still need.
// Returns whether this thread has a ray left to scan; `pre` marks the parked rays.
__device__ __forceinline__ bool adopt_first_hits(const UpdateParams& P, int row, const int (&px)[kPix], Lanes& L,
                                                 bool (&pre)[kPix]) {
    bool todo = false;
#pragma unroll
    for (int p = 0; p < kPix; ++p) {
        pre[p] = false;
        if (L.alive[p]) {
            const float2 fh = P.first_hit[(size_t)row * P.width + px[p]];
            const int code = __float_as_int(fh.y);
            if (code != -2) {
                L.best_t[p] = fh.x;
                L.best_i[p] = code;
                L.alive[p] = false;
                pre[p] = true;
            } else {
                arm_ray(L, p, origin_of(L, p), rtc::mk(L.dx[p], L.dy[p], L.dz[p]));
                todo = true;
            }
        }
    }
    return todo;
}
__device__ __forceinline__ void unpark(Lanes& L, const bool (&pre)[kPix]) {
#pragma unroll
    for (int p = 0; p < kPix; ++p)
        if (pre[p]) L.alive[p] = true;
}

__device__ __forceinline__ void flush_counters(const UpdateParams& P, uint32_t rays, uint32_t exact_tests, int lane) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        rays += __shfl_xor_sync(kFull, rays, o);
        exact_tests += __shfl_xor_sync(kFull, exact_tests, o);
    }
    if (lane == 0) {
        atomicAdd(&P.counters[0], (unsigned long long)rays);
        atomicAdd(&P.counters[1], (unsigned long long)exact_tests);
    }
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// Variant 2: scene resident in shared memory, persistent warps pulling 32x4 tiles.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlockThreads, 2) update_resident_kernel(const __grid_constant__ UpdateParams P) {
    extern __shared__ float4 s_geom[];
    const int n_pad = (int)((P.n_spheres + kGeomPad - 1) / kGeomPad * kGeomPad);
    for (int i = threadIdx.x; i < n_pad + kGeomTail; i += kBlockThreads) s_geom[i] = __ldg(&P.geom[i]);
    __syncthreads();

    const int lane = threadIdx.x & 31;
    const uint32_t seed_bits = rtc::f2u(P.cam.random_seed * 4294967296.0f);
    const uint32_t spp = rtc::f2u(P.cam.samples_per_pixel);
    const uint32_t max_depth = rtc::f2u(P.cam.max_depth);
    const bool reset = P.cam.camera_has_moved > 0.5f;
    uint32_t rays = 0, exact_tests = 0;

    for (;;) {
        uint32_t tile = 0;
        if (lane == 0) tile = atomicAdd(P.tile_counter, 1u);
        tile = __shfl_sync(kFull, tile, 0);
        if (tile >= P.n_tiles) break;

        Lanes L;
        PixelJob J;
        int row, px[kPix];
        begin_tile(P, tile, lane, seed_bits, spp, reset, L, J, row, px);

        for (uint32_t bounce = 0; bounce < max_depth; ++bounce) {
            bool mine = false;
#pragma unroll
            for (int p = 0; p < kPix; ++p) {
                mine |= L.alive[p];
                rays += L.alive[p] ? 1u : 0u;
            }
            if (!__any_sync(kFull, mine)) break;
            if (bounce == 0 && P.first_hit != nullptr) {
                bool pre[kPix];
                const bool todo = adopt_first_hits(P, row, px, L, pre);
                if (__any_sync(kFull, todo)) scan_spheres(s_geom, n_pad, 0, P.spheres, P.n_spheres, L, exact_tests);
                unpark(L, pre);
            } else {
                scan_spheres(s_geom, n_pad, 0, P.spheres, P.n_spheres, L, exact_tests);
            }
            after_traversal(P, bounce, bounce + 1u >= max_depth, L, J);
        }
        end_tile(P, L, J, row, px, reset);
    }
    flush_counters(P, rays, exact_tests, lane);
}

// ---------------------------------------------------------------------------------------------
// Common-origin primary pass (variant 4).  For a pinhole camera every
// primary ray starts at cam.center, so the broad phase is ONE 3-term dot product per (ray, sphere)
// against the per-frame record (C - O, -q) written by prep_primary_kernel: candidate <=> sign bit
// of h' clear (rtc::primary_h).  A thread owns a 1 x 8 pixel strip (4 FFMA2 pairs), a warp a
// 32 x 8 tile; per record: 12 FFMA2 + 4 LOP3 (sign bits AND-ed).  Candidates go through the
// reference quadratic in index order (narrow phase), so the (t, index) written to first_hit is
// exactly what sphere_list_hit returns for the primary ray.  Shading and the texel update stay in
// the update kernel, which skips its own first traversal when first_hit is present.
// ---------------------------------------------------------------------------------------------
// Linear RGBA8 of the running mean: q(x) = floor(clamp(x, 0, 1) * 255 + 0.5), alpha 255.
// (saturate maps NaN to 0 as the two-comparison clamp + NaN check of the oracle does; multiply and add round separately
// under -fmad=false; the conversion rounds toward -infinity = floor)
__device__ __forceinline__ unsigned quant1(float v) { return __float2uint_rd(__saturatef(v) * 255.0f + 0.5f); }
__device__ __forceinline__ unsigned quant_texel(float4 t) {
    return quant1(t.x) | (quant1(t.y) << 8) | (quant1(t.z) << 16) | 0xFF000000u;
}

namespace {

// The reference loop itself (compute_shader.wgsl:164-180) for a ray whose tree traversal met a NaN root.
__device__ __noinline__ float2 retrace_sequential(const GpuSphere* __restrict__ spheres, uint32_t n_spheres, float ox, float oy,
                                                  float oz, float dx, float dy, float dz, float a) {
    float bt = kTMax;
    int bi = -1;
    for (uint32_t si = 0; si < n_spheres; ++si) {
        const float4 g = __ldg(reinterpret_cast<const float4*>(spheres + si));
        float root;
        if (rtc::sphere_root(rtc::mk(ox, oy, oz), rtc::mk(dx, dy, dz), a, rtc::mk(g.x, g.y, g.z), g.w, kTMin, bt, &root)) {
            bt = root;
            bi = (int)si;
        }
    }
    return make_float2(bt, __int_as_float(bi));
}

constexpr int kPrimSlots = 8;   // shared-memory floats per ray: d.xyz, o.xyz, nearest hit (best_t, best_i)
static_assert((kPrimList + 4) * 6 >= kPrimRays * 32, "prim_setup parks one float per ray and lane in the list area");
static_assert(kPrimThreads % 32 == 0 && kPrimWork >= 160 + 64, "entry list (rtt::kMaxEntry) + growth");

constexpr int kPrimSched = 4;   // per-warp words behind the work stack: the current tile's row and starting clock (+ padding)
__device__ __forceinline__ uint32_t clock32() {
    uint32_t c;
    asm volatile("mov.u32 %0, %%clock;" : "=r"(c));
    return c;
}

struct PrimLanes {
    float2 ux[kPrimRays / 2], uy[kPrimRays / 2], uz[kPrimRays / 2];   // .x = row 2q, .y = row 2q + 1
    uint32_t alive;   // bit p: ray p is traced here
    float bt_max;     // the farthest nearest-hit among the traced rays (kTMax while one of them has none)
    float near_scale; // rtc::primary_near_scale of the smallest 1 / |d| among them
};

__device__ __forceinline__ float ray_half(const float2& v, int p) { return (p & 1) ? v.y : v.x; }

// Exact tests (the reference quadratic) of sphere `idx` for the rays in `cand`.  The tree hands the spheres over in
// no particular order, so the nearest hit is kept with rtc::nearest_update (ties to the lower index); a ray that
// meets a NaN root is handed to the sequential fallback (`untraced`), where the reference's order decides.
__device__ __forceinline__ void prim_exact(uint32_t cand, int idx, const float4 g /* centre, radius as uploaded */,
                                           float* __restrict__ s_d, PrimLanes& L, uint32_t& untraced, uint32_t& exact_tests) {
    const f3 center = rtc::mk(g.x, g.y, g.z);
    // one rolled copy of the quadratic for all eight rays (ray state lives in the thread's shared-memory slots):
    // the kernel's instruction footprint, not its instruction count, is what the profile shows to matter
#pragma unroll 1
    while (cand) {
        const int p = __ffs((int)cand) - 1;
        cand &= cand - 1u;
        ++exact_tests;
        float* dp = s_d + (p * kPrimSlots) * kPrimThreads;
        const f3 d = rtc::mk(dp[0], dp[kPrimThreads], dp[2 * kPrimThreads]);
        const f3 origin = rtc::mk(dp[3 * kPrimThreads], dp[4 * kPrimThreads], dp[5 * kPrimThreads]);
        float bt = dp[6 * kPrimThreads];
        int bi = __float_as_int(dp[7 * kPrimThreads]);
        if (!rtc::nearest_update(origin, d, rtc::dot(d, d), center, g.w, kTMin, kTMax, idx, &bt, &bi)) {
            untraced |= 1u << p;
            L.alive &= ~(1u << p);
        }
        dp[6 * kPrimThreads] = bt;
        dp[7 * kPrimThreads] = __int_as_float(bi);
    }
    if (!RT_PRIM_CULL) return;
    // the strip's farthest nearest-hit, for the "already beaten" cull of the spheres still to come
    float far = -kTMax;
#pragma unroll
    for (int p = 0; p < kPrimRays; ++p)
        if (L.alive & (1u << p)) far = fmaxf(far, s_d[(p * kPrimSlots + 6) * kPrimThreads]);
    L.bt_max = far;
}

// Primary rays of a strip (get_ray, compute_shader.wgsl:305-325): exact (d, o) parked in shared memory, unit
// directions in registers.  `untraced` collects sampling rays outside primary_arm's window.
template <int F>
__device__ __forceinline__ void prim_setup(const UpdateParams& P, int x, int row0, float* __restrict__ s_d,
                                           float* __restrict__ s_tmp, uint32_t seed_bits, uint32_t spp, bool reset, PrimLanes& L,
                                           uint32_t& untraced) {
    L.alive = 0;
    untraced = 0;
    float inv_min = __int_as_float(0x7f800000);
    const int y0 = abs_row(P, row0);   // (the tile's rows are consecutive image rows)
#pragma unroll 1
    for (int p = 0; p < kPrimRays; ++p) {
        const int row = row0 + p;
        f3 d = rtc::mk(0.0f, 0.0f, 0.0f), o = d;
        if ((uint32_t)x < P.width && (uint32_t)row < P.row_count) {
            uint32_t n = 0;
            if (!reset) n = rtc::f2u(P.accum[(size_t)row * P.width + x].w);
            if (n < spp) {
                const uint32_t seed = 1u + n + seed_bits;   // compute_shader.wgsl:353
                const rtc::Ray r = rtc::get_ray<F>(P.cam, x, y0 + p, seed, seed_bits);
                d = r.d;
                o = r.o;
                L.alive |= 1u << p;
            }
        }
        // unit direction of the common-origin test (rtc::primary_arm), here in the rolled loop: parked in the slots
        // of the nearest hit, which are initialised when the directions are picked up below
        float ux, uy, uz, inv = __int_as_float(0x7f800000);
        const bool ok = rtc::primary_arm<F>(d, rtc::dot(d, d), &ux, &uy, &uz, &inv);
        if (!ok && (L.alive & (1u << p))) {
            untraced |= 1u << p;
            L.alive &= ~(1u << p);
        }
        if (L.alive & (1u << p)) inv_min = fminf(inv_min, inv);
        float* dp = s_d + (p * kPrimSlots) * kPrimThreads;
        dp[0] = d.x; dp[kPrimThreads] = d.y; dp[2 * kPrimThreads] = d.z;
        dp[3 * kPrimThreads] = o.x; dp[4 * kPrimThreads] = o.y; dp[5 * kPrimThreads] = o.z;
        dp[6 * kPrimThreads] = ux; dp[7 * kPrimThreads] = uy;
        s_tmp[p * 32] = uz;   // (the warp's list area is idle during set-up)
    }
#pragma unroll
    for (int p = 0; p < kPrimRays; ++p) {
        float* dp = s_d + (p * kPrimSlots) * kPrimThreads;
        const float ux = dp[6 * kPrimThreads], uy = dp[7 * kPrimThreads], uz = s_tmp[p * 32];
        const int q = p >> 1;
        if (p & 1) { L.ux[q].y = ux; L.uy[q].y = uy; L.uz[q].y = uz; }
        else       { L.ux[q].x = ux; L.uy[q].x = uy; L.uz[q].x = uz; }
        dp[6 * kPrimThreads] = kTMax;
        dp[7 * kPrimThreads] = __int_as_float(-1);
    }
    L.bt_max = kTMax;
    L.near_scale = rtc::primary_near_scale(inv_min);   // (+inf without traced rays: nothing to cull for)
}

// Axis and chord bound of the warp's tile (rt_core.cuh, "Tile level"): bounding box of the traced rays' unit
// directions, reduced over the warp with min / max (exact, so the CPU emulation reproduces it bit for bit).
__device__ __forceinline__ void prim_tile_bound(const PrimLanes& L, f3& m, float& chord) {
    const float inf = __int_as_float(0x7f800000);
    f3 lo = rtc::mk(inf, inf, inf), hi = rtc::mk(-inf, -inf, -inf);
#pragma unroll
    for (int p = 0; p < kPrimRays; ++p) {
        if (L.alive & (1u << p)) {
            const float ux = ray_half(L.ux[p >> 1], p), uy = ray_half(L.uy[p >> 1], p), uz = ray_half(L.uz[p >> 1], p);
            lo = rtc::mk(fminf(lo.x, ux), fminf(lo.y, uy), fminf(lo.z, uz));
            hi = rtc::mk(fmaxf(hi.x, ux), fmaxf(hi.y, uy), fmaxf(hi.z, uz));
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lo = rtc::mk(fminf(lo.x, __shfl_xor_sync(kFull, lo.x, o)), fminf(lo.y, __shfl_xor_sync(kFull, lo.y, o)),
                     fminf(lo.z, __shfl_xor_sync(kFull, lo.z, o)));
        hi = rtc::mk(fmaxf(hi.x, __shfl_xor_sync(kFull, hi.x, o)), fmaxf(hi.y, __shfl_xor_sync(kFull, hi.y, o)),
                     fmaxf(hi.z, __shfl_xor_sync(kFull, hi.z, o)));
    }
    m = rtc::primary_tile_axis(lo, hi);
    float e2 = 0.0f;
#pragma unroll
    for (int p = 0; p < kPrimRays; ++p) {
        if (L.alive & (1u << p))
            e2 = fmaxf(e2, rtc::primary_tile_e2(ray_half(L.ux[p >> 1], p), ray_half(L.uy[p >> 1], p), ray_half(L.uz[p >> 1], p), m));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) e2 = fmaxf(e2, __shfl_xor_sync(kFull, e2, o));
    chord = rtc::primary_tile_chord(e2);
}

// The warp's tile against the tree of bounding spheres (rt_tree.h) with this frame's common-origin records
// (P.prim_tree, one per child slot, written by prep_primary_tree_kernel).
//   walk: the warp takes up to 4 nodes from its work stack, every lane tests ONE child record against the TILE
//         (primary_tile_candidate); surviving nodes go back on the stack, surviving spheres (children of leaf nodes)
//         into the warp's candidate list (ballot + popc compaction);
//   scan: every lane runs the per-ray sign test (3 FFMA per ray, as FFMA2 pairs) on the listed records, the next one
//         prefetched, and puts its candidates through the exact test.
// A 32 x 8 pixel tile of a 4K frame ends up with about 6 of 1024 spheres after ~10 warp steps; the work grows with the
// logarithm of the scene.  Records come straight from global memory (L1/L2): a tile touches a few hundred bytes.
struct PrimTree {
    const float4* rec;           // this frame's records, kTreeStride slots per node
    const uint32_t* link;        // kTreeWide per node
    const uint32_t* entry_node;  // the entry list (rt_tree.h): nodes ...
    const float4* entry_rec;     // ... and this frame's records of them, contiguous (behind the child slots)
    uint32_t first_leaf, n_entry;
};

// Nearest first: the warp sorts its listed spheres by rtc::primary_near (32 at a time, rank by counting), so that the
// hits found early beat most of what follows (prim_scan_list).  Any order gives the same image.
__device__ __forceinline__ void prim_sort_list(float4* __restrict__ s_rec, uint32_t* __restrict__ s_idx, float* __restrict__ s_near,
                                               int n_list, int lane) {
#pragma unroll 1
    for (int c0 = 0; c0 < n_list; c0 += 32) {
        const int m = min(32, n_list - c0);
        if (m > 1) {
            const bool mine = lane < m;
            const float4 rec = mine ? s_rec[c0 + lane] : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            const uint32_t idx = mine ? s_idx[c0 + lane] : 0u;
            const float key = mine ? s_near[c0 + lane] : 0.0f;
            int rank = 0;
#pragma unroll 1
            for (int j = 0; j < m; ++j) {
                const float kj = __shfl_sync(kFull, key, j);
                rank += (kj < key || (kj == key && j < lane)) ? 1 : 0;
            }
            __syncwarp();
            if (mine) {
                s_rec[c0 + rank] = rec;
                s_idx[c0 + rank] = idx;
                s_near[c0 + rank] = key;
            }
            __syncwarp();
        }
    }
}

__device__ __forceinline__ void prim_scan_list(float4* __restrict__ s_rec, uint32_t* __restrict__ s_idx, float* __restrict__ s_near,
                                               int n_list, float* __restrict__ s_d, const UpdateParams& P, PrimLanes& L,
                                               uint32_t& untraced, uint32_t& exact_tests, int lane) {
    if (n_list == 0) return;
    const bool sorted = RT_PRIM_CULL && n_list >= RT_PRIM_SORT_MIN && n_list <= RT_PRIM_SORT_MAX;
    if (sorted) prim_sort_list(s_rec, s_idx, s_near, n_list, lane);
    float4 rec = s_rec[0];
#pragma unroll 1
    for (int e = 0; e < n_list; ++e) {
        // already beaten for every ray of the strip (rt_core.cuh, primary_near); within a 32-entry run the bounds
        // ascend, so once the whole warp says so the rest of the run is skipped as well
        const bool beaten = RT_PRIM_CULL && s_near[e] * L.near_scale > L.bt_max;
        if (sorted && __all_sync(kFull, beaten)) {
            const int next = ((e >> 5) + 1) << 5;
            if (next >= n_list) break;
            e = next - 1;
            rec = s_rec[next];
            continue;
        }
        const float4 recn = s_rec[e + 1];   // (one readable slot past the list)
        // the sphere as uploaded (exact test operand): a broadcast load issued ahead of the sign tests, which cover its
        // latency in scenes past L1 (65,536 spheres are 2 MB)
        const uint32_t idx = s_idx[e];
        const float4 sph = __ldg(reinterpret_cast<const float4*>(P.spheres + idx));
        const float2 sx = make_float2(rec.x, rec.x), sy = make_float2(rec.y, rec.y), sz = make_float2(rec.z, rec.z),
                     sw = make_float2(rec.w, rec.w);
        float2 h[kPrimRays / 2];
        int acc = -1;   // AND of the rays' sign bits
#pragma unroll
        for (int q = 0; q < kPrimRays / 2; ++q) {
            h[q] = __ffma2_rn(sx, L.ux[q], __ffma2_rn(sy, L.uy[q], __ffma2_rn(sz, L.uz[q], sw)));
            acc &= __float_as_int(h[q].x) & __float_as_int(h[q].y);
        }
        if (acc >= 0) {
            uint32_t cand = 0;
#pragma unroll
            for (int p = 0; p < kPrimRays; ++p) cand |= rtc::primary_candidate(ray_half(h[p >> 1], p)) ? (1u << p) : 0u;
            cand &= L.alive;
            if (cand && !beaten) prim_exact(cand, (int)idx, sph, s_d, L, untraced, exact_tests);
        }
        rec = recn;
    }
}

__device__ __forceinline__ void prim_tree_walk(const PrimTree& T, uint32_t* __restrict__ s_work, float4* __restrict__ s_rec,
                                               uint32_t* __restrict__ s_idx, float* __restrict__ s_near, f3 m, float chord, float* __restrict__ s_d,
                                               const UpdateParams& P, PrimLanes& L, uint32_t& untraced, uint32_t& exact_tests,
                                               int lane) {
    const unsigned lt = (1u << lane) - 1u;
    int n_work = 0, n_list = 0;
    // the entry list first: independent iterations, 32 records per step, instead of a walk down from the root whose
    // every step waits for the one before
#pragma unroll 1
    for (uint32_t e0 = 0; e0 < T.n_entry; e0 += 32u) {
        const uint32_t e = e0 + (uint32_t)lane;
        bool keep = false;
        uint32_t node = 0;
        if (e < T.n_entry) {
            node = T.entry_node[e];
            const float4 r = __ldg(T.entry_rec + e);
            keep = rtc::primary_tile_candidate(r.x, r.y, r.z, r.w, m, chord);
        }
        const unsigned b = __ballot_sync(kFull, keep);
        if (keep) s_work[n_work + __popc(b & lt)] = node;
        n_work += __popc(b);
    }
    __syncwarp();
    while (n_work > 0) {
        // a nearly full stack is worked off one node at a time (at most 7 entries of growth per level below it)
        const int take = n_work > kPrimWork - 64 ? 1 : min(n_work, 4);
        n_work -= take;
        const int g = lane >> 3, k = lane & 7;
        bool keep = false, leaf = false;
        uint32_t lk = 0xFFFFFFFFu;
        float4 r = make_float4(0.0f, 0.0f, 0.0f, -1.0f);
        float near = 0.0f;   // rtc::primary_near of the child's record
        if (g < take) {
            const uint32_t node = s_work[n_work + g];
            // link and record are fetched side by side (an empty slot holds a readable never-record), not one after the other
            lk = __ldg(T.link + node * kTreeWide + k);
            r = __ldg(T.rec + node * kTreeStride + k);
            leaf = node >= T.first_leaf;
            keep = rtc::primary_tile_candidate(r.x, r.y, r.z, r.w, m, chord, &near) && lk != 0xFFFFFFFFu;
        }
        __syncwarp();   // the taken entries have been read before their slots are reused
        const unsigned bi = __ballot_sync(kFull, keep && !leaf), bl = __ballot_sync(kFull, keep && leaf);
        if (keep && !leaf) s_work[n_work + __popc(bi & lt)] = lk;
        n_work += __popc(bi);
        if (keep && leaf) {
            const int j = n_list + __popc(bl & lt);
            s_rec[j] = r;   // the sphere's record travels with it: the scan reads shared memory only
            s_idx[j] = lk;
            if (RT_PRIM_CULL) s_near[j] = near;
        }
        n_list += __popc(bl);
        __syncwarp();
        if (n_list > kPrimList - 32) {   // no room for another step's spheres: scan what is listed
            prim_scan_list(s_rec, s_idx, s_near, n_list, s_d, P, L, untraced, exact_tests, lane);
            n_list = 0;
            __syncwarp();
        }
    }
    prim_scan_list(s_rec, s_idx, s_near, n_list, s_d, P, L, untraced, exact_tests, lane);
    __syncwarp();   // the lists are rewritten by the next tile
}

// Running mean and texel store for a finished path (compute_shader.wgsl:356-363).
template <int F = 0>
__device__ __forceinline__ void finalize_texel(const UpdateParams& P, uint32_t texel_index, uint32_t n, f3 color, bool reset) {
    float4* texel = &P.accum[texel_index];
    f3 mean = rtc::mk(0.0f, 0.0f, 0.0f);
    if (!reset) {
        const float4 t = *texel;
        mean = rtc::mk(t.x, t.y, t.z);
    }
    mean = running_mean<F>(mean, color, n);
    *texel = make_float4(mean.x, mean.y, mean.z, (float)(n + 1u));
}

// Warp-aggregated append of one path per pushing lane.  Every lane of the warp must call this.
__device__ __forceinline__ void push_path(const UpdateParams& P, uint32_t next_bounce, bool push, f3 o, f3 d, f3 factor,
                                          uint32_t texel_index, uint32_t bseed, uint32_t n, int lane) {
    const unsigned m = __ballot_sync(kFull, push);
    if (m == 0u) return;
    const int leader = __ffs((int)m) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(&P.wf_counts[next_bounce], (unsigned)__popc(m));
    base = __shfl_sync(kFull, base, leader);
    if (push) {
        const uint32_t j = base + (uint32_t)__popc(m & ((1u << lane) - 1u));
        float4* q = P.wf_queue[next_bounce & 1u];
        q[j] = make_float4(o.x, o.y, o.z, d.x);
        q[(size_t)P.wf_capacity + j] = make_float4(d.y, d.z, factor.x, factor.y);
        q[2 * (size_t)P.wf_capacity + j] =
            make_float4(factor.z, __uint_as_float(texel_index), __uint_as_float(bseed), __uint_as_float(n));
    }
}

// What happens to a traced ray (compute_shader.wgsl:267-296): returns true when the path continues (then o, d,
// factor are the scattered ray), false when the texel has been finished.
template <int F = 0>
__device__ __forceinline__ bool resolve_hit(const UpdateParams& P, uint32_t bounce, bool last, int best_i, float best_t,
                                            f3& o, f3& d, f3& factor, uint32_t texel_index, uint32_t bseed, uint32_t n,
                                            bool reset, rtc::UnitVecCache* ruv = nullptr) {
    if (best_i < 0) {
        finalize_texel<F>(P, texel_index, n, rtc::sky_times<F>(factor, d), reset);
        return false;
    }
    const GpuSphere sp = load_sphere(P.spheres, best_i);
    rtc::Ray r;
    r.o = o;
    r.d = d;
    const bool ok = rtc::scatter<F>(&r, &factor, best_t, sp, rtc::hash(bseed + bounce * 1000u), ruv);   // compute_shader.wgsl:268
    if (!ok) {
        finalize_texel<F>(P, texel_index, n, rtc::mk(0.0f, 0.0f, 0.0f), reset);
        return false;
    }
    if (last) {
        finalize_texel<F>(P, texel_index, n, rtc::sky_times<F>(factor, r.d), reset);
        return false;
    }
    o = r.o;
    d = r.d;
    return true;
}

// kMode = kPrimHits:  write (t, index) per texel to P.first_hit; the update kernel shades and continues the paths
//         (in-kernel bounce loop, max_depth > kWfMaxDepth).
// kMode = kPrimShade (max_depth == 1): the whole `update` for the texel is finished here — scatter once for the
//         attenuation and the outgoing direction, sky colour, running mean, store (compute_shader.wgsl:267-296,
//         356-363) — and no second kernel runs.
// kMode = kPrimQueue (wavefront, 1 < max_depth <= kWfMaxDepth): the first hit is resolved here as well — texel finished
//         (miss, absorbed) or the scattered path appended to queue 1 (resolve_hit; a texel-per-thread pass of its own
//         over P.first_hit did this in round 1: 8 B written and read back per texel, every ray generated twice).
constexpr int kPrimHits = 0, kPrimShade = 1, kPrimQueue = 2;
template <int kMode, int F>
__device__ __forceinline__ void prim_finish(const UpdateParams& P, int x, int row0, float* __restrict__ s_d,
                                            const PrimLanes& L, uint32_t untraced, uint32_t seed_bits, bool reset,
                                            uint32_t& rays) {
    if constexpr (kMode == kPrimQueue) {
        const uint32_t sampling = L.alive | untraced;
        const uint32_t max_depth = rtc::f2u(P.cam.max_depth);
        const int lane = threadIdx.x & 31;
        rtc::UnitVecCache ruv_queue;   // the strip's rays share their bounce seed (rtc::UnitVecCache)
        ruv_queue.valid = false;
        // pass A: resolve every ray's first hit; a continuing path parks its scattered ray in the ray's own slots
        // (d, o; n in the best_t slot, the sphere index stays: the factor after one hit is that sphere's attenuation)
        uint32_t pmask = 0;
#pragma unroll 1
        for (int p = 0; p < kPrimRays; ++p) {
            const int row = row0 + p;
            if (!((uint32_t)x < P.width && (uint32_t)row < P.row_count)) continue;
            const uint32_t texel_index = (uint32_t)row * P.width + (uint32_t)x;
            uint32_t n = 0;
            if (!reset) n = rtc::f2u(P.accum[texel_index].w);
            if (sampling & (1u << p)) {
                float* dp = s_d + (p * kPrimSlots) * kPrimThreads;
                f3 d = rtc::mk(dp[0], dp[kPrimThreads], dp[2 * kPrimThreads]);
                f3 o = rtc::mk(dp[3 * kPrimThreads], dp[4 * kPrimThreads], dp[5 * kPrimThreads]);
                float best_t = dp[6 * kPrimThreads];
                int best_i = __float_as_int(dp[7 * kPrimThreads]);
                if (untraced & (1u << p)) {   // outside primary_arm's window: the reference loop itself
                    const float2 hit = retrace_sequential(P.spheres, P.n_spheres, o.x, o.y, o.z, d.x, d.y, d.z, rtc::dot(d, d));
                    best_t = hit.x;
                    best_i = __float_as_int(hit.y);
                    dp[7 * kPrimThreads] = hit.y;
                }
                ++rays;
                f3 factor = rtc::mk(1.0f, 1.0f, 1.0f);
                const uint32_t bseed = 2u + n + seed_bits;   // hash seed of bounce 0 (compute_shader.wgsl:353, 268)
                if (resolve_hit<F>(P, 0u, max_depth <= 1u, best_i, best_t, o, d, factor, texel_index, bseed, n, reset, &ruv_queue)) {
                    dp[0] = d.x; dp[kPrimThreads] = d.y; dp[2 * kPrimThreads] = d.z;
                    dp[3 * kPrimThreads] = o.x; dp[4 * kPrimThreads] = o.y; dp[5 * kPrimThreads] = o.z;
                    dp[6 * kPrimThreads] = __uint_as_float(n);
                    pmask |= 1u << p;
                }
            } else if (reset) {
                P.accum[texel_index] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);   // reset rule with nothing to sample
            }
        }
        // pass B: ONE reservation per tile, the tile's surviving paths appended row by row — a 64-ray batch of the first
        // bounce is then a 32 x 2 pixel block of one tile (appending row by row with a reservation each interleaves
        // the tiles of different warps in the queue and costs the bounce launches 3 %)
        uint32_t total = (uint32_t)__popc(pmask);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) total += __shfl_xor_sync(kFull, total, o);
        if (total == 0u) return;
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(&P.wf_counts[1], total);
        base = __shfl_sync(kFull, base, 0);
        float4* q = P.wf_queue[1];
#pragma unroll 1
        for (int p = 0; p < kPrimRays; ++p) {
            const bool mine = (pmask >> p) & 1u;
            const unsigned m = __ballot_sync(kFull, mine);
            if (mine) {
                const float* dp = s_d + (p * kPrimSlots) * kPrimThreads;
                const uint32_t j = base + (uint32_t)__popc(m & ((1u << lane) - 1u));
                const uint32_t n = __float_as_uint(dp[6 * kPrimThreads]);
                const float4 col = __ldg(reinterpret_cast<const float4*>(P.spheres + __float_as_int(dp[7 * kPrimThreads])) + 1);
                // attenuation of the one hit so far (rtc::scatter): albedo for lambertian and metal, 1 for a dielectric
                const bool glass = !(col.w <= 1.0f);   // (the else-branch of scatter, NaN included)
                const f3 factor = rtc::mk(glass ? 1.0f : col.x, glass ? 1.0f : col.y, glass ? 1.0f : col.z);
                const uint32_t texel_index = (uint32_t)(row0 + p) * P.width + (uint32_t)x;
                q[j] = make_float4(dp[3 * kPrimThreads], dp[4 * kPrimThreads], dp[5 * kPrimThreads], dp[0]);
                q[(size_t)P.wf_capacity + j] = make_float4(dp[kPrimThreads], dp[2 * kPrimThreads], factor.x, factor.y);
                q[2 * (size_t)P.wf_capacity + j] = make_float4(factor.z, __uint_as_float(texel_index),
                                                                __uint_as_float(2u + n + seed_bits), __uint_as_float(n));
            }
            base += (uint32_t)__popc(m);
        }
    } else if constexpr (kMode == kPrimHits) {
        // one (t, index) per texel: 8-byte stores, 256 B contiguous per warp row
#pragma unroll
        for (int p = 0; p < kPrimRays; ++p) {
            const int row = row0 + p;
            if ((uint32_t)x < P.width && (uint32_t)row < P.row_count) {
                const float* dp = s_d + (p * kPrimSlots) * kPrimThreads;
                const float code = (untraced & (1u << p)) ? __int_as_float(-2) : dp[7 * kPrimThreads];
                P.first_hit[(size_t)row * P.width + x] = make_float2(dp[6 * kPrimThreads], code);
            }
        }
    } else {
        // fused epilogue: one rolled pass over the strip (rays and their nearest hits are in the slots).  Tolerance modes:
        // a ray whose scatter sits inside a decision margin (rtc::scatter_impl, *risky) is left alone by the first
        // pass and finished with the exact arithmetic by a second one, which the hot loop so need not carry a call for
        const uint32_t sampling = L.alive | untraced;
        rtc::UnitVecCache ruv_main, ruv_redo;   // (the redo pass is exact arithmetic: its own cache)
        ruv_main.valid = false;
        ruv_redo.valid = false;
        auto shade_one = [&](int p, auto prec, rtc::UnitVecCache* ruv) -> bool {
            constexpr int FF = decltype(prec)::value;
            const int row = row0 + p;
            if (!((uint32_t)x < P.width && (uint32_t)row < P.row_count)) return false;
            float4* texel = &P.accum[(size_t)row * P.width + x];
            f3 mean = rtc::mk(0.0f, 0.0f, 0.0f);
            uint32_t n = 0;
            if (!reset) {
                const float4 t = *texel;
                mean = rtc::mk(t.x, t.y, t.z);
                n = rtc::f2u(t.w);
            }
            if (sampling & (1u << p)) {
                const float* dp = s_d + (p * kPrimSlots) * kPrimThreads;
                rtc::Ray r;
                r.d = rtc::mk(dp[0], dp[kPrimThreads], dp[2 * kPrimThreads]);
                r.o = rtc::mk(dp[3 * kPrimThreads], dp[4 * kPrimThreads], dp[5 * kPrimThreads]);
                float best_t = dp[6 * kPrimThreads];
                int best_i = __float_as_int(dp[7 * kPrimThreads]);
                if (untraced & (1u << p)) {
                    // outside primary_arm's window: the reference loop itself (compute_shader.wgsl:164-180)
                    const float2 hit = retrace_sequential(P.spheres, P.n_spheres, r.o.x, r.o.y, r.o.z, r.d.x, r.d.y, r.d.z,
                                                          rtc::dot(r.d, r.d));
                    best_t = hit.x;
                    best_i = __float_as_int(hit.y);
                }
                f3 factor = rtc::mk(1.0f, 1.0f, 1.0f);
                bool black = false;
                if (best_i >= 0) {
                    const GpuSphere sp = load_sphere(P.spheres, best_i);
                    const uint32_t seed = 1u + n + seed_bits;                       // compute_shader.wgsl:353
                    bool risky = false;
                    black = !rtc::scatter_impl<FF>(&r, &factor, best_t, sp, rtc::hash(seed + 1u), &risky, ruv);   // :268 with i = 0
                    if (FF != 0 && risky) return true;   // texel untouched: the exact pass takes it
                }
                ++rays;
                const f3 color = black ? rtc::mk(0.0f, 0.0f, 0.0f) : rtc::sky_times<FF>(factor, r.d);
                mean = running_mean<FF>(mean, color, n);
                n += 1u;
            }
            *texel = make_float4(mean.x, mean.y, mean.z, (float)n);
            // the frame's RGBA8 image in the same pass when asked for (rt_set_rgba8_output): 128 B contiguous per warp row
            // (rgba8_abs: a gather target holds the FULL image, texel (x, y) at y * width + x)
            if (P.rgba8_out)
                P.rgba8_out[(size_t)(P.rgba8_abs ? abs_row(P, row) : row) * P.width + x] = quant_texel(make_float4(mean.x, mean.y, mean.z, 0.0f));
            return false;
        };
        uint32_t redo = 0;
#pragma unroll 1
        for (int p = 0; p < kPrimRays; ++p)
            if (shade_one(p, std::integral_constant<int, F>(), &ruv_main)) redo |= 1u << p;
        if constexpr (F != 0) {
#pragma unroll 1
            while (redo) {
                const int p = __ffs((int)redo) - 1;
                redo &= redo - 1u;
                shade_one(p, std::integral_constant<int, 0>(), &ruv_redo);
            }
        }
    }
}

__device__ __forceinline__ void prim_flush(const UpdateParams& P, uint32_t exact_tests, uint32_t rays, int lane) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        exact_tests += __shfl_xor_sync(kFull, exact_tests, o);
        rays += __shfl_xor_sync(kFull, rays, o);
    }
    if (lane == 0 && exact_tests) atomicAdd(&P.counters[1], (unsigned long long)exact_tests);
    if (lane == 0 && rays) atomicAdd(&P.counters[0], (unsigned long long)rays);
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// Bulk copies into shared memory (cp.async.bulk + mbarrier), used by the bounce kernels below.
// ---------------------------------------------------------------------------------------------
namespace {

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_copy_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
    } while (!done);
}

}  // namespace

// One kernel for any scene size: persistent CTAs, every WARP pulls 32 x 8 pixel tiles from an atomic counter on its
// own (no block barrier anywhere), sets up its rays, walks the tree with the tile, scans the surviving spheres and
// finishes the texels.  Shared memory per warp: the per-ray slots (8 KB) and the warp's work stack and candidate
// lists (2.1 KB); 2 CTAs of 11 warps per SM (rt_kernels.cuh).
template <int kMode, int F>
__global__ void __launch_bounds__(kPrimThreads, kPrimCtas) primary_hits_kernel(const __grid_constant__ UpdateParams P) {
    extern __shared__ float4 s_dyn[];
    float* s_d = reinterpret_cast<float*>(s_dyn) + threadIdx.x;                       // ray slots, this thread's column
    // per warp: [kPrimList + 4] float4 records of the listed spheres, [kPrimList + 4] their indices, as many
    // nearest-distance bounds, [kPrimWork] nodes
    uint32_t* s_warp = reinterpret_cast<uint32_t*>(s_dyn) + kPrimRays * kPrimSlots * kPrimThreads +
                       (threadIdx.x >> 5) * (6 * (kPrimList + 4) + kPrimWork + kPrimSched);
    float4* s_rec = reinterpret_cast<float4*>(s_warp);
    uint32_t* s_idx = s_warp + 4 * (kPrimList + 4);
    float* s_near = reinterpret_cast<float*>(s_idx + (kPrimList + 4));
    uint32_t* s_work = s_idx + 2 * (kPrimList + 4);
    uint32_t* s_sched = s_work + kPrimWork;   // [0] the current tile's row, [1] the clock at its start
    const int lane = threadIdx.x & 31;

    const uint32_t seed_bits = rtc::f2u(P.cam.random_seed * 4294967296.0f);
    const uint32_t spp = rtc::f2u(P.cam.samples_per_pixel);
    const bool reset = P.cam.camera_has_moved > 0.5f;
    const uint32_t tiles_x = (P.width + kPrimTileW - 1) / kPrimTileW;
    const uint32_t n_tiles = tiles_x * ((P.row_count + kPrimTileH - 1) / kPrimTileH);
    PrimTree T;
    {
        const rtt::ImageLayout lay = rtt::image_layout(P.cl_n_nodes, P.cl_n_entry);
        const unsigned char* blob = static_cast<const unsigned char*>(P.cl_blob);
        T.rec = P.prim_tree;
        T.link = reinterpret_cast<const uint32_t*>(blob + lay.off_link);
        T.first_leaf = P.cl_first_leaf;
        T.n_entry = P.cl_n_entry;
        T.entry_node = reinterpret_cast<const uint32_t*>(blob + lay.off_entry_node);
        T.entry_rec = P.prim_tree + (size_t)P.cl_n_nodes * kTreeStride;
    }
    uint32_t exact_tests = 0, rays = 0;
#if RT_PRIM_PROBE
    uint32_t probe_busy = 0;
    // modes 2 / 3: when the warps end / start, in ns after the first warp's arrival (the frame's update-kernel ticket word
    // is free during a primary pass and holds that reference): rt_exact_tests = the sum over warps, rt_rays_traced = the max
    uint32_t probe_ref = 0, probe_start = 0;
    if (RT_PRIM_PROBE >= 2) {
        unsigned long long g;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g));
        if (lane == 0) {
            const uint32_t mine = (uint32_t)g | 1u;
            const uint32_t seen = atomicCAS(P.tile_counter, 0u, mine);
            probe_ref = seen ? seen : mine;
        }
        probe_ref = __shfl_sync(kFull, probe_ref, 0);
        probe_start = (uint32_t)g - probe_ref;
        if ((int32_t)probe_start < 0) probe_start = 0;
    }
#endif

    // A warp's first ticket is its number in the grid — thousands of warps drawing from one counter in the launch's first
    // microsecond queue up behind each other — and the counter hands out the tickets behind those.
    uint32_t ticket = blockIdx.x * (uint32_t)(kPrimThreads / 32) + (threadIdx.x >> 5);
#if !RT_PRIM_STATIC_FIRST
    if (lane == 0) ticket = atomicAdd(P.tile_counter + 1, 1u);   // (A/B arm: every ticket drawn)
#endif
    for (;;) {
        uint32_t tile = __shfl_sync(kFull, ticket, 0);
        if (tile >= n_tiles) break;
#if RT_PRIM_TILE_ORDER == 1
        tile = n_tiles - 1u - tile;   // bottom-up: see RT_PRIM_TILE_ORDER
#elif RT_PRIM_TILE_ORDER == 2
        {   // middle-out by tile rows: mid, mid - 1, mid + 1, mid - 2, ... (a bijection of the rows for any count)
            const uint32_t k = tile / tiles_x, mid = (n_tiles / tiles_x) / 2u;
            const uint32_t row_of = (k & 1u) ? mid - (k + 1u) / 2u : mid + k / 2u;
            tile = row_of * tiles_x + tile % tiles_x;
        }
#endif
        // ticket -> tile: the ticket's tile row goes through the frame's row order (dearest rows of the last frame first)
        // (the tile's row and starting time wait in the warp's shared memory, not in registers the tile's work is short of)
        uint32_t tile_row = tile / tiles_x;
        if (P.row_order != nullptr) tile_row = P.row_order[tile_row];
        if (lane == 0) {
            s_sched[0] = tile_row;
            s_sched[1] = clock32();
        }
        const int x = (int)((tile % tiles_x) * kPrimTileW) + lane;
        const int row0 = (int)(tile_row * kPrimTileH);
        PrimLanes L;
        uint32_t untraced;
        prim_setup<F>(P, x, row0, s_d, reinterpret_cast<float*>(s_rec) + lane, seed_bits, spp, reset, L, untraced);
        if (__any_sync(kFull, L.alive != 0u)) {
            f3 m;
            float chord;
            prim_tile_bound(L, m, chord);
            prim_tree_walk(T, s_work, s_rec, s_idx, s_near, m, chord, s_d, P, L, untraced, exact_tests, lane);
        }
        prim_finish<kMode, F>(P, x, row0, s_d, L, untraced, seed_bits, reset, rays);
        if (lane == 0) {
            const uint32_t spent = clock32() - s_sched[1];
            if (P.row_cost_cur != nullptr) atomicAdd(P.row_cost_cur + s_sched[0], spent >> 6);
#if RT_PRIM_PROBE
            probe_busy += spent;
#endif
            ticket = (RT_PRIM_STATIC_FIRST ? gridDim.x * (uint32_t)(kPrimThreads / 32) : 0u) + atomicAdd(P.tile_counter + 1, 1u);
        }
    }
#if RT_PRIM_PROBE
    exact_tests = lane == 0 ? probe_busy : 0u;   // developer probe: the counter reports the warps' busy cycles instead
    if (RT_PRIM_PROBE >= 2) {
        unsigned long long g;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g));
        const uint32_t when = RT_PRIM_PROBE == 2 ? (uint32_t)g - probe_ref : probe_start;
        if (lane == 0) {
            atomicAdd(&P.counters[1], (unsigned long long)when);
            atomicMax(&P.counters[0], (unsigned long long)when);
        }
        return;
    }
#endif
    prim_flush(P, exact_tests, rays, lane);
}

// ---------------------------------------------------------------------------------------------
// Wavefront bounces (max_depth > 1 behind a primary pass).  Paths that survive
// a bounce are compacted into a queue, so every traversal launch works on dense warps of live rays
// instead of dragging finished pixels along: the primary pass (kPrimQueue) turns first hits into colours or
// queued paths, the bounce kernel traces queue b (general 7-FFMA broad phase, records streamed through
// shared memory by bulk copies) and appends the survivors to queue b + 1.  Bounce rays are incoherent,
// so the exact tests are DEFERRED: a thread notes the groups in which one of its rays passed the broad
// test and the warp drains those notes together, every lane working on its own candidates, instead of
// the whole warp stalling on one lane's candidate.  Per-ray order of exact tests is still ascending
// index, so the result is the reference's.
// ---------------------------------------------------------------------------------------------
namespace {

// narrow_phase with the broad test re-evaluated by the same packed operations as the scan (half the instructions
// of the scalar form; the values are identical, each half of an FFMA2 being the IEEE fused operation).
template <int NP>
__device__ __forceinline__ void narrow_phase_packed(const float4* __restrict__ s_geom, int slot, int count, int index0,
                                                    const GpuSphere* __restrict__ spheres, uint32_t n_spheres, Lanes& L,
                                                    uint32_t& exact_tests) {
#pragma unroll 1
    for (int u = 0; u < count; ++u) {
        const int idx = index0 + u;
        if ((uint32_t)idx >= n_spheres) break;   // padding
        const float4 sp = s_geom[slot + u];
        bool cand[2 * NP];
        bool some = false;
#pragma unroll
        for (int q = 0; q < NP; ++q) {
            const float2 g = broad_g2(sp, L, q);
            cand[2 * q] = L.alive[2 * q] && !(g.x < L.thr[2 * q]);
            cand[2 * q + 1] = L.alive[2 * q + 1] && !(g.y < L.thr[2 * q + 1]);
            some |= cand[2 * q] | cand[2 * q + 1];
        }
        if (!some) continue;
        const float radius = __ldg(&spheres[idx].radius);
        const f3 center = rtc::mk(sp.x, sp.y, sp.z);
#pragma unroll
        for (int p = 0; p < 2 * NP; ++p) {
            if (cand[p]) {
                ++exact_tests;
                float root;
                if (rtc::sphere_root(origin_of(L, p), rtc::mk(L.dx[p], L.dy[p], L.dz[p]), L.a[p], center, radius, kTMin,
                                     L.best_t[p], &root)) {
                    L.best_t[p] = root;
                    L.best_i[p] = idx;
                }
            }
        }
    }
}

// scan_spheres with the exact tests deferred (see the section header).  s_note: this thread's column of
// [kWfDefer][kBlockThreads] group offsets.
template <int NP>
__device__ __forceinline__ void drain_notes(const float4* __restrict__ s_geom, int index0, const unsigned short* __restrict__ s_note,
                                            int& n_notes, const GpuSphere* __restrict__ spheres, uint32_t n_spheres, Lanes& L,
                                            uint32_t& exact_tests) {
    for (int e = 0; __any_sync(kFull, e < n_notes); ++e) {
        if (e < n_notes) {
            const int s = (int)s_note[e * kBlockThreads];
            narrow_phase_packed<NP>(s_geom, s, kUnroll, index0 + s, spheres, n_spheres, L, exact_tests);
        }
    }
    n_notes = 0;
}

template <int NP>
__device__ __forceinline__ void scan_spheres_deferred(const float4* __restrict__ s_geom, int count, int index0,
                                                      unsigned short* __restrict__ s_note, const GpuSphere* __restrict__ spheres,
                                                      uint32_t n_spheres, Lanes& L, uint32_t& exact_tests) {
    uint32_t addr = (uint32_t)__cvta_generic_to_shared(s_geom);
    int n_notes = 0;
    float4 cur[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) cur[u] = lds128(addr + 16u * u);
#pragma unroll 1
    for (int s = 0; s < count; s += kUnroll) {
        addr += 16u * kUnroll;
        float4 nxt[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) nxt[u] = lds128(addr + 16u * u);
        bool any = false;
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
#pragma unroll
            for (int q = 0; q < NP; ++q) {
                const float2 g = broad_g2(cur[u], L, q);
                any |= !(g.x < L.thr[2 * q]);
                any |= !(g.y < L.thr[2 * q + 1]);
            }
        }
        s_note[n_notes * kBlockThreads] = (unsigned short)s;   // slot n_notes < kWfDefer is free; kept only if `any`
        n_notes += any ? 1 : 0;
        if (__any_sync(kFull, n_notes >= kWfDefer)) drain_notes<NP>(s_geom, index0, s_note, n_notes, spheres, n_spheres, L, exact_tests);
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) cur[u] = nxt[u];
    }
    drain_notes<NP>(s_geom, index0, s_note, n_notes, spheres, n_spheres, L, exact_tests);
}

}  // namespace

namespace {

// Rays j0 + p (p < NR) of the queue — consecutive entries, paths of neighbouring pixels, so the two rays of an FFMA2
// pair tend to meet the same clusters — into the thread's first NR lanes; the others idle.
template <int NR, int F = 0>
__device__ __forceinline__ void wf_load_rays(const UpdateParams& P, const float4* __restrict__ qin, uint32_t count,
                                             uint32_t j0, Lanes& L, uint32_t& rays) {
#pragma unroll
    for (int p = 0; p < kPix; ++p) {
        const uint32_t j = j0 + (uint32_t)p;
        if (p < NR && j < count) {
            const float4 A = __ldcg(qin + j), B = __ldcg(qin + (size_t)P.wf_capacity + j);
            arm_ray<F>(L, p, rtc::mk(A.x, A.y, A.z), rtc::mk(A.w, B.x, B.y));
            ++rays;
        } else {
            idle_ray(L, p);
        }
    }
}

// Scatter / finish / re-queue the traced rays of the thread.  Called by whole warps.
template <int NR>
__device__ __forceinline__ void wf_resolve(const UpdateParams& P, const float4* __restrict__ qin, uint32_t count, uint32_t j0,
                                           const Lanes& L, uint32_t bounce, bool last, bool reset, int lane) {
#pragma unroll 1
    for (int p = 0; p < NR; ++p) {
        const uint32_t j = j0 + (uint32_t)p;
        bool push = false;
        f3 o = rtc::mk(0.0f, 0.0f, 0.0f), d = o, factor = o;
        uint32_t texel_index = 0, bseed = 0, n = 0;
        if (j < count) {
            const float4 B = __ldcg(qin + (size_t)P.wf_capacity + j), C = __ldcg(qin + 2 * (size_t)P.wf_capacity + j);
            factor = rtc::mk(B.z, B.w, C.x);
            texel_index = __float_as_uint(C.y);
            bseed = __float_as_uint(C.z);
            n = __float_as_uint(C.w);
            float best_t = kTMax;
            int best_i = -1;
            // p is a run-time index here: pick this ray's state without dynamic register indexing
#pragma unroll
            for (int k = 0; k < kPix; ++k) {
                if (k == p) {
                    o = origin_of(L, k);
                    d = rtc::mk(L.dx[k], L.dy[k], L.dz[k]);
                    best_t = L.best_t[k];
                    best_i = L.best_i[k];
                }
            }
            push = resolve_hit(P, bounce, last, best_i, best_t, o, d, factor, texel_index, bseed, n, reset);
        }
        push_path(P, bounce + 1u, push, o, d, factor, texel_index, bseed, n, lane);
    }
}


// The same for a 64-ray batch of the tree kernel (2 rays per thread: queue entries j0 and j0 + 1), with ONE reservation
// per batch and the survivors appended in their queue order.  (Two appends per batch, one per p, put the even entries'
// survivors before the odd ones — a perfect unshuffle per bounce, so the rays of an FFMA2 pair drift 2^k entries apart —
// and let other warps' appends fall between.)  Ray 0's outcome waits in `s_park`, 12 words of the thread's idle candidate
// column (its OWN column: a wider view of the area would land in other warps' cells, which are in use), while the
// rolled loop resolves ray 1.
// Per-thread columns of the bounce kernels (node stack, candidate codes): row k of thread t.  Plain layout: element
// k * kBlockThreads + t — for 16-bit entries lanes 2j and 2j + 1 share a bank, and once their rows differ (divergent
// stack depths) every access is a 2-way conflict (16 % of the shared-memory wavefronts of a bounce launch).  Paired
// layout (RT_WF_PAIR_ROWS): rows 2m and 2m + 1 of a thread share ONE 32-bit word, element (m * kBlockThreads + t) * 2 +
// (k & 1), so lane t always hits bank t whatever its row.
#ifndef RT_WF_PAIR_ROWS
#define RT_WF_PAIR_ROWS 0
#endif
#ifndef RT_WF_COOP_GLOBAL
#define RT_WF_COOP_GLOBAL 1   /* trees walked in GLOBAL memory use the warp-cooperative traversal (trace_tree_coop) */
#endif
// A word of a warp's cooperative queue area used as parking space by wf_resolve_batch: column stride 32, not a CTA's.
enum CoopWord : uint32_t {};
template <typename T>
__device__ __forceinline__ int col_at(int k) {
    if constexpr (std::is_same<T, CoopWord>::value) return k * 32;   // a column of the WARP's queue area
    else if constexpr (sizeof(T) == 2 && RT_WF_PAIR_ROWS) return (k >> 1) * (2 * kBlockThreads) + (k & 1);
    else return k * kBlockThreads;
}
template <typename T>
__device__ __forceinline__ T* col_base(T* block_base) {
    if constexpr (sizeof(T) == 2 && RT_WF_PAIR_ROWS) return block_base + 2 * threadIdx.x;
    else return block_base + threadIdx.x;
}

template <typename CandT>
__device__ __forceinline__ void park_word(CandT* s_cand, int k, uint32_t v) {
    if constexpr (sizeof(CandT) == 4) {
        s_cand[col_at<CandT>(k)] = (CandT)v;
    } else {
        s_cand[col_at<CandT>(2 * k)] = (CandT)(v & 0xffffu);
        s_cand[col_at<CandT>(2 * k + 1)] = (CandT)(v >> 16);
    }
}
template <typename CandT>
__device__ __forceinline__ uint32_t parked_word(const CandT* s_cand, int k) {
    if constexpr (sizeof(CandT) == 4) return (uint32_t)s_cand[col_at<CandT>(k)];
    else return (uint32_t)s_cand[col_at<CandT>(2 * k)] | ((uint32_t)s_cand[col_at<CandT>(2 * k + 1)] << 16);
}
template <int F, typename CandT>
__device__ __forceinline__ void wf_resolve_batch(const UpdateParams& P, const float4* __restrict__ qin, uint32_t count, uint32_t j0,
                                                 const Lanes& L, uint32_t bounce, bool last, bool reset, int lane,
                                                 CandT* s_park) {
    static_assert(kWfCand * sizeof(CandT) >= 12 * sizeof(uint32_t), "12 parked words per thread");
    bool push = false, push0 = false;
    f3 o = rtc::mk(0.0f, 0.0f, 0.0f), d = o, factor = o;
    uint32_t texel_index = 0, bseed = 0, n = 0;
    rtc::UnitVecCache ruv_batch;   // the thread's two paths usually share their bounce seed
    ruv_batch.valid = false;
    __syncwarp();
#pragma unroll 1
    for (int p = 0; p < 2; ++p) {
        const uint32_t j = j0 + (uint32_t)p;
        push = false;
        if (j < count) {
            const float4 B = __ldcg(qin + (size_t)P.wf_capacity + j), C = __ldcg(qin + 2 * (size_t)P.wf_capacity + j);
            factor = rtc::mk(B.z, B.w, C.x);
            texel_index = __float_as_uint(C.y);
            bseed = __float_as_uint(C.z);
            n = __float_as_uint(C.w);
            const bool second = p != 0;   // picked with selects: no dynamic register indexing
            o = second ? origin_of(L, 1) : origin_of(L, 0);
            d = second ? rtc::mk(L.dx[1], L.dy[1], L.dz[1]) : rtc::mk(L.dx[0], L.dy[0], L.dz[0]);
            const float best_t = second ? L.best_t[1] : L.best_t[0];
            const int best_i = second ? L.best_i[1] : L.best_i[0];
            push = resolve_hit<F>(P, bounce, last, best_i, best_t, o, d, factor, texel_index, bseed, n, reset, &ruv_batch);
        }
        if (p == 0) {
            push0 = push;
            if (push) {
                park_word(s_park, 0, __float_as_uint(o.x)); park_word(s_park, 1, __float_as_uint(o.y));
                park_word(s_park, 2, __float_as_uint(o.z)); park_word(s_park, 3, __float_as_uint(d.x));
                park_word(s_park, 4, __float_as_uint(d.y)); park_word(s_park, 5, __float_as_uint(d.z));
                park_word(s_park, 6, __float_as_uint(factor.x)); park_word(s_park, 7, __float_as_uint(factor.y));
                park_word(s_park, 8, __float_as_uint(factor.z)); park_word(s_park, 9, texel_index);
                park_word(s_park, 10, bseed); park_word(s_park, 11, n);
            }
        }
    }
    const unsigned m0 = __ballot_sync(kFull, push0), m1 = __ballot_sync(kFull, push);
    const uint32_t total = (uint32_t)(__popc(m0) + __popc(m1));
    if (total != 0u) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(&P.wf_counts[bounce + 1u], total);
        base = __shfl_sync(kFull, base, 0);
        const unsigned lt = (1u << lane) - 1u;
        const uint32_t rank0 = base + (uint32_t)(__popc(m0 & lt) + __popc(m1 & lt));
        float4* q = P.wf_queue[(bounce + 1u) & 1u];
        if (push0) {
            q[rank0] = make_float4(__uint_as_float(parked_word(s_park, 0)), __uint_as_float(parked_word(s_park, 1)),
                                   __uint_as_float(parked_word(s_park, 2)), __uint_as_float(parked_word(s_park, 3)));
            q[(size_t)P.wf_capacity + rank0] =
                make_float4(__uint_as_float(parked_word(s_park, 4)), __uint_as_float(parked_word(s_park, 5)),
                            __uint_as_float(parked_word(s_park, 6)), __uint_as_float(parked_word(s_park, 7)));
            q[2 * (size_t)P.wf_capacity + rank0] =
                make_float4(__uint_as_float(parked_word(s_park, 8)), __uint_as_float(parked_word(s_park, 9)),
                            __uint_as_float(parked_word(s_park, 10)), __uint_as_float(parked_word(s_park, 11)));
        }
        if (push) {
            const uint32_t j = rank0 + (push0 ? 1u : 0u);
            q[j] = make_float4(o.x, o.y, o.z, d.x);
            q[(size_t)P.wf_capacity + j] = make_float4(d.y, d.z, factor.x, factor.y);
            q[2 * (size_t)P.wf_capacity + j] =
                make_float4(factor.z, __uint_as_float(texel_index), __uint_as_float(bseed), __uint_as_float(n));
        }
    }
    __syncwarp();   // the parked words are candidate codes again from the next batch on
}

// ---------------------------------------------------------------------------------------------
// Tree traversal for incoherent rays (rt_tree.h).  The scene image in shared memory is
// [child records, 9 slots per node | child links, 8 per node | largest child radius per node].  A thread traces one
// FFMA2 pair of rays: it keeps a private stack of node numbers (a column of shared memory), pops a node, runs the
// 7-FFMA2 test against its 8 child records branch-free (empty slots never pass) — with the ray's lowered threshold
// (rtc::cluster_thr) for bounding spheres, with the ordinary one for the spheres under a leaf node — and pushes the
// children either ray passes, or, under a leaf node, appends (sphere, ray) candidates to its list.  The warp drains the
// candidate lists together, one exact test per lane and iteration.  The exact tests do not come in index order, so the
// nearest hit is kept with rtc::nearest_update (ties to the lower index); a NaN root makes the reference's result
// order dependent, and such a ray is retraced sequentially.
// ---------------------------------------------------------------------------------------------
struct TreeView {
    const float4* rec;
    const uint32_t* link;
    const float* node_r;
    const float4* entry_rec;            // the entry list (rt_tree.h): records ...
    const uint32_t* entry_node;         // ... and the nodes they stand for
    int first_leaf, n_entry_pad, stack_cap;
    float entry_r;
};

__device__ __forceinline__ TreeView tree_view(const void* blob, uint32_t n_nodes, uint32_t first_leaf, uint32_t n_entry,
                                              float entry_r, uint32_t stack_cap) {
    const unsigned char* b = static_cast<const unsigned char*>(blob);
    const rtt::ImageLayout lay = rtt::image_layout(n_nodes, n_entry);
    TreeView tv;
    tv.rec = reinterpret_cast<const float4*>(b);
    tv.link = reinterpret_cast<const uint32_t*>(b + lay.off_link);
    tv.node_r = reinterpret_cast<const float*>(b + lay.off_r);
    tv.entry_rec = reinterpret_cast<const float4*>(b + lay.off_entry_rec);
    tv.entry_node = reinterpret_cast<const uint32_t*>(b + lay.off_entry_node);
    tv.first_leaf = (int)first_leaf;
    tv.n_entry_pad = (int)lay.n_entry_pad;
    tv.stack_cap = (int)stack_cap;
    tv.entry_r = entry_r;
    return tv;
}

// One exact test per lane and iteration: every lane works on its own (sphere, ray) candidate, so the reference
// quadratic runs on full warps instead of on the one lane that happened to find a candidate.
template <typename CandT>
__device__ __forceinline__ void drain_candidates(const CandT* __restrict__ s_cand, int& n_cand,
                                                 const GpuSphere* __restrict__ spheres, Lanes& L, uint32_t& exact_tests,
                                                 uint32_t& nan_rays) {
    for (int e = 0; __any_sync(kFull, e < n_cand); ++e) {
        if (e < n_cand) {
            constexpr int kRayBit = (int)sizeof(CandT) * 8 - 1;   // top bit: the thread's second ray
            const uint32_t code = (uint32_t)s_cand[col_at<CandT>(e)];
            const int idx = (int)(code & ((1u << kRayBit) - 1u));
            const bool second = (code >> kRayBit) != 0u;
            const float4 gs = __ldg(reinterpret_cast<const float4*>(spheres + idx));   // centre, radius as uploaded
            // the ray is picked with selects: one copy of the quadratic for both rays
            const f3 o = second ? origin_of(L, 1) : origin_of(L, 0);
            const f3 d = second ? rtc::mk(L.dx[1], L.dy[1], L.dz[1]) : rtc::mk(L.dx[0], L.dy[0], L.dz[0]);
            const float a = second ? L.a[1] : L.a[0];
            float bt = second ? L.best_t[1] : L.best_t[0];
            int bi = second ? L.best_i[1] : L.best_i[0];
            ++exact_tests;
            if (!rtc::nearest_update_from_inside(o, d, a, rtc::mk(gs.x, gs.y, gs.z), gs.w, kTMin, kTMax, idx, &bt, &bi))
                nan_rays |= second ? 2u : 1u;
            if (second) {
                L.best_t[1] = bt;
                L.best_i[1] = bi;
            } else {
                L.best_t[0] = bt;
                L.best_i[0] = bi;
            }
        }
    }
    n_cand = 0;
}

// Nearest hits of the thread's rays 0 and 1 through the tree.  s_stack / s_cand: this thread's columns.
template <typename CandT>
__device__ __forceinline__ void trace_tree(const TreeView& tv, unsigned short* __restrict__ s_stack,
                                           CandT* __restrict__ s_cand, const GpuSphere* __restrict__ spheres,
                                           uint32_t n_spheres, Lanes& L, uint32_t& exact_tests) {
    const float inf = __int_as_float(0x7f800000);
    const float thr0 = L.alive[0] ? L.thr[0] : inf, thr1 = L.alive[1] ? L.thr[1] : inf;
    const f3 o0 = origin_of(L, 0), o1 = origin_of(L, 1);
    const float m0 = rtc::cluster_m(rtc::dot(o0, o0)), m1 = rtc::cluster_m(rtc::dot(o1, o1));
    uint32_t nan_rays = 0;
    int n_stack = 0, n_cand = 0;
    auto traverse = [&]() {   // empties every lane's stack
    while (__any_sync(kFull, n_stack > 0)) {
        if (n_stack > 0) {
            const int node = (int)s_stack[col_at<unsigned short>(--n_stack)];
            const bool leaf = node >= tv.first_leaf;
            float t0 = thr0, t1 = thr1;
            if (!leaf) {
                const float r = tv.node_r[node];
                t0 = rtc::cluster_thr(thr0, m0, r);
                t1 = rtc::cluster_thr(thr1, m1, r);
            }
            const float4* rec = tv.rec + node * kTreeStride;
            uint32_t pass = 0;   // bit 2k: ray 0 passes child k, bit 2k + 1: ray 1
#pragma unroll
            for (int k = 0; k < kTreeWide; ++k) {
                const float2 g = broad_g2(rec[k], L, 0);
                pass |= (!(g.x < t0) ? 1u : 0u) << (2 * k);
                pass |= (!(g.y < t1) ? 1u : 0u) << (2 * k + 1);
            }
            const uint32_t* link = tv.link + node * kTreeWide;
            if (leaf) {
                while (pass) {
                    const int bit = __ffs((int)pass) - 1;
                    pass &= pass - 1u;
                    const unsigned idx = link[bit >> 1];
                    if (idx != 0xFFFFFFFFu) {   // (an empty slot can only get here through a NaN threshold)
                        s_cand[col_at<CandT>(n_cand)] = (CandT)(idx | ((unsigned)(bit & 1) << ((int)sizeof(CandT) * 8 - 1)));
                        ++n_cand;
                    }
                }
            } else {
                pass = (pass | (pass >> 1)) & 0x5555u;   // either ray
                while (pass) {
                    const int bit = __ffs((int)pass) - 1;
                    pass &= pass - 1u;
                    const unsigned child = link[bit >> 1];
                    if (child != 0xFFFFFFFFu) {
                        s_stack[col_at<unsigned short>(n_stack)] = (unsigned short)child;
                        ++n_stack;
                    }
                }
            }
        }
        // a leaf node adds at most 2 * kTreeWide candidates
        if (__any_sync(kFull, n_cand > kWfCand - 2 * kTreeWide)) drain_candidates(s_cand, n_cand, spheres, L, exact_tests, nan_rays);
    }
    };
    // the entry list, all lanes on the same record: the nodes either ray passes start the lane's stack
    {
        const float e0 = rtc::cluster_thr(thr0, m0, tv.entry_r), e1 = rtc::cluster_thr(thr1, m1, tv.entry_r);
#pragma unroll 1
        for (int c0 = 0; c0 < tv.n_entry_pad; c0 += 4) {
            float4 rec[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) rec[u] = tv.entry_rec[c0 + u];
            const uint4 nodes4 = *reinterpret_cast<const uint4*>(tv.entry_node + c0);
            const uint32_t node_of[4] = {nodes4.x, nodes4.y, nodes4.z, nodes4.w};
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const float2 g = broad_g2(rec[u], L, 0);
                const bool pass = !(g.x < e0) || !(g.y < e1);
                s_stack[col_at<unsigned short>(n_stack)] = (unsigned short)node_of[u];   // slot n_stack is free; kept only if `pass`
                n_stack += (pass && node_of[u] != 0xFFFFFFFFu) ? 1 : 0;
            }
            // the stack keeps room for 4 more entries and for the growth of the traversal below them (rt_api.cu)
            if (__any_sync(kFull, n_stack > kTreeBacklog)) traverse();
        }
    }
    traverse();
    drain_candidates(s_cand, n_cand, spheres, L, exact_tests, nan_rays);
    if (nan_rays) {
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            if (nan_rays & (1u << p)) {
                const float2 hit = retrace_sequential(spheres, n_spheres, L.ox[p], L.oy[p], L.oz[p], L.dx[p], L.dy[p], L.dz[p], L.a[p]);
                L.best_t[p] = hit.x;
                L.best_i[p] = __float_as_int(hit.y);
                exact_tests += n_spheres;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Warp-cooperative traversal (trees walked in GLOBAL memory, RT_WF_COOP_GLOBAL).  trace_tree above keeps a stack per THREAD, and the warp runs as long as its
// busiest lane: from the second bounce on the 64 rays of a batch come from all over the image and 15-17 of 32 lanes are
// active per instruction (profiles/r2_bounce_ncu.md).  Here the work belongs to the WARP: the entry scan (coherent, all
// lanes on the same record, two rays per lane in an FFMA2 pair — unchanged) emits (ray, node) items into a queue in the
// warp's shared memory; each step the 32 top items go one to a lane, which tests the node's 8 child records for THAT ray
// (scalar FFMA: the ray's parameters come from shared memory) and pushes (ray, child) items or, under a leaf node,
// (ray, sphere) candidates through a warp prefix sum; candidates are drained 32 at a time, one reference quadratic per
// lane.  A lane is busy whenever the warp has 32 items, whatever ray they belong to, and rays that ended (to-completion
// batches of finish_paths) cost nothing.  The nearest hit per ray is the 64-bit minimum of (bits of t, sphere index) —
// t > tmin > 0, so the float's bits order as the value does — kept with atomicMin in shared memory: the lexicographic
// (t, index) minimum of rtc::nearest_update, whatever the order of the tests.  A NaN root flags its ray for the
// sequential retrace as before.  Same candidate set, same exact test, same result as trace_tree, bit for bit.
// Measured (B200, depth-8 frames): where the tree image sits in SHARED memory the per-thread traversal wins — 1,024
// spheres at 4K: 2.55 ms against 3.60 cooperative, depth 30: 3.78 against 4.65 (an FFMA2 serves two rays per
// instruction there and a step's queue bookkeeping costs as much as its 56 FMAs) — so that path keeps trace_tree.  Where
// the tree is walked in GLOBAL memory the walk is latency-bound and full lanes hide it: 65,536 spheres at 1080p,
// 8.0 -> 5.3 ms.  The kGlobal instantiations use this traversal.
//   per warp: 64 rays x 4 float4 (u, h0 | -2o, thr | o, a | d, m), 64 x u64 nearest keys, kCoopItems node items
//   (ray << 16 | node), kCoopCands candidates (ray << 26 | sphere), 2 words of NaN flags.
// ---------------------------------------------------------------------------------------------
constexpr int kCoopItems = 768, kCoopCands = 512;
constexpr int kCoopStepGrowth = 32 * (kTreeWide - 1);   // a full step pops 32 items and pushes at most 32 x 8
constexpr int kCoopDfsReserve = 7 * 16;                 // one-node-at-a-time descent: + 7 items per level
constexpr size_t kCoopWarpBytes = 64 * 4 * sizeof(float4) + 64 * sizeof(unsigned long long) + (kCoopItems + kCoopCands) * sizeof(uint32_t) + 16;
static_assert(kCoopWarpBytes % 16 == 0, "warp regions stay 16-byte aligned");

struct CoopWarp {
    float4* ray;
    unsigned long long* best;
    uint32_t* items;
    uint32_t* cands;
    uint32_t* nan;
};
__device__ __forceinline__ CoopWarp coop_warp(void* cta_base, int warp) {
    unsigned char* b = static_cast<unsigned char*>(cta_base) + (size_t)warp * kCoopWarpBytes;
    CoopWarp w;
    w.ray = reinterpret_cast<float4*>(b);
    w.best = reinterpret_cast<unsigned long long*>(b + 64 * 4 * sizeof(float4));
    w.items = reinterpret_cast<uint32_t*>(b + 64 * 4 * sizeof(float4) + 64 * sizeof(unsigned long long));
    w.cands = w.items + kCoopItems;
    w.nan = w.cands + kCoopCands;
    return w;
}

// Inclusive warp prefix sum of a packed pair of counts (low 16 bits | high 16 bits; sums stay below 2^16).
__device__ __forceinline__ uint32_t warp_scan_incl(uint32_t v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(kFull, v, o);
        if (lane >= o) v += t;
    }
    return v;
}

__device__ __forceinline__ void coop_drain(const CoopWarp& w, int& n_cand, int keep_below, const GpuSphere* __restrict__ spheres,
                                           uint32_t n_spheres, uint32_t& exact_tests, int lane) {
    // full steps of 32 while more than `keep_below` candidates wait (0: everything)
    while (n_cand > keep_below) {
        const int take = min(32, n_cand);
        n_cand -= take;
        if (lane < take) {
            const uint32_t c = w.cands[n_cand + lane];
            const uint32_t ray = c >> 26, idx = c & 0x3FFFFFFu;
            if (idx < n_spheres) {   // (an empty slot can only get here through a NaN threshold)
                const float4 q2 = w.ray[ray * 4 + 2], q3 = w.ray[ray * 4 + 3];
                const float4 gs = __ldg(reinterpret_cast<const float4*>(spheres + idx));   // centre, radius as uploaded
                float bt = kTMax;
                int bi = -1;
                ++exact_tests;
                if (!rtc::nearest_update_from_inside(rtc::mk(q2.x, q2.y, q2.z), rtc::mk(q3.x, q3.y, q3.z), q2.w, rtc::mk(gs.x, gs.y, gs.z),
                                                     gs.w, kTMin, kTMax, (int)idx, &bt, &bi))
                    atomicOr(&w.nan[ray >> 5], 1u << (ray & 31u));
                else if (bi >= 0)
                    atomicMin(&w.best[ray], ((unsigned long long)__float_as_uint(bt) << 32) | (unsigned long long)idx);
            }
        }
        __syncwarp();
    }
}

// Works the node queue down to `keep_below` items (0: empty).
__device__ __forceinline__ void coop_nodes(const TreeView& tv, const CoopWarp& w, int& n_items, int& n_cand, int keep_below,
                                           const GpuSphere* __restrict__ spheres, uint32_t n_spheres, uint32_t& exact_tests, int lane) {
    while (n_items > keep_below) {
        if (n_cand > kCoopCands - 32 * kTreeWide) coop_drain(w, n_cand, 31, spheres, n_spheres, exact_tests, lane);
        // a nearly full queue is worked off one node at a time, depth first: it then grows by at most 7 per level
        const int take = (n_items + kCoopStepGrowth + kCoopDfsReserve <= kCoopItems) ? min(32, n_items) : 1;
        n_items -= take;
        uint32_t pass = 0, ray = 0, node = 0;
        bool leaf = false;
        if (lane < take) {
            const uint32_t it = w.items[n_items + lane];
            ray = it >> 16;
            node = it & 0xFFFFu;
            if (node != 0xFFFFu) {   // (the link of an empty slot)
                leaf = (int)node >= tv.first_leaf;
                const float4 q0 = w.ray[ray * 4], q1 = w.ray[ray * 4 + 1];
                float thr = q1.w;
                if (!leaf) thr = rtc::cluster_thr(thr, w.ray[ray * 4 + 3].w, tv.node_r[node]);
                const float4* rec = tv.rec + node * kTreeStride;
#pragma unroll
                for (int k = 0; k < kTreeWide; ++k) {
                    const float4 r = rec[k];
                    const float g = rtc::broad_g(r.x, r.y, r.z, r.w, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z);
                    pass |= (!(g < thr) ? 1u : 0u) << k;
                }
            }
        }
        __syncwarp();   // the taken items have been read before their slots are reused
        const uint32_t c = (uint32_t)__popc(pass);
        const uint32_t mine = leaf ? (c << 16) : c;
        const uint32_t incl = warp_scan_incl(mine, lane);
        const uint32_t total = __shfl_sync(kFull, incl, 31);
        const uint32_t excl = incl - mine;
        const uint32_t* link = tv.link + node * kTreeWide;
        if (leaf) {
            uint32_t pos = (uint32_t)n_cand + (excl >> 16);
            while (pass) {
                const int k = __ffs((int)pass) - 1;
                pass &= pass - 1u;
                w.cands[pos++] = (ray << 26) | (link[k] & 0x3FFFFFFu);
            }
        } else {
            uint32_t pos = (uint32_t)n_items + (excl & 0xFFFFu);
            while (pass) {
                const int k = __ffs((int)pass) - 1;
                pass &= pass - 1u;
                w.items[pos++] = (ray << 16) | (link[k] & 0xFFFFu);
            }
        }
        n_items += (int)(total & 0xFFFFu);
        n_cand += (int)(total >> 16);
        __syncwarp();
    }
}

// Nearest hits of the thread's rays 0 and 1 (rays 2 lane and 2 lane + 1 of the warp's batch), cooperatively.
__device__ __forceinline__ void trace_tree_coop(const TreeView& tv, const CoopWarp& w, const GpuSphere* __restrict__ spheres,
                                                uint32_t n_spheres, Lanes& L, uint32_t& exact_tests) {
    const int lane = threadIdx.x & 31;
    const float inf = __int_as_float(0x7f800000);
    const float thr0 = L.alive[0] ? L.thr[0] : inf, thr1 = L.alive[1] ? L.thr[1] : inf;
    const f3 o0 = origin_of(L, 0), o1 = origin_of(L, 1);
    const float m0 = rtc::cluster_m(rtc::dot(o0, o0)), m1 = rtc::cluster_m(rtc::dot(o1, o1));
    __syncwarp();   // (the queue area may have held the previous batch's parked words)
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        float4* q = w.ray + (2 * lane + p) * 4;
        q[0] = make_float4(half(L.ux[0], p), half(L.uy[0], p), half(L.uz[0], p), half(L.h0[0], p));
        q[1] = make_float4(half(L.mox[0], p), half(L.moy[0], p), half(L.moz[0], p), p ? thr1 : thr0);
        q[2] = make_float4(L.ox[p], L.oy[p], L.oz[p], L.a[p]);
        q[3] = make_float4(L.dx[p], L.dy[p], L.dz[p], p ? m1 : m0);
        w.best[2 * lane + p] = ((unsigned long long)__float_as_uint(kTMax) << 32) | 0xFFFFFFFFull;
    }
    if (lane < 2) w.nan[lane] = 0u;
    __syncwarp();
    int n_items = 0, n_cand = 0;
    // the entry list, all lanes on the same record: one pass bit per (ray, entry), 32 entries at a time, then the set
    // bits become (ray, node) items, one per lane and round
    const float e0 = rtc::cluster_thr(thr0, m0, tv.entry_r), e1 = rtc::cluster_thr(thr1, m1, tv.entry_r);
#pragma unroll 1
    for (int w0 = 0; w0 < tv.n_entry_pad; w0 += 32) {
        const int wn = min(32, tv.n_entry_pad - w0);   // (a multiple of 4)
        uint32_t bits0 = 0u, bits1 = 0u;
#pragma unroll 4
        for (int i = 0; i < wn; ++i) {
            const float2 g = broad_g2(tv.entry_rec[w0 + i], L, 0);
            bits0 = (bits0 >> 1) | (!(g.x < e0) ? 0x80000000u : 0u);
            bits1 = (bits1 >> 1) | (!(g.y < e1) ? 0x80000000u : 0u);
        }
        bits0 >>= (32 - wn);
        bits1 >>= (32 - wn);
        while (__any_sync(kFull, (bits0 | bits1) != 0u)) {
            if (n_items > kCoopItems - kCoopStepGrowth - kCoopDfsReserve - 32)
                coop_nodes(tv, w, n_items, n_cand, 31, spheres, n_spheres, exact_tests, lane);
            const bool first = bits0 != 0u;
            const uint32_t bits = first ? bits0 : bits1;
            const bool has = bits != 0u;
            const int e = __ffs((int)bits) - 1;
            if (first) bits0 &= bits0 - 1u; else bits1 &= bits1 - 1u;
            const unsigned m = __ballot_sync(kFull, has);
            if (has) {
                const uint32_t node = tv.entry_node[w0 + e];
                w.items[n_items + __popc(m & ((1u << lane) - 1u))] = ((uint32_t)(2 * lane + (first ? 0 : 1)) << 16) | (node & 0xFFFFu);
            }
            n_items += __popc(m);
            __syncwarp();
        }
    }
    coop_nodes(tv, w, n_items, n_cand, 0, spheres, n_spheres, exact_tests, lane);
    coop_drain(w, n_cand, 0, spheres, n_spheres, exact_tests, lane);
    __syncwarp();
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        const unsigned long long key = w.best[2 * lane + p];
        L.best_t[p] = __uint_as_float((uint32_t)(key >> 32));
        L.best_i[p] = (int)(uint32_t)(key & 0xFFFFFFFFull);
        if ((w.nan[(2 * lane + p) >> 5] >> ((2 * lane + p) & 31)) & 1u) {
            const float2 hit = retrace_sequential(spheres, n_spheres, L.ox[p], L.oy[p], L.oz[p], L.dx[p], L.dy[p], L.dz[p], L.a[p]);
            L.best_t[p] = hit.x;
            L.best_i[p] = __float_as_int(hit.y);
            exact_tests += n_spheres;
        }
    }
    __syncwarp();   // the queues may become parking space (wf_resolve_batch)
}

// Few paths left (less than RT_WF_COMPLETE_BELOW rays per resident warp): finish them inside ONE launch, every
// remaining bounce, instead of paying a launch and a pass latency per bounce for a nearly empty machine.  Nothing is
// queued, so the later bounce launches find empty queues and return.  Kept out of line: inlined, its live state
// costs the full-queue loops of bounce_kernel 3 % (measured), and it runs when the machine is mostly idle anyway.
#ifndef RT_WF_COMPLETE_BELOW
#define RT_WF_COMPLETE_BELOW 64u   /* one 64-ray batch per resident warp (A/B on B200: 16 / 32 / 48 / 64 / 80 -> 4.52 / 4.49 / 4.41 / 4.31 / 4.60 ms per depth-30 4K frame) */
#endif
template <int F, bool kCoop, typename CandT>
__device__ __noinline__ void finish_paths(const UpdateParams& P, const float4* __restrict__ qin, uint32_t count,
                                          const TreeView& tv, unsigned short* __restrict__ s_stack, CandT* __restrict__ s_cand,
                                          void* coop_base, unsigned int* ticket, uint32_t bounce, uint32_t max_depth, bool reset, int lane,
                                          uint32_t batch_rays, uint32_t stop_at,
                                          uint32_t& exact_tests_out, uint32_t& rays_out, uint32_t& rays_later_out) {
    // batch_rays: paths a warp takes at a time — 64 (two per thread, consecutive queue entries) or, for a queue that has
    // been thinned out and re-dealt (bounce_tail_kernel), an even share of at most 32, one per thread.
    // stop_at: 0, or the bounce at which the surviving paths go back to the queue (its counter is P.wf_counts[stop_at])
    // instead of being traced on: stop_at - bounce is odd, so that queue is not the one this call loads from.
    const uint32_t n_batches = (count + batch_rays - 1u) / batch_rays;
    const bool paired = batch_rays > 32u;
    unsigned int* traced = &P.wf_counts[2u * (kWfMaxDepth + 1u)];
    uint32_t exact_tests = 0, rays = 0, rays_later = 0;
    for (;;) {
        uint32_t batch = 0;
        if (lane == 0) batch = atomicAdd(ticket, 1u);
        batch = __shfl_sync(kFull, batch, 0);
        if (batch >= n_batches) break;
        const uint32_t batch_end = min(count, (batch + 1u) * batch_rays);   // (a thin batch must not reach into the next one)
        const uint32_t j0 = batch * batch_rays + (paired ? 2u : 1u) * (uint32_t)lane;
        Lanes L;
        if (paired) wf_load_rays<2, F>(P, qin, batch_end, j0, L, rays);
        else wf_load_rays<1, F>(P, qin, batch_end, j0, L, rays);
        f3 factor[2];
        uint32_t texel_index[2], bseed[2], n_samples[2];
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            const uint32_t j = j0 + (uint32_t)p;
            factor[p] = rtc::mk(0.0f, 0.0f, 0.0f);
            texel_index[p] = bseed[p] = n_samples[p] = 0u;
            if (j < batch_end && (paired || p == 0)) {
                const float4 B = __ldcg(qin + (size_t)P.wf_capacity + j), C = __ldcg(qin + 2 * (size_t)P.wf_capacity + j);
                factor[p] = rtc::mk(B.z, B.w, C.x);
                texel_index[p] = __float_as_uint(C.y);
                bseed[p] = __float_as_uint(C.z);
                n_samples[p] = __float_as_uint(C.w);
            }
        }
        for (uint32_t b = bounce;; ++b) {
            if constexpr (kCoop) trace_tree_coop(tv, coop_warp(coop_base, threadIdx.x >> 5), P.spheres, P.n_spheres, L, exact_tests);
            else trace_tree(tv, s_stack, s_cand, P.spheres, P.n_spheres, L, exact_tests);
#pragma unroll
            for (int p = 0; p < 2; ++p) {
                if (!L.alive[p]) continue;
                f3 o = origin_of(L, p), d = rtc::mk(L.dx[p], L.dy[p], L.dz[p]);
                const bool go = resolve_hit<F>(P, b, b + 1u >= max_depth, L.best_i[p], L.best_t[p], o, d, factor[p],
                                               texel_index[p], bseed[p], n_samples[p], reset);
                if (go) {
                    arm_ray<F>(L, p, o, d);
                } else {
                    idle_ray(L, p);
                }
            }
            const unsigned m0 = __ballot_sync(kFull, L.alive[0]), m1 = __ballot_sync(kFull, L.alive[1]);
            if ((m0 | m1) == 0u) break;
            if (b + 1u == stop_at) {
                // back to the queue: the survivors are dealt out again, evenly, by the next pass over queue stop_at
#pragma unroll
                for (int p = 0; p < 2; ++p)
                    push_path(P, stop_at, L.alive[p], origin_of(L, p), rtc::mk(L.dx[p], L.dy[p], L.dz[p]), factor[p], texel_index[p],
                              bseed[p], n_samples[p], lane);
                break;
            }
            if (lane == 0) atomicAdd(&traced[b + 1u], (unsigned)(__popc(m0) + __popc(m1)));
            const uint32_t mine = (L.alive[0] ? 1u : 0u) + (L.alive[1] ? 1u : 0u);
            rays += mine;
            rays_later += mine;
        }
    }
    exact_tests_out += exact_tests;
    rays_out += rays;
    rays_later_out += rays_later;
}

}  // namespace

// Rays traced at this bounce (the to-completion path adds its later bounces itself) and the frame counters.
__device__ __forceinline__ void wf_flush(const UpdateParams& P, uint32_t bounce, uint32_t exact_tests, uint32_t rays,
                                         uint32_t rays_later, int lane) {
    uint32_t mine = rays - rays_later;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mine += __shfl_xor_sync(kFull, mine, o);
    if (lane == 0 && mine) atomicAdd(&P.wf_counts[2u * (kWfMaxDepth + 1u) + bounce], mine);
    prim_flush(P, exact_tests, rays, lane);
}

// Traces queue P.wf_bounce through the tree image of the scene (rt_tree.h), which is copied into shared memory once
// per CTA; every WARP then pulls batches of 64 rays (2 per thread, one FFMA2 pair) on its own — no block barrier in
// the loop, so a warp with deep stacks or long candidate queues delays nobody.  Two rays per thread rather than four:
// the traversal runs as long as the busiest lane's, and a thread's work grows with its rays.
// This kernel serves the early bounces, where nearly all the rays are: three CTAs per SM (80 registers, ~74 KB of shared
// memory for 1024 spheres).  The late bounces, whose queues may be short enough to finish from registers, go to
// bounce_tail_kernel, which carries the finishing code (it would spill at this kernel's register budget).
// kGlobal: the instantiation for scenes whose image does not fit shared memory (cl_bytes == 0): the same traversal
// reads the tree where it lies in global memory (L1 / L2; the upper levels stay hot), and the candidate codes are
// 32 bits wide, so the sphere count is not limited by the code's 15 index bits either.
template <bool kGlobal, int F>
__global__ void __launch_bounds__(kBlockThreads, (kGlobal && RT_WF_COOP_GLOBAL) ? 2 : 3) bounce_clustered_kernel(const __grid_constant__ UpdateParams P) {
    extern __shared__ float4 s_dyn[];
    __shared__ __align__(8) uint64_t s_bar;
    using CandT = typename std::conditional<kGlobal, uint32_t, unsigned short>::type;
    const uint32_t image_slots = kGlobal ? 0u : (P.cl_bytes + 15u) / 16u;
    // [kWfCand][kBlockThreads] candidate codes, then [cl_stack_cap][kBlockThreads] node numbers
    CandT* s_cand = col_base(reinterpret_cast<CandT*>(s_dyn + image_slots));

    unsigned short* s_stack = col_base(reinterpret_cast<unsigned short*>(reinterpret_cast<CandT*>(s_dyn + image_slots) + kWfCand * kBlockThreads));
    const int lane = threadIdx.x & 31;
    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_bar);
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const uint32_t bounce = P.wf_bounce;
    const uint32_t count = P.wf_counts[bounce];
    const uint32_t max_depth = rtc::f2u(P.cam.max_depth);
    const bool last = bounce + 1u >= max_depth;
    const bool reset = P.cam.camera_has_moved > 0.5f;
    const float4* qin = P.wf_queue[bounce & 1u];
    unsigned int* ticket = &P.wf_counts[kWfMaxDepth + 1u + bounce];
    uint32_t exact_tests = 0, rays = 0, rays_later = 0;
    if (count == 0u) return;

    if constexpr (!kGlobal) {
        if (threadIdx.x == 0) {
            mbar_expect_tx(bar, P.cl_bytes);
            bulk_copy_g2s((uint32_t)__cvta_generic_to_shared(s_dyn), P.cl_blob, P.cl_bytes, bar);
        }
        mbar_wait(bar, 0u);
    }
    const TreeView tv = tree_view(kGlobal ? P.cl_blob : (const void*)s_dyn, P.cl_n_nodes, P.cl_first_leaf, P.cl_n_entry, P.cl_entry_r,
                                  P.cl_stack_cap);
    const uint32_t n_batches = (count + 63u) / 64u;
    for (;;) {
        uint32_t batch = 0;
        if (lane == 0) batch = atomicAdd(ticket, 1u);
        batch = __shfl_sync(kFull, batch, 0);
        if (batch >= n_batches) break;
        const uint32_t j0 = batch * 64u + 2u * (uint32_t)lane;
        Lanes L;
        wf_load_rays<2, F>(P, qin, count, j0, L, rays);
        if constexpr (kGlobal && RT_WF_COOP_GLOBAL) {
            const CoopWarp cw = coop_warp(s_dyn + image_slots, threadIdx.x >> 5);
            trace_tree_coop(tv, cw, P.spheres, P.n_spheres, L, exact_tests);
            wf_resolve_batch<F>(P, qin, count, j0, L, bounce, last, reset, lane, reinterpret_cast<CoopWord*>(cw.items) + lane);   // (the queue is idle here)
        } else {
            trace_tree(tv, s_stack, s_cand, P.spheres, P.n_spheres, L, exact_tests);
            wf_resolve_batch<F>(P, qin, count, j0, L, bounce, last, reset, lane, s_cand);   // (the thread's candidate column is idle here)
        }
    }
    wf_flush(P, bounce, exact_tests, rays, rays_later, lane);
}

// Every bounce from P.wf_bounce to the end of the paths in ONE cooperative launch: the grid traces queue b exactly as
// bounce_clustered_kernel does, meets at a grid barrier, reads the length of queue b + 1 and goes on — until a queue is
// short enough to be finished from registers (finish_paths) or empty.  The late bounces of a frame are launch-bound, not
// work-bound: at 1280x720 twenty-odd of the 29 bounce launches found a handful of rays or none, each costing a launch
// and a drain.  Queue lengths come through ld.cg (the counters are bumped by atomics in L2; a plain load could be served
// from an L1 line fetched for the neighbouring counter one bounce earlier), and so do the queue planes themselves
// (wf_load_rays / wf_resolve_batch / finish_paths use __ldcg): queue b & 1 was read by this SM two bounces ago and has
// been rewritten by others since.
template <bool kGlobal, int F>
__global__ void __launch_bounds__(kBlockThreads, 2) bounce_tail_kernel(const __grid_constant__ UpdateParams P) {
    extern __shared__ float4 s_dyn[];
    __shared__ __align__(8) uint64_t s_bar;
    using CandT = typename std::conditional<kGlobal, uint32_t, unsigned short>::type;
    const uint32_t image_slots = kGlobal ? 0u : (P.cl_bytes + 15u) / 16u;
    CandT* s_cand = col_base(reinterpret_cast<CandT*>(s_dyn + image_slots));
    unsigned short* s_stack = col_base(reinterpret_cast<unsigned short*>(reinterpret_cast<CandT*>(s_dyn + image_slots) + kWfCand * kBlockThreads));
    const int lane = threadIdx.x & 31;
    cooperative_groups::grid_group grid = cooperative_groups::this_grid();
    if constexpr (!kGlobal) {
        const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_bar);
        if (threadIdx.x == 0) {
            mbar_init(bar, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            mbar_expect_tx(bar, P.cl_bytes);
            bulk_copy_g2s((uint32_t)__cvta_generic_to_shared(s_dyn), P.cl_blob, P.cl_bytes, bar);
        }
        mbar_wait(bar, 0u);
    }
    const TreeView tv = tree_view(kGlobal ? P.cl_blob : (const void*)s_dyn, P.cl_n_nodes, P.cl_first_leaf, P.cl_n_entry, P.cl_entry_r,
                                  P.cl_stack_cap);
    const uint32_t max_depth = rtc::f2u(P.cam.max_depth);
    const bool reset = P.cam.camera_has_moved > 0.5f;
    const uint32_t resident_warps = gridDim.x * (kBlockThreads / 32);
    uint32_t deals = 0;   // passes over a short queue so far
    for (uint32_t bounce = P.wf_bounce, next = 0; bounce < max_depth; bounce = next) {
        const uint32_t count = __ldcg(&P.wf_counts[bounce]);   // grid-uniform: written before the barrier below
        if (count == 0u) break;
        const float4* qin = P.wf_queue[bounce & 1u];
        unsigned int* ticket = &P.wf_counts[kWfMaxDepth + 1u + bounce];
        bool last = bounce + 1u >= max_depth;
        next = bounce + 1u;
        uint32_t exact_tests = 0, rays = 0, rays_later = 0;
        if (count >= resident_warps * RT_WF_COMPLETE_BELOW) {
            const uint32_t n_batches = (count + 63u) / 64u;
            for (;;) {
                uint32_t batch = 0;
                if (lane == 0) batch = atomicAdd(ticket, 1u);
                batch = __shfl_sync(kFull, batch, 0);
                if (batch >= n_batches) break;
                const uint32_t j0 = batch * 64u + 2u * (uint32_t)lane;
                Lanes L;
                wf_load_rays<2, F>(P, qin, count, j0, L, rays);
                if constexpr (kGlobal && RT_WF_COOP_GLOBAL) {
                    const CoopWarp cw = coop_warp(s_dyn + image_slots, threadIdx.x >> 5);
                    trace_tree_coop(tv, cw, P.spheres, P.n_spheres, L, exact_tests);
                    wf_resolve_batch<F>(P, qin, count, j0, L, bounce, last, reset, lane, reinterpret_cast<CoopWord*>(cw.items) + lane);
                } else {
                    trace_tree(tv, s_stack, s_cand, P.spheres, P.n_spheres, L, exact_tests);
                    wf_resolve_batch<F>(P, qin, count, j0, L, bounce, last, reset, lane, s_cand);
                }
            }
        } else if (const uint32_t keep = deals == 0u ? RT_WF_REDEAL_AFTER : deals == 1u ? RT_WF_REDEAL_AGAIN : 0u;
                   keep != 0u && bounce + keep < max_depth) {
            // Short queue, first pass: the warps keep their 64 paths for a few bounces — most paths end there — and hand
            // the survivors back.  What is left in the end are the paths that go on to max_depth (in the reference's
            // default scene 0.7 % of the pixels bounce on, bounce after bounce, until the limit); finished from the
            // registers of the warps that happened to load them, the frame waited for the warp that drew the most of them
            // (measured: the slowest of 2,002 batches took 2.5 x the mean).
            next = bounce + keep;
            uint32_t share = 64u;
            if (deals != 0u) share = min(64u, max(2u, ((count + resident_warps - 1u) / resident_warps + 1u) & ~1u));
            finish_paths<F, (kGlobal && RT_WF_COOP_GLOBAL)>(P, qin, count, tv, s_stack, s_cand, s_dyn + image_slots, ticket, bounce, max_depth, reset, lane, share, next, exact_tests, rays, rays_later);
            deals += 1u;
        } else {
            // short queue, every path to its end: nothing is queued, so the next length read is 0 for the whole grid.
            // After the first pass the survivors are dealt out evenly over all the resident warps (one per thread up to 32).
            uint32_t share = 64u;
            if (deals != 0u) share = min(64u, max(2u, ((count + resident_warps - 1u) / resident_warps + 1u) & ~1u));
            finish_paths<F, (kGlobal && RT_WF_COOP_GLOBAL)>(P, qin, count, tv, s_stack, s_cand, s_dyn + image_slots, ticket, bounce, max_depth, reset, lane, share, 0u, exact_tests, rays, rays_later);
            last = true;
        }
        wf_flush(P, bounce, exact_tests, rays, rays_later, lane);
        if (last) break;
        grid.sync();   // queue bounce + 1 and its length are complete (the barrier fences)
    }
}

// Traces queue P.wf_bounce for scenes without a clustered image (more spheres than its shared-memory budget): 4 rays
// per thread, a CTA takes batches of 1024 rays and streams the plain records through two stages as
// primary_hits_kernel does.
__global__ void __launch_bounds__(kBlockThreads, 2) bounce_streamed_kernel(const __grid_constant__ UpdateParams P) {
    extern __shared__ float4 s_dyn[];
    constexpr int kStage = kWfChunk + kGeomTail;
    __shared__ __align__(8) uint64_t s_bar[2];
    __shared__ uint32_t s_batch;
    float4* s_stage = s_dyn;                                                                                // [2][kStage]
    unsigned short* s_note = reinterpret_cast<unsigned short*>(s_dyn + 2 * kStage) + threadIdx.x;           // [kWfDefer][kBlockThreads]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t bar0 = (uint32_t)__cvta_generic_to_shared(&s_bar[0]);
    const uint32_t stage0 = (uint32_t)__cvta_generic_to_shared(s_stage);
    if (threadIdx.x == 0) {
        mbar_init(bar0, 1);
        mbar_init(bar0 + 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const uint32_t bounce = P.wf_bounce;
    const uint32_t count = P.wf_counts[bounce];
    const uint32_t max_depth = rtc::f2u(P.cam.max_depth);
    const bool last = bounce + 1u >= max_depth;
    const bool reset = P.cam.camera_has_moved > 0.5f;
    const float4* qin = P.wf_queue[bounce & 1u];
    unsigned int* ticket = &P.wf_counts[kWfMaxDepth + 1u + bounce];
    const int n_pad = (int)((P.n_spheres + kGeomPad - 1) / kGeomPad * kGeomPad);
    const int n_chunks = (n_pad + kWfChunk - 1) / kWfChunk;
    uint32_t exact_tests = 0, rays = 0, parity = 0;
    if (count == 0u) return;

    auto request = [&](int c) {   // thread 0: chunk c -> stage c & 1
        const int base = c * kWfChunk;
        const uint32_t bytes = (uint32_t)(min(kWfChunk, n_pad - base) + kGeomTail) * (uint32_t)sizeof(float4);
        const uint32_t bar = bar0 + 8u * (uint32_t)(c & 1);
        mbar_expect_tx(bar, bytes);
        bulk_copy_g2s(stage0 + (uint32_t)(c & 1) * (uint32_t)(kStage * sizeof(float4)), P.geom + base, bytes, bar);
    };

    const uint32_t n_batches = (count + 8u * 32u * kPix - 1u) / (8u * 32u * kPix);
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) s_batch = atomicAdd(ticket, 1u);
        __syncthreads();
        const uint32_t batch = s_batch;
        if (batch >= n_batches) break;
        if (threadIdx.x == 0) {
            request(0);
            if (n_chunks > 1) request(1);
        }
        const uint32_t j0 = batch * (8u * 32u * kPix) + (uint32_t)warp * (32u * kPix) + (uint32_t)(kPix * lane);
        Lanes L;
        wf_load_rays<kPix>(P, qin, count, j0, L, rays);
        for (int c = 0; c < n_chunks; ++c) {
            const int st = c & 1, base = c * kWfChunk;
            mbar_wait(bar0 + 8u * (uint32_t)st, (parity >> st) & 1u);
            parity ^= 1u << st;
            scan_spheres_deferred<kPairs>(s_stage + st * kStage, min(kWfChunk, n_pad - base), base, s_note, P.spheres,
                                          P.n_spheres, L, exact_tests);
            __syncthreads();   // every warp has left stage st
            if (threadIdx.x == 0 && c + 2 < n_chunks) request(c + 2);
        }
        wf_resolve<kPix>(P, qin, count, j0, L, bounce, last, reset, lane);
    }
    wf_flush(P, bounce, exact_tests, rays, 0u, lane);
}

// Per-frame common-origin records of the tree (rt_tree.h): one thread per child slot.  A sphere under a leaf node gets
// its own rtc::primary_record; a group gets the record of its bounding sphere, radius Rc' + m_cam (m_cam = 2^-7 (|Oc| +
// rho): the camera's share of the members' kappa S inflation, so that every member's inflated sphere lies inside the
// group's); children hanging under always-passing records stay always candidates; empty slots never are.
__global__ void __launch_bounds__(kPrepThreads) prep_primary_tree_kernel(const float4* __restrict__ tree_rec, const uint32_t* __restrict__ link,
                                         const float* __restrict__ child_r, const GpuSphere* __restrict__ spheres,
                                         const uint32_t* __restrict__ entry_slot, uint32_t n_entry,
                                         float4* __restrict__ out, uint32_t n_nodes, uint32_t first_leaf, float ox, float oy,
                                         float oz, double rho, float m_cam, unsigned int* __restrict__ tile_counter,
                                         unsigned int* __restrict__ wf_counts, uint32_t record_blocks,
                                         const unsigned int* __restrict__ row_cost_prev, unsigned int* __restrict__ row_cost_cur,
                                         uint32_t* __restrict__ row_order, uint32_t tile_rows) {
    if (blockIdx.x >= record_blocks) {
        // The blocks behind the record blocks rank the tile rows by what they cost in the last frame, dearest first
        // (ties: top-down, so a first frame is handed out top-down), and clear the record this frame adds to.  A
        // thread per row counts the rows that go before its own: tile_rows <= kOrderMaxRows broadcast reads of a record no
        // kernel in flight writes (the frame that filled it is two frames back), so the ranks are a permutation.
        const uint32_t i = (blockIdx.x - record_blocks) * blockDim.x + threadIdx.x;
        if (i >= tile_rows) return;
        const unsigned int mine = __ldg(row_cost_prev + i);
        uint32_t before = 0;
#pragma unroll 8
        for (uint32_t j = 0; j < tile_rows; ++j) {
            const unsigned int c = __ldg(row_cost_prev + j);
            before += (c > mine || (c == mine && j < i)) ? 1u : 0u;
        }
        row_order[before] = i;
        row_cost_cur[i] = 0u;
        return;
    }
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    // the frame's tile tickets and (deep frames) queue lengths, batch tickets and per-bounce counters start at zero:
    // done here rather than by memset nodes of their own, the kernels that use them are ordered behind this one
    if (t < 2u) tile_counter[t] = 0u;
    if (wf_counts != nullptr && t < kWfWords) wf_counts[t] = 0u;
    // threads past the child slots write the entry list's records once more, contiguously, behind them
    if (t >= n_nodes * kTreeStride + n_entry) return;
    uint32_t slot = t;
    if (t >= n_nodes * kTreeStride) {
        slot = entry_slot[t - n_nodes * kTreeStride];
        if (slot == 0xFFFFFFFFu) {   // the root itself: always a candidate
            out[t] = make_float4(0.0f, 0.0f, 0.0f, 1.0f);
            return;
        }
    }
    const uint32_t node = slot / kTreeStride, k = slot % kTreeStride;
    float r[4];
    rtc::primary_record_pad(r);
    if (k < (uint32_t)kTreeWide) {
        const uint32_t lk = link[node * kTreeWide + k];
        if (lk != 0xFFFFFFFFu) {
            if (node >= first_leaf) {
                rtc::primary_record(spheres[lk], rtc::mk(ox, oy, oz), rho, r);
            } else {
                const float4 b = tree_rec[slot];
                if (b.w < -1.0e38f) {   // rtt::kAlways
                    r[0] = 0.0f; r[1] = 0.0f; r[2] = 0.0f; r[3] = 1.0f;
                } else {
                    rtc::primary_group_record(b.x, b.y, b.z, child_r[node * kTreeWide + k], m_cam, rtc::mk(ox, oy, oz), rho, r);
                }
            }
        }
    }
    out[t] = make_float4(r[0], r[1], r[2], r[3]);
}

// ---------------------------------------------------------------------------------------------
// Variant 1: baseline.  The shader as written: 8x8 workgroups, one invocation per pixel, every
// sphere read from the storage buffer and put through the full quadratic (src/lib.rs:416,
// compute_shader.wgsl:164-221).  Kept as the A/B arm and as an independent second GPU path.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(64) update_baseline_kernel(const __grid_constant__ UpdateParams P) {
    const int x = blockIdx.x * 8 + (threadIdx.x & 7);
    const int row = blockIdx.y * 8 + (threadIdx.x >> 3);
    uint32_t rays = 0;
    if ((uint32_t)x < P.width && (uint32_t)row < P.row_count) {
        const int y_abs = abs_row(P, row);
        const uint32_t seed_bits = rtc::f2u(P.cam.random_seed * 4294967296.0f);
        const uint32_t spp = rtc::f2u(P.cam.samples_per_pixel);
        const uint32_t max_depth = rtc::f2u(P.cam.max_depth);
        float4* texel = &P.accum[(size_t)row * P.width + x];
        float4 t = *texel;
        f3 mean = rtc::mk(t.x, t.y, t.z);
        uint32_t n = rtc::f2u(t.w);
        if (P.cam.camera_has_moved > 0.5f) {
            mean = rtc::mk(0.0f, 0.0f, 0.0f);
            n = 0u;
        }
        if (n < spp) {
            const uint32_t seed = 1u + n + seed_bits;
            rtc::Ray r = rtc::get_ray(P.cam, x, y_abs, seed, seed_bits);
            f3 factor = rtc::mk(1.0f, 1.0f, 1.0f);
            bool black = false;
            for (uint32_t i = 0; i < max_depth; ++i) {
                ++rays;
                const float a = rtc::dot(r.d, r.d);
                float closest = kTMax;
                int best = -1;
                for (uint32_t s = 0; s < P.n_spheres; ++s) {
                    const float4 g = __ldg(reinterpret_cast<const float4*>(P.spheres + s));
                    float root;
                    if (rtc::sphere_root(r.o, r.d, a, rtc::mk(g.x, g.y, g.z), g.w, kTMin, closest, &root)) {
                        closest = root;
                        best = (int)s;
                    }
                }
                if (best < 0) break;
                const GpuSphere sp = load_sphere(P.spheres, best);
                if (!rtc::scatter(&r, &factor, closest, sp, rtc::hash(seed + 1u + i * 1000u))) {
                    black = true;
                    break;
                }
            }
            f3 color = black ? rtc::mk(0.0f, 0.0f, 0.0f) : rtc::sky_times(factor, r.d);
            mean = rtc::add(mean, rtc::divs(rtc::sub(color, mean), (float)(n + 1u)));
            n += 1u;
        }
        *texel = make_float4(mean.x, mean.y, mean.z, (float)n);
    }
    // one atomic per warp
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) rays += __shfl_xor_sync(kFull, rays, o);
    if (lane == 0 && rays) atomicAdd(&P.counters[0], (unsigned long long)rays);
}

// ---------------------------------------------------------------------------------------------
// Scene preparation: per-sphere broad-phase record (C, |C|^2 - r^2 - kappa (|C|^2 + r^2)),
// evaluated in binary64 and rounded toward -inf so the slack can only grow.  Entries past n up
// to the padded length can never be candidates.
// ---------------------------------------------------------------------------------------------
__global__ void prep_geometry_kernel(const GpuSphere* __restrict__ spheres, float4* __restrict__ geom, uint32_t n,
                                     uint32_t n_pad) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pad) return;
    if (i >= n) {
        geom[i] = make_float4(0.0f, 0.0f, 0.0f, 3.0e38f);
        return;
    }
    const GpuSphere s = spheres[i];
    geom[i] = make_float4(s.position.x, s.position.y, s.position.z, rtc::broad_record_k(s));
}

__global__ void quantize_kernel(const float4* __restrict__ accum, uint32_t* __restrict__ rgba8, uint64_t n_texels) {
    // four texels per thread: 64 B in, one 16 B store out
    const uint64_t i4 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i4 + 3 < n_texels) {
        uint4 o;
        o.x = quant_texel(accum[i4 + 0]);
        o.y = quant_texel(accum[i4 + 1]);
        o.z = quant_texel(accum[i4 + 2]);
        o.w = quant_texel(accum[i4 + 3]);
        *reinterpret_cast<uint4*>(rgba8 + i4) = o;
    } else {
        for (uint64_t i = i4; i < n_texels; ++i) rgba8[i] = quant_texel(accum[i]);
    }
}

// ---------------------------------------------------------------------------------------------
// Variant 3: scenes larger than shared memory.  A CTA owns 8 consecutive warp tiles (1024 pixels);
// the broad-phase records stream through a double-buffered 2 x 32 KB stage, one barrier per chunk.
// Per-ray nearest-hit state stays in registers across chunks, so index order (and with it the
// reference's tie rule) is preserved.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlockThreads, 2) update_streamed_kernel(const __grid_constant__ UpdateParams P) {
    extern __shared__ float4 s_stage[];   // [2][kChunkSpheres + kGeomTail]
    constexpr int kStage = kChunkSpheres + kGeomTail;
    __shared__ uint32_t s_block_tile;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t seed_bits = rtc::f2u(P.cam.random_seed * 4294967296.0f);
    const uint32_t spp = rtc::f2u(P.cam.samples_per_pixel);
    const uint32_t max_depth = rtc::f2u(P.cam.max_depth);
    const bool reset = P.cam.camera_has_moved > 0.5f;
    const int n_pad = (int)((P.n_spheres + kGeomPad - 1) / kGeomPad * kGeomPad);
    const int n_chunks = (n_pad + kChunkSpheres - 1) / kChunkSpheres;
    const uint32_t n_block_tiles = (P.n_tiles + 7u) / 8u;
    uint32_t rays = 0, exact_tests = 0;

    for (;;) {
        __syncthreads();   // everyone is done with s_block_tile and the stages of the previous tile
        if (threadIdx.x == 0) s_block_tile = atomicAdd(P.tile_counter, 1u);
        __syncthreads();
        const uint32_t bt = s_block_tile;
        if (bt >= n_block_tiles) break;

        Lanes L;
        PixelJob J;
        int row, px[kPix];
        begin_tile(P, bt * 8u + (uint32_t)warp, lane, seed_bits, spp, reset, L, J, row, px);

        for (uint32_t bounce = 0; bounce < max_depth; ++bounce) {
            bool mine = false;
#pragma unroll
            for (int p = 0; p < kPix; ++p) {
                mine |= L.alive[p];
                rays += L.alive[p] ? 1u : 0u;
            }
            if (!__syncthreads_or(mine ? 1 : 0)) break;   // also fences the stages against the previous bounce
            bool pre[kPix];
            bool todo = mine;
            const bool adopt = bounce == 0 && P.first_hit != nullptr;
            if (adopt) {
                todo = adopt_first_hits(P, row, px, L, pre);
                if (!__syncthreads_or(todo ? 1 : 0)) {   // block-uniform: nothing left to scan
                    unpark(L, pre);
                    after_traversal(P, bounce, bounce + 1u >= max_depth, L, J);
                    continue;
                }
            }
            const bool warp_alive = __any_sync(kFull, todo);

            // prologue: chunk 0 -> stage 0
            {
                const int cnt = min(kChunkSpheres, n_pad);
                for (int i = threadIdx.x; i < cnt + kGeomTail; i += kBlockThreads) s_stage[i] = __ldg(&P.geom[i]);
            }
            __syncthreads();
            for (int c = 0; c < n_chunks; ++c) {
                const int base = c * kChunkSpheres;
                const int cnt = min(kChunkSpheres, n_pad - base);
                if (c + 1 < n_chunks) {
                    const int nbase = base + kChunkSpheres;
                    const int ncnt = min(kChunkSpheres, n_pad - nbase);
                    float4* dst = s_stage + ((c + 1) & 1) * kStage;
                    for (int i = threadIdx.x; i < ncnt + kGeomTail; i += kBlockThreads) dst[i] = __ldg(&P.geom[nbase + i]);
                }
                if (warp_alive)
                    scan_spheres(s_stage + (c & 1) * kStage, cnt, base, P.spheres, P.n_spheres, L, exact_tests);
                __syncthreads();
            }
            if (adopt) unpark(L, pre);
            after_traversal(P, bounce, bounce + 1u >= max_depth, L, J);
        }
        end_tile(P, L, J, row, px, reset);
    }
    flush_counters(P, rays, exact_tests, lane);
}

// ---------------------------------------------------------------------------------------------
// Launchers
// ---------------------------------------------------------------------------------------------
static size_t prim_smem_bytes() {
    return (size_t)kPrimRays * kPrimSlots * kPrimThreads * sizeof(float) +
           (size_t)(kPrimThreads / 32) * (6 * (kPrimList + 4) + kPrimWork + kPrimSched) * sizeof(uint32_t);
}

static size_t bounce_smem_bytes() {
    return (size_t)2 * (kWfChunk + kGeomTail) * sizeof(float4) + (size_t)kWfDefer * kBlockThreads * sizeof(unsigned short);
}
static size_t clustered_smem_bytes(uint32_t image_bytes, uint32_t stack_cap) {

    return (size_t)(image_bytes + 15u) / 16u * 16u + (size_t)(((stack_cap + 1u) & ~1u) + kWfCand) * kBlockThreads * sizeof(unsigned short);
}
static size_t clustered_global_smem_bytes(uint32_t stack_cap) {   // tree in global memory, 32-bit candidate codes
    if (RT_WF_COOP_GLOBAL) return (size_t)(kBlockThreads / 32) * kCoopWarpBytes;
    return (size_t)kWfCand * kBlockThreads * sizeof(uint32_t) + (size_t)((stack_cap + 1u) & ~1u) * kBlockThreads * sizeof(unsigned short);
}

template <typename K>
static cudaError_t set_smem(K kernel, size_t bytes) {
    return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
}

cudaError_t configure_kernels() {
    cudaError_t e = set_smem(update_resident_kernel, (kResidentMaxSpheres + kGeomTail) * sizeof(float4));
    if (e != cudaSuccess) return e;
    const size_t prim = prim_smem_bytes();
    if ((e = set_smem(primary_hits_kernel<kPrimHits, 0>, prim)) != cudaSuccess) return e;
    if ((e = set_smem(primary_hits_kernel<kPrimShade, 0>, prim)) != cudaSuccess) return e;
    if ((e = set_smem(primary_hits_kernel<kPrimQueue, 0>, prim)) != cudaSuccess) return e;
    if ((e = set_smem(primary_hits_kernel<kPrimShade, 1>, prim)) != cudaSuccess) return e;
    if ((e = set_smem(primary_hits_kernel<kPrimQueue, 1>, prim)) != cudaSuccess) return e;
    if ((e = set_smem(primary_hits_kernel<kPrimShade, 2>, prim)) != cudaSuccess) return e;
    if ((e = set_smem(primary_hits_kernel<kPrimQueue, 2>, prim)) != cudaSuccess) return e;
    if ((e = set_smem(bounce_streamed_kernel, bounce_smem_bytes())) != cudaSuccess) return e;
    const size_t cl = clustered_smem_bytes((uint32_t)kClusterImageMax, kTreeStackMax), clg = clustered_global_smem_bytes(kTreeStackMax);
    if ((e = set_smem(bounce_clustered_kernel<false, 0>, cl)) != cudaSuccess) return e;
    if ((e = set_smem(bounce_tail_kernel<false, 0>, cl)) != cudaSuccess) return e;
    if ((e = set_smem(bounce_clustered_kernel<true, 0>, clg)) != cudaSuccess) return e;
    if ((e = set_smem(bounce_tail_kernel<true, 0>, clg)) != cudaSuccess) return e;
    if ((e = set_smem(bounce_clustered_kernel<false, 1>, cl)) != cudaSuccess) return e;
    if ((e = set_smem(bounce_tail_kernel<false, 1>, cl)) != cudaSuccess) return e;
    if ((e = set_smem(bounce_clustered_kernel<true, 1>, clg)) != cudaSuccess) return e;
    if ((e = set_smem(bounce_tail_kernel<true, 1>, clg)) != cudaSuccess) return e;
    return set_smem(update_streamed_kernel, 2 * (kChunkSpheres + kGeomTail) * sizeof(float4));
}

bool fused_shade_applies(const UpdateParams& p, int variant) {
    return (variant == 0 || variant >= 4) && primary_pass_applies(p) && rtc::f2u(p.cam.max_depth) == 1u;
}

bool primary_pass_applies(const UpdateParams& p) {
    return p.prim_tree != nullptr && p.cl_n_nodes != 0u && p.first_hit != nullptr && p.n_spheres > 0 && rtc::primary_rho(p.cam) >= 0.0 &&
           rtc::f2u(p.cam.max_depth) > 0u;
}

// One bounce launch of the wavefront: which instantiation depends on where the tree lives, whether this launch may finish
// a short queue to the end, and the precision mode.
template <int F>
static void launch_bounce(const UpdateParams& p, uint32_t full_grid, uint32_t finish_grid, cudaStream_t stream) {
    if (p.cl_bytes == 0u && p.cl_global != 0u) {   // the tree stays in global memory
        bounce_clustered_kernel<true, F><<<full_grid, kBlockThreads, clustered_global_smem_bytes(p.cl_stack_cap), stream>>>(p);
    } else if (p.cl_bytes == 0u) {   // no tree the bounce kernel can walk (node numbers beyond 16 bits): exact mode only
        bounce_streamed_kernel<<<finish_grid, kBlockThreads, bounce_smem_bytes(), stream>>>(p);
    } else {
        bounce_clustered_kernel<false, F><<<full_grid, kBlockThreads, clustered_smem_bytes(p.cl_bytes, p.cl_stack_cap), stream>>>(p);
    }
}
// Every remaining bounce in one cooperative launch (bounce_tail_kernel); grid = the co-resident CTAs.
template <int F>
static cudaError_t launch_bounce_tail(const UpdateParams& p, uint32_t grid, cudaStream_t stream) {
    void* args[] = {const_cast<UpdateParams*>(&p)};
    if (p.cl_bytes == 0u)
        return cudaLaunchCooperativeKernel((const void*)bounce_tail_kernel<true, F>, dim3(grid), dim3(kBlockThreads), args,
                                           clustered_global_smem_bytes(p.cl_stack_cap), stream);
    return cudaLaunchCooperativeKernel((const void*)bounce_tail_kernel<false, F>, dim3(grid), dim3(kBlockThreads), args,
                                       clustered_smem_bytes(p.cl_bytes, p.cl_stack_cap), stream);
}

cudaError_t launch_update(const UpdateParams& p_in, int variant, int sm_count, cudaStream_t stream, int* launches,
                          const PrepAsync* prep_async, bool* prep_was_async) {
    UpdateParams p = p_in;
    if (prep_was_async) *prep_was_async = false;
    int n_launch = 0;
    const bool primary = (variant == 0 || variant >= 4) && primary_pass_applies(p);
    if (!primary) p.first_hit = nullptr;
    if (variant >= 4) variant = 0;
    if (variant == 0) variant = (p.n_spheres <= kResidentMaxSpheres) ? 2 : 3;
    if (variant == 2 && p.n_spheres > kResidentMaxSpheres) variant = 3;
    if (variant == 1) {
        dim3 grid((p.width + 7) / 8, (p.row_count + 7) / 8);
        update_baseline_kernel<<<grid, 64, 0, stream>>>(p);
        if (launches) *launches = 1;
        return cudaGetLastError();
    }
    cudaError_t e;
    const uint32_t resident = (uint32_t)sm_count * 2u;
    if (primary) {
        const uint32_t tiles = ((p.width + kPrimTileW - 1) / kPrimTileW) * ((p.row_count + kPrimTileH - 1) / kPrimTileH);
        constexpr uint32_t warps = kPrimThreads / 32;
        uint32_t blocks = (tiles + warps - 1) / warps;
        if (blocks > (uint32_t)sm_count * kPrimCtas) blocks = (uint32_t)sm_count * kPrimCtas;
        const uint32_t depth = rtc::f2u(p.cam.max_depth);
        const bool shade = depth == 1u;
        const bool wavefront = !shade && p.wf_queue[0] != nullptr && p.wf_counts != nullptr && depth <= kWfMaxDepth;
        // tolerance mode exists for the fused depth-1 pass and the wavefront; the other paths are exact whatever the mode
        const uint32_t fast = (shade || wavefront) ? p.precision : 0u;
        // per-frame records; the same launch zeroes the tile tickets and (wavefront) the frame's queue counters
        if (shade && prep_async != nullptr) {
            // nothing the prep kernel reads or writes belongs to the frame before this one: it runs beside it
            e = cudaStreamWaitEvent(prep_async->stream, prep_async->buffers_free, 0);
            if (e == cudaSuccess && prep_async->also_wait != nullptr) e = cudaStreamWaitEvent(prep_async->stream, prep_async->also_wait, 0);
            if (e == cudaSuccess) e = launch_prep_primary_tree(p, false, prep_async->stream);
            if (e == cudaSuccess) e = cudaEventRecord(prep_async->prep_done, prep_async->stream);
            if (e == cudaSuccess) e = cudaStreamWaitEvent(stream, prep_async->prep_done, 0);
            if (e == cudaSuccess && prep_was_async) *prep_was_async = true;
        } else {
            e = launch_prep_primary_tree(p, wavefront, stream);
        }
        if (e != cudaSuccess) return e;
        if (shade) {
            if (fast >= 2u) primary_hits_kernel<kPrimShade, 2><<<blocks, kPrimThreads, prim_smem_bytes(), stream>>>(p);
            else if (fast == 1u) primary_hits_kernel<kPrimShade, 1><<<blocks, kPrimThreads, prim_smem_bytes(), stream>>>(p);
            else primary_hits_kernel<kPrimShade, 0><<<blocks, kPrimThreads, prim_smem_bytes(), stream>>>(p);
        } else if (wavefront) {
            if (fast >= 2u) primary_hits_kernel<kPrimQueue, 2><<<blocks, kPrimThreads, prim_smem_bytes(), stream>>>(p);
            else if (fast == 1u) primary_hits_kernel<kPrimQueue, 1><<<blocks, kPrimThreads, prim_smem_bytes(), stream>>>(p);
            else primary_hits_kernel<kPrimQueue, 0><<<blocks, kPrimThreads, prim_smem_bytes(), stream>>>(p);
        } else {
            primary_hits_kernel<kPrimHits, 0><<<blocks, kPrimThreads, prim_smem_bytes(), stream>>>(p);
        }
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
        n_launch += 2;
        if (shade) {   // the texels are finished: no update kernel
            if (launches) *launches = n_launch;
            return cudaSuccess;
        }
        if (wavefront) {
            // the primary pass has resolved the first hits and queued the survivors (kPrimQueue): one launch per bounce
            const uint64_t n_texels = (uint64_t)p.width * p.row_count;
            // First bounce launched with the finishing code compiled in: not before the queue can be short enough to
            // use it, assuming it shrinks by no more than a fifth per bounce (texels * 0.8^b < 8 * the short-queue
            // threshold).  A scene that sheds paths faster only delays the switch by a few cheap launches.
            const double short_queue = (double)resident * (kBlockThreads / 32) * RT_WF_COMPLETE_BELOW;
            uint32_t finish_from = kWfFinishFrom;
            for (double q = (double)n_texels; q > 8.0 * short_queue && finish_from < depth; q *= 0.8) ++finish_from;
            finish_from = finish_from > kWfFinishFrom ? finish_from - 1 : kWfFinishFrom;
            const bool walkable = p.cl_bytes != 0u || p.cl_global != 0u;
            for (uint32_t b = 1; b < depth; ++b) {
                p.wf_bounce = b;
                if (walkable && b >= finish_from) {
                    // from here on the queues can be short: the rest of the frame is one cooperative launch
                    e = fast ? launch_bounce_tail<1>(p, resident, stream) : launch_bounce_tail<0>(p, resident, stream);
                    if (e != cudaSuccess) return e;
                    n_launch += 1;
                    break;
                }
                if (fast) launch_bounce<1>(p, (uint32_t)sm_count * 3u, resident, stream);   // (bounce rays: one fast level)
                else launch_bounce<0>(p, (uint32_t)sm_count * 3u, resident, stream);
                n_launch += 1;
            }
            if (launches) *launches = n_launch;
            return cudaGetLastError();
        }
    }
    e = cudaMemsetAsync(p.tile_counter, 0, 2 * sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    const uint32_t n_pad = (p.n_spheres + kGeomPad - 1) / kGeomPad * kGeomPad;
    uint32_t blocks = (p.n_tiles + 7) / 8;
    if (blocks > resident) blocks = resident;
    if (blocks == 0) blocks = 1;
    if (variant == 2) {
        update_resident_kernel<<<blocks, kBlockThreads, (n_pad + kGeomTail) * sizeof(float4), stream>>>(p);
    } else {
        update_streamed_kernel<<<blocks, kBlockThreads, 2 * (kChunkSpheres + kGeomTail) * sizeof(float4), stream>>>(p);
    }
    if (launches) *launches = n_launch + 1;
    return cudaGetLastError();
}

cudaError_t launch_prep_primary_tree(const UpdateParams& p, bool zero_wavefront, cudaStream_t stream) {
    const rtt::ImageLayout lay = rtt::image_layout(p.cl_n_nodes, p.cl_n_entry);
    const unsigned char* blob = static_cast<const unsigned char*>(p.cl_blob);
    const double rho = rtc::primary_rho(p.cam);
    const float m_cam = rtc::primary_m_cam(p.cam, rho);
    uint32_t n_threads = p.cl_n_nodes * kTreeStride + p.cl_n_entry;
    if (n_threads < kWfWords) n_threads = kWfWords;
    const uint32_t record_blocks = (n_threads + kPrepThreads - 1) / kPrepThreads;
    const uint32_t tile_rows = p.row_order != nullptr ? (p.row_count + kPrimTileH - 1) / kPrimTileH : 0u;
    prep_primary_tree_kernel<<<record_blocks + (tile_rows + kPrepThreads - 1) / kPrepThreads, kPrepThreads, 0, stream>>>(
        reinterpret_cast<const float4*>(blob), reinterpret_cast<const uint32_t*>(blob + lay.off_link),
        reinterpret_cast<const float*>(blob + lay.off_child_r), p.spheres, reinterpret_cast<const uint32_t*>(blob + lay.off_entry_slot),
        p.cl_n_entry, const_cast<float4*>(p.prim_tree), p.cl_n_nodes,
        p.cl_first_leaf, p.cam.center.x, p.cam.center.y, p.cam.center.z, rho, m_cam, p.tile_counter,
        zero_wavefront ? p.wf_counts : nullptr, record_blocks, p.row_cost_prev, p.row_cost_cur, p.row_order, tile_rows);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// Completion flags of a fused gather (rt_stream_signal / rt_stream_wait): one release store, resp. a bounded spin of
// acquire loads, at system scope — the flag may live in a peer GPU's memory.
// ---------------------------------------------------------------------------------------------
__global__ void signal_kernel(unsigned int* flag, unsigned int value) {
    // the stores of the kernels ahead of this one in the stream have been performed; the fence orders them before the
    // flag for an observer on another device
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(value) : "memory");
}

__global__ void wait_flags_kernel(const unsigned int* flags, unsigned int n, unsigned int value, unsigned long long timeout_ns,
                                  unsigned int* status) {
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (unsigned int i = threadIdx.x; i < n; i += blockDim.x) {
        for (;;) {
            unsigned int v;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + i) : "memory");
            if ((int)(v - value) >= 0) break;
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if (t - t0 > timeout_ns) {   // never hang the device on a peer that does not come
                atomicAdd(status, 1u);
                break;
            }
            __nanosleep(200);
        }
    }
    __threadfence_system();
}

cudaError_t launch_signal(void* flag, uint32_t value, cudaStream_t stream) {
    signal_kernel<<<1, 1, 0, stream>>>(static_cast<unsigned int*>(flag), value);
    return cudaGetLastError();
}
cudaError_t launch_wait_flags(const void* flags, uint32_t n, uint32_t value, uint32_t timeout_ms, unsigned int* status,
                              cudaStream_t stream) {
    wait_flags_kernel<<<1, 32, 0, stream>>>(static_cast<const unsigned int*>(flags), n, value, (unsigned long long)timeout_ms * 1000000ull, status);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// rt_update_static (include/rt_abi.h): `update` for callers that may not allocate — code under CUDA graph capture, tight
// serving loops.  It is update_resident_kernel (variant 2: every sphere's broad-phase record resident in shared memory, any
// max_depth in the kernel's own bounce loop) with its scratch — the tile ticket, the ray / test counters and the
// per-sphere records prep_geometry_kernel derives from the caller's sphere buffer — in ONE file-scope device array
// instead of cudaMalloc'ed buffers of a renderer handle.  Every GPU that loads the library holds the array (66,048
// bytes) until the process exits; every call on a device shares it.
//   [0, 8)    tile ticket (2 x u32)   cleared per call (must start at zero)
//   [16, 32)  rays, exact tests (2 x u64)   cleared per call
//   [256, ..) float4 records for kStaticMaxSpheres + padding + prefetch tail, rewritten per call up to the length the
//             kernel reads (prep_geometry_kernel writes the padding and the tail as never-candidates)
// In host code the array's NAME is a host-side stand-in, not a device address: the address comes from
// cudaGetSymbolAddress, once per device with that device current, and that pointer is what the kernels get.
// ---------------------------------------------------------------------------------------------
constexpr uint32_t kStaticGeomOffset = 256;
constexpr size_t kStaticScratchBytes = kStaticGeomOffset + (size_t)(kStaticMaxSpheres + kGeomPad + kGeomTail) * sizeof(float4);
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];
static_assert(kStaticMaxSpheres <= kResidentMaxSpheres, "the records must fit the resident kernel's shared memory");

constexpr int kStaticMaxDevices = 64;
static unsigned char* g_static_ptr[kStaticMaxDevices];   // device address of g_static_scratch, per device
static int g_static_sms[kStaticMaxDevices];

// Set-up of the calling device (current): symbol address, SM count, shared-memory opt-in of the kernel.  No allocation.
static cudaError_t static_setup(int device) {
    if (g_static_ptr[device] != nullptr) return cudaSuccess;
    void* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_scratch);
    if (e != cudaSuccess) return e;
    int sms = 0;
    e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    if (e != cudaSuccess) return e;
    e = set_smem(update_resident_kernel, (kResidentMaxSpheres + kGeomTail) * sizeof(float4));
    if (e != cudaSuccess) return e;
    g_static_sms[device] = sms;
    g_static_ptr[device] = static_cast<unsigned char*>(ptr);
    return cudaSuccess;
}

cudaError_t launch_update_static(int device, const GpuSphere* device_spheres, uint32_t n_spheres, const SceneCamera& cam,
                                 uint32_t width, uint32_t height, float4* device_accum, cudaStream_t stream) {
    if (device < 0 || device >= kStaticMaxDevices || n_spheres > kStaticMaxSpheres) return cudaErrorInvalidValue;
    cudaError_t e = static_setup(device);
    if (e != cudaSuccess) return e;
    unsigned char* scratch = g_static_ptr[device];
    UpdateParams p;
    memset(&p, 0, sizeof(p));   // no primary pass, no wavefront, no tree, no RGBA8 output: null pointers and zero counts
    p.cam = cam;
    p.accum = device_accum;
    p.geom = reinterpret_cast<const float4*>(scratch + kStaticGeomOffset);
    p.spheres = device_spheres;
    p.n_spheres = n_spheres;
    p.width = width; p.height = height;
    p.row_begin = 0; p.row_count = height;
    p.tiles_x = (width + kWarpTileW - 1) / kWarpTileW;
    p.n_tiles = p.tiles_x * ((height + kWarpTileH - 1) / kWarpTileH);
    p.tile_counter = reinterpret_cast<unsigned int*>(scratch);
    p.counters = reinterpret_cast<unsigned long long*>(scratch + 16);
    // the array keeps the last call's contents: ticket and counters must start at zero
    e = cudaMemsetAsync(scratch, 0, 32, stream);
    if (e != cudaSuccess) return e;
    const uint32_t n_pad = (n_spheres + kGeomPad - 1) / kGeomPad * kGeomPad;
    prep_geometry_kernel<<<(n_pad + kGeomTail + 255) / 256, 256, 0, stream>>>(device_spheres, reinterpret_cast<float4*>(scratch + kStaticGeomOffset),
                                                                               n_spheres, n_pad + kGeomTail);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    uint32_t blocks = (p.n_tiles + 7) / 8;
    const uint32_t resident = (uint32_t)g_static_sms[device] * 2u;
    if (blocks > resident) blocks = resident;
    if (blocks == 0) blocks = 1;
    update_resident_kernel<<<blocks, kBlockThreads, (n_pad + kGeomTail) * sizeof(float4), stream>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_prep_geometry(const GpuSphere* spheres, float4* geom, uint32_t n, cudaStream_t stream) {
    const uint32_t n_pad = (n + kGeomPad - 1) / kGeomPad * kGeomPad + kGeomTail;
    prep_geometry_kernel<<<(n_pad + 255) / 256, 256, 0, stream>>>(spheres, geom, n, n_pad);
    return cudaGetLastError();
}

cudaError_t launch_quantize(const float4* accum, uchar4* rgba8, uint64_t n_texels, cudaStream_t stream) {
    if (n_texels == 0) return cudaSuccess;
    const uint64_t threads = (n_texels + 3) / 4;
    quantize_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, stream>>>(accum, reinterpret_cast<uint32_t*>(rgba8),
                                                                         n_texels);
    return cudaGetLastError();
}

}  // namespace rtk
