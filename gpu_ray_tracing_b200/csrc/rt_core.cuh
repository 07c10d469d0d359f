// rt_core.cuh — per-ray arithmetic of the ray-trace pass, shared by every kernel variant.
//
// Follows /root/reference/assets/compute_shader.wgsl (cited per function).  All arithmetic is
// IEEE-754 binary32 with ONE rounding per WGSL operator: this translation unit must be built
// with -fmad=false (nvcc) so that a*b+c is never contracted; the only fused operations in the
// product are the explicit fmaf() calls of the broad phase (rt_kernels.cu), which never decide
// a result on their own.  Transcendentals come from include/rt_pinned_math.h.
//
// The functions are __host__ __device__ so that tests can run the very same code on the CPU
// (tests/emu) when no GPU is present; the product never executes them on the host.
#pragma once

#include <stdint.h>
#include <math.h>
#include <string.h>

#include "../../include/rt_abi.h"
#include "../../include/rt_pinned_math.h"

#if defined(__CUDACC__)
#define RT_DEV __host__ __device__ __forceinline__
#else
#define RT_DEV static inline
#endif

namespace rtc {

struct f3 {
    float x, y, z;
};
RT_DEV f3 mk(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
RT_DEV f3 mk(const RtVec3& v) { return mk(v.x, v.y, v.z); }
RT_DEV f3 add(f3 a, f3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
RT_DEV f3 sub(f3 a, f3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
RT_DEV f3 mul(f3 a, f3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
RT_DEV f3 scale(float s, f3 a) { return mk(s * a.x, s * a.y, s * a.z); }
RT_DEV f3 divs(f3 a, float s) { return mk(a.x / s, a.y / s, a.z / s); }
RT_DEV f3 neg(f3 a) { return mk(-a.x, -a.y, -a.z); }
RT_DEV float dot(f3 a, f3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
RT_DEV f3 normalize(f3 a) { return divs(a, sqrtf(dot(a, a))); }
// An IEEE square root and three IEEE divisions expand to ~45 instructions plus slow-path calls, and the kernels'
// instruction footprint is what limits them: the metal and dielectric branches of scatter,
// which a minority of the rays take, share ONE out-of-line copy on the device.  (Every ray's own divisions — sky
// colour, running mean, outward normal — stay inline: there the call costs more than the footprint.)
#if defined(__CUDACC__) && !defined(RT_INLINE_DIVS)
#define RT_DEV_SHARED static __host__ __device__ __noinline__
#else
#define RT_DEV_SHARED RT_DEV
#endif
RT_DEV_SHARED f3 normalize_shared(f3 a) {
    const float s = sqrtf(dot(a, a));
    return mk(a.x / s, a.y / s, a.z / s);
}

// ---------------------------------------------------------------------------------------------
// Tolerance modes (rt_set_precision).  The functions below that take an `int F` template
// argument have up to three instantiations: F = 0 is the bit-exact restatement (one IEEE rounding per WGSL operator,
// pinned sin / cos); F >= 1 evaluates scatter, sky colour, running mean and the broad phases' unit directions with
// the hardware approximations the port target names — MUFU rcp / rsqrt / sqrt / sin / cos (about 1-2 ulp, sin / cos
// 2^-21 absolute) and explicit FMA contraction — which is what a WGSL implementation is allowed to do as well (builtin
// precision is implementation defined); F >= 2 does the same to get_ray.  What decides WHICH sphere a ray hits
// (discriminant sign, root order, rtc::sphere_root) stays IEEE in every mode.  With F = 1 the primary rays are the
// reference's bit for bit, so a depth-1 image differs from the exact one by colour rounding only (and by the odd
// material-branch decision that sits on its threshold); with F = 2 a primary ray is displaced by ~1e-7 relative and a
// sample that grazes a silhouette may land on the other side of it.
// ---------------------------------------------------------------------------------------------
namespace fastm {
RT_DEV float rcp(float x) {
#if defined(__CUDA_ARCH__)
    float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y;
#else
    return 1.0f / x;
#endif
}
RT_DEV float rsq(float x) {
#if defined(__CUDA_ARCH__)
    float y; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y;
#else
    return 1.0f / sqrtf(x);
#endif
}
RT_DEV float sqr(float x) {
#if defined(__CUDA_ARCH__)
    float y; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y;
#else
    return sqrtf(x);
#endif
}
// sin and cos of an angle in [0, 2 pi] (the shader only forms u * 2 pi): MUFU.SIN / MUFU.COS are accurate to 2^-21
// absolute on [-pi, pi], so the argument is shifted by pi and the signs flipped.
RT_DEV void sincos2pi(float a, float* s_out, float* c_out) {
#if defined(__CUDA_ARCH__)
    const float b = a - 3.14159274101257324f;
    float s, c;
    asm("sin.approx.ftz.f32 %0, %1;" : "=f"(s) : "f"(b));
    asm("cos.approx.ftz.f32 %0, %1;" : "=f"(c) : "f"(b));
    *s_out = -s; *c_out = -c;
#else
    *s_out = sinf(a); *c_out = cosf(a);
#endif
}
// 1 / sqrt(a) to ~2^-23: MUFU.RSQ and one Newton step (the broad phases assume unit directions good to 2^-22)
RT_DEV float rsq_refined(float a) {
    const float y = rsq(a);
    const float h = 0.5f * a * y;
    return fmaf(y, fmaf(-h, y, 0.5f), y);
}
}  // namespace fastm

template <int F> RT_DEV float dot_t(f3 a, f3 b) {
    if constexpr (F) return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z));
    else return dot(a, b);
}
template <int F> RT_DEV f3 axpy_t(float s, f3 a, f3 b) {   // s * a + b
    if constexpr (F) return mk(fmaf(s, a.x, b.x), fmaf(s, a.y, b.y), fmaf(s, a.z, b.z));
    else return add(b, scale(s, a));
}
// normalize() of the metal / dielectric branches: the shared out-of-line IEEE copy, or v * rsqrt(v.v) inline
template <int F> RT_DEV f3 normalize_branch(f3 a) {
    if constexpr (F) return scale(fastm::rsq(dot_t<1>(a, a)), a);
    else return normalize_shared(a);
}

// WGSL u32(f32): truncate, saturate, NaN -> 0
RT_DEV uint32_t f2u(float f) {
#if defined(__CUDA_ARCH__)
    return __float2uint_rz(f);
#else
    if (!(f > 0.0f)) return 0u;
    if (f >= 4294967296.0f) return 0xFFFFFFFFu;
    return (uint32_t)f;
#endif
}

// compute_shader.wgsl:50-59 (naga precedence: shift before xor)
RT_DEV uint32_t hash(uint32_t s) {
    s ^= 2747636419u;
    s *= 2654435769u;
    s ^= s >> 16;
    s *= 2654435769u;
    s ^= s >> 16;
    s *= 2654435769u;
    return s;
}
// compute_shader.wgsl:61-63 — the f32 literal 4294967295.0 is 2^32
RT_DEV float random_float(uint32_t v) { return (float)hash(v) * 2.3283064365386963e-10f; }

// compute_shader.wgsl:234-243
template <int F = 0>
RT_DEV f3 random_unit_vector(uint32_t seed) {
    float z = 2.0f * random_float(seed) - 1.0f;
    float a = random_float(seed + 1u) * 6.283185307f;
    float sn, cs;
    if constexpr (F) {
        const float r = fastm::sqr(fmaxf(fmaf(-z, z, 1.0f), 0.0f));
        fastm::sincos2pi(a, &sn, &cs);
        return mk(r * cs, r * sn, z);
    } else {
        float r = sqrtf(1.0f - z * z);
        rt_sincos(a, &sn, &cs);
        return mk(r * cs, r * sn, z);
    }
}

// ray_color's seed does not depend on the pixel (compute_shader.wgsl:268, 353: hash(seed + i * 1000) with seed = 1 + n +
// u32(random_seed * 2^32)), so every path of a frame with the same sample count n draws the SAME random unit vector at
// bounce i.  A caller that shades several rays in a row (a thread's strip of the primary pass) keeps the last one: two
// hashes, a square root and a sin / cos per thread instead of per ray.  Pure memoisation: same bits.  One cache per
// precision level (the levels' values differ).
struct UnitVecCache {
    uint32_t seed;
    f3 v;
    bool valid;
};
template <int F>
RT_DEV f3 random_unit_vector_cached(uint32_t seed, UnitVecCache* c) {
    if (c == nullptr) return random_unit_vector<F>(seed);
    if (!(c->valid && c->seed == seed)) {
        c->v = random_unit_vector<F>(seed);
        c->seed = seed;
        c->valid = true;
    }
    return c->v;
}

struct Ray {
    f3 o, d;
};

// compute_shader.wgsl:305-331 (get_ray, sample_square, defocus_disk_sample)
template <int F = 0>
RT_DEV Ray get_ray(const SceneCamera& cam, int x, int y, uint32_t sample_index, uint32_t seed_bits) {
    uint32_t seed = hash(hash((uint32_t)x * 73u) ^ hash((uint32_t)y * 51u) ^ (sample_index * 25u + seed_bits));
    float ox = random_float(seed) - 0.5f;
    float oy = random_float(seed * seed) - 0.5f;
    float sx = ((float)x + 0.5f) + ox;
    float sy = ((float)y + 0.5f) + oy;
    if constexpr (F >= 2) {   // RT_PRECISION_FAST_ALL: the primary ray itself is approximate (silhouette pixels may flip)
        Ray r;
        const f3 pc = axpy_t<1>(sy, mk(cam.pixel_delta_v), axpy_t<1>(sx, mk(cam.pixel_delta_u), mk(cam.viewport_upper_left)));
        r.o = mk(cam.center);
        if (cam.defocus_angle > 0.0f) {
            // (cos, sin) of the approximation is unit to 2^-20: the reference's normalize of it is a no-op at that level
            float sn, cs;
            fastm::sincos2pi((2.0f * 3.1415926f) * random_float(seed + 1u), &sn, &cs);
            r.o = axpy_t<1>(sn, mk(cam.defocus_disk_v), axpy_t<1>(cs, mk(cam.defocus_disk_u), r.o));
        }
        r.d = sub(pc, r.o);
        return r;
    }
    f3 pc = add(add(mk(cam.viewport_upper_left), mk(cam.pixel_delta_u.x * sx, cam.pixel_delta_u.y * sx, cam.pixel_delta_u.z * sx)),
                mk(cam.pixel_delta_v.x * sy, cam.pixel_delta_v.y * sy, cam.pixel_delta_v.z * sy));
    Ray r;
    if (cam.defocus_angle > 0.0f) {
        float angle = (2.0f * 3.1415926f) * random_float(seed + 1u);
        float sn, cs;
        rt_sincos(angle, &sn, &cs);
        float len = sqrtf(cs * cs + sn * sn);
        float px = cs / len, py = sn / len;
        r.o = add(add(mk(cam.center), scale(px, mk(cam.defocus_disk_u))), scale(py, mk(cam.defocus_disk_v)));
    } else {
        r.o = mk(cam.center);
    }
    r.d = sub(pc, r.o);
    return r;
}

// The quadratic of sphere_hit (compute_shader.wgsl:183-201) up to the accepted root.
// `a` = dot(d, d) is loop invariant and passed in.  Returns true and the root when the sphere is
// hit inside (tmin, tmax) with the reference's <= / <= rejection.
RT_DEV bool sphere_root(f3 o, f3 d, float a, f3 center, float radius, float tmin, float tmax, float* root_out) {
    f3 oc = sub(center, o);
    float h = dot(oc, d);
    float c = dot(oc, oc) - radius * radius;
    float disc = h * h - a * c;
    if (disc < 0.0f) return false;
    float sq = sqrtf(disc);
    float root = (h - sq) / a;
    if (root <= tmin || tmax <= root) {
        root = (h + sq) / a;
        if (root <= tmin || tmax <= root) return false;
    }
    *root_out = root;
    return true;
}

// Sky gradient of ray_color (compute_shader.wgsl:293-296) times the accumulated factor.
template <int F = 0>
RT_DEV f3 sky_times(f3 factor, f3 dir) {
    float uy;
    if constexpr (F) uy = dir.y * fastm::rsq(dot_t<1>(dir, dir));
    else uy = normalize(dir).y;
    float a = 0.5f * (uy + 1.0f);
    float ia = 1.0f - a;
    f3 sky = mk(ia * 1.0f + a * 0.5f, ia * 1.0f + a * 0.7f, ia * 1.0f + a * 1.0f);
    return mul(factor, sky);
}

// Hit record construction (compute_shader.wgsl:204-218, 158-162) followed by the material
// dispatch of ray_color (compute_shader.wgsl:268-286).  On return *ray is the scattered ray and
// *factor has been multiplied by the attenuation.  Returns false when the sample is absorbed
// (ray_color returns black, compute_shader.wgsl:274,278,282).
//
// Tolerance modes (F >= 1) evaluate the same formulas with approximations, and the branch DECISIONS in here — front
// face, lambertian degenerate direction, metal absorb, dielectric reflect / refract — sit on thresholds: a grazing hit
// has dot(d, n) ~ 0, and which side it falls on follows from the last bits of p.  Each decision is therefore taken with
// a margin several times the approximation's error; a ray inside a margin sets *risky, and scatter<F> then repeats it
// with the exact arithmetic (one shared out-of-line copy).  The tolerance modes so make the reference's decisions, ray
// for ray, and differ from the exact mode by the rounding of the values only.
template <int F>
RT_DEV bool scatter_impl(Ray* ray, f3* factor, float t, const GpuSphere& sp, uint32_t bounce_seed, bool* risky,
                         UnitVecCache* ruv_cache = nullptr) {
    f3 o = ray->o, d = ray->d;
    f3 center = mk(sp.position);
    f3 p, outward;
    if constexpr (F) {
        p = axpy_t<1>(t, d, o);
        const float inv_r = fastm::rcp(sp.radius);
        outward = scale(inv_r, sub(p, center));
        // |dot(d, outward)| against the rounding of p (a few 2^-24 of |o| + |t d| + |c|, divided by r) times |d|
        const float dd = dot_t<1>(d, d);
        const float len_d = dd * fastm::rsq(dd);
        const float mag = (fabsf(o.x) + fabsf(o.y) + fabsf(o.z)) + fabsf(t) * (fabsf(d.x) + fabsf(d.y) + fabsf(d.z)) +
                          (fabsf(center.x) + fabsf(center.y) + fabsf(center.z));
        const float tol = len_d * fmaf(mag * 7.62939453125e-06f /* 2^-17 */, fabsf(inv_r), 7.62939453125e-06f);
        *risky = !(fabsf(dot_t<1>(d, outward)) > tol);
    } else {
        p = add(o, scale(t, d));
        outward = divs(sub(p, center), sp.radius);
    }
    bool front = dot_t<F>(d, outward) < 0.0f;
    f3 n = front ? outward : neg(outward);
    float w = sp.color[3];
    f3 albedo = mk(sp.color[0], sp.color[1], sp.color[2]);
    Ray out;
    out.o = p;
    if (w < -1.0f) {
        // lambertian_scatter, compute_shader.wgsl:84-93
        f3 dir = add(n, random_unit_vector_cached<F>(bounce_seed, ruv_cache));
        if constexpr (F) *risky |= !(dot_t<1>(dir, dir) > 1e-3f);   // nearly opposite vectors: the direction is ill-conditioned
        if (dot_t<F>(dir, dir) < 1e-6f) dir = n;
        out.d = dir;
        *factor = mul(*factor, albedo);
        *ray = out;
        return true;
    } else if (w <= 1.0f) {
        // metal_scatter, compute_shader.wgsl:95-100; reflect(e1,e2) = e1 - 2 dot(e2,e1) e2
        f3 refl = sub(d, scale(2.0f * dot_t<F>(n, d), n));
        f3 reflected = add(normalize_branch<F>(refl), scale(w, random_unit_vector_cached<F>(bounce_seed, ruv_cache)));
        out.d = normalize_branch<F>(reflected);
        *factor = mul(*factor, albedo);
        *ray = out;
        if constexpr (F) *risky |= !(fabsf(dot_t<1>(reflected, n)) > 1e-4f) || !(dot_t<1>(reflected, reflected) > 1e-3f);
        return dot_t<F>(reflected, n) > 0.0f;
    } else {
        // dielectric_scatter + reflectance, compute_shader.wgsl:102-141
        float ior = sp.color[0];
        float ratio, sin_theta, r0;
        f3 ud = normalize_branch<F>(d);
        float cos_theta = fminf(dot_t<F>(neg(ud), n), 1.0f);
        if constexpr (F) {
            ratio = front ? fastm::rcp(ior) : ior;
            sin_theta = fastm::sqr(fmaxf(fmaf(-cos_theta, cos_theta, 1.0f), 0.0f));
            r0 = (1.0f - ratio) * fastm::rcp(1.0f + ratio);
        } else {
            ratio = front ? (1.0f / ior) : ior;
            sin_theta = sqrtf(1.0f - cos_theta * cos_theta);
            r0 = (1.0f - ratio) / (1.0f + ratio);
        }
        bool cannot_refract = ratio * sin_theta > 1.0f;
        r0 = r0 * r0;
        float refl_prob = r0 + (1.0f - r0) * rt_pow5(1.0f - cos_theta);
        bool should_reflect = cannot_refract || refl_prob > random_float(bounce_seed);
        if constexpr (F)
            *risky |= !(fabsf(ratio * sin_theta - 1.0f) > 1e-4f) || !(fabsf(refl_prob - random_float(bounce_seed)) > 1e-4f);
        f3 dir;
        if (should_reflect) {
            dir = sub(ud, scale(2.0f * dot_t<F>(n, ud), n));
        } else {
            // WGSL refract(e1, e2, e3)
            float dn = dot_t<F>(n, ud);
            float k = 1.0f - (ratio * ratio) * (1.0f - dn * dn);
            if (k < 0.0f) {
                dir = mk(0.0f, 0.0f, 0.0f);
            } else {
                const float sk = F ? fastm::sqr(k) : sqrtf(k);
                dir = sub(scale(ratio, ud), scale(ratio * dn + sk, n));
            }
        }
        out.d = normalize_branch<F>(dir);
        // attenuation = vec3(1.0): color_factor *= 1.0 is exact, skipped
        *ray = out;
        return true;
    }
}
RT_DEV_SHARED bool scatter_exact_shared(Ray* ray, f3* factor, float t, const GpuSphere& sp, uint32_t bounce_seed) {
    bool unused = false;
    return scatter_impl<0>(ray, factor, t, sp, bounce_seed, &unused);
}
template <int F = 0>
RT_DEV bool scatter(Ray* ray, f3* factor, float t, const GpuSphere& sp, uint32_t bounce_seed, UnitVecCache* ruv_cache = nullptr) {
    bool risky = false;
    if constexpr (F == 0) {
        return scatter_impl<0>(ray, factor, t, sp, bounce_seed, &risky, ruv_cache);
    } else {
        Ray r = *ray;
        f3 f = *factor;
        const bool ok = scatter_impl<F>(&r, &f, t, sp, bounce_seed, &risky, ruv_cache);
        if (risky) return scatter_exact_shared(ray, factor, t, sp, bounce_seed);
        *ray = r;
        *factor = f;
        return ok;
    }
}


// ---------------------------------------------------------------------------------------------
// Broad phase.  Conservative sign test of the discriminant with fused
// arithmetic: 7 FFMA per (ray, sphere).  Shared by the kernels and the CPU emulation in tests/emu.
// ---------------------------------------------------------------------------------------------
#define RT_KAPPA 1.52587890625e-05 /* 2^-16 */
#define RT_BROAD_MAG 1073741824.0 /* 2^30: |C_i|, |r| window of the general records */

struct BroadRay {
    float ux, uy, uz;     // d / |d|
    float h0;             // -dot(O, u)
    float mox, moy, moz;  // -2 O
    float thr;            // |O|^2 (1 - kappa)
};

template <int F = 0>
RT_DEV BroadRay broad_arm(f3 o, f3 d, float a) {
    BroadRay b;
    float inv;
    if constexpr (F) inv = fastm::rsq_refined(a);
    else inv = 1.0f / sqrtf(a);
    b.ux = d.x * inv; b.uy = d.y * inv; b.uz = d.z * inv;
    b.h0 = -((o.x * b.ux + o.y * b.uy) + o.z * b.uz);
    b.mox = -2.0f * o.x; b.moy = -2.0f * o.y; b.moz = -2.0f * o.z;
    float oo = dot(o, o);
    b.thr = fmaf(-(float)RT_KAPPA, oo, oo);
    return b;
}

// g >= thr  <=>  candidate.  (sx, sy, sz, sk) is the per-sphere record of broad_record().
RT_DEV float broad_g(float sx, float sy, float sz, float sk, float ux, float uy, float uz, float h0, float mox,
                     float moy, float moz) {
    float h = fmaf(sx, ux, fmaf(sy, uy, fmaf(sz, uz, h0)));
    float c = fmaf(sx, mox, fmaf(sy, moy, fmaf(sz, moz, sk)));
    return fmaf(h, h, -c);
}

// k = |C|^2 - r^2 - kappa (|C|^2 + r^2), evaluated in binary64 and rounded toward -infinity.  A sphere outside the
// magnitude window (any of |C_i|, |r| beyond 2^30, or not finite) gets k = -infinity, which passes for every ray:
// beyond the window the float test itself would overflow (c = +inf rejects a sphere whose reference discriminant
// is NaN, i.e. a hit), and the reference's arithmetic is no longer covered by the kappa bound anyway.
RT_DEV float broad_record_k(const GpuSphere& s) {
    const double cx = s.position.x, cy = s.position.y, cz = s.position.z, r = s.radius;
    if (!(fabs(cx) <= RT_BROAD_MAG && fabs(cy) <= RT_BROAD_MAG && fabs(cz) <= RT_BROAD_MAG && fabs(r) <= RT_BROAD_MAG))
        return -INFINITY;
    const double cc = cx * cx + cy * cy + cz * cz, rr = r * r;
    const double k = cc - rr - RT_KAPPA * (cc + rr);
    float f = (float)k;
    if ((double)f > k) f = nextafterf(f, -INFINITY);
    return f;
}

// ---------------------------------------------------------------------------------------------
// Common-origin broad phase.  Primary rays start at cam.center = Oc
// (pinhole, compute_shader.wgsl:319) or on the rim of the defocus disk around it (:327-331), i.e.
// at O = Oc + e with |e| <= rho.  A ray from O that comes within R of a sphere centre C at some
// t > 0 has its parallel through Oc come within R + rho of C at the same t, so with
//     oc = C - Oc,  u = d / |d|,  R^2 = r^2 + kappa S  (kappa S absorbs every rounding of the
//     reference's own discriminant, S = |oc|^2 + r^2 + rho^2),  c' = |oc|^2 - (R + rho)^2 > 0,
// a hit needs  oc.u >= sqrt(c').  The test is ONE 3-term dot product against a per-sphere,
// per-frame record (oc, -q), q = sqrt(c') minus the slack for the broad phase's own rounding:
//     candidate  <=>  sign bit of fma(ocx, ux, fma(ocy, uy, fma(ocz, uz, -q))) clear.
// 3 FFMA per (ray, sphere).  Spheres the bound does not cover (Oc inside or within the slack of
// the inflated sphere, non-finite or extreme magnitudes) get the record (0, 0, 0, +1): always a
// candidate.  Rays outside the magnitude window of primary_arm() are never candidates and must be
// traced exactly by the caller.
// ---------------------------------------------------------------------------------------------
#define RT_PRIMARY_MAG 1073741824.0 /* 2^30: |oc_i|, |r|, |Oc_i|, rho window */

RT_DEV float rt_double2float_rd(double v) {
#if defined(__CUDA_ARCH__)
    return __double2float_rd(v);
#else
    float f = (float)v;
    if ((double)f > v) f = nextafterf(f, -INFINITY);
    return f;
#endif
}

// Upper bound of |O - cam.center| over the primary rays of this camera, or a negative value when
// the primary pass must not be used (non-finite or extreme camera).  Pinhole: 0.  Defocus:
// e = px ddu + py ddv with |(px, py)| = 1 up to rounding, so |e|^2 <= lambda_max of the Gram matrix
// of (ddu, ddv); the float evaluation of O adds a few ulps of |Oc| + |e|.
RT_DEV double primary_rho(const SceneCamera& cam) {
    const double lim = RT_PRIMARY_MAG;
    const double cx = cam.center.x, cy = cam.center.y, cz = cam.center.z;
    if (!(fabs(cx) <= lim && fabs(cy) <= lim && fabs(cz) <= lim)) return -1.0;
    if (!(cam.defocus_angle > 0.0f)) return (cam.defocus_angle <= 0.0f) ? 0.0 : -1.0;   // NaN angle: WGSL select() takes the pinhole branch, but stay out
    const double ux = cam.defocus_disk_u.x, uy = cam.defocus_disk_u.y, uz = cam.defocus_disk_u.z;
    const double vx = cam.defocus_disk_v.x, vy = cam.defocus_disk_v.y, vz = cam.defocus_disk_v.z;
    const double A = ux * ux + uy * uy + uz * uz, B = vx * vx + vy * vy + vz * vz, C = ux * vx + uy * vy + uz * vz;
    const double half = 0.5 * (A - B);
    const double lam = 0.5 * (A + B) + sqrt(half * half + C * C);
    const double cmax = fmax(fmax(fabs(cx), fabs(cy)), fabs(cz));
    const double rho = sqrt(lam) * (1.0 + 1e-5) + 4.76837158203125e-07 /* 2^-21 */ * (cmax + sqrt(lam));
    if (!(rho <= lim)) return -1.0;
    return rho;
}

// Record of sphere `s` for rays starting within `rho` of `o`.  out = (ocx, ocy, ocz, -q).
RT_DEV void primary_record(const GpuSphere& s, f3 o, double rho, float out[4]) {
    const double ocx = (double)s.position.x - (double)o.x, ocy = (double)s.position.y - (double)o.y,
                 ocz = (double)s.position.z - (double)o.z, r = fabs((double)s.radius);
    const double cc = ocx * ocx + ocy * ocy + ocz * ocz, rr = r * r, S = cc + rr + rho * rho;
    const double R = sqrt(rr + RT_KAPPA * S) + rho;
    const double cp = cc - R * R * (1.0 + 9.5367431640625e-07 /* 2^-20 */);
    const double m = fmax(fmax(fabs(ocx), fabs(ocy)), fmax(fabs(ocz), r));
    // written so that a NaN anywhere fails the test
    bool ok = (m <= RT_PRIMARY_MAG) && (S >= 8.673617379884035e-19 /* 2^-60 */) && (cp > 0.0);
    double q = 0.0;
    if (ok) {
        q = sqrt(cp) * (1.0 - 9.5367431640625e-07) - 3.814697265625e-06 /* 2^-18 */ * sqrt(S);
        ok = q > 0.0;
    }
    if (!ok) {
        out[0] = 0.0f; out[1] = 0.0f; out[2] = 0.0f; out[3] = 1.0f;
        return;
    }
    out[0] = (float)ocx; out[1] = (float)ocy; out[2] = (float)ocz;
    out[3] = -rt_double2float_rd(q);
}

// Record of a padding slot: never a candidate.
RT_DEV void primary_record_pad(float out[4]) { out[0] = 0.0f; out[1] = 0.0f; out[2] = 0.0f; out[3] = -1.0f; }

// Unit direction of a ray for the common-origin test.  Returns false (and u = 0) when |d|^2 is
// outside [2^-40, 2^40] or not finite: such a ray has to be traced with the exact test only.
template <int F = 0>
RT_DEV bool primary_arm(f3 d, float a, float* ux, float* uy, float* uz, float* inv_len = nullptr) {
    const bool ok = (a >= 9.094947017729282e-13f) && (a <= 1099511627776.0f);
    if (!ok) {
        *ux = 0.0f; *uy = 0.0f; *uz = 0.0f;
        return false;
    }
    float inv;
    if constexpr (F) inv = fastm::rsq_refined(a);
    else inv = 1.0f / sqrtf(a);
    *ux = d.x * inv; *uy = d.y * inv; *uz = d.z * inv;
    if (inv_len) *inv_len = inv;   // 1 / |d|
    return true;
}

// h' of one (ray, record): candidate <=> sign bit clear.
RT_DEV float primary_h(float rx, float ry, float rz, float rw, float ux, float uy, float uz) {
    return fmaf(rx, ux, fmaf(ry, uy, fmaf(rz, uz, rw)));
}
RT_DEV bool primary_candidate(float h) {
#if defined(__CUDA_ARCH__)
    return __float_as_int(h) >= 0;
#else
    int32_t b;
    memcpy(&b, &h, 4);
    return b >= 0;
#endif
}

// ---------------------------------------------------------------------------------------------
// Tile level of the common-origin test.  The per-ray test  oc.u - q >= 0  says: the angle between u and oc is at
// most alpha, cos(alpha) = q / |oc| (alpha = the angular radius of the inflated sphere seen from the camera).  If
// every ray direction of a pixel tile is within the angle eps of a tile axis m (|m| = 1), a record can only be a
// candidate of one of them when the angle between m and oc is at most alpha + eps:
//     oc.m >= |oc| cos(alpha + eps) = q cos(eps) - R sin(eps),    R = |oc| sin(alpha) = sqrt(|oc|^2 - q^2).
// ONE such test per (tile, record) discards the record for all the tile's rays.  eps enters through the chord
// k = max |u - m| (cos(eps) = 1 - k^2/2, sin(eps) <= k for k < 1).  Floating point: the per-ray test is evaluated in
// floats (primary_h), h_fl >= 0 implies oc.u - q >= -2^-21 |oc|; the dot product here carries the same error, u and
// m are unit to 2^-22.  k is rounded up by 2^-10 relative + 2^-18 absolute, R^2 by 2^-18 |oc|^2, and 2^-18 |oc| is
// added to the left side — each several times the error it covers — so the tile test keeps every record that any
// per-ray test of the tile keeps: the set of exact tests, and with it the result, is unchanged.
// ---------------------------------------------------------------------------------------------
// The camera's share of a group's bounding radius in the primary pass over the tree (rt_tree.h): every member's
// inflated sphere (radius sqrt(r^2 + kappa S) + rho, S <= (|C| + |Oc| + r + rho)^2) must lie inside the group's, whose
// stored radius covers the |C| and r terms; m_cam = 2^-7 (|Oc| + rho), rounded up, covers the rest twice over.
RT_DEV float primary_m_cam(const SceneCamera& cam, double rho) {
    const double cx = cam.center.x, cy = cam.center.y, cz = cam.center.z;
    const double m = 0.0078125 * (sqrt(cx * cx + cy * cy + cz * cz) + rho) * (1.0 + 0.0009765625);
    float f = (float)m;
    if ((double)f < m) f = nextafterf(f, INFINITY);
    return f;
}
// Record of a group of the tree for this frame: bounding sphere (centre, radius + m_cam) through primary_record.
RT_DEV void primary_group_record(float cx, float cy, float cz, float radius, float m_cam, f3 o, double rho, float out[4]) {
    GpuSphere g;
    g.position.x = cx; g.position.y = cy; g.position.z = cz;
    g.radius = radius + m_cam;
    g.color[0] = 0.0f; g.color[1] = 0.0f; g.color[2] = 0.0f; g.color[3] = 0.0f;
    primary_record(g, o, rho, out);
}

// Axis of a tile from the bounding box [lo, hi] of its rays' unit directions.
RT_DEV f3 primary_tile_axis(f3 lo, f3 hi) {
    const f3 m = mk(0.5f * (lo.x + hi.x), 0.5f * (lo.y + hi.y), 0.5f * (lo.z + hi.z));
    const float inv = 1.0f / sqrtf(dot(m, m));
    return mk(m.x * inv, m.y * inv, m.z * inv);
}
RT_DEV float primary_tile_e2(float ux, float uy, float uz, f3 m) {
    const float dx = ux - m.x, dy = uy - m.y, dz = uz - m.z;
    return fmaf(dx, dx, fmaf(dy, dy, dz * dz));
}
// e2max: the largest primary_tile_e2 over the tile's traced rays.  Returns the chord bound k.
RT_DEV float primary_tile_chord(float e2max) { return sqrtf(e2max) * 1.0009765625f + 3.814697265625e-06f; }
// A NaN anywhere (degenerate axis) and k >= 1 (a tile wider than 60 degrees) keep the record.
// *near_out (optional): primary_near of the record (below), from the same |oc| and R.
RT_DEV bool primary_tile_candidate(float rx, float ry, float rz, float rw, f3 m, float k, float* near_out = nullptr) {
    const float q = -rw;
    const float t = fmaf(rx, m.x, fmaf(ry, m.y, rz * m.z));
    const float n2 = fmaf(rx, rx, fmaf(ry, ry, rz * rz));
    const float R = sqrtf(fmaxf(fmaf(n2, 1.000003814697265625f, -((q * q) * 0.999996185302734375f)), 0.0f));
    const float len = sqrtf(n2);
    if (near_out) *near_out = fmaf(-R, 1.0009765625f, len * 0.99609375f);
    if (!(k < 1.0f)) return true;
    const float ce = fmaf(-0.5f * k, k, 1.0f);
    const float lhs = fmaf(len, 3.814697265625e-06f, fmaf(R, k, fmaf(-q, ce, t)));
    return !(lhs < 0.0f);
}

// Nearest-first order and the "already beaten" cull of the primary pass.  From a sphere's common-origin record
// (oc, -q): |oc| - R, R = sqrt(|oc|^2 - q^2) >= r + rho (primary_record), is a lower bound of the distance from any
// origin on the defocus disk to the sphere's surface, so the reference's accepted root of a ray with direction d is at
// least (|oc| - R) / |d| — in exact arithmetic.  The reference's float root (h - sqrt(h^2 - a c)) / a is below the
// exact one by at most sqrt(delta) / a, delta the rounding of its discriminant: h, a and c each carry a few 2^-24 of
// |oc||d|, |d|^2 and |oc|^2 + r^2, so delta <= 2^-20 |oc|^2 |d|^2 and the root falls short by at most
// 2^-10 |oc| / |d| (its other roundings are of order 2^-22); primary_near takes 2^-8 |oc| and 2^-10 R off, four times
// that.  A sphere
// with  primary_near * (smallest 1/|d| of the rays concerned) * (1 - 2^-12)  >  the nearest hit they all hold
// cannot win for any of them (first root > best_t >= tmin: it loses or is rejected at tmax; the second root is larger
// still) and is not tested.  Negative for a camera inside the inflated sphere and 0 for an always-record: never culled.
RT_DEV float primary_near(float rx, float ry, float rz, float rw) {
    const float q = -rw;
    const float n2 = fmaf(rx, rx, fmaf(ry, ry, rz * rz));
    const float R = sqrtf(fmaxf(fmaf(n2, 1.000003814697265625f, -((q * q) * 0.999996185302734375f)), 0.0f));
    return fmaf(-R, 1.0009765625f, sqrtf(n2) * 0.99609375f);
}
RT_DEV float primary_near_scale(float inv_len_min) { return inv_len_min * 0.999755859375f; }

// ---------------------------------------------------------------------------------------------
// Group level of the general broad phase (rt_tree.h).  A group's bounding sphere has an ordinary record
// (Cc, k(Rc')); the ray's own share of the bound, m = 2^-7 |O|, enters through the threshold:
// (Rc' + m)^2 <= Rc'^2 + 2 Rmax m + m^2, so the group is descended into when g >= thr - (2 Rmax m + m^2).
// m = cluster_m(dot(O, O)); extra_r = Rmax, the largest bounding radius among the children of the node being visited.
// ---------------------------------------------------------------------------------------------
RT_DEV float cluster_m(float oo) { return sqrtf(oo) * 0.00782012939453125f; }   // 2^-7 (1 + 2^-10) |O|
RT_DEV float cluster_thr(float thr, float m, float extra_r) {
    const float extra = fmaf(m, m, (2.0f * extra_r) * m) * 1.0009765625f;
    return thr - extra;
}

// Nearest-hit update that does not depend on the order of the tests: the reference's sequential loop returns the
// lexicographic minimum of (first valid root, index) when no root is NaN.  Returns
// false when the root is NaN: the caller must then retrace the ray with the sequential loop.
RT_DEV bool nearest_update(f3 o, f3 d, float a, f3 center, float radius, float tmin, float tmax0, int idx, float* best_t,
                           int* best_i) {
    float root;
    if (!sphere_root(o, d, a, center, radius, tmin, tmax0, &root)) return true;
    if (root != root) return false;
    if (root < *best_t || (root == *best_t && idx < *best_i)) {
        *best_t = root;
        *best_i = idx;
    }
    return true;
}

// The same for rays that start inside the scene (bounce rays): their candidates come from a LINE test, so half of them
// lie behind the origin, and the reference spends a square root and two divisions on each to reject both roots.
// With h = oc.d < 0 and c = |oc|^2 - r^2 > 0 as the reference computes them, and a finite discriminant, that outcome
// is certain without them: fl(a c) >= 0, so disc <= fl(h^2) and sq = fl(sqrt(disc)) <= fl(sqrt(fl(h^2))) = |h| (square
// root and rounding are monotone), hence h + sq <= 0: the second root is <= 0 < tmin and the first is no larger.
// (h^2 = inf gives sq = inf and a root >= tmax, rejected as well; only a NaN discriminant, which the reference ACCEPTS
// with a NaN root, must take the full path — it fails `disc == disc` below.)  Same results, bit for bit.
RT_DEV bool nearest_update_from_inside(f3 o, f3 d, float a, f3 center, float radius, float tmin, float tmax0, int idx,
                                       float* best_t, int* best_i) {
    f3 oc = sub(center, o);
    float h = dot(oc, d);
    float c = dot(oc, oc) - radius * radius;
    float disc = h * h - a * c;
    if (disc < 0.0f) return true;
    if (h < 0.0f && c > 0.0f && disc == disc && tmin >= 0.0f) return true;
    float sq = sqrtf(disc);
    float root = (h - sq) / a;
    if (root <= tmin || tmax0 <= root) {
        root = (h + sq) / a;
        if (root <= tmin || tmax0 <= root) return true;
    }
    if (root != root) return false;
    if (root < *best_t || (root == *best_t && idx < *best_i)) {
        *best_t = root;
        *best_i = idx;
    }
    return true;
}

}  // namespace rtc
