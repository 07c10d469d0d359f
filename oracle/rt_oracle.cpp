// rt_oracle.cpp — CPU restatement of the reference's ray-trace pass.  TEST INFRASTRUCTURE ONLY.
//
// This file is the parity oracle (and the timed CPU baseline) for the CUDA kernels.  It restates
// /root/reference/assets/compute_shader.wgsl function by function in scalar binary32 arithmetic,
// one IEEE-754 rounding per WGSL operator (build: g++ -O2 -ffp-contract=off, no fast-math).
// Nothing in the product path (gpu_ray_tracing_b200/) may include, link or call it: only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do.
//
// PARITY UNPINNED BY THE REFERENCE: the reference repository holds no tests, golden vectors or
// fixtures for this path and its shader cannot be executed here or on
// the GPU box (WGSL through wgpu needs a graphics stack and a Rust toolchain; both unavailable).
// The pins this oracle is checked against are derived from the shader text itself:
// tests/golden/*.json (hash / random_float / camera known-answer values, sphere_hit hand cases).
//
// Third-party semantics restated (sources not vendored in the reference):
//   naga 23.1.0 (Cargo.lock:2629) — WGSL front end.  `state ^ state >> 16u` parses with shift
//   binding tighter than xor (the other reading makes hash() identically 0).
//   WGSL builtins: reflect, refract, normalize, select, u32(f32) as in the WGSL specification.
//   sin / cos / pow have implementation-defined precision in WGSL: pinned in
//   include/rt_pinned_math.h (or libm with -DRT_ORACLE_LIBM).
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../include/rt_abi.h"
#include "../include/rt_pinned_math.h"

namespace {

struct V3 {
    float x, y, z;
};
inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
inline V3 v3(const RtVec3& a) { return V3{a.x, a.y, a.z}; }
inline V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator-(V3 a) { return V3{-a.x, -a.y, -a.z}; }
inline V3 operator*(V3 a, V3 b) { return V3{a.x * b.x, a.y * b.y, a.z * b.z}; }
inline V3 operator*(float s, V3 a) { return V3{s * a.x, s * a.y, s * a.z}; }
inline V3 operator*(V3 a, float s) { return V3{a.x * s, a.y * s, a.z * s}; }
inline V3 operator/(V3 a, float s) { return V3{a.x / s, a.y / s, a.z / s}; }
// WGSL dot / length / normalize, evaluation order pinned left to right.
inline float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline float length(V3 a) { return sqrtf(dot(a, a)); }
inline V3 normalize(V3 a) { return a / length(a); }
// WGSL reflect(e1, e2) = e1 - 2 * dot(e2, e1) * e2
inline V3 reflect(V3 e1, V3 e2) { return e1 - (2.0f * dot(e2, e1)) * e2; }
// WGSL refract(e1, e2, e3): k = 1 - e3^2 (1 - dot(e2,e1)^2); k < 0 -> 0 else e3*e1 - (e3*dot(e2,e1) + sqrt(k))*e2
inline V3 refract(V3 e1, V3 e2, float e3) {
    float dn = dot(e2, e1);
    float k = 1.0f - (e3 * e3) * (1.0f - dn * dn);
    if (k < 0.0f) return v3(0.0f, 0.0f, 0.0f);
    return e3 * e1 - (e3 * dn + sqrtf(k)) * e2;
}
// WGSL u32(f32): truncate toward zero, saturate, NaN -> 0.
inline uint32_t u32_of(float f) {
    if (!(f > 0.0f)) return 0u;
    if (f >= 4294967296.0f) return 0xFFFFFFFFu;
    return (uint32_t)f;
}

inline void wgsl_sincos(float a, float* s, float* c) {
#ifdef RT_ORACLE_LIBM
    *s = sinf(a);
    *c = cosf(a);
#else
    rt_sincos(a, s, c);
#endif
}
inline float wgsl_pow5(float x) {
#ifdef RT_ORACLE_LIBM
    return powf(x, 5.0f);
#else
    return rt_pow5(x);
#endif
}

// compute_shader.wgsl:50-59
inline uint32_t hash(uint32_t value) {
    uint32_t state = value;
    state = state ^ 2747636419u;
    state = state * 2654435769u;
    state = state ^ (state >> 16u);
    state = state * 2654435769u;
    state = state ^ (state >> 16u);
    state = state * 2654435769u;
    return state;
}

// compute_shader.wgsl:61-63.  The literal 4294967295.0 is not representable in f32 and rounds to 2^32.
inline float random_float(uint32_t value) { return (float)hash(value) / 4294967296.0f; }

struct Ray {
    V3 origin, direction;
};
struct Material {
    float albedo[4];
};
// compute_shader.wgsl:143-149
struct HitRecord {
    float t;
    V3 p, normal;
    bool front_face;
    Material material;
};

struct Ctx {
    const SceneCamera* camera;
    const GpuSphere* spheres;
    uint32_t sphere_count;
};

// compute_shader.wgsl:234-243
inline V3 random_unit_vector(uint32_t seed) {
    float z = 2.0f * random_float(seed) - 1.0f;
    float a = random_float(seed + 1u) * 6.283185307f;
    float r = sqrtf(1.0f - z * z);
    float sn, cs;
    wgsl_sincos(a, &sn, &cs);
    float x = r * cs;
    float y = r * sn;
    return v3(x, y, z);
}

// compute_shader.wgsl:84-93
inline bool lambertian_scatter(const Material& m, const Ray&, const HitRecord& rec, V3* attenuation,
                               Ray* scattered, uint32_t seed) {
    V3 dir = rec.normal + random_unit_vector(seed);
    if (dot(dir, dir) < 1e-6f) dir = rec.normal;
    *scattered = Ray{rec.p, dir};
    *attenuation = v3(m.albedo[0], m.albedo[1], m.albedo[2]);
    return true;
}

// compute_shader.wgsl:95-100
inline bool metal_scatter(const Material& m, const Ray& ray, const HitRecord& rec, V3* attenuation,
                          Ray* scattered, uint32_t seed) {
    V3 reflected = normalize(reflect(ray.direction, rec.normal)) + m.albedo[3] * random_unit_vector(seed);
    *scattered = Ray{rec.p, normalize(reflected)};
    *attenuation = v3(m.albedo[0], m.albedo[1], m.albedo[2]);
    return dot(reflected, rec.normal) > 0.0f;
}

// compute_shader.wgsl:137-141
inline float reflectance(float cos_theta, float refractive_index) {
    float r0 = (1.0f - refractive_index) / (1.0f + refractive_index);
    r0 *= r0;
    return r0 + (1.0f - r0) * wgsl_pow5(1.0f - cos_theta);
}

// compute_shader.wgsl:102-135
inline bool dielectric_scatter(const Material& m, const Ray& ray, const HitRecord& rec, V3* attenuation,
                               Ray* scattered, uint32_t seed) {
    *attenuation = v3(1.0f, 1.0f, 1.0f);
    // select(f, t, cond): 1/ior when front_face
    float refraction_ratio = rec.front_face ? (1.0f / m.albedo[0]) : m.albedo[0];
    V3 unit_direction = normalize(ray.direction);
    float cos_theta = fminf(dot(-unit_direction, rec.normal), 1.0f);
    float sin_theta = sqrtf(1.0f - cos_theta * cos_theta);
    bool cannot_refract = refraction_ratio * sin_theta > 1.0f;
    bool should_reflect = cannot_refract || reflectance(cos_theta, refraction_ratio) > random_float(seed);
    // select() evaluates both operands; only the chosen one is observable.
    V3 direction = should_reflect ? reflect(unit_direction, rec.normal)
                                  : refract(unit_direction, rec.normal, refraction_ratio);
    *scattered = Ray{rec.p, normalize(direction)};
    return true;
}

// compute_shader.wgsl:158-162
inline void hit_record_set_face_normal(HitRecord* rec, const Ray& r, V3 outward_normal) {
    bool front_face = dot(r.direction, outward_normal) < 0.0f;
    rec->normal = front_face ? outward_normal : -outward_normal;
    rec->front_face = front_face;
}

// compute_shader.wgsl:182-221
inline bool sphere_hit(const GpuSphere& sphere, const Ray& r, float ray_tmin, float ray_tmax, HitRecord* rec) {
    V3 center = v3(sphere.position);
    V3 oc = center - r.origin;
    float a = dot(r.direction, r.direction);
    float h = dot(oc, r.direction);
    float c = dot(oc, oc) - sphere.radius * sphere.radius;
    float discriminant = h * h - a * c;
    if (discriminant < 0.0f) return false;
    float square_root = sqrtf(discriminant);
    float root = (h - square_root) / a;
    if (root <= ray_tmin || ray_tmax <= root) {
        root = (h + square_root) / a;
        if (root <= ray_tmin || ray_tmax <= root) return false;
    }
    V3 hit_point = r.origin + root * r.direction;
    V3 outward_normal = (hit_point - center) / sphere.radius;
    rec->t = root;
    rec->p = hit_point;
    rec->normal = outward_normal;
    rec->front_face = false;
    memcpy(rec->material.albedo, sphere.color, sizeof(float) * 4);
    hit_record_set_face_normal(rec, r, outward_normal);
    return true;
}

// compute_shader.wgsl:164-180
inline bool sphere_list_hit(const Ctx& ctx, const Ray& r, float ray_tmin, float ray_tmax, HitRecord* rec) {
    HitRecord temp_rec{};
    bool hit_anything = false;
    float closest_so_far = ray_tmax;
    for (uint32_t i = 0; i < ctx.sphere_count; i++) {
        if (sphere_hit(ctx.spheres[i], r, ray_tmin, closest_so_far, &temp_rec)) {
            hit_anything = true;
            closest_so_far = temp_rec.t;
            *rec = temp_rec;
        }
    }
    return hit_anything;
}

// compute_shader.wgsl:261-297
inline V3 ray_color(const Ctx& ctx, Ray r, uint32_t seed, uint64_t* rays) {
    V3 color_factor = v3(1.0f, 1.0f, 1.0f);
    uint32_t max_depth = u32_of(ctx.camera->max_depth);
    for (uint32_t i = 0; i < max_depth; i++) {
        HitRecord rec{};
        ++*rays;
        if (sphere_list_hit(ctx, r, 0.001f, 3.4e35f, &rec)) {
            uint32_t s = hash(seed + i * 1000u);
            Ray scattered{v3(0, 0, 0), v3(0, 0, 0)};
            V3 attenuation = v3(0, 0, 0);
            bool ok;
            if (rec.material.albedo[3] < -1.0f) {
                ok = lambertian_scatter(rec.material, r, rec, &attenuation, &scattered, s);
            } else if (rec.material.albedo[3] <= 1.0f) {
                ok = metal_scatter(rec.material, r, rec, &attenuation, &scattered, s);
            } else {
                ok = dielectric_scatter(rec.material, r, rec, &attenuation, &scattered, s);
            }
            if (!ok) return v3(0.0f, 0.0f, 0.0f);
            color_factor = color_factor * attenuation;
            r = scattered;
        } else {
            break;
        }
    }
    V3 unit_direction = normalize(r.direction);
    float a = 0.5f * (unit_direction.y + 1.0f);
    V3 sky_color = (1.0f - a) * v3(1.0f, 1.0f, 1.0f) + a * v3(0.5f, 0.7f, 1.0f);
    return color_factor * sky_color;
}

// compute_shader.wgsl:299-303
inline V3 sample_square(uint32_t seed) {
    float x = random_float(seed) - 0.5f;
    float y = random_float(seed * seed) - 0.5f;
    return v3(x, y, 0.0f);
}

// compute_shader.wgsl:327-331
inline V3 defocus_disk_sample(const SceneCamera& cam, uint32_t seed) {
    float angle = (2.0f * 3.1415926f) * random_float(seed);
    float sn, cs;
    wgsl_sincos(angle, &sn, &cs);
    float len = sqrtf(cs * cs + sn * sn);  // normalize(vec2(cos, sin))
    float px = cs / len, py = sn / len;
    return (v3(cam.center) + px * v3(cam.defocus_disk_u)) + py * v3(cam.defocus_disk_v);
}

// compute_shader.wgsl:305-325
inline Ray get_ray(const SceneCamera& cam, int32_t lx, int32_t ly, uint32_t sample_index) {
    uint32_t seed = hash(hash((uint32_t)lx * 73u) ^ hash((uint32_t)ly * 51u) ^
                         (sample_index * 25u + u32_of(cam.random_seed * 4294967296.0f)));
    V3 offset = sample_square(seed);
    V3 pixel_center = (v3(cam.viewport_upper_left) + v3(cam.pixel_delta_u) * (((float)lx + 0.5f) + offset.x)) +
                      v3(cam.pixel_delta_v) * (((float)ly + 0.5f) + offset.y);
    V3 ray_origin = (cam.defocus_angle > 0.0f) ? defocus_disk_sample(cam, seed + 1u) : v3(cam.center);
    V3 ray_direction = pixel_center - ray_origin;
    return Ray{ray_origin, ray_direction};
}

// compute_shader.wgsl:333-364, one invocation
inline void update_pixel(const Ctx& ctx, int32_t x, int32_t y, float* texel, uint64_t* rays) {
    const SceneCamera& cam = *ctx.camera;
    V3 color_until_now = v3(texel[0], texel[1], texel[2]);
    uint32_t samples_until_now = u32_of(texel[3]);
    uint32_t samples_per_pixel = u32_of(cam.samples_per_pixel);
    bool reset = cam.camera_has_moved > 0.5f;
    if (reset) {
        color_until_now = v3(0.0f, 0.0f, 0.0f);
        samples_until_now = 0u;
    }
    if (samples_until_now < samples_per_pixel) {
        uint32_t seed = 1u + samples_until_now + u32_of(cam.random_seed * 4294967296.0f);
        Ray ray = get_ray(cam, x, y, seed);
        V3 color = ray_color(ctx, ray, seed + 1u, rays);
        color_until_now = color_until_now + (color - color_until_now) / (float)(samples_until_now + 1u);
        samples_until_now += 1u;
    }
    texel[0] = color_until_now.x;
    texel[1] = color_until_now.y;
    texel[2] = color_until_now.z;
    texel[3] = (float)samples_until_now;
}

}  // namespace

// ------------------------------------------------------------------------------------------
// C entry points used by tests/, smoke() and bench.py (ctypes)
// ------------------------------------------------------------------------------------------
extern "C" {

#define ORACLE_API __attribute__((visibility("default")))

ORACLE_API int oracle_uses_libm(void) {
#ifdef RT_ORACLE_LIBM
    return 1;
#else
    return 0;
#endif
}

ORACLE_API int oracle_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

ORACLE_API uint32_t oracle_hash(uint32_t v) { return hash(v); }
ORACLE_API float oracle_random_float(uint32_t v) { return random_float(v); }
ORACLE_API void oracle_sincos(float a, float* s, float* c) { wgsl_sincos(a, s, c); }
ORACLE_API float oracle_pow5(float x) { return wgsl_pow5(x); }
ORACLE_API void oracle_random_unit_vector(uint32_t seed, float* out3) {
    V3 v = random_unit_vector(seed);
    out3[0] = v.x; out3[1] = v.y; out3[2] = v.z;
}

// sphere_hit on explicit inputs.  out = (t, p.xyz, normal.xyz, front_face)
ORACLE_API int oracle_sphere_hit(const GpuSphere* sphere, const float* origin3, const float* dir3, float tmin,
                                 float tmax, float* out8) {
    Ray r{v3(origin3[0], origin3[1], origin3[2]), v3(dir3[0], dir3[1], dir3[2])};
    HitRecord rec{};
    bool hit = sphere_hit(*sphere, r, tmin, tmax, &rec);
    if (hit && out8) {
        out8[0] = rec.t;
        out8[1] = rec.p.x; out8[2] = rec.p.y; out8[3] = rec.p.z;
        out8[4] = rec.normal.x; out8[5] = rec.normal.y; out8[6] = rec.normal.z;
        out8[7] = rec.front_face ? 1.0f : 0.0f;
    }
    return hit ? 1 : 0;
}

// sphere_list_hit: returns hit flag; out_index = index of the sphere whose record was kept
// (found by matching t against a second pass), out8 as above.
ORACLE_API int oracle_sphere_list_hit(const GpuSphere* spheres, uint32_t count, const float* origin3,
                                      const float* dir3, float tmin, float tmax, float* out8, int32_t* out_index) {
    Ray r{v3(origin3[0], origin3[1], origin3[2]), v3(dir3[0], dir3[1], dir3[2])};
    HitRecord temp{}, rec{};
    bool hit_anything = false;
    float closest = tmax;
    int32_t idx = -1;
    for (uint32_t i = 0; i < count; i++) {
        if (sphere_hit(spheres[i], r, tmin, closest, &temp)) {
            hit_anything = true;
            closest = temp.t;
            rec = temp;
            idx = (int32_t)i;
        }
    }
    if (out_index) *out_index = idx;
    if (hit_anything && out8) {
        out8[0] = rec.t;
        out8[1] = rec.p.x; out8[2] = rec.p.y; out8[3] = rec.p.z;
        out8[4] = rec.normal.x; out8[5] = rec.normal.y; out8[6] = rec.normal.z;
        out8[7] = rec.front_face ? 1.0f : 0.0f;
    }
    return hit_anything ? 1 : 0;
}

// get_ray for one pixel: out6 = origin.xyz, direction.xyz
ORACLE_API void oracle_get_ray(const SceneCamera* cam, int32_t x, int32_t y, uint32_t sample_index, float* out6) {
    Ray r = get_ray(*cam, x, y, sample_index);
    out6[0] = r.origin.x; out6[1] = r.origin.y; out6[2] = r.origin.z;
    out6[3] = r.direction.x; out6[4] = r.direction.y; out6[5] = r.direction.z;
}

// WGSL `init` over a row window: zero the texels.
ORACLE_API void oracle_init(uint32_t width, uint32_t row_count, float* accum_rgba32f) {
    memset(accum_rgba32f, 0, (size_t)width * row_count * 4 * sizeof(float));
}

// WGSL `update` over rows [row_begin, row_begin + row_count) of a width x height image.
// accum_rgba32f holds row_count * width texels (the window only), updated in place.
// threads <= 0: use every OpenMP thread.  Returns the number of sphere_list_hit traversals.
ORACLE_API uint64_t oracle_update(const SceneCamera* camera, const GpuSphere* spheres, uint32_t count,
                                  uint32_t width, uint32_t height, uint32_t row_begin, uint32_t row_count,
                                  float* accum_rgba32f, int threads) {
    (void)height;
    Ctx ctx{camera, spheres, count};
    uint64_t rays_total = 0;
#ifdef _OPENMP
    int nt = threads > 0 ? threads : omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1) num_threads(nt) reduction(+ : rays_total)
#endif
    for (int64_t row = 0; row < (int64_t)row_count; ++row) {
        uint64_t rays = 0;
        int32_t y = (int32_t)(row_begin + (uint32_t)row);
        float* line = accum_rgba32f + (size_t)row * width * 4;
        for (uint32_t x = 0; x < width; ++x) update_pixel(ctx, (int32_t)x, y, line + (size_t)x * 4, &rays);
        rays_total += rays;
    }
    return rays_total;
}

// Linear RGBA8 quantisation used on both sides of the parity check:
// q(x) = floor(clamp(x, 0, 1) * 255 + 0.5), alpha = 255.
ORACLE_API void oracle_quantize(const float* accum_rgba32f, uint64_t n_texels, uint8_t* out_rgba8) {
    for (uint64_t i = 0; i < n_texels; ++i) {
        for (int c = 0; c < 3; ++c) {
            float v = accum_rgba32f[i * 4 + c];
            v = v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v);   // NaN -> stays NaN -> cast below gives 0
            float q = floorf(v * 255.0f + 0.5f);
            out_rgba8[i * 4 + c] = (q == q) ? (uint8_t)q : 0;
        }
        out_rgba8[i * 4 + 3] = 255;
    }
}

// SceneCamera::from(&CameraSettings) (src/camera.rs:293-351) with crate::SIZE replaced by
// (width, height) and rand::random() replaced by the caller's random_seed.  glam semantics:
// Vec3::normalize = self * (1 / length) (glam 0.29 `normalize` multiplies by length_recip),
// cross = (y*oz - oy*z, z*ox - oz*x, x*oy - ox*y).
ORACLE_API void oracle_camera_from_settings(const CameraSettings* s, uint32_t width, uint32_t height,
                                            float random_seed, SceneCamera* out) {
    auto glam_normalize = [](V3 a) {
        float inv = 1.0f / sqrtf(dot(a, a));
        return a * inv;
    };
    auto cross = [](V3 a, V3 b) {
        return v3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
    };
    // f32::to_radians: self * (PI / 180) with the constant rounded to f32
    const float rads_per_deg = 0.017453292519943295769236907684886f;
    float aspect_ratio = (float)width / (float)height;
    V3 look_from = v3(s->look_from), look_at = v3(s->look_at), vup = v3(s->vup);
    float theta = s->field_of_view * rads_per_deg;
    float h = tanf(theta / 2.0f);
    float viewport_height = 2.0f * h * s->focus_distance;
    float viewport_width = viewport_height * aspect_ratio;
    V3 w = glam_normalize(look_from - look_at);
    V3 u = glam_normalize(cross(vup, w));
    V3 v = cross(w, u);
    V3 viewport_u = viewport_width * u;
    V3 viewport_v = (-viewport_height) * v;
    V3 pixel_delta_u = viewport_u / (float)width;
    V3 pixel_delta_v = viewport_v / (float)height;
    V3 viewport_upper_left = ((look_from - s->focus_distance * w) - viewport_u / 2.0f) - viewport_v / 2.0f;
    float defocus_radius = s->focus_distance * tanf((s->defocus_angle / 2.0f) * rads_per_deg);
    V3 ddu = u * defocus_radius, ddv = v * defocus_radius;
    auto put = [](RtVec3* d, V3 a) { d->x = a.x; d->y = a.y; d->z = a.z; };
    memset(out, 0, sizeof(*out));
    put(&out->center, look_from);
    out->viewport_height = viewport_height;
    put(&out->viewport_upper_left, viewport_upper_left);
    out->viewport_width = viewport_width;
    put(&out->pixel_delta_u, pixel_delta_u);
    out->defocus_angle = s->defocus_angle;
    put(&out->pixel_delta_v, pixel_delta_v);
    out->aspect_ratio = aspect_ratio;
    put(&out->defocus_disk_u, ddu);
    put(&out->viewport_u, viewport_u);
    put(&out->defocus_disk_v, ddv);
    out->max_depth = (float)s->max_depth;
    put(&out->look_from, look_from);
    out->samples_per_pixel = (float)s->samples_per_pixel;
    put(&out->look_at, look_at);
    out->camera_has_moved = s->camera_has_moved ? 1.0f : 0.0f;
    put(&out->vup, vup);
    out->random_seed = random_seed;
    put(&out->viewport_v, viewport_v);
    out->defocus_radius = defocus_radius;
}

}  // extern "C"
