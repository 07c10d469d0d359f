// rt_host.cpp — host-side helpers on either side of the hot path: the camera uniform derivation
// (src/camera.rs:293-351), the default scene recipe (src/scene/sphere.rs:45-153) with a seeded
// generator, the benchmark scenes of SURVEY.md section 8(d) and a scripted orbit of the keyboard
// yaw controls (src/camera.rs:125-253).  Plain C++ (no CUDA); built with -ffp-contract=off so the
// derivation rounds like the reference's scalar glam code.
#include <cmath>
#include <cstdint>
#include <cstring>

#include "../../include/rt_abi.h"
#include "rt_internal_host.h"

namespace {

struct Vec3 {
    float x, y, z;
};
inline Vec3 V(const RtVec3& a) { return Vec3{a.x, a.y, a.z}; }
inline RtVec3 R(Vec3 a) { return RtVec3{a.x, a.y, a.z}; }
inline Vec3 operator+(Vec3 a, Vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline Vec3 operator-(Vec3 a, Vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline Vec3 operator*(Vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline Vec3 operator*(float s, Vec3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline Vec3 operator/(Vec3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline float dot(Vec3 a, Vec3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
// glam 0.29 Vec3::normalize: self * length_recip()
inline Vec3 normalize(Vec3 a) { return a * (1.0f / std::sqrt(dot(a, a))); }
// glam Vec3::cross
inline Vec3 cross(Vec3 a, Vec3 b) {
    return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y};
}
inline float to_radians(float deg) { return deg * 0.017453292519943295769236907684886f; }

// splitmix64; f32 draws take the top 24 bits like rand 0.9's StandardUniform for f32 ([0, 1)).
struct Rng {
    uint64_t s;
    explicit Rng(uint64_t seed) : s(seed) {}
    uint64_t next() {
        uint64_t z = (s += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    float f32() { return (float)(next() >> 40) * (1.0f / 16777216.0f); }
};


// glam 0.29 Quat as the controller uses it (camera.rs:168-232): from_rotation_y / from_axis_angle build
// (axis * sin(angle/2), cos(angle/2)); mul_vec3 is glam's scalar formula
//   v * (w*w - b.b) + b * (2 * v.b) + (b x v) * (2 * w),  b = (x, y, z).
// glam is a Cargo dependency whose source is not under /root/reference, so this follows its published definition;
// the last bits of a pose may differ from a real build (sin/cos come from libm here).
struct Quat {
    float x, y, z, w;
};
inline Quat quat_axis_angle(Vec3 axis, float angle) {
    const float s = std::sin(angle * 0.5f), c = std::cos(angle * 0.5f);
    return Quat{axis.x * s, axis.y * s, axis.z * s, c};
}
inline Vec3 quat_mul_vec3(Quat q, Vec3 v) {
    const Vec3 b{q.x, q.y, q.z};
    const float b2 = dot(b, b);
    return (v * (q.w * q.w - b2) + b * (dot(v, b) * 2.0f)) + cross(b, v) * (q.w * 2.0f);
}
inline float length(Vec3 a) { return std::sqrt(dot(a, a)); }

// One yaw step of the arrow keys: look_from swings about look_at, distance kept (camera.rs:167-186).
inline void yaw_about_target(CameraSettings* c, float angle) {
    const Quat rot = quat_axis_angle(Vec3{0.0f, 1.0f, 0.0f}, angle);
    const Vec3 view = V(c->look_from) - V(c->look_at);
    const float len = length(view);
    c->look_from = R(V(c->look_at) + normalize(quat_mul_vec3(rot, view)) * len);
}

// One pitch step of the 1 / 2 keys, refused within ~18 degrees of the poles (camera.rs:209-243).
inline bool pitch_about_target(CameraSettings* c, float angle) {
    const Vec3 view = V(c->look_from) - V(c->look_at);
    const float len = length(view);
    const Vec3 forward = normalize(view);
    const Vec3 right = normalize(cross(forward, Vec3{0.0f, 1.0f, 0.0f}));
    const Vec3 dir = normalize(quat_mul_vec3(quat_axis_angle(right, angle), forward));
    if (!(std::fabs(dir.y) < 0.95f)) return false;
    c->look_from = R(V(c->look_at) + dir * len);
    return true;
}

inline GpuSphere sphere(Vec3 c, float r, float cx, float cy, float cz, float cw) {
    GpuSphere s;
    s.position = R(c);
    s.radius = r;
    s.color[0] = cx; s.color[1] = cy; s.color[2] = cz; s.color[3] = cw;
    return s;
}

// Material draw shared by every scene recipe (sphere.rs:70-108).
inline GpuSphere small_sphere(Rng& rng, float choose_mat, Vec3 center, float radius) {
    if (choose_mat < 0.8f) {
        float ax = rng.f32() * rng.f32(), ay = rng.f32() * rng.f32(), az = rng.f32() * rng.f32();
        return sphere(center, radius, ax, ay, az, -2.0f);
    } else if (choose_mat < 0.95f) {
        float ax = 0.5f * (1.0f + rng.f32()), ay = 0.5f * (1.0f + rng.f32()), az = 0.5f * (1.0f + rng.f32());
        float fuzz = 0.5f * rng.f32();
        return sphere(center, radius, ax, ay, az, fuzz);
    }
    return sphere(center, radius, 1.5f, 0.0f, 0.0f, 2.0f);
}

}  // namespace

extern "C" {

void rt_camera_settings_default(CameraSettings* out) {
    if (!out) return;
    out->samples_per_pixel = 500;
    out->max_depth = 30;
    out->camera_has_moved = 1;
    out->field_of_view = 20.0f;
    out->look_from = RtVec3{13.0f, 2.0f, 3.0f};
    out->look_at = RtVec3{0.0f, 0.0f, 0.0f};
    out->vup = RtVec3{0.0f, 1.0f, 0.0f};
    out->defocus_angle = 0.6f;
    out->focus_distance = 10.0f;
}

int rt_camera_from_settings(const CameraSettings* settings, uint32_t width, uint32_t height, float random_seed,
                            SceneCamera* out) {
    if (!settings || !out) return rti::fail(RT_ERR_INVALID_ARG, "null argument");
    if (width == 0 || height == 0) return rti::fail(RT_ERR_INVALID_ARG, "empty image");
    const CameraSettings& c = *settings;
    const float aspect_ratio = (float)width / (float)height;
    const Vec3 look_from = V(c.look_from), look_at = V(c.look_at), vup = V(c.vup);
    const float theta = to_radians(c.field_of_view);
    const float h = std::tan(theta / 2.0f);
    const float viewport_height = 2.0f * h * c.focus_distance;
    const float viewport_width = viewport_height * aspect_ratio;
    const Vec3 w = normalize(look_from - look_at);
    const Vec3 u = normalize(cross(vup, w));
    const Vec3 v = cross(w, u);
    const Vec3 viewport_u = viewport_width * u;
    const Vec3 viewport_v = (-viewport_height) * v;
    const Vec3 pixel_delta_u = viewport_u / (float)width;
    const Vec3 pixel_delta_v = viewport_v / (float)height;
    const Vec3 viewport_upper_left = look_from - (c.focus_distance * w) - viewport_u / 2.0f - viewport_v / 2.0f;
    const float defocus_radius = c.focus_distance * std::tan(to_radians(c.defocus_angle / 2.0f));
    SceneCamera s;
    memset(&s, 0, sizeof(s));
    s.center = R(look_from);
    s.viewport_height = viewport_height;
    s.viewport_upper_left = R(viewport_upper_left);
    s.viewport_width = viewport_width;
    s.pixel_delta_u = R(pixel_delta_u);
    s.defocus_angle = c.defocus_angle;
    s.pixel_delta_v = R(pixel_delta_v);
    s.aspect_ratio = aspect_ratio;
    s.defocus_disk_u = R(u * defocus_radius);
    s.viewport_u = R(viewport_u);
    s.defocus_disk_v = R(v * defocus_radius);
    s.max_depth = (float)c.max_depth;
    s.look_from = R(look_from);
    s.samples_per_pixel = (float)c.samples_per_pixel;
    s.look_at = R(look_at);
    s.camera_has_moved = c.camera_has_moved ? 1.0f : 0.0f;
    s.vup = R(vup);
    s.random_seed = random_seed;
    s.viewport_v = R(viewport_v);
    s.defocus_radius = defocus_radius;
    *out = s;
    return RT_OK;
}

int rt_create_default_spheres(uint64_t seed, GpuSphere* out, uint32_t capacity, uint32_t* count, uint32_t* buffer_len) {
    if (!out || !count) return rti::fail(RT_ERR_INVALID_ARG, "null argument");
    const uint32_t kMaxSpheres = 100;   // sphere.rs:11 — only a pad-up-to floor
    Rng rng(seed);
    uint32_t n = 0;
    auto push = [&](const GpuSphere& s) -> bool {
        if (n >= capacity) return false;
        out[n++] = s;
        return true;
    };
    bool ok = push(sphere(Vec3{0.0f, -1000.0f, 0.0f}, 1000.0f, 0.5f, 0.5f, 0.5f, -2.0f));
    for (int a = -7; a < 7 && ok; ++a) {
        for (int b = -7; b < 7 && ok; ++b) {
            const float choose_mat = rng.f32();
            const Vec3 center{(float)a + 0.9f * rng.f32(), 0.2f, (float)b + 0.9f * rng.f32()};
            const Vec3 d = center - Vec3{4.0f, 0.2f, 0.0f};
            if (std::sqrt(dot(d, d)) > 0.9f) ok = push(small_sphere(rng, choose_mat, center, 0.2f));
        }
    }
    ok = ok && push(sphere(Vec3{0.0f, 1.0f, 0.0f}, 1.0f, 1.5f, 0.0f, 0.0f, 2.0f));
    ok = ok && push(sphere(Vec3{-4.0f, 1.0f, 0.0f}, 1.0f, 0.4f, 0.2f, 0.1f, -2.0f));
    ok = ok && push(sphere(Vec3{4.0f, 1.0f, 0.0f}, 1.0f, 0.7f, 0.6f, 0.5f, 0.0f));
    if (!ok) return rti::fail(RT_ERR_INVALID_ARG, "capacity %u too small for the default scene (needs up to 200)", capacity);
    *count = n;
    uint32_t len = n;
    while (len < kMaxSpheres && len < capacity) out[len++] = sphere(Vec3{0, 0, 0}, 0.0f, 0, 0, 0, 0);
    if (buffer_len) *buffer_len = len;
    return RT_OK;
}

int rt_create_random_spheres(uint64_t seed, uint32_t n, GpuSphere* out) {
    if (!out || n == 0) return rti::fail(RT_ERR_INVALID_ARG, "null or empty output");
    Rng rng(seed);
    out[0] = sphere(Vec3{0.0f, -1000.0f, 0.0f}, 1000.0f, 0.5f, 0.5f, 0.5f, -2.0f);
    for (uint32_t i = 1; i < n; ++i) {
        const float choose_mat = rng.f32();
        const float r = 0.1f + 0.3f * rng.f32();
        const Vec3 center{-11.0f + 22.0f * rng.f32(), r, -11.0f + 22.0f * rng.f32()};
        out[i] = small_sphere(rng, choose_mat, center, r);
    }
    return RT_OK;
}

int rt_create_grid_spheres(uint64_t seed, uint32_t n, GpuSphere* out) {
    if (!out || n == 0) return rti::fail(RT_ERR_INVALID_ARG, "null or empty output");
    Rng rng(seed);
    out[0] = sphere(Vec3{0.0f, -1000.0f, 0.0f}, 1000.0f, 0.5f, 0.5f, 0.5f, -2.0f);
    uint32_t side = 1;
    while ((uint64_t)side * side < (uint64_t)(n - 1)) ++side;
    const int half = (int)(side / 2);
    uint32_t i = 1;
    for (int a = -half; a < (int)side - half && i < n; ++a) {
        for (int b = -half; b < (int)side - half && i < n; ++b) {
            const float choose_mat = rng.f32();
            const Vec3 center{(float)a + 0.9f * rng.f32(), 0.2f, (float)b + 0.9f * rng.f32()};
            out[i++] = small_sphere(rng, choose_mat, center, 0.2f);
        }
    }
    return RT_OK;
}

int rt_orbit_camera(uint32_t k, uint32_t n, float radius, float height, float fov_deg, uint32_t width,
                    uint32_t height_px, SceneCamera* out) {
    if (!out || n == 0) return rti::fail(RT_ERR_INVALID_ARG, "null output or zero poses");
    CameraSettings c;
    rt_camera_settings_default(&c);
    const float phi = 6.283185307179586f * ((float)k / (float)n);
    c.look_from = RtVec3{radius * std::cos(phi), height, radius * std::sin(phi)};
    c.look_at = RtVec3{0.0f, 0.0f, 0.0f};
    c.field_of_view = fov_deg;
    c.defocus_angle = 0.0f;
    c.focus_distance = 10.0f;
    c.max_depth = 1;
    c.samples_per_pixel = 1;
    c.camera_has_moved = 1;
    return rt_camera_from_settings(&c, width, height_px, (float)k / (float)n, out);
}

// camera_movement_system (src/camera.rs:125-253) for one frame of `dt` seconds with the keys in `keys` held.
// Kept as the reference has it, including: "forward" points from look_at to look_from, so W moves AWAY from the
// target; the arrow-left/right block appears twice in the reference, so a held arrow key yaws by 2 * dt radians per
// frame; camera_has_moved is set on any movement and cleared by the first idle frame.
int rt_camera_step(CameraSettings* c, uint32_t keys, float dt) {
    if (!c) return rti::fail(RT_ERR_INVALID_ARG, "settings is null");
    const float move_speed = 2.0f, rotate_speed = 1.0f, vertical_speed = 1.0f;   // camera.rs:115-117
    bool moved = false;
    const Vec3 forward = normalize(V(c->look_from) - V(c->look_at));
    const Vec3 right = normalize(cross(forward, Vec3{0.0f, 1.0f, 0.0f}));
    const Vec3 up{0.0f, 1.0f, 0.0f};
    if (keys & RT_KEY_W) { c->look_from = R(V(c->look_from) + forward * move_speed * dt); moved = true; }
    if (keys & RT_KEY_S) { c->look_from = R(V(c->look_from) - forward * move_speed * dt); moved = true; }
    if (keys & RT_KEY_A) { c->look_from = R(V(c->look_from) - right * move_speed * dt); moved = true; }
    if (keys & RT_KEY_D) { c->look_from = R(V(c->look_from) + right * move_speed * dt); moved = true; }
    if (keys & RT_KEY_ARROW_UP) { c->look_from = R(V(c->look_from) + up * vertical_speed * dt); moved = true; }
    if (keys & RT_KEY_ARROW_DOWN) { c->look_from = R(V(c->look_from) - up * vertical_speed * dt); moved = true; }
    for (int pass = 0; pass < 2; ++pass) {   // camera.rs:166-186 and its verbatim repeat at :187-207
        if (keys & RT_KEY_ARROW_LEFT) { yaw_about_target(c, rotate_speed * dt); moved = true; }
        if (keys & RT_KEY_ARROW_RIGHT) { yaw_about_target(c, -rotate_speed * dt); moved = true; }
    }
    if (keys & RT_KEY_DIGIT1) moved |= pitch_about_target(c, rotate_speed * dt);
    if (keys & RT_KEY_DIGIT2) moved |= pitch_about_target(c, -rotate_speed * dt);
    c->camera_has_moved = moved ? 1u : 0u;   // camera.rs:245-252
    return RT_OK;
}

// A scripted fly-through: frame f applies rt_camera_step(keys[f], dt) and derives the 176-byte uniform the render
// world would extract that frame (lib.rs:117; random_seed = seeds[f], or f / n_frames when seeds is null).
int rt_camera_script(CameraSettings* c, const uint32_t* keys, uint32_t n_frames, float dt, const float* seeds,
                     uint32_t width, uint32_t height_px, SceneCamera* out) {
    if (!c || (n_frames && (!keys || !out))) return rti::fail(RT_ERR_INVALID_ARG, "null argument");
    for (uint32_t f = 0; f < n_frames; ++f) {
        int rc = rt_camera_step(c, keys[f], dt);
        if (rc != RT_OK) return rc;
        rc = rt_camera_from_settings(c, width, height_px, seeds ? seeds[f] : (float)f / (float)n_frames, out + f);
        if (rc != RT_OK) return rc;
    }
    return RT_OK;
}

}  // extern "C"
