/*
 * rt_abi.h — C ABI of the B200-native per-pixel ray-trace pass.
 *
 * This header is the drop-in boundary for the ONE hot path of Sur091/GPU-Ray-Tracing:
 * the `init` / `update` compute entry points of assets/compute_shader.wgsl, together with
 * the two byte layouts that feed them.  Everything is plain C: fixed-size structs, raw
 * pointers and sizes, `int` status codes.  No C++ or torch types cross this boundary.
 *
 * Reference interfaces replaced (paths relative to the reference repo):
 *   SceneCamera  176-byte uniform ........ src/camera.rs:256-291 == compute_shader.wgsl:7-40
 *   GpuSphere    32-byte storage stride .. src/scene/sphere.rs:14-26 == compute_shader.wgsl:80-82,151-155
 *   accumulation texel rgba32float ....... src/lib.rs:63-77 == compute_shader.wgsl:4-5,339-341,362-363
 *   `init` dispatch ...................... src/lib.rs:398-407 -> compute_shader.wgsl:65-70
 *   `update` dispatch .................... src/lib.rs:408-417 -> compute_shader.wgsl:333-364
 *   per-frame buffer upload .............. src/lib.rs:151-207 (camera, spheres, count)
 */
#ifndef RT_ABI_H_
#define RT_ABI_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define RT_API __declspec(dllexport)
#else
#define RT_API __attribute__((visibility("default")))
#endif

/* ---------------------------------------------------------------------------------------
 * Binary layouts (bit-identical to the reference's #[repr(C)] Pod structs)
 * ------------------------------------------------------------------------------------- */

typedef struct RtVec3 {
    float x, y, z;
} RtVec3;

/* src/camera.rs:256-291.  11 rows of (vec3, f32).  Fields the shader reads are marked *. */
typedef struct SceneCamera {
    RtVec3 center;              float viewport_height;   /* off   0* / 12  */
    RtVec3 viewport_upper_left; float viewport_width;    /* off  16* / 28  */
    RtVec3 pixel_delta_u;       float defocus_angle;     /* off  32* / 44* */
    RtVec3 pixel_delta_v;       float aspect_ratio;      /* off  48* / 60  */
    RtVec3 defocus_disk_u;      float _padding0;         /* off  64* / 76  */
    RtVec3 viewport_u;          float _padding1;         /* off  80  / 92  */
    RtVec3 defocus_disk_v;      float max_depth;         /* off  96* / 108* */
    RtVec3 look_from;           float samples_per_pixel; /* off 112  / 124* */
    RtVec3 look_at;             float camera_has_moved;  /* off 128  / 140* */
    RtVec3 vup;                 float random_seed;       /* off 144  / 156* */
    RtVec3 viewport_v;          float defocus_radius;    /* off 160  / 172  */
} SceneCamera;

/* src/scene/sphere.rs:14-26.  color.w is the material selector (compute_shader.wgsl:272,276):
 *   w < -1  lambertian (albedo = color.xyz)
 *  -1..1    metal      (albedo = color.xyz, fuzz = w)
 *   w >  1  dielectric (index of refraction = color.x) */
typedef struct GpuSphere {
    RtVec3 position;
    float  radius;
    float  color[4];
} GpuSphere;

/* src/camera.rs:9-28 (CameraSettings) — the host-side camera model SceneCamera derives from. */
typedef struct CameraSettings {
    float    field_of_view;      /* vertical, degrees */
    uint32_t samples_per_pixel;
    uint32_t camera_has_moved;   /* bool in the reference */
    uint32_t max_depth;
    RtVec3   vup;
    RtVec3   look_from;
    RtVec3   look_at;
    float    defocus_angle;      /* degrees */
    float    focus_distance;
} CameraSettings;

#if defined(__cplusplus)
static_assert(sizeof(RtVec3) == 12, "Vec3 is 3 x f32, align 4 (glam 0.29)");
static_assert(sizeof(SceneCamera) == 176, "SceneCamera uniform is 176 bytes");
static_assert(offsetof(SceneCamera, viewport_upper_left) == 16, "layout");
static_assert(offsetof(SceneCamera, defocus_angle) == 44, "layout");
static_assert(offsetof(SceneCamera, pixel_delta_v) == 48, "layout");
static_assert(offsetof(SceneCamera, defocus_disk_u) == 64, "layout");
static_assert(offsetof(SceneCamera, defocus_disk_v) == 96, "layout");
static_assert(offsetof(SceneCamera, max_depth) == 108, "layout");
static_assert(offsetof(SceneCamera, samples_per_pixel) == 124, "layout");
static_assert(offsetof(SceneCamera, camera_has_moved) == 140, "layout");
static_assert(offsetof(SceneCamera, random_seed) == 156, "layout");
static_assert(offsetof(SceneCamera, defocus_radius) == 172, "layout");
static_assert(sizeof(GpuSphere) == 32, "GpuSphere stride is 32 bytes");
static_assert(offsetof(GpuSphere, radius) == 12, "layout");
static_assert(offsetof(GpuSphere, color) == 16, "layout");
#endif

/* Status codes.  Nothing aborts or throws across this boundary; the reference's panic!/unwrap
 * sites (src/lib.rs:216-217,356-358,399-401) map to a negative status + rt_last_error(). */
enum {
    RT_OK                 = 0,
    RT_ERR_INVALID_ARG    = -1,
    RT_ERR_CUDA           = -2,
    RT_ERR_NO_DEVICE      = -3,
    RT_ERR_NO_SCENE       = -4,
    RT_ERR_OUT_OF_MEMORY  = -5,
    RT_ERR_STATE          = -6
};

typedef struct RtRenderer RtRenderer;          /* opaque: one device, one image (or row window) */
typedef struct RtMultiRenderer RtMultiRenderer; /* opaque: one image sharded over several devices */

/* ---------------------------------------------------------------------------------------
 * Library / device
 * ------------------------------------------------------------------------------------- */
RT_API const char* rt_last_error(void);              /* thread-local, never NULL */
RT_API const char* rt_version(void);
RT_API int rt_device_count(int* count);              /* RT_ERR_NO_DEVICE when CUDA is unusable */

/* ---------------------------------------------------------------------------------------
 * Renderer handle == the render-world state the Bevy node owns (src/lib.rs:138-229):
 * the accumulation image, the sphere storage buffer + count and the camera uniform.
 * The reference ping-pongs two textures (src/lib.rs:218-228) only because a WebGPU storage
 * texture cannot be read and written in one pass; each pixel touches its own texel only
 * (compute_shader.wgsl:339,363), so here the accumulation is one in-place float4 image.
 *
 * A renderer may own a ROW WINDOW [row_begin, row_begin+row_count) of a width x height image:
 * that is the multi-GPU shard.  Pixel seeds use absolute (x, y) (compute_shader.wgsl:309-310)
 * so a window renders exactly the texels the full image would hold.
 * ------------------------------------------------------------------------------------- */
RT_API int rt_create(int device, uint32_t width, uint32_t height, RtRenderer** out);
RT_API int rt_create_window(int device, uint32_t width, uint32_t height,
                            uint32_t row_begin, uint32_t row_count, RtRenderer** out);
/* An INTERLEAVED shard: the renderer owns bands index, index + n_shards, index + 2 n_shards, ... of band_rows rows each
 * (band_rows: a positive multiple of 8).  Contiguous row blocks balance only while every row costs the same; the culled
 * kernels spend more on rows that hold geometry than on sky, and bounce rays (max_depth > 1) cluster on the lower rows,
 * so a sharded image is balanced by dealing narrow bands round-robin.  Window row r is image row
 * (r / band_rows * n_shards + index) * band_rows + r % band_rows; rt_read_accum / rt_read_rgba8 return the window's rows
 * in window order (rt_shard_bands gives their number).  Everything else behaves as for rt_create_window. */
RT_API int rt_create_bands(int device, uint32_t width, uint32_t height, uint32_t band_rows, uint32_t n_shards,
                           uint32_t index, RtRenderer** out);
RT_API int rt_shard_bands(uint32_t height, uint32_t band_rows, uint32_t n_shards, uint32_t index, uint32_t* row_count);
RT_API int rt_bands(RtRenderer* r, uint32_t* band_rows, uint32_t* n_shards, uint32_t* index);   /* band_rows 0: contiguous */
RT_API int rt_destroy(RtRenderer* r);

/* prepare_sphere_buffer (src/lib.rs:177-207): `spheres` points at `buffer_len` structs of which
 * the first `count` are traced (the reference pads the Vec to MAX_SPHERES, sphere.rs:142-150).
 * Borrowed for the call only. */
RT_API int rt_set_scene(RtRenderer* r, const GpuSphere* spheres, uint32_t buffer_len, uint32_t count);

/* WGSL `init` (compute_shader.wgsl:65-70): zero every accumulation texel. */
RT_API int rt_init(RtRenderer* r);

/* WGSL `update` (compute_shader.wgsl:333-364): one dispatch over every pixel of the window with
 * this 176-byte uniform (prepare_camera_bind_group, src/lib.rs:151-175).  Asynchronous on the
 * renderer's stream; any read-back call synchronises. */
RT_API int rt_update(RtRenderer* r, const SceneCamera* camera);

/* `n` consecutive `update` dispatches, cameras[i] used for dispatch i — the per-frame loop of
 * ComputeShaderNode (src/lib.rs:369-374) with the per-frame random_seed re-roll of
 * src/camera.rs:346 supplied by the caller.  Identical to calling rt_update n times. */
RT_API int rt_update_n(RtRenderer* r, const SceneCamera* cameras, uint32_t n);

RT_API int rt_synchronize(RtRenderer* r);

/* Read / restore the accumulation window: row_count*width texels of (mean_r, mean_g, mean_b, n). */
RT_API int rt_read_accum(RtRenderer* r, float* host_rgba32f);
RT_API int rt_write_accum(RtRenderer* r, const float* host_rgba32f);
/* Linear RGBA8 of the running mean: q(x) = floor(clamp(x,0,1)*255 + 0.5), alpha = 255. */
RT_API int rt_read_rgba8(RtRenderer* r, uint8_t* host_rgba8);
/* The same without blocking: the frame as it stands after the work queued so far is quantised into one of two device
 * buffers and copied to `host_rgba8` on a separate copy stream, so the next rt_update overlaps the transfer (the role
 * the render graph's frame pipelining plays for the reference's sprite, src/lib.rs:79-102).  `host_rgba8` must stay
 * valid (and should be pinned) until rt_read_wait returns; at most two reads are in flight, a third waits for the
 * first on the device. */
RT_API int rt_read_rgba8_async(RtRenderer* r, uint8_t* host_rgba8);
RT_API int rt_read_wait(RtRenderer* r);   /* every asynchronous read issued so far has landed */
/* enable != 0: updates whose shading is finished inside the primary pass (max_depth == 1) also store the frame's RGBA8
 * image, into the buffer the next rt_read_rgba8_async sends, so that read needs no quantise pass of its own (one kernel
 * less per frame, 166 MB less HBM traffic at 4K).  Other updates and reads behave as before.  Off by default. */
RT_API int rt_set_rgba8_output(RtRenderer* r, int enable);
/* Fused gather.  device_rgba8 != NULL names texel (0, 0) of a FULL width x height RGBA8 image in device memory of this
 * GPU or of a peer (mapped with rt_ipc_open, or any pointer of a peer with access enabled): updates whose shading is
 * finished inside the primary pass (max_depth == 1) store every texel of the renderer's window straight to its place
 * there, device_rgba8[y * width + x] with y the image row.  The rows of a sharded image so land in the owner's frame
 * WHILE the kernel computes, 128 contiguous bytes per warp row over NVLink, and no copy follows the kernel ("the final
 * image is gathered by peer copy over NVLink", fused into the producing pass).  NULL: off.
 * rt_read_rgba8* keep working (they quantise into the renderer's own buffers). */
RT_API int rt_set_rgba8_target(RtRenderer* r, void* device_rgba8);
/* Completion protocol of such a gather, enqueued on the renderer's stream (no host synchronisation):
 * rt_stream_signal: after everything queued so far — including the fused stores above — *device_flag = value, released at
 * system scope (device_flag may be peer memory).  rt_stream_wait: the stream does not proceed until every one of
 * device_flags[0..n) is >= value (wrap-safe), or `timeout_ms` of device time has passed, in which case the wait gives up
 * and rt_stream_wait_timed_out reports it.  Together: rank i signals arrived[i] = frame + 1 behind its update, the owner
 * waits for all of them and then knows the frame is whole. */
RT_API int rt_stream_signal(RtRenderer* r, void* device_flag, uint32_t value);
RT_API int rt_stream_wait(RtRenderer* r, const void* device_flags, uint32_t n, uint32_t value, uint32_t timeout_ms);
RT_API int rt_stream_wait_timed_out(RtRenderer* r, uint32_t* count);   /* synchronises; number of waits that gave up */
/* The PIPELINED gather: rt_send_rgba8_async puts the window's RGBA8 rows (as rt_read_rgba8_async would read them: the frame
 * as it stands after the work queued so far, from one of two device buffers) into their places in a FULL width x height
 * image at device_full_rgba8 — this GPU's memory or a peer's — with a copy engine on the renderer's copy stream: one copy
 * for a contiguous window, one strided 2-D copy for an interleaved one.  The next update overlaps the transfer, so the
 * link carries frame f while the SMs render f + 1 (one frame of latency; the fused target above has none but its stores
 * share the link with the kernel that issues them).  rt_copy_signal / rt_copy_wait are rt_stream_signal / rt_stream_wait
 * on that copy stream: a flag behind the transfer, and back-pressure before it.  rt_read_wait drains the copy stream. */
RT_API int rt_send_rgba8_async(RtRenderer* r, void* device_full_rgba8);
RT_API int rt_copy_signal(RtRenderer* r, void* device_flag, uint32_t value);
RT_API int rt_copy_wait(RtRenderer* r, const void* device_flags, uint32_t n, uint32_t value, uint32_t timeout_ms);
/* Device buffers that other processes can write into.  A multi-process run (one process per GPU) keeps its gather
 * target in one process (rt_device_alloc + rt_ipc_export), the others map it (rt_ipc_open, peer access enabled
 * lazily) and hand the mapped address to rt_read_rgba8_async, whose destination may be pinned host memory or device
 * memory of this or a peer GPU: the finished frame then travels by one peer copy over NVLink, the analogue of the
 * "final image is gathered by peer copy" step for frames sharded over processes. */
RT_API int rt_device_alloc(int device, uint64_t bytes, void** ptr);
RT_API int rt_device_free(int device, void* ptr);
RT_API int rt_ipc_export(int device, void* ptr, uint8_t handle[64]);
RT_API int rt_ipc_open(int device, const uint8_t handle[64], void** ptr);
RT_API int rt_ipc_close(int device, void* ptr);

/* Number of sphere_list_hit traversals (compute_shader.wgsl:266) executed since the last
 * rt_reset_counters: the "rays" of the Mrays/s metric. */
RT_API int rt_rays_traced(RtRenderer* r, uint64_t* rays);
/* Number of (ray, sphere) pairs that went through the exact quadratic after the broad phase. */
RT_API int rt_exact_tests(RtRenderer* r, uint64_t* tests);
RT_API int rt_reset_counters(RtRenderer* r);
/* Rays the wavefront traced per bounce in the last deep (max_depth > 1) update: counts[b] for b >= 1 (counts[0] is
 * unused: bounce 0 is the primary pass).  *n_out = entries written (0 before any deep frame). */
RT_API int rt_wavefront_counts(RtRenderer* r, uint32_t* counts, uint32_t capacity, uint32_t* n_out);
/* The scene's tree of bounding spheres as of the last update (zeros before the first one that built it): node count,
 * levels, and how bounce rays (max_depth > 1) meet it — 0 = no tree, 1 = image copied into shared memory per CTA,
 * 2 = walked where it lies in global memory (scenes past the shared-memory budget), 3 = not walked: plain records
 * streamed (node numbers beyond 16 bits, or kernel variant 5). */
RT_API int rt_tree_info(RtRenderer* r, uint32_t* n_nodes, uint32_t* depth, uint32_t* bounce_path);
/* Host time (ms) of the last tree build + pack + upload; the build runs inside the first update after rt_set_scene
 * brought new sphere bytes (the reference scene is static, src/scene/sphere.rs:167-170, so this is paid once). */
RT_API int rt_tree_build_ms(RtRenderer* r, double* ms);

/* Device-side timing of the update kernels on the renderer's own stream (CUDA events). */
RT_API int rt_timer_begin(RtRenderer* r);
RT_API int rt_timer_end(RtRenderer* r, float* elapsed_ms);   /* synchronises */
RT_API int rt_kernel_launches(RtRenderer* r, uint64_t* launches);

/* Interop for a host that already lives on the GPU (torch, another CUDA library). */
RT_API int rt_accum_device_ptr(RtRenderer* r, void** ptr);   /* float4[row_count*width] */
RT_API int rt_rgba8_device_ptr(RtRenderer* r, void** ptr);   /* uchar4[row_count*width], filled by rt_quantize */
RT_API int rt_quantize(RtRenderer* r);                       /* accum -> rgba8, on the stream */
/* Same, written straight into a caller-owned DEVICE buffer of row_count*width uchar4 (e.g. the send
 * buffer of an NCCL gather).  Asynchronous on the renderer's stream. */
RT_API int rt_quantize_into(RtRenderer* r, void* device_rgba8);
/* The renderer's stream: every kernel of rt_update runs on it or is ordered into it by events (a depth-1 frame's
 * per-frame record kernel runs on an internal stream beside the frame before it and the frame's main kernel waits for
 * it), so work a caller enqueues on this stream after rt_update sees the finished frame. */
RT_API int rt_stream(RtRenderer* r, void** cuda_stream);
RT_API int rt_window(RtRenderer* r, uint32_t* width, uint32_t* height,
                     uint32_t* row_begin, uint32_t* row_count);

/* Kernel selection for experiments/tests: 0 = auto, 1 = baseline (one thread per pixel, exact
 * quadratic per sphere), 2 = tiled broad phase + exact narrow phase with the scene resident in shared
 * memory, 3 = the same streaming the scene through shared memory, 4 = common-origin primary pass
 * (primary rays of pinhole and defocus cameras, records streamed through shared memory) followed by 2 or 3
 * for the bounces, 5 = as 4 with the bounce rays forced through the streamed-record kernel (the path of scenes whose
 * tree the bounce kernel cannot walk).  0 picks 4 when it applies, else 2 or 3. */
RT_API int rt_set_kernel_variant(RtRenderer* r, int variant);

/* Arithmetic of the per-ray O(1) work (camera ray, scatter, sky colour, running mean).
 *   RT_PRECISION_EXACT (default): one IEEE-754 rounding per WGSL operator, pinned sin / cos — bit-identical to the CPU
 *                                 oracle of this repository.
 *   RT_PRECISION_FAST_SHADING:    scatter, sky colour and running mean through the hardware approximations (MUFU rcp /
 *                                 rsqrt / sqrt / sin / cos, FMA contraction) — what a WGSL implementation may do too, its
 *                                 builtin precision being implementation defined.  Primary rays are generated exactly
 *                                 and which sphere a ray hits is decided by the reference's IEEE quadratic, so a depth-1
 *                                 image differs from the exact one by colour rounding only (<= 1/255 after RGBA8
 *                                 quantisation; bench.py and the tests count the pixels over it).
 *   RT_PRECISION_FAST_ALL:        get_ray as well.  A primary ray is then displaced by ~1e-7 relative, and a sample that
 *                                 grazes a silhouette can land on the other side of it (a few pixels per million).
 * Applies to the primary pass with fused shading (max_depth == 1) and to the wavefront (1 < max_depth <= 64), whose
 * bounce rays know one fast level; other kernel variants ignore it. */
enum { RT_PRECISION_EXACT = 0, RT_PRECISION_FAST_SHADING = 1, RT_PRECISION_FAST_ALL = 2 };
RT_API int rt_set_precision(RtRenderer* r, int precision);

/* ---------------------------------------------------------------------------------------
 * `update` for callers that cannot allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc.
 *
 * One `update` dispatch (compute_shader.wgsl:333-364) over a whole width x height image, in place, without a renderer
 * handle: the caller passes DEVICE pointers for the input (n_spheres GpuSphere structs as the reference lays them out)
 * and for the input / output accumulation image (width*height float4 texels (mean.rgb, n); zero it for `init`), the
 * 176-byte camera uniform by host pointer (read during the call) and the CUDA stream to enqueue on (NULL = the default
 * stream).  Asynchronous on that stream.  The call performs no cudaMalloc, cudaFree or other allocating routine: the
 * kernel's scratch (per-sphere broad-phase records, tile ticket, counters) lives in a file-scope device array of 66,048
 * bytes that every GPU which loads the library holds until the process exits.
 * LIMIT: n_spheres <= RT_STATIC_MAX_SPHERES (4096); more is rejected with RT_ERR_INVALID_ARG and a message.
 * SHARING: every call on a device uses the same array, which keeps the previous call's contents (what must start at zero is
 * cleared by each call).  Calls that OVERLAP on one GPU — different streams, or different threads — race on it: serialise
 * them (one stream, or events).  Calls on different GPUs are independent.
 * The kernel is the resident-scene variant (rt_set_kernel_variant 2: every sphere's record in shared memory, any max_depth
 * in the kernel's own bounce loop); results are bit-identical to rt_update's. */
#define RT_STATIC_MAX_SPHERES 4096u
RT_API int rt_update_static(int device, const GpuSphere* device_spheres, uint32_t n_spheres, const SceneCamera* camera,
                            uint32_t width, uint32_t height, float* device_accum_rgba32f, void* cuda_stream);

/* ---------------------------------------------------------------------------------------
 * Headless one-shot entries (new surface asked for by the port target; the reference has
 * only the windowed app, src/lib.rs:28-58)
 * ------------------------------------------------------------------------------------- */
/* render(scene, camera, w, h): zeroed accumulators, `samples` update dispatches with this
 * camera (random_seed as given), result as rgba32f (w*h*4 floats). */
RT_API int rt_render(int device, const GpuSphere* spheres, uint32_t count,
                     const SceneCamera* camera, uint32_t width, uint32_t height,
                     uint32_t samples, float* out_rgba32f);
/* One frame per camera (fly-through); out_rgba8 holds n_frames*w*h*4 bytes. */
RT_API int rt_render_frames(int device, const GpuSphere* spheres, uint32_t count,
                            const SceneCamera* cameras, uint32_t n_frames,
                            uint32_t width, uint32_t height, uint8_t* out_rgba8);

/* ---------------------------------------------------------------------------------------
 * One image sharded by row tiles over several devices of one host, scene replicated,
 * result gathered to devices[0] by peer copy over NVLink.
 * ------------------------------------------------------------------------------------- */
/* The row block shard `index` of `n_shards` owns: contiguous, multiples of the 4-row warp tile except the
 * last.  Pure host arithmetic (usable without a device; also what a multi-process host uses per rank). */
RT_API int rt_shard_rows(uint32_t height, uint32_t n_shards, uint32_t index, uint32_t* row_begin, uint32_t* row_count);
RT_API int rt_multi_create(const int* devices, uint32_t n_devices, uint32_t width, uint32_t height,
                           RtMultiRenderer** out);
/* The same with interleaved bands of band_rows rows (rt_create_bands) instead of contiguous blocks: balanced whatever
 * the image holds where. */
RT_API int rt_multi_create_bands(const int* devices, uint32_t n_devices, uint32_t width, uint32_t height,
                                 uint32_t band_rows, RtMultiRenderer** out);
RT_API int rt_multi_destroy(RtMultiRenderer* m);
RT_API int rt_multi_set_scene(RtMultiRenderer* m, const GpuSphere* spheres, uint32_t buffer_len, uint32_t count);
RT_API int rt_multi_init(RtMultiRenderer* m);
RT_API int rt_multi_update(RtMultiRenderer* m, const SceneCamera* camera);
RT_API int rt_multi_gather_rgba8(RtMultiRenderer* m, uint8_t* host_rgba8);   /* may be NULL: gather on device only */
RT_API int rt_multi_gather_accum(RtMultiRenderer* m, float* host_rgba32f);
RT_API int rt_multi_rays_traced(RtMultiRenderer* m, uint64_t* rays);
RT_API int rt_multi_shard(RtMultiRenderer* m, uint32_t index, RtRenderer** shard);

/* ---------------------------------------------------------------------------------------
 * Host helpers: the CPU code on either side of the path
 * ------------------------------------------------------------------------------------- */
/* CameraSettings::default() (src/camera.rs:30-46). */
RT_API void rt_camera_settings_default(CameraSettings* out);
/* SceneCamera::from(&CameraSettings) (src/camera.rs:293-351) generalised from crate::SIZE to
 * (width, height); random_seed is supplied (the reference rolls rand::random::<f32>() there). */
RT_API int rt_camera_from_settings(const CameraSettings* settings, uint32_t width, uint32_t height,
                                   float random_seed, SceneCamera* out);
/* create_default_spheres (src/scene/sphere.rs:45-153) with a seeded generator instead of the
 * thread RNG.  Writes at most `capacity` spheres, returns the number in *count and the padded
 * buffer length (>= MAX_SPHERES = 100) in *buffer_len. */
RT_API int rt_create_default_spheres(uint64_t seed, GpuSphere* out, uint32_t capacity,
                                     uint32_t* count, uint32_t* buffer_len);
/* Benchmark scenes: ground sphere + (n-1) random spheres. */
RT_API int rt_create_random_spheres(uint64_t seed, uint32_t n, GpuSphere* out);
/* Ground sphere + (n-1) small spheres on a square grid, the reference recipe scaled up. */
RT_API int rt_create_grid_spheres(uint64_t seed, uint32_t n, GpuSphere* out);
/* Orbit fly-through: pose k of n around look_at (camera.rs:125-253 yaw controls, scripted). */
RT_API int rt_orbit_camera(uint32_t k, uint32_t n, float radius, float height, float fov_deg,
                           uint32_t width, uint32_t height_px, SceneCamera* out);

/* Keys of camera_movement_system (src/camera.rs:125-253), as a bit mask of the keys held during a frame. */
#define RT_KEY_W           0x001u  /* look_from += forward * 2 dt (forward = from look_at TO look_from) */
#define RT_KEY_S           0x002u
#define RT_KEY_A           0x004u  /* strafe along forward x Y */
#define RT_KEY_D           0x008u
#define RT_KEY_ARROW_UP    0x010u  /* look_from.y +/- dt */
#define RT_KEY_ARROW_DOWN  0x020u
#define RT_KEY_ARROW_LEFT  0x040u  /* yaw look_from about look_at; applied twice per frame, as in the reference */
#define RT_KEY_ARROW_RIGHT 0x080u
#define RT_KEY_DIGIT1      0x100u  /* pitch about look_at, refused when |dir.y| >= 0.95 */
#define RT_KEY_DIGIT2      0x200u
/* One frame of the keyboard controller: updates look_from and camera_has_moved in place. */
RT_API int rt_camera_step(CameraSettings* settings, uint32_t keys, float dt);
/* n_frames controller frames, each followed by the uniform derivation: out[f] is what `update` would see at frame f.
 * seeds may be null (random_seed = f / n_frames). */
RT_API int rt_camera_script(CameraSettings* settings, const uint32_t* keys, uint32_t n_frames, float dt,
                            const float* seeds, uint32_t width, uint32_t height_px, SceneCamera* out);

#ifdef __cplusplus
}
#endif
#endif /* RT_ABI_H_ */
