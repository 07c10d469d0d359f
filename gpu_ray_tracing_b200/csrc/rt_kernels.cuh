// rt_kernels.cuh — launch interface between the C-ABI host code (rt_api.cu) and the kernels.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/rt_abi.h"
#include "rt_tree_layout.h"


namespace rtk {

constexpr int kPix = 4;             // pixels per thread
constexpr int kWarpTileW = 32;      // a warp owns a 32 x 4 pixel tile: lane -> (lane & 7) + 8 p, lane >> 3
constexpr int kWarpTileH = 4;
constexpr int kBlockThreads = 256;  // 8 warps
constexpr int kChunkSpheres = 2048; // streamed kernel: spheres per shared-memory stage (32 KB)
constexpr uint32_t kResidentMaxSpheres = 4096;  // 64 KB of shared memory per CTA at most
// common-origin primary pass: a warp owns a 32 x 8 pixel tile, a thread a 1 x 8 strip
constexpr int kPrimRays = 8;
constexpr int kPrimTileW = 32;
constexpr int kPrimTileH = 8;
#ifndef RT_PRIM_THREADS
#define RT_PRIM_THREADS 352
#endif
#ifndef RT_PRIM_CTAS
#define RT_PRIM_CTAS 2
#endif
#ifndef RT_PRIM_LIST
#define RT_PRIM_LIST 48
#endif
#ifndef RT_PRIM_SORT_MIN
#define RT_PRIM_SORT_MIN 2
#endif
#ifndef RT_PRIM_CULL
#define RT_PRIM_CULL 1
#endif
#ifndef RT_PRIM_SORT_MAX
#define RT_PRIM_SORT_MAX 16
#endif
// Order in which the dynamic tile tickets walk the window: 0 = top-down (row-major), 1 = bottom-up, 2 = middle-out by
// tile rows (the window's edges last: sky above and the near ground below are the cheapest tiles of most frames).  A warp's last
// tile decides the kernel's tail; tiles that hold geometry cost several times a sky tile, and with the horizon in the
// upper part of the frame (the reference's default camera looks slightly down, camera.rs:38-40) top-down hands out the
// expensive tiles LAST.
#ifndef RT_PRIM_TILE_ORDER
#define RT_PRIM_TILE_ORDER 0
#endif
// 1: the primary pass hands out its tile rows dearest first, by the last frame's measured cost (UpdateParams::row_order);
// 0: in RT_PRIM_TILE_ORDER's fixed order (the A/B arm)
#ifndef RT_PRIM_ROW_ORDER
#define RT_PRIM_ROW_ORDER 1
#endif
// 1: a depth-1 frame's prep kernel runs on a stream of its own, beside the frame before it (PrepAsync); 0: in line (A/B arm)
#ifndef RT_PREP_ASYNC
#define RT_PREP_ASYNC 1
#endif
// 1: a warp's first tile of a primary pass is its number in the grid, tickets are drawn from the second tile on; 0: every
// tile is drawn (the A/B arm)
#ifndef RT_PRIM_STATIC_FIRST
#define RT_PRIM_STATIC_FIRST 1
#endif
// Bounces a warp of the tail launch keeps its 64 paths of a short queue before the survivors are dealt out again, evenly
// (an ODD number: the hand-back must not go to the queue the pass loads from); 0: every path is finished where it was
// first loaded (the A/B arm)
#ifndef RT_WF_REDEAL_AFTER
#define RT_WF_REDEAL_AFTER 1u
#endif
// ... and the bounces the evenly dealt survivors are kept before they are dealt out once more (odd; 0: to their end)
#ifndef RT_WF_REDEAL_AGAIN
#define RT_WF_REDEAL_AGAIN 0u
#endif
#ifndef RT_PRIM_PROBE
#define RT_PRIM_PROBE 0   // developer probe: rt_exact_tests reports the primary pass's busy cycles summed over its warps
#endif
#ifndef RT_PRIM_WORK
#define RT_PRIM_WORK 224
#endif
// CTA shape of the primary pass: no block barrier anywhere, so the CTA is only a container of warps; 2 x 11 warps
// per SM is what 80 registers (ptxas' cap for 2 x 352 threads: the file is allotted per 4 warps) and 10.2 KB of shared
// memory per warp (8 KB ray slots + lists) allow
constexpr int kPrimThreads = RT_PRIM_THREADS;
constexpr int kPrimCtas = RT_PRIM_CTAS;
constexpr int kPrimWork = RT_PRIM_WORK;   // per-warp stack of tree nodes awaiting their tile test (entry list <= 160 + growth)
constexpr int kPrimList = RT_PRIM_LIST;   // per-warp list of spheres that passed the tile test, scanned when it fills
constexpr uint32_t kWfMaxDepth = 64; // deeper paths fall back to the in-kernel bounce loop of variants 2 / 3
constexpr uint32_t kWfWords = 3 * (kWfMaxDepth + 1);   // length of wf_counts
constexpr int kTreeWide = rtt::kWide;                  // children per node of the scene's tree (rt_tree_layout.h)
constexpr int kTreeStride = rtt::kStride;              // record slots per node in the image
constexpr int kTreeBacklog = 16;                       // entry nodes a lane may stack up before the warp traverses
constexpr uint32_t kTreeStackMax = 96;                 // deepest per-thread node stack the bounce kernel is configured for
constexpr uint32_t kWfFinishFrom = 3;                   // first bounce whose launch may finish a short queue to the end
constexpr int kWfChunk = 2048;      // wavefront bounce kernel: general records per stage (32 KB)
constexpr size_t kClusterImageMax = (size_t)2 * (kWfChunk + 8) * 16;   // the bounce kernel's two record stages
constexpr int kWfCand = 32;         // per-thread queue of (member, ray) candidates awaiting their exact tests (clustered traversal)
constexpr int kWfDefer = 24;        // per-thread queue of broad-phase groups awaiting their exact tests

struct UpdateParams {
    SceneCamera cam;            // 176 B, lands in the kernel parameter constant bank
    float4* accum;              // window base: row_count * width texels (mean.rgb, n)
    const float4* geom;         // n_spheres x (cx, cy, cz, |C|^2 - r^2 - slack)
    const GpuSphere* spheres;   // the caller's buffer as uploaded (32 B stride)
    uint32_t n_spheres;
    uint32_t width, height;     // full image
    uint32_t row_begin, row_count;  // window: rows [row_begin, row_begin + row_count) of the image, or (band_rows != 0) the
                                    // bands row_begin / band_rows + k * band_stride / band_rows, row_count rows in all
    uint32_t band_rows, band_stride;
    uint32_t tiles_x, n_tiles;  // warp tiles over the window
    unsigned int* tile_counter; // dynamic tile schedulers, zeroed before the launch: [0] update, [1] primary pass
    const float4* prim_tree;    // per-frame common-origin records (oc, -q) of the tree's child slots (cl_n_nodes * kTreeStride);
                                // null = no primary pass
    // wavefront bounces (max_depth > 1 after a primary pass): compacted path queues, three float4 planes of
    // `wf_capacity` entries each: (o.xyz, d.x) (d.yz, factor.xy) (factor.z, bits(texel), bits(seed), bits(n))
    float4* wf_queue[2];
    unsigned int* wf_counts;    // 3 x [kWfMaxDepth + 1]: queue length per bounce, batch tickets, rays traced per bounce (kWfWords)
    uint32_t wf_capacity;
    uint32_t wf_bounce;         // bounce index this launch traces
    float2* first_hit;          // window texels x (t, bits(sphere index)): nearest hit of the primary ray found by
                                // the primary pass; index -1 = miss, -2 = not traced (the update kernel scans it)
    // tree image of the scene (rt_tree.h), built after rt_set_scene, in global memory; cl_n_nodes == 0: none.
    // cl_bytes: the part the bounce kernel copies into shared memory, 0 when it does not fit (streamed bounce kernel)
    const void* cl_blob;
    // cl_global: cl_bytes == 0 but the bounce kernel can walk the tree in global memory (node numbers fit 16 bits)
    uint32_t cl_n_nodes, cl_first_leaf, cl_bytes, cl_stack_cap, cl_n_entry, cl_global;
    float cl_entry_r;
    uint32_t* rgba8_out;            // non-null: the fused shading pass (max_depth 1) also stores the texel's RGBA8 here:
                                    // window-relative texel index, in this device's memory or in a peer's (NVLink stores)
    uint32_t rgba8_abs;             // rgba8_out holds the full image (gather target): index by absolute row
    uint32_t precision;             // RT_PRECISION_* (rt_set_precision)
    unsigned long long* counters;   // [0] rays (sphere_list_hit traversals), [1] narrow-phase tests
    // Order in which the primary pass hands out its tile rows: every tile adds its duration to
    // row_cost_cur[its tile row]; the next frame's prep kernel ranks the rows of row_cost_prev, dearest first, into
    // row_order, so the frame ends on its cheapest tiles.  All three null (or tile_rows > kOrderMaxRows): top-down.
    unsigned int* row_cost_cur = nullptr;
    const unsigned int* row_cost_prev = nullptr;
    uint32_t* row_order = nullptr;
};
constexpr uint32_t kOrderMaxRows = 2048;   // tile rows the prep kernel ranks (16,384 image rows); taller windows go top-down

// variant: 1 = baseline (thread per pixel, exact test per sphere from global memory),
//          2 = tiled broad phase with the scene resident in shared memory,
//          3 = tiled broad phase streaming sphere chunks through shared memory.
// Where a depth-1 frame's prep kernel may run ahead of the frame before it: on `stream`, once `buffers_free` (the frame
// that last used this frame's record buffers) has passed; the primary pass waits for `prep_done`.  Null: on the frame's stream.
struct PrepAsync {
    cudaStream_t stream;
    cudaEvent_t buffers_free, also_wait, prep_done;   // also_wait: null, or a second event the prep kernel must follow
};
cudaError_t launch_update(const UpdateParams& p, int variant, int sm_count, cudaStream_t stream, int* launches = nullptr,
                          const PrepAsync* prep_async = nullptr, bool* prep_was_async = nullptr);
bool primary_pass_applies(const UpdateParams& p);
// max_depth == 1 behind a primary pass: primary_hits_kernel<true> finishes the texels (and honours p.rgba8_out)
bool fused_shade_applies(const UpdateParams& p, int variant);
// variant 4 = common-origin primary pass followed by the update kernel (chosen by variant 0 whenever the camera is
// a pinhole, defocus_angle <= 0); p.prim / p.first_hit must be set.
cudaError_t launch_prep_primary_tree(const UpdateParams& p, bool zero_wavefront, cudaStream_t stream);
constexpr int kPrepThreads = 64;   // small blocks of few registers and no shared memory: they fit beside the primary pass's CTAs
cudaError_t launch_prep_geometry(const GpuSphere* spheres, float4* geom, uint32_t n, cudaStream_t stream);
cudaError_t launch_quantize(const float4* accum, uchar4* rgba8, uint64_t n_texels, cudaStream_t stream);
cudaError_t launch_signal(void* flag, uint32_t value, cudaStream_t stream);
cudaError_t launch_wait_flags(const void* flags, uint32_t n, uint32_t value, uint32_t timeout_ms, unsigned int* status,
                              cudaStream_t stream);
// rt_update_static: variant 2 with its scratch in a file-scope device array (rt_kernels.cu); allocates nothing
constexpr uint32_t kStaticMaxSpheres = 4096;   // == RT_STATIC_MAX_SPHERES of include/rt_abi.h
cudaError_t launch_update_static(int device, const GpuSphere* device_spheres, uint32_t n_spheres, const SceneCamera& cam,
                                 uint32_t width, uint32_t height, float4* device_accum, cudaStream_t stream);
cudaError_t configure_kernels();   // one-off cudaFuncSetAttribute calls

}  // namespace rtk
