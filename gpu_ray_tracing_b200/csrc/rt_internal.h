// rt_internal.h — state behind the opaque handles of include/rt_abi.h.
#pragma once

#include <vector>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/rt_abi.h"

struct RtRenderer {
    int device = 0;
    int sm_count = 0;
    uint32_t width = 0, height = 0, row_begin = 0, row_count = 0;
    uint32_t band_rows = 0, band_stride = 0;   // interleaved window (rt_create_bands); 0 = contiguous
    uint64_t n_texels = 0;
    cudaStream_t stream = nullptr;
    float4* accum = nullptr;              // (mean.rgb, n) per texel of the window
    uchar4* rgba8 = nullptr;
    GpuSphere* spheres = nullptr;         // the caller's storage buffer, as uploaded
    uint32_t sphere_capacity = 0;
    float4* geom = nullptr;               // broad-phase records
    uint32_t geom_capacity = 0;
    float4* prim_tree = nullptr;          // per-frame common-origin records of the tree's child slots: two buffers of
    size_t prim_tree_capacity = 0;        // prim_tree_slots records (capacity in bytes per buffer), by frame parity
    size_t prim_tree_slots = 0;
    float2* first_hit = nullptr;          // (t, sphere index) per texel, written by the primary pass
    float4* wf_queue[2] = {nullptr, nullptr};   // wavefront path queues (3 planes of n_texels float4 each), on first deep frame
    unsigned int* wf_counts = nullptr;
    // tree image of the scene for the bounce rays (rt_tree.h): built lazily, on the first deep frame after rt_set_scene,
    // from the host copy of the sphere list kept for that purpose
    std::vector<GpuSphere> host_spheres;
    bool clusters_stale = true;
    unsigned char* cl_blob = nullptr;
    size_t cl_capacity = 0;
    uint32_t cl_n_nodes = 0, cl_first_leaf = 0, cl_bytes = 0, cl_stack_cap = 0, cl_n_entry = 0, cl_global = 0, cl_depth = 0;
    float cl_entry_r = 0.0f;
    uint32_t n_spheres = 0;
    bool has_scene = false;
    // Per-frame records are double-buffered by frame parity so that a depth-1 frame's prep kernel can run on prep_stream
    // while the frame before it still renders: tile_counter = [2][2] tickets, then
    // (order_rows != 0) [3][order_rows] tile-row costs (frame g adds to g % 3, ranks (g - 2) % 3), [2][order_rows] order.
    unsigned int* tile_counter = nullptr;
    uint32_t order_rows = 0;               // tile rows of the primary pass with a cost / order record; 0 = none
    uint32_t frame_index = 0;              // updates issued so far
    bool last_prep_async = false;          // the last update's prep kernel ran on prep_stream
    cudaStream_t prep_stream = nullptr;
    cudaEvent_t ev_frame[2] = {nullptr, nullptr};   // recorded on `stream` behind the last kernel of a frame of that parity
    cudaEvent_t ev_prep[2] = {nullptr, nullptr};    // recorded on prep_stream behind that parity's prep kernel
    unsigned long long* counters = nullptr;
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr;
    // asynchronous read-back (rt_read_rgba8_async): second RGBA8 buffer, copy stream, per-slot events
    cudaStream_t copy_stream = nullptr;
    uchar4* rgba8_alt = nullptr;
    bool rgba8_output = false;   // rt_set_rgba8_output: fused updates also write the next async slot
    bool rgba8_fresh = false;    // ... and the last update did, nothing has touched the accumulators since
    cudaEvent_t ev_quantized[2] = {nullptr, nullptr}, ev_copied[2] = {nullptr, nullptr};
    uint32_t async_reads = 0;
    uint64_t launches = 0;
    int variant = 0;
    int precision = 0;                    // RT_PRECISION_* (rt_set_precision)
    // rt_set_rgba8_target: where fused depth-1 updates store the window's RGBA8 texels instead of the read-back slots —
    // this device's memory or a peer's (the gather target of a sharded image); target_fresh: the last update did
    uint32_t* rgba8_target = nullptr;
    bool target_fresh = false;
    double tree_build_ms = 0.0;           // host time of the last tree build + pack (rt_tree_build_ms)
    unsigned int* wait_status = nullptr;  // device word a timed-out rt_stream_wait sets
};

namespace rti {
int fail(int code, const char* fmt, ...);
int fail_cuda(cudaError_t e, const char* what);
const char* last_error();
}  // namespace rti
