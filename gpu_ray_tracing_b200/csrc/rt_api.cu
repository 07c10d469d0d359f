// rt_api.cu — the extern "C" boundary declared in include/rt_abi.h (single-device part).
//
// Mirrors the render-world side of the reference plugin (src/lib.rs:138-229, 326-422): the
// renderer owns what the Bevy node owned — accumulation image, sphere storage buffer + count,
// camera uniform — and `rt_init` / `rt_update` are the two dispatches of ComputeShaderNode::run.
// There is no CPU fallback: without a usable CUDA device every entry returns RT_ERR_NO_DEVICE.
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <chrono>
#include <cstring>
#include <new>
#include <vector>

#include "../../include/rt_abi.h"
#include "rt_kernels.cuh"
#include "rt_internal.h"
#include "rt_tree.h"

namespace rti {

static thread_local char g_error[512] = "";

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
    return code;
}
int fail_cuda(cudaError_t e, const char* what) {
    int code = (e == cudaErrorMemoryAllocation) ? RT_ERR_OUT_OF_MEMORY
             : (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) ? RT_ERR_NO_DEVICE
             : RT_ERR_CUDA;
    return fail(code, "%s: %s (%s)", what, cudaGetErrorString(e), cudaGetErrorName(e));
}
const char* last_error() { return g_error; }

}  // namespace rti

using rti::fail;
using rti::fail_cuda;

#define RT_CUDA(call)                                   \
    do {                                                \
        cudaError_t e_ = (call);                        \
        if (e_ != cudaSuccess) return fail_cuda(e_, #call); \
    } while (0)

#define RT_CHECK_HANDLE(r)                                           \
    do {                                                             \
        if (!(r)) return fail(RT_ERR_INVALID_ARG, "null renderer");  \
        RT_CUDA(cudaSetDevice((r)->device));                         \
    } while (0)

extern "C" {

const char* rt_last_error(void) { return rti::last_error(); }
const char* rt_version(void) { return "gpu-ray-tracing_b200 0.1 (sm_100a)"; }

int rt_device_count(int* count) {
    if (!count) return fail(RT_ERR_INVALID_ARG, "count is null");
    *count = 0;
    cudaError_t e = cudaGetDeviceCount(count);
    if (e != cudaSuccess) {
        *count = 0;
        return fail_cuda(e, "cudaGetDeviceCount");
    }
    if (*count == 0) return fail(RT_ERR_NO_DEVICE, "no CUDA device");
    return RT_OK;
}

int rt_create_window(int device, uint32_t width, uint32_t height, uint32_t row_begin, uint32_t row_count,
                     RtRenderer** out) {
    if (!out) return fail(RT_ERR_INVALID_ARG, "out is null");
    *out = nullptr;
    if (width == 0 || height == 0 || row_count == 0) return fail(RT_ERR_INVALID_ARG, "empty image %ux%u rows %u", width, height, row_count);
    if ((uint64_t)row_begin + row_count > height) return fail(RT_ERR_INVALID_ARG, "row window [%u,+%u) outside height %u", row_begin, row_count, height);
    if (width > (1u << 20) || height > (1u << 20)) return fail(RT_ERR_INVALID_ARG, "image too large");
    int n = 0;
    int rc = rt_device_count(&n);
    if (rc != RT_OK) return rc;
    if (device < 0 || device >= n) return fail(RT_ERR_INVALID_ARG, "device %d out of range (have %d)", device, n);
    RT_CUDA(cudaSetDevice(device));
    RT_CUDA(rtk::configure_kernels());

    RtRenderer* r = new (std::nothrow) RtRenderer();
    if (!r) return fail(RT_ERR_OUT_OF_MEMORY, "host allocation failed");
    r->device = device;
    r->width = width; r->height = height; r->row_begin = row_begin; r->row_count = row_count;
    r->n_texels = (uint64_t)width * row_count;
    cudaDeviceProp prop;
    cudaError_t e = cudaGetDeviceProperties(&prop, device);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&r->stream, cudaStreamNonBlocking);
    r->sm_count = prop.multiProcessorCount;
    if (e == cudaSuccess) e = cudaMalloc(&r->accum, r->n_texels * sizeof(float4));
    if (e == cudaSuccess) e = cudaMalloc(&r->rgba8, r->n_texels * sizeof(uchar4));
    if (e == cudaSuccess) e = cudaMalloc(&r->first_hit, r->n_texels * sizeof(float2));
    {   // the tile tickets and, behind them, the primary pass's per-tile-row costs (two frames) and row order
        const uint32_t tile_rows = (row_count + rtk::kPrimTileH - 1) / rtk::kPrimTileH;
        r->order_rows = (RT_PRIM_ROW_ORDER && tile_rows <= rtk::kOrderMaxRows) ? tile_rows : 0u;
        const size_t words = 4 + 5 * (size_t)r->order_rows;
        if (e == cudaSuccess) e = cudaMalloc(&r->tile_counter, words * sizeof(unsigned int));
        if (e == cudaSuccess) e = cudaMemsetAsync(r->tile_counter, 0, words * sizeof(unsigned int), r->stream);
    }
    if (e == cudaSuccess) e = cudaMalloc(&r->counters, 2 * sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMalloc(&r->wait_status, sizeof(unsigned int));
    if (e == cudaSuccess) e = cudaMemsetAsync(r->wait_status, 0, sizeof(unsigned int), r->stream);
    if (e == cudaSuccess) e = cudaEventCreate(&r->ev_begin);
    if (e == cudaSuccess) e = cudaEventCreate(&r->ev_end);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&r->copy_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&r->prep_stream, cudaStreamNonBlocking);
    for (int b = 0; b < 2 && e == cudaSuccess; ++b) {
        e = cudaEventCreateWithFlags(&r->ev_frame[b], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&r->ev_prep[b], cudaEventDisableTiming);
    }
    for (int b = 0; b < 2 && e == cudaSuccess; ++b) {
        e = cudaEventCreateWithFlags(&r->ev_quantized[b], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&r->ev_copied[b], cudaEventDisableTiming);
    }
    if (e == cudaSuccess) e = cudaMemsetAsync(r->accum, 0, r->n_texels * sizeof(float4), r->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(r->counters, 0, 2 * sizeof(unsigned long long), r->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(r->stream);
    if (e != cudaSuccess) {
        int code = fail_cuda(e, "rt_create");
        rt_destroy(r);
        return code;
    }
    *out = r;
    return RT_OK;
}

int rt_create(int device, uint32_t width, uint32_t height, RtRenderer** out) {
    return rt_create_window(device, width, height, 0, height, out);
}

int rt_shard_bands(uint32_t height, uint32_t band_rows, uint32_t n_shards, uint32_t index, uint32_t* row_count) {
    if (!row_count) return fail(RT_ERR_INVALID_ARG, "null argument");
    if (band_rows == 0 || band_rows % 8u != 0u) return fail(RT_ERR_INVALID_ARG, "band_rows %u is not a positive multiple of 8", band_rows);
    if (n_shards == 0 || index >= n_shards) return fail(RT_ERR_INVALID_ARG, "shard %u of %u", index, n_shards);
    const uint32_t bands = (height + band_rows - 1) / band_rows;
    if (bands <= index) { *row_count = 0; return RT_OK; }
    const uint32_t mine = (bands - index + n_shards - 1) / n_shards;        // bands index, index + n, ...
    uint32_t rows = mine * band_rows;
    const uint32_t last_band = index + (mine - 1) * n_shards;
    if (last_band == bands - 1) rows -= bands * band_rows - height;         // the image's last band may be short
    *row_count = rows;
    return RT_OK;
}

int rt_create_bands(int device, uint32_t width, uint32_t height, uint32_t band_rows, uint32_t n_shards, uint32_t index,
                    RtRenderer** out) {
    if (!out) return fail(RT_ERR_INVALID_ARG, "out is null");
    *out = nullptr;
    uint32_t rows = 0;
    int rc = rt_shard_bands(height, band_rows, n_shards, index, &rows);
    if (rc != RT_OK) return rc;
    if (rows == 0) return fail(RT_ERR_INVALID_ARG, "shard %u of %u owns no band of a %u-row image", index, n_shards, height);
    // created as a window of `rows` rows at the top, then re-labelled: the kernels map window rows to image rows (abs_row)
    rc = rt_create_window(device, width, height, 0, rows, out);
    if (rc != RT_OK) return rc;
    (*out)->row_begin = index * band_rows;
    (*out)->band_rows = band_rows;
    (*out)->band_stride = n_shards * band_rows;
    return RT_OK;
}

int rt_bands(RtRenderer* r, uint32_t* band_rows, uint32_t* n_shards, uint32_t* index) {
    if (!r) return fail(RT_ERR_INVALID_ARG, "null renderer");
    if (band_rows) *band_rows = r->band_rows;
    if (n_shards) *n_shards = r->band_rows ? r->band_stride / r->band_rows : 1u;
    if (index) *index = r->band_rows ? r->row_begin / r->band_rows : 0u;
    return RT_OK;
}

int rt_destroy(RtRenderer* r) {
    if (!r) return RT_OK;
    cudaSetDevice(r->device);
    if (r->stream) cudaStreamSynchronize(r->stream);
    if (r->copy_stream) cudaStreamSynchronize(r->copy_stream);
    if (r->prep_stream) cudaStreamSynchronize(r->prep_stream);
    cudaFree(r->rgba8_alt);
    for (int b = 0; b < 2; ++b) {
        if (r->ev_frame[b]) cudaEventDestroy(r->ev_frame[b]);
        if (r->ev_prep[b]) cudaEventDestroy(r->ev_prep[b]);
    }
    if (r->prep_stream) cudaStreamDestroy(r->prep_stream);
    for (int b = 0; b < 2; ++b) {
        if (r->ev_quantized[b]) cudaEventDestroy(r->ev_quantized[b]);
        if (r->ev_copied[b]) cudaEventDestroy(r->ev_copied[b]);
    }
    if (r->copy_stream) cudaStreamDestroy(r->copy_stream);
    cudaFree(r->accum);
    cudaFree(r->rgba8);
    cudaFree(r->spheres);
    cudaFree(r->geom);
    cudaFree(r->prim_tree);
    cudaFree(r->first_hit);
    cudaFree(r->wf_queue[0]);
    cudaFree(r->wf_queue[1]);
    cudaFree(r->wf_counts);
    cudaFree(r->cl_blob);
    cudaFree(r->tile_counter);
    cudaFree(r->counters);
    cudaFree(r->wait_status);
    if (r->ev_begin) cudaEventDestroy(r->ev_begin);
    if (r->ev_end) cudaEventDestroy(r->ev_end);
    if (r->stream) cudaStreamDestroy(r->stream);
    delete r;
    return RT_OK;
}

int rt_set_scene(RtRenderer* r, const GpuSphere* spheres, uint32_t buffer_len, uint32_t count) {
    RT_CHECK_HANDLE(r);
    if (count > buffer_len) return fail(RT_ERR_INVALID_ARG, "count %u exceeds buffer_len %u", count, buffer_len);
    if (buffer_len > 0 && !spheres) return fail(RT_ERR_INVALID_ARG, "spheres is null");
    if (count > (1u << 26)) return fail(RT_ERR_INVALID_ARG, "too many spheres");
    // storage buffers cannot be empty in wgpu either; keep one slot so pointers stay valid
    const uint32_t cap_needed = buffer_len > 0 ? buffer_len : 1;
    const uint32_t pad_needed = (count + 7) / 8 * 8 + 8;   // rtk: kGeomPad rounding + kGeomTail
    if (cap_needed > r->sphere_capacity || pad_needed > r->geom_capacity) {
        RT_CUDA(cudaStreamSynchronize(r->stream));
        cudaFree(r->spheres); r->spheres = nullptr; r->sphere_capacity = 0;
        cudaFree(r->geom); r->geom = nullptr; r->geom_capacity = 0;
        RT_CUDA(cudaMalloc(&r->spheres, (size_t)cap_needed * sizeof(GpuSphere)));
        r->sphere_capacity = cap_needed;
        RT_CUDA(cudaMalloc(&r->geom, (size_t)pad_needed * sizeof(float4)));
        r->geom_capacity = pad_needed;
    }
    if (buffer_len > 0)
        RT_CUDA(cudaMemcpyAsync(r->spheres, spheres, (size_t)buffer_len * sizeof(GpuSphere), cudaMemcpyHostToDevice, r->stream));
    r->n_spheres = count;
    // The tree of bounding spheres (rt_tree.h) is built on the host from a copy of the list, on the next update that
    // needs it; an upload of the same bytes (the reference re-creates its buffers every frame, src/lib.rs:177-207)
    // keeps the tree it has.
    const bool same = r->host_spheres.size() == count &&
                      (count == 0 || memcmp(r->host_spheres.data(), spheres, (size_t)count * sizeof(GpuSphere)) == 0);
    if (!same) {
        r->host_spheres.assign(spheres, spheres + count);
        r->clusters_stale = true;
        r->cl_n_nodes = 0;
        r->cl_bytes = 0;
        r->cl_global = 0;
    }
    RT_CUDA(rtk::launch_prep_geometry(r->spheres, r->geom, count, r->stream));
    r->launches += 1;
    // the source is borrowed for the call only (pageable memory: the copy above has been staged,
    // but make the contract unconditional)
    RT_CUDA(cudaStreamSynchronize(r->stream));
    r->has_scene = true;
    return RT_OK;
}

int rt_init(RtRenderer* r) {
    RT_CHECK_HANDLE(r);
    RT_CUDA(cudaMemsetAsync(r->accum, 0, r->n_texels * sizeof(float4), r->stream));
    r->rgba8_fresh = false;
    r->target_fresh = false;
    return RT_OK;
}

static int update_one(RtRenderer* r, const SceneCamera* camera) {
    rtk::UpdateParams p;
    memcpy(&p.cam, camera, sizeof(SceneCamera));
    p.accum = r->accum;
    p.geom = r->geom;
    p.spheres = r->spheres;
    p.n_spheres = r->n_spheres;
    p.width = r->width; p.height = r->height;
    p.row_begin = r->row_begin; p.row_count = r->row_count;
    p.band_rows = r->band_rows; p.band_stride = r->band_stride;
    p.tiles_x = (r->width + rtk::kWarpTileW - 1) / rtk::kWarpTileW;
    p.n_tiles = p.tiles_x * ((r->row_count + rtk::kWarpTileH - 1) / rtk::kWarpTileH);
    const uint32_t frame = r->frame_index++;
    const uint32_t parity = frame & 1u;
    p.tile_counter = r->tile_counter + 2 * parity;
    if (r->order_rows) {
        unsigned int* cost = r->tile_counter + 4;
        p.row_cost_cur = cost + (size_t)(frame % 3u) * r->order_rows;
        p.row_cost_prev = cost + (size_t)((frame + 1u) % 3u) * r->order_rows;   // the frame before the last one
        p.row_order = cost + (size_t)(3u + parity) * r->order_rows;
    }
    p.counters = r->counters;
    p.first_hit = r->first_hit;
    if (r->clusters_stale && (r->variant == 0 || r->variant >= 4) && r->n_spheres > 0) {
        // first update of this scene that can use the primary pass: the tree of bounding spheres, kept in global
        // memory for the primary pass; the bounce kernel copies its image into shared memory when it fits there
        r->clusters_stale = false;
        rtt::Tree tree;
        const auto t_build = std::chrono::steady_clock::now();
        if (rtt::build(r->host_spheres.data(), r->n_spheres, tree)) {
            const size_t rec_slots = (size_t)tree.n_nodes * rtk::kTreeStride + tree.n_entry_pad;
            const size_t rec_bytes = rec_slots * sizeof(float4);
            if (tree.bytes() > r->cl_capacity || rec_bytes > r->prim_tree_capacity) {
                RT_CUDA(cudaStreamSynchronize(r->stream));
                cudaFree(r->cl_blob); r->cl_blob = nullptr; r->cl_capacity = 0;
                cudaFree(r->prim_tree); r->prim_tree = nullptr; r->prim_tree_capacity = 0;
                RT_CUDA(cudaMalloc(&r->cl_blob, tree.bytes()));
                r->cl_capacity = tree.bytes();
                RT_CUDA(cudaMalloc(&r->prim_tree, 2 * rec_bytes));
                r->prim_tree_capacity = rec_bytes;
            }
            std::vector<unsigned char> image(tree.bytes());
            tree.pack(image.data());
            RT_CUDA(cudaMemcpyAsync(r->cl_blob, image.data(), image.size(), cudaMemcpyHostToDevice, r->stream));
            RT_CUDA(cudaStreamSynchronize(r->stream));   // `image` is about to go away
            r->prim_tree_slots = rec_slots;
            r->cl_n_nodes = tree.n_nodes;
            r->cl_first_leaf = tree.first_leaf;
            r->cl_n_entry = tree.n_entry;
            r->cl_entry_r = tree.entry_r;
            r->cl_depth = tree.depth;
            // per-thread node stack of the bounce kernel: the entry backlog, 4 more entries, and 7 per level of
            // depth-first growth below them
            r->cl_stack_cap = (uint32_t)rtk::kTreeBacklog + 4u + 7u * tree.depth;
            const bool fits = r->n_spheres < 32768u && tree.n_nodes < 65536u && tree.bounce_bytes() <= rtk::kClusterImageMax &&
                              r->cl_stack_cap <= rtk::kTreeStackMax;
            r->cl_bytes = fits ? (uint32_t)tree.bounce_bytes() : 0u;
            // too large for shared memory: the bounce kernel walks the tree where it lies (16-bit node numbers on its stack)
            // (sphere indices travel in 26 bits of the cooperative traversal's candidate codes)
            r->cl_global = (!fits && tree.n_nodes < 65536u && r->n_spheres < (1u << 26) && r->cl_stack_cap <= rtk::kTreeStackMax) ? 1u : 0u;
        }
        r->tree_build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_build).count();
    }
    p.prim_tree = r->cl_n_nodes ? r->prim_tree + parity * r->prim_tree_slots : nullptr;
    p.cl_blob = r->cl_blob;
    p.cl_n_nodes = r->cl_n_nodes;
    p.cl_first_leaf = r->cl_first_leaf;
    p.cl_bytes = r->cl_bytes;
    p.cl_global = r->cl_bytes ? 0u : r->cl_global;
    if (r->variant == 5) p.cl_bytes = p.cl_global = 0u;   // bounce rays through the streamed kernel whatever the scene size
    p.cl_stack_cap = r->cl_stack_cap;
    p.cl_n_entry = r->cl_n_entry;
    p.cl_entry_r = r->cl_entry_r;
    if (camera->max_depth > 1.0f && !r->wf_counts && rtk::primary_pass_applies(p) && (r->variant == 0 || r->variant >= 4)) {
        // first deep frame: path queues for the wavefront bounces (two queues x three float4 planes)
        for (int q = 0; q < 2; ++q) RT_CUDA(cudaMalloc(&r->wf_queue[q], 3 * r->n_texels * sizeof(float4)));
        RT_CUDA(cudaMalloc(&r->wf_counts, rtk::kWfWords * sizeof(unsigned int)));
    }
    p.wf_queue[0] = r->wf_queue[0];
    p.wf_queue[1] = r->wf_queue[1];
    p.wf_counts = r->wf_counts;
    p.wf_capacity = (uint32_t)r->n_texels;
    p.wf_bounce = 0;
    p.rgba8_out = nullptr;
    p.rgba8_abs = 0u;
    p.precision = (uint32_t)r->precision;
    r->rgba8_fresh = false;
    r->target_fresh = false;
    if (r->rgba8_target && rtk::fused_shade_applies(p, r->variant)) {
        p.rgba8_out = r->rgba8_target;   // fused gather: the texels go into the caller's full image, possibly on a peer GPU
        p.rgba8_abs = 1u;
        r->target_fresh = true;
    } else if (r->rgba8_output && rtk::fused_shade_applies(p, r->variant)) {
        // the slot the next rt_read_rgba8_async will send: its previous copy must have left it
        const int b = (int)(r->async_reads & 1u);
        if (b && !r->rgba8_alt) RT_CUDA(cudaMalloc(&r->rgba8_alt, r->n_texels * sizeof(uchar4)));
        if (r->async_reads >= 2) RT_CUDA(cudaStreamWaitEvent(r->stream, r->ev_copied[b], 0));
        p.rgba8_out = reinterpret_cast<uint32_t*>(b ? r->rgba8_alt : r->rgba8);
    }
    int launched = 0;
    // a depth-1 frame's prep kernel goes to prep_stream, behind the frame that last used this parity's buffers (two
    // frames back) — and behind the last frame too if that one's prep kernel ran in line (it shares the cost records)
    const rtk::PrepAsync prep_async{r->prep_stream, r->ev_frame[parity], r->last_prep_async ? nullptr : r->ev_frame[parity ^ 1u],
                                    r->ev_prep[parity]};
    bool was_async = false;
    RT_CUDA(rtk::launch_update(p, r->variant, r->sm_count, r->stream, &launched, RT_PREP_ASYNC ? &prep_async : nullptr, &was_async));
    r->last_prep_async = was_async;
    RT_CUDA(cudaEventRecord(r->ev_frame[parity], r->stream));
    r->launches += (uint64_t)launched;
    r->rgba8_fresh = p.rgba8_out != nullptr && !r->target_fresh;
    return RT_OK;
}

int rt_update(RtRenderer* r, const SceneCamera* camera) {
    RT_CHECK_HANDLE(r);
    if (!camera) return fail(RT_ERR_INVALID_ARG, "camera is null");
    if (!r->has_scene) return fail(RT_ERR_NO_SCENE, "rt_update before rt_set_scene");
    return update_one(r, camera);
}

int rt_update_n(RtRenderer* r, const SceneCamera* cameras, uint32_t n) {
    RT_CHECK_HANDLE(r);
    if (n > 0 && !cameras) return fail(RT_ERR_INVALID_ARG, "cameras is null");
    if (!r->has_scene) return fail(RT_ERR_NO_SCENE, "rt_update_n before rt_set_scene");
    for (uint32_t i = 0; i < n; ++i) {
        int rc = update_one(r, cameras + i);
        if (rc != RT_OK) return rc;
    }
    return RT_OK;
}

int rt_synchronize(RtRenderer* r) {
    RT_CHECK_HANDLE(r);
    RT_CUDA(cudaStreamSynchronize(r->stream));
    return RT_OK;
}

int rt_read_accum(RtRenderer* r, float* host) {
    RT_CHECK_HANDLE(r);
    if (!host) return fail(RT_ERR_INVALID_ARG, "host buffer is null");
    RT_CUDA(cudaMemcpyAsync(host, r->accum, r->n_texels * sizeof(float4), cudaMemcpyDeviceToHost, r->stream));
    RT_CUDA(cudaStreamSynchronize(r->stream));
    return RT_OK;
}

int rt_write_accum(RtRenderer* r, const float* host) {
    RT_CHECK_HANDLE(r);
    if (!host) return fail(RT_ERR_INVALID_ARG, "host buffer is null");
    r->rgba8_fresh = false;
    r->target_fresh = false;
    RT_CUDA(cudaMemcpyAsync(r->accum, host, r->n_texels * sizeof(float4), cudaMemcpyHostToDevice, r->stream));
    RT_CUDA(cudaStreamSynchronize(r->stream));
    return RT_OK;
}

int rt_quantize(RtRenderer* r) {
    RT_CHECK_HANDLE(r);
    // a fused update (rt_set_rgba8_output) may have written this very buffer already: slot 0 of the asynchronous reads
    if (r->rgba8_fresh && (r->async_reads & 1u) == 0u) return RT_OK;
    // r->rgba8 is slot 0 of the asynchronous reads as well: an earlier rt_read_rgba8_async may still be copying out of it
    if (r->async_reads > 0) RT_CUDA(cudaStreamWaitEvent(r->stream, r->ev_copied[0], 0));
    RT_CUDA(rtk::launch_quantize(r->accum, r->rgba8, r->n_texels, r->stream));
    r->launches += 1;
    return RT_OK;
}

int rt_quantize_into(RtRenderer* r, void* device_rgba8) {
    RT_CHECK_HANDLE(r);
    if (!device_rgba8) return fail(RT_ERR_INVALID_ARG, "device buffer is null");
    RT_CUDA(rtk::launch_quantize(r->accum, reinterpret_cast<uchar4*>(device_rgba8), r->n_texels, r->stream));
    r->launches += 1;
    return RT_OK;
}

int rt_read_rgba8(RtRenderer* r, uint8_t* host) {
    RT_CHECK_HANDLE(r);
    if (!host) return fail(RT_ERR_INVALID_ARG, "host buffer is null");
    int rc = rt_quantize(r);
    if (rc != RT_OK) return rc;
    RT_CUDA(cudaMemcpyAsync(host, r->rgba8, r->n_texels * sizeof(uchar4), cudaMemcpyDeviceToHost, r->stream));
    RT_CUDA(cudaStreamSynchronize(r->stream));
    return RT_OK;
}

// The window's rows (window order, `elem` bytes per texel) to their places in a full image: one copy for a contiguous
// window, a strided 2-D copy over the whole bands (+ one for a short last band) for an interleaved one.
static cudaError_t scatter_window(const RtRenderer* r, void* full, const void* src, size_t elem, cudaStream_t stream) {
    const size_t row_bytes = (size_t)r->width * elem;
    unsigned char* dst = static_cast<unsigned char*>(full);
    const unsigned char* from = static_cast<const unsigned char*>(src);
    if (r->band_rows == 0)
        return cudaMemcpyAsync(dst + (size_t)r->row_begin * row_bytes, from, (size_t)r->row_count * row_bytes, cudaMemcpyDefault, stream);
    const size_t band_bytes = (size_t)r->band_rows * row_bytes;
    const uint32_t whole = r->row_count / r->band_rows, rest = r->row_count % r->band_rows;
    cudaError_t e = cudaSuccess;
    if (whole)
        e = cudaMemcpy2DAsync(dst + (size_t)r->row_begin * row_bytes, (size_t)r->band_stride * row_bytes, from, band_bytes, band_bytes,
                              whole, cudaMemcpyDefault, stream);
    if (e == cudaSuccess && rest)
        e = cudaMemcpyAsync(dst + ((size_t)r->row_begin + (size_t)whole * r->band_stride) * row_bytes, from + (size_t)whole * band_bytes,
                            (size_t)rest * row_bytes, cudaMemcpyDefault, stream);
    return e;
}

// Common part of the asynchronous reads: the frame as it stands into device slot b (unless a fused update wrote it
// already), the copy stream ordered behind it.  Returns the slot's buffer.
static int stage_async_slot(RtRenderer* r, uchar4** dev_out, int* slot_out) {
    if (!r->rgba8_alt) RT_CUDA(cudaMalloc(&r->rgba8_alt, r->n_texels * sizeof(uchar4)));
    const int b = (int)(r->async_reads & 1u);
    uchar4* dev = b ? r->rgba8_alt : r->rgba8;
    // slot b was last used two reads ago: its copy must have left the buffer before it is overwritten
    if (!r->rgba8_fresh) {   // (a fused update has already written slot b, after the same wait)
        if (r->async_reads >= 2) RT_CUDA(cudaStreamWaitEvent(r->stream, r->ev_copied[b], 0));
        RT_CUDA(rtk::launch_quantize(r->accum, dev, r->n_texels, r->stream));
        r->launches += 1;
    }
    r->rgba8_fresh = false;
    RT_CUDA(cudaEventRecord(r->ev_quantized[b], r->stream));
    RT_CUDA(cudaStreamWaitEvent(r->copy_stream, r->ev_quantized[b], 0));
    *dev_out = dev;
    *slot_out = b;
    return RT_OK;
}

int rt_send_rgba8_async(RtRenderer* r, void* device_full_rgba8) {
    RT_CHECK_HANDLE(r);
    if (!device_full_rgba8) return fail(RT_ERR_INVALID_ARG, "destination is null");
    uchar4* dev = nullptr;
    int b = 0;
    int rc = stage_async_slot(r, &dev, &b);
    if (rc != RT_OK) return rc;
    RT_CUDA(scatter_window(r, device_full_rgba8, dev, sizeof(uchar4), r->copy_stream));
    RT_CUDA(cudaEventRecord(r->ev_copied[b], r->copy_stream));
    r->async_reads += 1;
    return RT_OK;
}

int rt_copy_signal(RtRenderer* r, void* device_flag, uint32_t value) {
    RT_CHECK_HANDLE(r);
    if (!device_flag) return fail(RT_ERR_INVALID_ARG, "flag is null");
    RT_CUDA(rtk::launch_signal(device_flag, value, r->copy_stream));
    r->launches += 1;
    return RT_OK;
}

int rt_copy_wait(RtRenderer* r, const void* device_flags, uint32_t n, uint32_t value, uint32_t timeout_ms) {
    RT_CHECK_HANDLE(r);
    if (!device_flags || n == 0) return fail(RT_ERR_INVALID_ARG, "no flags");
    RT_CUDA(rtk::launch_wait_flags(device_flags, n, value, timeout_ms ? timeout_ms : 1u, r->wait_status, r->copy_stream));
    r->launches += 1;
    return RT_OK;
}

int rt_read_rgba8_async(RtRenderer* r, uint8_t* host) {
    RT_CHECK_HANDLE(r);
    if (!host) return fail(RT_ERR_INVALID_ARG, "host buffer is null");
    if (!r->rgba8_alt) RT_CUDA(cudaMalloc(&r->rgba8_alt, r->n_texels * sizeof(uchar4)));
    const int b = (int)(r->async_reads & 1u);
    uchar4* dev = b ? r->rgba8_alt : r->rgba8;
    // slot b was last used two reads ago: its copy must have left the buffer before it is overwritten
    if (!r->rgba8_fresh) {   // (a fused update has already written slot b, after the same wait)
        if (r->async_reads >= 2) RT_CUDA(cudaStreamWaitEvent(r->stream, r->ev_copied[b], 0));
        RT_CUDA(rtk::launch_quantize(r->accum, dev, r->n_texels, r->stream));
        r->launches += 1;
    }
    r->rgba8_fresh = false;
    RT_CUDA(cudaEventRecord(r->ev_quantized[b], r->stream));
    RT_CUDA(cudaStreamWaitEvent(r->copy_stream, r->ev_quantized[b], 0));
    // cudaMemcpyDefault: the destination may be pinned host memory or device memory of this or a peer GPU
    RT_CUDA(cudaMemcpyAsync(host, dev, r->n_texels * sizeof(uchar4), cudaMemcpyDefault, r->copy_stream));
    RT_CUDA(cudaEventRecord(r->ev_copied[b], r->copy_stream));
    r->async_reads += 1;
    return RT_OK;
}

int rt_set_rgba8_output(RtRenderer* r, int enable) {
    RT_CHECK_HANDLE(r);
    r->rgba8_output = enable != 0;
    r->rgba8_fresh = false;
    return RT_OK;
}

int rt_set_rgba8_target(RtRenderer* r, void* device_rgba8) {
    RT_CHECK_HANDLE(r);
    r->rgba8_target = static_cast<uint32_t*>(device_rgba8);
    r->target_fresh = false;
    r->rgba8_fresh = false;
    return RT_OK;
}

int rt_stream_signal(RtRenderer* r, void* device_flag, uint32_t value) {
    RT_CHECK_HANDLE(r);
    if (!device_flag) return fail(RT_ERR_INVALID_ARG, "flag is null");
    RT_CUDA(rtk::launch_signal(device_flag, value, r->stream));
    r->launches += 1;
    return RT_OK;
}

int rt_stream_wait(RtRenderer* r, const void* device_flags, uint32_t n, uint32_t value, uint32_t timeout_ms) {
    RT_CHECK_HANDLE(r);
    if (!device_flags || n == 0) return fail(RT_ERR_INVALID_ARG, "no flags");
    RT_CUDA(rtk::launch_wait_flags(device_flags, n, value, timeout_ms ? timeout_ms : 1u, r->wait_status, r->stream));
    r->launches += 1;
    return RT_OK;
}

int rt_stream_wait_timed_out(RtRenderer* r, uint32_t* count) {
    RT_CHECK_HANDLE(r);
    if (!count) return fail(RT_ERR_INVALID_ARG, "count is null");
    unsigned int v = 0;
    RT_CUDA(cudaMemcpyAsync(&v, r->wait_status, sizeof(v), cudaMemcpyDeviceToHost, r->stream));
    RT_CUDA(cudaStreamSynchronize(r->stream));
    *count = v;
    return RT_OK;
}

int rt_set_precision(RtRenderer* r, int precision) {
    if (!r) return fail(RT_ERR_INVALID_ARG, "null renderer");
    if (precision < RT_PRECISION_EXACT || precision > RT_PRECISION_FAST_ALL) return fail(RT_ERR_INVALID_ARG, "precision %d", precision);
    r->precision = precision;
    return RT_OK;
}

int rt_tree_build_ms(RtRenderer* r, double* ms) {
    if (!r || !ms) return fail(RT_ERR_INVALID_ARG, "null argument");
    *ms = r->tree_build_ms;
    return RT_OK;
}

int rt_read_wait(RtRenderer* r) {
    RT_CHECK_HANDLE(r);
    RT_CUDA(cudaStreamSynchronize(r->copy_stream));
    return RT_OK;
}

// Device buffers other processes can write into: the gather target of a multi-process run (bench.py) lives in one
// process and is mapped by the others through CUDA IPC, so that rt_read_rgba8_async can copy a finished frame
// straight into it over NVLink.
int rt_device_alloc(int device, uint64_t bytes, void** out) {
    if (!out || bytes == 0) return fail(RT_ERR_INVALID_ARG, "null output or zero size");
    RT_CUDA(cudaSetDevice(device));
    RT_CUDA(cudaMalloc(out, (size_t)bytes));
    return RT_OK;
}
int rt_device_free(int device, void* ptr) {
    RT_CUDA(cudaSetDevice(device));
    RT_CUDA(cudaFree(ptr));
    return RT_OK;
}
int rt_ipc_export(int device, void* ptr, uint8_t handle[64]) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handles are 64 bytes");
    if (!ptr || !handle) return fail(RT_ERR_INVALID_ARG, "null argument");
    RT_CUDA(cudaSetDevice(device));
    cudaIpcMemHandle_t h;
    RT_CUDA(cudaIpcGetMemHandle(&h, ptr));
    memcpy(handle, &h, 64);
    return RT_OK;
}
int rt_ipc_open(int device, const uint8_t handle[64], void** out) {
    if (!handle || !out) return fail(RT_ERR_INVALID_ARG, "null argument");
    RT_CUDA(cudaSetDevice(device));
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    RT_CUDA(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
    return RT_OK;
}
int rt_ipc_close(int device, void* ptr) {
    RT_CUDA(cudaSetDevice(device));
    RT_CUDA(cudaIpcCloseMemHandle(ptr));
    return RT_OK;
}

static int read_counter(RtRenderer* r, int which, uint64_t* out) {
    RT_CHECK_HANDLE(r);
    if (!out) return fail(RT_ERR_INVALID_ARG, "out is null");
    unsigned long long v = 0;
    RT_CUDA(cudaMemcpyAsync(&v, r->counters + which, sizeof(v), cudaMemcpyDeviceToHost, r->stream));
    RT_CUDA(cudaStreamSynchronize(r->stream));
    *out = v;
    return RT_OK;
}
int rt_rays_traced(RtRenderer* r, uint64_t* rays) { return read_counter(r, 0, rays); }
int rt_exact_tests(RtRenderer* r, uint64_t* tests) { return read_counter(r, 1, tests); }
int rt_reset_counters(RtRenderer* r) {
    RT_CHECK_HANDLE(r);
    RT_CUDA(cudaMemsetAsync(r->counters, 0, 2 * sizeof(unsigned long long), r->stream));
    return RT_OK;
}

int rt_tree_info(RtRenderer* r, uint32_t* n_nodes, uint32_t* depth, uint32_t* bounce_path) {
    RT_CHECK_HANDLE(r);
    if (!n_nodes || !depth || !bounce_path) return fail(RT_ERR_INVALID_ARG, "null argument");
    *n_nodes = r->cl_n_nodes;
    *depth = r->cl_n_nodes ? r->cl_depth : 0u;
    *bounce_path = !r->cl_n_nodes ? 0u : r->variant == 5 ? 3u : r->cl_bytes ? 1u : r->cl_global ? 2u : 3u;
    return RT_OK;
}

int rt_wavefront_counts(RtRenderer* r, uint32_t* counts, uint32_t capacity, uint32_t* n_out) {
    RT_CHECK_HANDLE(r);
    if (!counts || !n_out) return fail(RT_ERR_INVALID_ARG, "null argument");
    *n_out = 0;
    if (!r->wf_counts) return RT_OK;   // no deep frame has run
    unsigned int host[rtk::kWfMaxDepth + 1];
    RT_CUDA(cudaMemcpyAsync(host, r->wf_counts + 2 * (rtk::kWfMaxDepth + 1), sizeof(host), cudaMemcpyDeviceToHost, r->stream));
    RT_CUDA(cudaStreamSynchronize(r->stream));
    uint32_t n = rtk::kWfMaxDepth + 1;
    if (n > capacity) n = capacity;
    for (uint32_t i = 0; i < n; ++i) counts[i] = host[i];
    *n_out = n;
    return RT_OK;
}

int rt_timer_begin(RtRenderer* r) {
    RT_CHECK_HANDLE(r);
    RT_CUDA(cudaEventRecord(r->ev_begin, r->stream));
    return RT_OK;
}
int rt_timer_end(RtRenderer* r, float* elapsed_ms) {
    RT_CHECK_HANDLE(r);
    if (!elapsed_ms) return fail(RT_ERR_INVALID_ARG, "elapsed_ms is null");
    RT_CUDA(cudaEventRecord(r->ev_end, r->stream));
    RT_CUDA(cudaEventSynchronize(r->ev_end));
    RT_CUDA(cudaEventElapsedTime(elapsed_ms, r->ev_begin, r->ev_end));
    return RT_OK;
}
int rt_kernel_launches(RtRenderer* r, uint64_t* launches) {
    if (!r || !launches) return fail(RT_ERR_INVALID_ARG, "null argument");
    *launches = r->launches;
    return RT_OK;
}

int rt_accum_device_ptr(RtRenderer* r, void** ptr) {
    if (!r || !ptr) return fail(RT_ERR_INVALID_ARG, "null argument");
    *ptr = r->accum;
    return RT_OK;
}
int rt_rgba8_device_ptr(RtRenderer* r, void** ptr) {
    if (!r || !ptr) return fail(RT_ERR_INVALID_ARG, "null argument");
    *ptr = r->rgba8;
    return RT_OK;
}
int rt_stream(RtRenderer* r, void** s) {
    if (!r || !s) return fail(RT_ERR_INVALID_ARG, "null argument");
    *s = (void*)r->stream;
    return RT_OK;
}
int rt_window(RtRenderer* r, uint32_t* width, uint32_t* height, uint32_t* row_begin, uint32_t* row_count) {
    if (!r) return fail(RT_ERR_INVALID_ARG, "null renderer");
    if (width) *width = r->width;
    if (height) *height = r->height;
    if (row_begin) *row_begin = r->row_begin;
    if (row_count) *row_count = r->row_count;
    return RT_OK;
}
int rt_set_kernel_variant(RtRenderer* r, int variant) {
    if (!r) return fail(RT_ERR_INVALID_ARG, "null renderer");
    if (variant < 0 || variant > 5) return fail(RT_ERR_INVALID_ARG, "variant %d not in 0..5", variant);
    r->variant = variant;
    return RT_OK;
}

// ---------------------------------------------------------------------------------------------
// rt_update_static: one `update` without a renderer handle and without any allocation (include/rt_abi.h).  Calls no
// cudaMalloc / cudaFree / allocating library routine: the kernel's scratch is a file-scope device array in rt_kernels.cu.
// ---------------------------------------------------------------------------------------------
int rt_update_static(int device, const GpuSphere* device_spheres, uint32_t n_spheres, const SceneCamera* camera,
                     uint32_t width, uint32_t height, float* device_accum_rgba32f, void* cuda_stream) {
    if (!camera || !device_accum_rgba32f) return fail(RT_ERR_INVALID_ARG, "null camera or accumulation image");
    if (n_spheres > 0 && !device_spheres) return fail(RT_ERR_INVALID_ARG, "spheres is null");
    if (n_spheres > RT_STATIC_MAX_SPHERES)
        return fail(RT_ERR_INVALID_ARG, "rt_update_static: %u spheres exceed the static scratch's maximum of %u (use a renderer handle)",
                    n_spheres, (unsigned)RT_STATIC_MAX_SPHERES);
    if (width == 0 || height == 0 || width > (1u << 20) || height > (1u << 20)) return fail(RT_ERR_INVALID_ARG, "image %ux%u", width, height);
    if (device < 0 || device >= 64) return fail(RT_ERR_INVALID_ARG, "device %d out of range", device);
    RT_CUDA(cudaSetDevice(device));
    RT_CUDA(rtk::launch_update_static(device, device_spheres, n_spheres, *camera, width, height,
                                      reinterpret_cast<float4*>(device_accum_rgba32f), static_cast<cudaStream_t>(cuda_stream)));
    return RT_OK;
}

// ---------------------------------------------------------------------------------------------
// Headless one-shot entries
// ---------------------------------------------------------------------------------------------
int rt_render(int device, const GpuSphere* spheres, uint32_t count, const SceneCamera* camera, uint32_t width,
              uint32_t height, uint32_t samples, float* out_rgba32f) {
    if (!camera || !out_rgba32f) return fail(RT_ERR_INVALID_ARG, "null argument");
    RtRenderer* r = nullptr;
    int rc = rt_create(device, width, height, &r);
    if (rc != RT_OK) return rc;
    rc = rt_set_scene(r, spheres, count, count);
    if (rc == RT_OK) rc = rt_init(r);
    for (uint32_t i = 0; rc == RT_OK && i < samples; ++i) {
        SceneCamera cam = *camera;
        if (i > 0) cam.camera_has_moved = 0.0f;   // the reset flag holds for the first frame only (camera.rs:245-252)
        rc = rt_update(r, &cam);
    }
    if (rc == RT_OK) rc = rt_read_accum(r, out_rgba32f);
    rt_destroy(r);
    return rc;
}

int rt_render_frames(int device, const GpuSphere* spheres, uint32_t count, const SceneCamera* cameras,
                     uint32_t n_frames, uint32_t width, uint32_t height, uint8_t* out_rgba8) {
    if ((n_frames && !cameras) || !out_rgba8) return fail(RT_ERR_INVALID_ARG, "null argument");
    RtRenderer* r = nullptr;
    int rc = rt_create(device, width, height, &r);
    if (rc != RT_OK) return rc;
    rc = rt_set_scene(r, spheres, count, count);
    const size_t frame_bytes = (size_t)width * height * 4;
    for (uint32_t f = 0; rc == RT_OK && f < n_frames; ++f) {
        rc = rt_init(r);
        if (rc == RT_OK) rc = rt_update(r, cameras + f);
        if (rc == RT_OK) rc = rt_read_rgba8(r, out_rgba8 + f * frame_bytes);
    }
    rt_destroy(r);
    return rc;
}

}  // extern "C"
