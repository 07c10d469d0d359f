// rt_multi.cu — one image sharded by contiguous row blocks over several devices of one host, scene
// replicated, result gathered into devices[0] over NVLink.  Depth-1 frames are gathered BY THE SHADING
// PASS: every shard's kernel stores its RGBA8 texels straight into its rows of the full frame on
// devices[0] (rt_set_rgba8_target on a peer pointer), so the gather is finished when the kernels are;
// deeper frames (and shards without peer access) are quantised and sent with cudaMemcpyPeerAsync on
// the shard's own stream.  No collective is needed: the shader has no inter-pixel dependency
// (compute_shader.wgsl:333-364 touches only its own texel).
#include <cuda_runtime.h>

#include <new>
#include <vector>

#include "../../include/rt_abi.h"
#include "rt_internal.h"

using rti::fail;
using rti::fail_cuda;

struct RtMultiRenderer {
    std::vector<RtRenderer*> shards;
    std::vector<char> direct;        // shard i can store into devices[0]'s memory (same device or peer access)
    uint32_t width = 0, height = 0;
    uint32_t band_rows = 0;          // 0: contiguous row blocks (rt_shard_rows); else interleaved bands (rt_create_bands)
    uchar4* full_rgba8 = nullptr;    // on devices[0]
    float4* full_accum = nullptr;    // on devices[0], allocated on first use
};

#define RTM_CUDA(call)                                      \
    do {                                                    \
        cudaError_t e_ = (call);                            \
        if (e_ != cudaSuccess) return fail_cuda(e_, #call); \
    } while (0)

extern "C" {

int rt_shard_rows(uint32_t height, uint32_t n_shards, uint32_t index, uint32_t* row_begin, uint32_t* row_count) {
    if (!row_begin || !row_count) return fail(RT_ERR_INVALID_ARG, "null argument");
    if (n_shards == 0 || index >= n_shards) return fail(RT_ERR_INVALID_ARG, "shard %u of %u", index, n_shards);
    if (height < n_shards) return fail(RT_ERR_INVALID_ARG, "fewer rows (%u) than shards (%u)", height, n_shards);
    const uint32_t tiles = (height + 3) / 4;
    uint32_t r0, r1;
    if (tiles >= n_shards) {
        r0 = (uint32_t)((uint64_t)tiles * index / n_shards) * 4;
        r1 = (index + 1 == n_shards) ? height : (uint32_t)((uint64_t)tiles * (index + 1) / n_shards) * 4;
    } else {   // fewer 4-row tiles than shards: fall back to single rows
        r0 = (uint32_t)((uint64_t)height * index / n_shards);
        r1 = (uint32_t)((uint64_t)height * (index + 1) / n_shards);
    }
    if (r1 > height) r1 = height;
    *row_begin = r0;
    *row_count = r1 - r0;
    return RT_OK;
}

int rt_multi_destroy(RtMultiRenderer* m) {
    if (!m) return RT_OK;
    if (!m->shards.empty()) {
        cudaSetDevice(m->shards[0]->device);
        cudaFree(m->full_rgba8);
        cudaFree(m->full_accum);
    }
    for (RtRenderer* s : m->shards) rt_destroy(s);
    delete m;
    return RT_OK;
}

// One shard's rows (window order) into its places in a full image on devices[0]: one copy for a contiguous block, one
// strided 2-D copy over the whole bands (+ one for a short last band) for an interleaved shard.  elem = bytes per texel.
static cudaError_t scatter_rows(const RtMultiRenderer* m, const RtRenderer* s, void* full, const void* src, size_t elem) {
    const size_t row_bytes = (size_t)m->width * elem;
    unsigned char* dst = static_cast<unsigned char*>(full);
    const unsigned char* from = static_cast<const unsigned char*>(src);
    if (s->band_rows == 0)
        return cudaMemcpyAsync(dst + (size_t)s->row_begin * row_bytes, from, (size_t)s->row_count * row_bytes, cudaMemcpyDefault, s->stream);
    const size_t band_bytes = (size_t)s->band_rows * row_bytes;
    const uint32_t whole = s->row_count / s->band_rows, rest = s->row_count % s->band_rows;
    cudaError_t e = cudaSuccess;
    if (whole)
        e = cudaMemcpy2DAsync(dst + (size_t)s->row_begin * row_bytes, (size_t)s->band_stride * row_bytes, from, band_bytes, band_bytes,
                              whole, cudaMemcpyDefault, s->stream);
    if (e == cudaSuccess && rest)
        e = cudaMemcpyAsync(dst + ((size_t)s->row_begin + (size_t)whole * s->band_stride) * row_bytes, from + (size_t)whole * band_bytes,
                            (size_t)rest * row_bytes, cudaMemcpyDefault, s->stream);
    return e;
}

static int multi_create(const int* devices, uint32_t n_devices, uint32_t width, uint32_t height, uint32_t band_rows,
                        RtMultiRenderer** out);

int rt_multi_create(const int* devices, uint32_t n_devices, uint32_t width, uint32_t height, RtMultiRenderer** out) {
    return multi_create(devices, n_devices, width, height, 0, out);
}

int rt_multi_create_bands(const int* devices, uint32_t n_devices, uint32_t width, uint32_t height, uint32_t band_rows,
                          RtMultiRenderer** out) {
    if (band_rows == 0 || band_rows % 8u != 0u) return fail(RT_ERR_INVALID_ARG, "band_rows %u is not a positive multiple of 8", band_rows);
    return multi_create(devices, n_devices, width, height, band_rows, out);
}

static int multi_create(const int* devices, uint32_t n_devices, uint32_t width, uint32_t height, uint32_t band_rows,
                        RtMultiRenderer** out) {
    if (!out) return fail(RT_ERR_INVALID_ARG, "out is null");
    *out = nullptr;
    if (!devices || n_devices == 0) return fail(RT_ERR_INVALID_ARG, "no devices");
    if (height < n_devices) return fail(RT_ERR_INVALID_ARG, "fewer rows (%u) than devices (%u)", height, n_devices);
    RtMultiRenderer* m = new (std::nothrow) RtMultiRenderer();
    if (!m) return fail(RT_ERR_OUT_OF_MEMORY, "host allocation failed");
    m->width = width; m->height = height; m->band_rows = band_rows;
    if (band_rows && (height + band_rows - 1) / band_rows < n_devices) { delete m; return fail(RT_ERR_INVALID_ARG, "fewer bands than devices"); }
    for (uint32_t i = 0; i < n_devices; ++i) {
        uint32_t r0 = 0, rows = 0;
        int rc = RT_OK;
        RtRenderer* s = nullptr;
        if (band_rows) {
            rc = rt_create_bands(devices[i], width, height, band_rows, n_devices, i, &s);
        } else {
            rc = rt_shard_rows(height, n_devices, i, &r0, &rows);
            if (rc == RT_OK) rc = rt_create_window(devices[i], width, height, r0, rows, &s);
        }
        if (rc != RT_OK) { rt_multi_destroy(m); return rc; }
        m->shards.push_back(s);
    }
    // peer access devices[0] <-> devices[i]; cudaMemcpyPeerAsync falls back to staging when absent
    m->direct.assign(n_devices, 0);
    m->direct[0] = 1;
    for (uint32_t i = 1; i < n_devices; ++i) {
        if (devices[i] == devices[0]) { m->direct[i] = 1; continue; }
        int can = 0;
        if (cudaDeviceCanAccessPeer(&can, devices[0], devices[i]) == cudaSuccess && can) {
            cudaSetDevice(devices[0]);
            cudaError_t e = cudaDeviceEnablePeerAccess(devices[i], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { rt_multi_destroy(m); return fail_cuda(e, "cudaDeviceEnablePeerAccess"); }
            cudaSetDevice(devices[i]);
            e = cudaDeviceEnablePeerAccess(devices[0], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { rt_multi_destroy(m); return fail_cuda(e, "cudaDeviceEnablePeerAccess"); }
            cudaGetLastError();
            m->direct[i] = 1;
        }
    }
    // the gather sends RGBA8 blocks: depth-1 updates write them in the shading pass (no quantise kernel per shard)
    for (RtRenderer* s : m->shards) rt_set_rgba8_output(s, 1);
    cudaError_t e = cudaSetDevice(devices[0]);
    if (e == cudaSuccess) e = cudaMalloc(&m->full_rgba8, (size_t)width * height * sizeof(uchar4));
    if (e != cudaSuccess) { rt_multi_destroy(m); return fail_cuda(e, "rt_multi_create"); }
    // fused gather: the shading pass of a depth-1 update stores the shard's rows straight into the full frame
    for (uint32_t i = 0; i < n_devices; ++i)
        if (m->direct[i]) rt_set_rgba8_target(m->shards[i], m->full_rgba8);
    *out = m;
    return RT_OK;
}

int rt_multi_set_scene(RtMultiRenderer* m, const GpuSphere* spheres, uint32_t buffer_len, uint32_t count) {
    if (!m) return fail(RT_ERR_INVALID_ARG, "null renderer");
    for (RtRenderer* s : m->shards) {
        int rc = rt_set_scene(s, spheres, buffer_len, count);
        if (rc != RT_OK) return rc;
    }
    return RT_OK;
}

int rt_multi_init(RtMultiRenderer* m) {
    if (!m) return fail(RT_ERR_INVALID_ARG, "null renderer");
    for (RtRenderer* s : m->shards) {
        int rc = rt_init(s);
        if (rc != RT_OK) return rc;
    }
    return RT_OK;
}

int rt_multi_update(RtMultiRenderer* m, const SceneCamera* camera) {
    if (!m) return fail(RT_ERR_INVALID_ARG, "null renderer");
    for (RtRenderer* s : m->shards) {   // asynchronous: all devices run concurrently
        int rc = rt_update(s, camera);
        if (rc != RT_OK) return rc;
    }
    return RT_OK;
}

int rt_multi_gather_rgba8(RtMultiRenderer* m, uint8_t* host) {
    if (!m) return fail(RT_ERR_INVALID_ARG, "null renderer");
    const int dev0 = m->shards[0]->device;
    for (RtRenderer* s : m->shards) {
        if (s->target_fresh) continue;   // the last update's shading pass has already stored these rows in place
        int rc = rt_quantize(s);
        if (rc != RT_OK) return rc;
        RTM_CUDA(cudaSetDevice(s->device));
        RTM_CUDA(scatter_rows(m, s, m->full_rgba8, s->rgba8, sizeof(uchar4)));
    }
    for (RtRenderer* s : m->shards) {
        RTM_CUDA(cudaSetDevice(s->device));
        RTM_CUDA(cudaStreamSynchronize(s->stream));
    }
    if (host) {
        RTM_CUDA(cudaSetDevice(dev0));
        RTM_CUDA(cudaMemcpy(host, m->full_rgba8, (size_t)m->width * m->height * sizeof(uchar4), cudaMemcpyDeviceToHost));
    }
    return RT_OK;
}

int rt_multi_gather_accum(RtMultiRenderer* m, float* host) {
    if (!m) return fail(RT_ERR_INVALID_ARG, "null renderer");
    const int dev0 = m->shards[0]->device;
    if (!m->full_accum) {
        RTM_CUDA(cudaSetDevice(dev0));
        RTM_CUDA(cudaMalloc(&m->full_accum, (size_t)m->width * m->height * sizeof(float4)));
    }
    for (RtRenderer* s : m->shards) {
        RTM_CUDA(cudaSetDevice(s->device));
        RTM_CUDA(scatter_rows(m, s, m->full_accum, s->accum, sizeof(float4)));
    }
    for (RtRenderer* s : m->shards) {
        RTM_CUDA(cudaSetDevice(s->device));
        RTM_CUDA(cudaStreamSynchronize(s->stream));
    }
    if (host) {
        RTM_CUDA(cudaSetDevice(dev0));
        RTM_CUDA(cudaMemcpy(host, m->full_accum, (size_t)m->width * m->height * sizeof(float4), cudaMemcpyDeviceToHost));
    }
    return RT_OK;
}

int rt_multi_rays_traced(RtMultiRenderer* m, uint64_t* rays) {
    if (!m || !rays) return fail(RT_ERR_INVALID_ARG, "null argument");
    uint64_t total = 0;
    for (RtRenderer* s : m->shards) {
        uint64_t v = 0;
        int rc = rt_rays_traced(s, &v);
        if (rc != RT_OK) return rc;
        total += v;
    }
    *rays = total;
    return RT_OK;
}

int rt_multi_shard(RtMultiRenderer* m, uint32_t index, RtRenderer** shard) {
    if (!m || !shard) return fail(RT_ERR_INVALID_ARG, "null argument");
    if (index >= m->shards.size()) return fail(RT_ERR_INVALID_ARG, "shard %u of %zu", index, m->shards.size());
    *shard = m->shards[index];
    return RT_OK;
}

}  // extern "C"
