// rt_cluster.h — spatial clusters of the sphere list for incoherent (bounce) rays.  Host side, plain C++;
// shared by rt_api.cu (rt_set_scene) and the CPU emulation in tests/emu.
//
// sphere_list_hit (compute_shader.wgsl:164-180) tests every sphere for every ray.  Without NaNs its result is the
// lexicographic minimum of (first valid root, index) over the spheres, so the order of
// the exact tests is free as long as ties go to the lower index — and spheres the ray's line cannot reach need no test
// at all.  The spheres are therefore split at medians (k-d fashion) into clusters of up to kSlots members; a
// cluster is represented by a bounding sphere whose record has the SAME form as a sphere's own broad-phase record
// (rtc::broad_record_k), so the kernels test "does the ray's line come near this cluster" with the 7-FFMA test they
// already have and descend into the members only then.
//
// Why the bound is safe.  An exact hit of member i (the reference's float discriminant >= 0 or NaN) implies, in real
// numbers, dist(line, C_i)^2 <= r_i^2 + kappa S_i with S_i <= 2 (|C_i|^2 + r_i^2 + |O|^2) (kappa
// = 2^-16 absorbs every rounding of the reference's own discriminant).  Hence
//     dist(line, Cc) <= |C_i - Cc| + r_i + 2^-7.5 (sqrt(|C_i|^2 + r_i^2) + |O|)
// and the cluster radius is  Rc' = max_i(|C_i - Cc| + r_i) + 2^-7 A_c  (A_c = max_i sqrt(|C_i|^2 + r_i^2)), with the
// ray's own share m = 2^-7 |O| added at run time (rtc::cluster_thr: the threshold drops by 2 Rmax m + m^2, Rmax the
// largest Rc').  Spheres that would spoil this — non-finite or beyond 2^30, or much larger than a cluster (a ground
// sphere) — are not clustered: they go to "always" clusters whose record passes for every ray, and are then tested
// by their own records exactly as before.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "rt_core.cuh"

namespace rtcl {

constexpr int kSlots = RT_CLUSTER_SLOTS;   // member slots per cluster
constexpr int kStride = kSlots + 1;        // record slots per cluster in the image: the odd stride spreads the clusters over the shared-memory banks
constexpr uint32_t kMaxAlways = 64;     // at most this many oversized spheres are pulled out of the clusters
constexpr double kMargin = 0.0078125;   // 2^-7
constexpr uint16_t kNoSphere = 0xFFFF;

struct Clusters {
    uint32_t n_top = 0;        // clusters, the always-scanned ones last
    uint32_t n_top_pad = 0;    // top-level records, padded to a multiple of 8 with never-passing records
    float extra_r = 0.0f;      // largest bounding radius Rc' over the bounded clusters
    std::vector<float> top;        // n_top_pad x (Cc.xyz, k)
    std::vector<float> mem;        // n_top * kStride x (C.xyz, k): the members' own broad-phase records
    std::vector<uint16_t> order;   // n_top * kStride: original sphere index of a member slot, kNoSphere for padding
    std::vector<uint8_t> count;    // n_top_pad: members per cluster

    // Device image: [top | mem | order | count], every part padded to 16 bytes.
    static size_t pad16(size_t b) { return (b + 15) / 16 * 16; }
    size_t off_mem() const { return (size_t)n_top_pad * 16; }
    size_t off_order() const { return off_mem() + (size_t)n_top * kStride * 16; }
    size_t off_count() const { return off_order() + pad16((size_t)n_top * kStride * 2); }
    size_t bytes() const { return off_count() + pad16(n_top_pad); }
    void pack(unsigned char* blob) const {
        memset(blob, 0, bytes());
        memcpy(blob, top.data(), top.size() * 4);
        memcpy(blob + off_mem(), mem.data(), mem.size() * 4);
        memcpy(blob + off_order(), order.data(), order.size() * 2);
        memcpy(blob + off_count(), count.data(), count.size());
    }
};

inline float round_up(double v) {
    float f = (float)v;
    if ((double)f < v) f = nextafterf(f, INFINITY);
    return f;
}

inline double coord(const GpuSphere& g, int axis) { return axis == 0 ? g.position.x : (axis == 1 ? g.position.y : g.position.z); }

// Median splits along the longest axis of the centres' bounding box until a part fits one cluster: balanced,
// compact leaves (a space-filling curve cut into runs leaves a few clusters that span the scene).
inline void split(const GpuSphere* s, std::vector<uint32_t>& ids, size_t lo_i, size_t hi_i,
                  std::vector<std::vector<uint32_t>>& groups) {
    const size_t cnt = hi_i - lo_i;
    if (cnt <= (size_t)kSlots) {
        if (cnt) groups.emplace_back(ids.begin() + lo_i, ids.begin() + hi_i);
        return;
    }
    double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (size_t j = lo_i; j < hi_i; ++j)
        for (int k = 0; k < 3; ++k) {
            lo[k] = std::min(lo[k], coord(s[ids[j]], k));
            hi[k] = std::max(hi[k], coord(s[ids[j]], k));
        }
    int axis = 0;
    for (int k = 1; k < 3; ++k)
        if (hi[k] - lo[k] > hi[axis] - lo[axis]) axis = k;
    // split so that the left part is a whole number of clusters when possible (fuller leaves)
    size_t half = cnt / 2;
    if (cnt > 2 * (size_t)kSlots) half = (half + kSlots - 1) / kSlots * kSlots;
    std::nth_element(ids.begin() + lo_i, ids.begin() + lo_i + half, ids.begin() + hi_i, [&](uint32_t a, uint32_t b) {
        const double ca = coord(s[a], axis), cb = coord(s[b], axis);
        return ca != cb ? ca < cb : a < b;
    });
    split(s, ids, lo_i, lo_i + half, groups);
    split(s, ids, lo_i + half, hi_i, groups);
}

// Clusters of `ids`; returns each cluster's geometric radius max(|C_i - Cc| + r_i) around its bounding-box centre.
inline void cut(const GpuSphere* s, std::vector<uint32_t> ids, std::vector<std::vector<uint32_t>>& groups,
                std::vector<double>& radii) {
    groups.clear();
    radii.clear();
    split(s, ids, 0, ids.size(), groups);
    for (std::vector<uint32_t>& g : groups) {
        double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (uint32_t i : g)
            for (int k = 0; k < 3; ++k) {
                lo[k] = std::min(lo[k], coord(s[i], k));
                hi[k] = std::max(hi[k], coord(s[i], k));
            }
        double r = 0.0;
        for (uint32_t i : g) {
            const double dx = s[i].position.x - 0.5 * (lo[0] + hi[0]), dy = s[i].position.y - 0.5 * (lo[1] + hi[1]),
                         dz = s[i].position.z - 0.5 * (lo[2] + hi[2]);
            r = std::max(r, std::sqrt(dx * dx + dy * dy + dz * dz) + std::fabs((double)s[i].radius));
        }
        std::sort(g.begin(), g.end());   // members in index order
        radii.push_back(r);
    }
}

// Builds the clusters of spheres[0, n).  Returns false (out unusable) when n does not fit 16-bit member indices.
inline bool build(const GpuSphere* s, uint32_t n, Clusters& out) {
    out = Clusters();
    if (n == 0 || n >= kNoSphere) return false;
    const double lim = 1073741824.0;   // 2^30, the magnitude window of the bound
    std::vector<uint32_t> bounded, always;
    for (uint32_t i = 0; i < n; ++i) {
        const double m = std::max(std::max(std::fabs((double)s[i].position.x), std::fabs((double)s[i].position.y)),
                                  std::max(std::fabs((double)s[i].position.z), std::fabs((double)s[i].radius)));
        if (m <= lim) bounded.push_back(i);   // false for NaN
        else always.push_back(i);
    }
    std::vector<std::vector<uint32_t>> groups;
    std::vector<double> radii;
    for (int pass = 0; pass < 2; ++pass) {
        cut(s, bounded, groups, radii);
        if (pass == 1 || groups.empty()) break;
        // spheres larger than a typical cluster would inflate their cluster for everyone: pull the largest out
        std::vector<double> sorted = radii;
        std::nth_element(sorted.begin(), sorted.begin() + sorted.size() / 2, sorted.end());
        const double typical = sorted[sorted.size() / 2];
        std::vector<uint32_t> big;
        for (uint32_t i : bounded)
            if (std::fabs((double)s[i].radius) > typical) big.push_back(i);
        if (big.empty()) break;
        std::sort(big.begin(), big.end(), [&](uint32_t a, uint32_t b) {
            const double ra = std::fabs((double)s[a].radius), rb = std::fabs((double)s[b].radius);
            return ra != rb ? ra > rb : a < b;
        });
        if (big.size() + always.size() > kMaxAlways) big.resize(kMaxAlways > always.size() ? kMaxAlways - always.size() : 0);
        if (big.empty()) break;
        std::vector<uint8_t> out_flag(n, 0);
        for (uint32_t i : big) { out_flag[i] = 1; always.push_back(i); }
        std::vector<uint32_t> rest;
        for (uint32_t i : bounded)
            if (!out_flag[i]) rest.push_back(i);
        bounded.swap(rest);
    }
    if (bounded.empty()) groups.clear();
    std::sort(always.begin(), always.end());
    const uint32_t n_bounded = (uint32_t)groups.size();
    const uint32_t n_always = (uint32_t)((always.size() + kSlots - 1) / kSlots);
    out.n_top = n_bounded + n_always;
    out.n_top_pad = (out.n_top + 7) / 8 * 8;
    out.top.assign((size_t)out.n_top_pad * 4, 0.0f);
    out.mem.assign((size_t)out.n_top * kStride * 4, 0.0f);
    out.order.assign((size_t)out.n_top * kStride, kNoSphere);
    out.count.assign(out.n_top_pad, 0);
    for (uint32_t c = 0; c < out.n_top_pad; ++c) out.top[(size_t)c * 4 + 3] = 3.0e38f;          // never passes
    for (size_t j = 0; j < (size_t)out.n_top * kStride; ++j) out.mem[j * 4 + 3] = 3.0e38f;
    auto put_members = [&](uint32_t c, const uint32_t* ids, uint32_t cnt) {
        out.count[c] = (uint8_t)cnt;
        for (uint32_t u = 0; u < cnt; ++u) {
            const GpuSphere& sp = s[ids[u]];
            float* rec = &out.mem[((size_t)c * kStride + u) * 4];
            rec[0] = sp.position.x; rec[1] = sp.position.y; rec[2] = sp.position.z;
            rec[3] = rtc::broad_record_k(sp);
            out.order[(size_t)c * kStride + u] = (uint16_t)ids[u];
        }
    };
    double extra = 0.0;
    for (uint32_t c = 0; c < n_bounded; ++c) {
        const std::vector<uint32_t>& g = groups[c];
        double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (uint32_t i : g) {
            const double p[3] = {s[i].position.x, s[i].position.y, s[i].position.z};
            for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], p[k]); hi[k] = std::max(hi[k], p[k]); }
        }
        GpuSphere bound;
        memset(&bound, 0, sizeof(bound));
        bound.position.x = (float)(0.5 * (lo[0] + hi[0]));
        bound.position.y = (float)(0.5 * (lo[1] + hi[1]));
        bound.position.z = (float)(0.5 * (lo[2] + hi[2]));
        double rc = 0.0, ac = 0.0;
        for (uint32_t i : g) {   // against the centre as rounded to float
            const double dx = (double)s[i].position.x - bound.position.x, dy = (double)s[i].position.y - bound.position.y,
                         dz = (double)s[i].position.z - bound.position.z, r = std::fabs((double)s[i].radius);
            rc = std::max(rc, std::sqrt(dx * dx + dy * dy + dz * dz) + r);
            const double cx = s[i].position.x, cy = s[i].position.y, cz = s[i].position.z;
            ac = std::max(ac, std::sqrt(cx * cx + cy * cy + cz * cz + r * r));
        }
        bound.radius = round_up(rc * (1.0 + 9.5367431640625e-07) + kMargin * ac + 1e-30);
        extra = std::max(extra, (double)bound.radius);
        float* rec = &out.top[(size_t)c * 4];
        rec[0] = bound.position.x; rec[1] = bound.position.y; rec[2] = bound.position.z;
        rec[3] = rtc::broad_record_k(bound);
        put_members(c, g.data(), (uint32_t)g.size());
    }
    for (uint32_t c = 0; c < n_always; ++c) {
        float* rec = &out.top[(size_t)(n_bounded + c) * 4];
        rec[0] = 0.0f; rec[1] = 0.0f; rec[2] = 0.0f;
        rec[3] = -INFINITY;   // c = -inf, g = +inf: passes every threshold
        const uint32_t first = c * kSlots, cnt = std::min<uint32_t>(kSlots, (uint32_t)always.size() - first);
        put_members(n_bounded + c, always.data() + first, cnt);
    }
    out.extra_r = round_up(extra);
    return true;
}

}  // namespace rtcl
