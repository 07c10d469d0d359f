// rt_tree.h — an 8-wide tree of bounding spheres over the sphere list, for incoherent (bounce) rays.  Host side,
// plain C++; shared by rt_api.cu (built lazily after rt_set_scene) and the CPU emulation in tests/emu.
//
// sphere_list_hit (compute_shader.wgsl:164-180) tests every sphere for every ray.  Without NaNs its result is the
// lexicographic minimum of (first valid root, index) over the spheres, so the order of
// the exact tests is free as long as ties go to the lower index — and spheres the ray's line cannot reach need no test
// at all.  The spheres are split at medians (k-d fashion) into leaves of up to 8; 8 leaves form a node, 8 nodes the
// next level, and so on.  Every child of a node is represented by a bounding sphere whose record has the SAME form as a
// sphere's own broad-phase record (rtc::broad_record_k), so the kernels test "does the ray's line come near this
// subtree" with the 7-FFMA test they already have and descend only then.  The children of a leaf node are the
// spheres themselves, with their ordinary records.
//
// Why the bound is safe.  An exact hit of sphere i (the reference's float discriminant >= 0 or NaN) implies, in real
// numbers, dist(line, C_i)^2 <= r_i^2 + kappa S_i, S_i = |C_i|^2 + r_i^2 + |O|^2 (kappa = 2^-16
// absorbs every rounding of the reference's own discriminant).  Hence for a group with bounding centre Cc
//     dist(line, Cc) <= |C_i - Cc| + r_i + 2^-8 (sqrt(|C_i|^2 + r_i^2) + |O|)
// and the group's radius is  Rc' = max_i(|C_i - Cc| + r_i) + 2^-7 A  (A = max_i sqrt(|C_i|^2 + r_i^2)), with the ray's
// own share m = 2^-7 |O| added at run time (rtc::cluster_thr: the threshold drops by 2 R m + m^2, R the largest Rc'
// among the node's children).  Spheres that would spoil this — non-finite or beyond 2^30, or much larger than a leaf
// (a ground sphere) — are kept out of the tree proper: they hang under "always" records that pass for every ray, and
// are then tested by their own records exactly as before.
#pragma once
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "rt_core.cuh"
#include "rt_tree_layout.h"

namespace rtt {

constexpr uint32_t kMaxEntry = 160;     // longest entry list (coherent scan) before the per-lane traversal takes over
constexpr uint32_t kMaxAlways = 64;     // at most this many oversized spheres are pulled out of the tree
constexpr double kMargin = 0.0078125;   // 2^-7
constexpr uint32_t kNone = 0xFFFFFFFFu;
constexpr float kAlways = -3.0e38f;     // record constant of a child that every ray descends into
constexpr float kNever = 3.0e38f;       // ... and of an empty slot

struct Tree {
    uint32_t n_nodes = 0;
    uint32_t first_leaf = 0;       // nodes [first_leaf, n_nodes) have spheres as children, the others have nodes
    uint32_t depth = 0;            // levels, root included
    std::vector<float> rec;        // n_nodes * kStride x (C.xyz, k)
    std::vector<uint32_t> link;    // n_nodes * kWide: child node, or sphere index under a leaf node; kNone = empty
    std::vector<float> child_r;    // n_nodes * kWide: bounding radius Rc' of an internal node's child (0 elsewhere); the
                                   // primary pass derives its per-frame records of the groups from it
    std::vector<float> node_r;     // n_nodes: largest bounding radius among the node's children (0 for leaf nodes)
    // Entry list: the nodes of the deepest level that still has at most kMaxEntry nodes (plus leaves above it), each
    // with the record its parent holds for it.  A warp scans this list coherently, all lanes on the same record, and
    // only the traversal below it is per lane.
    uint32_t n_entry = 0, n_entry_pad = 0;   // padded to a multiple of 4 with never-passing records
    float entry_r = 0.0f;                    // largest bounding radius among the entry records
    std::vector<float> entry_rec;            // n_entry_pad x (C.xyz, k)
    std::vector<uint32_t> entry_node;        // n_entry_pad
    std::vector<uint32_t> entry_slot;        // n_entry_pad: the child slot (node * kStride + k) the entry's record came from

    // Device image: [rec | link | node_r | entry_rec | entry_node | child_r | entry_slot], every part padded to 16 bytes.  The
    // bounce kernel copies the image up to child_r (bounce_bytes) into shared memory.
    // offsets: rt_tree_layout.h, the one definition the kernels use as well
    ImageLayout layout() const { return image_layout(n_nodes, n_entry); }
    size_t off_link() const { return layout().off_link; }
    size_t off_r() const { return layout().off_r; }
    size_t off_entry_rec() const { return layout().off_entry_rec; }
    size_t off_entry_node() const { return layout().off_entry_node; }
    size_t off_child_r() const { return layout().off_child_r; }
    size_t bounce_bytes() const { return layout().off_child_r; }
    size_t off_entry_slot() const { return layout().off_entry_slot; }
    size_t bytes() const { return layout().bytes; }
    void pack(unsigned char* blob) const {
        memset(blob, 0, bytes());
        memcpy(blob, rec.data(), rec.size() * 4);
        memcpy(blob + off_link(), link.data(), link.size() * 4);
        memcpy(blob + off_r(), node_r.data(), node_r.size() * 4);
        memcpy(blob + off_entry_rec(), entry_rec.data(), entry_rec.size() * 4);
        memcpy(blob + off_entry_node(), entry_node.data(), entry_node.size() * 4);
        memcpy(blob + off_child_r(), child_r.data(), child_r.size() * 4);
        memcpy(blob + off_entry_slot(), entry_slot.data(), entry_slot.size() * 4);
    }
};

inline float round_up(double v) {
    float f = (float)v;
    if ((double)f < v) f = nextafterf(f, INFINITY);
    return f;
}
inline double coord(const GpuSphere& g, int axis) { return axis == 0 ? g.position.x : (axis == 1 ? g.position.y : g.position.z); }

// Bounding record of the spheres ids[0, cnt): centre of the centres' bounding box, radius as derived above.
inline std::array<float, 4> bound_of(const GpuSphere* s, const uint32_t* ids, size_t cnt, float* radius) {
    double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (size_t j = 0; j < cnt; ++j)
        for (int k = 0; k < 3; ++k) {
            lo[k] = std::min(lo[k], coord(s[ids[j]], k));
            hi[k] = std::max(hi[k], coord(s[ids[j]], k));
        }
    GpuSphere b;
    memset(&b, 0, sizeof(b));
    b.position.x = (float)(0.5 * (lo[0] + hi[0]));
    b.position.y = (float)(0.5 * (lo[1] + hi[1]));
    b.position.z = (float)(0.5 * (lo[2] + hi[2]));
    double rc = 0.0, ac = 0.0;
    for (size_t j = 0; j < cnt; ++j) {   // against the centre as rounded to float
        const GpuSphere& g = s[ids[j]];
        const double dx = (double)g.position.x - b.position.x, dy = (double)g.position.y - b.position.y,
                     dz = (double)g.position.z - b.position.z, r = std::fabs((double)g.radius);
        rc = std::max(rc, std::sqrt(dx * dx + dy * dy + dz * dz) + r);
        const double cx = g.position.x, cy = g.position.y, cz = g.position.z;
        ac = std::max(ac, std::sqrt(cx * cx + cy * cy + cz * cz + r * r));
    }
    b.radius = round_up(rc * (1.0 + 9.5367431640625e-07) + kMargin * ac + 1e-30);
    *radius = b.radius;
    return {b.position.x, b.position.y, b.position.z, rtc::broad_record_k(b)};
}

struct Builder {
    struct Node {
        bool leaf = false;
        std::vector<uint32_t> spheres;                // leaf: members, in index order
        std::vector<int> kids;                        // internal: child nodes (indices into `nodes`)
        std::vector<std::array<float, 4>> kid_rec;    // internal: one record per child
        std::vector<float> kid_r;                     // internal: bounding radius per child (0 under always records)
        float rmax = 0.0f;
    };
    const GpuSphere* s;
    std::vector<uint32_t> ids;
    std::vector<Node> nodes;

    // Median splits of ids[lo, hi) along the longest axis into `parts` runs of `cap` spheres (the last one shorter).
    void split(size_t lo, size_t hi, size_t parts, size_t cap, std::vector<std::pair<size_t, size_t>>& out) {
        if (parts <= 1) {
            if (hi > lo) out.push_back({lo, hi});
            return;
        }
        double blo[3] = {INFINITY, INFINITY, INFINITY}, bhi[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (size_t j = lo; j < hi; ++j)
            for (int k = 0; k < 3; ++k) {
                blo[k] = std::min(blo[k], coord(s[ids[j]], k));
                bhi[k] = std::max(bhi[k], coord(s[ids[j]], k));
            }
        int axis = 0;
        for (int k = 1; k < 3; ++k)
            if (bhi[k] - blo[k] > bhi[axis] - blo[axis]) axis = k;
        const size_t left_parts = (parts + 1) / 2, left = std::min(hi - lo, left_parts * cap);
        const GpuSphere* sp = s;
        std::nth_element(ids.begin() + lo, ids.begin() + lo + left, ids.begin() + hi, [sp, axis](uint32_t a, uint32_t b) {
            const double ca = coord(sp[a], axis), cb = coord(sp[b], axis);
            return ca != cb ? ca < cb : a < b;
        });
        split(lo, lo + left, left_parts, cap, out);
        split(lo + left, hi, parts - left_parts, cap, out);
    }
    int make_leaf(const uint32_t* first, size_t cnt) {
        Node nd;
        nd.leaf = true;
        nd.spheres.assign(first, first + cnt);
        std::sort(nd.spheres.begin(), nd.spheres.end());
        nodes.push_back(std::move(nd));
        return (int)nodes.size() - 1;
    }
    // Subtree over ids[lo, hi); every child holds `cap` spheres, the smallest power of 8 with 8 * cap >= the count,
    // so the lower levels are full.
    int make_subtree(size_t lo, size_t hi) {
        const size_t cnt = hi - lo;
        if (cnt <= (size_t)kWide) return make_leaf(ids.data() + lo, cnt);
        size_t cap = kWide;
        while (cap * kWide < cnt) cap *= kWide;
        std::vector<std::pair<size_t, size_t>> parts;
        split(lo, hi, (cnt + cap - 1) / cap, cap, parts);
        Node nd;
        for (auto& pt : parts) {
            float r;
            nd.kid_rec.push_back(bound_of(s, ids.data() + pt.first, pt.second - pt.first, &r));
            nd.kid_r.push_back(r);
            nd.rmax = std::max(nd.rmax, r);
        }
        for (auto& pt : parts) nd.kids.push_back(make_subtree(pt.first, pt.second));   // (may reallocate `nodes`)
        nodes.push_back(std::move(nd));
        return (int)nodes.size() - 1;
    }
    int make_always(const std::vector<uint32_t>& always) {
        if (always.size() <= (size_t)kWide) return make_leaf(always.data(), always.size());
        Node nd;
        for (size_t lo = 0; lo < always.size(); lo += kWide) {
            nd.kids.push_back(make_leaf(always.data() + lo, std::min(always.size() - lo, (size_t)kWide)));
            nd.kid_rec.push_back({0.0f, 0.0f, 0.0f, kAlways});
            nd.kid_r.push_back(0.0f);
        }
        nodes.push_back(std::move(nd));
        return (int)nodes.size() - 1;
    }
    uint32_t depth_of(int n) const {
        uint32_t d = 0;
        for (int k : nodes[n].kids) d = std::max(d, depth_of(k));
        return d + 1;
    }
};

// Radius of a typical leaf of a plain cut of `ids` into runs of 8, to spot oversized spheres.
inline double typical_leaf_radius(const GpuSphere* s, const std::vector<uint32_t>& ids) {
    Builder b{s, ids, {}};
    std::vector<std::pair<size_t, size_t>> leaves;
    b.split(0, ids.size(), (ids.size() + kWide - 1) / kWide, kWide, leaves);
    std::vector<double> radii;
    for (auto& lf : leaves) {
        float r;
        bound_of(s, b.ids.data() + lf.first, lf.second - lf.first, &r);
        radii.push_back(r);
    }
    if (radii.empty()) return INFINITY;
    std::nth_element(radii.begin(), radii.begin() + radii.size() / 2, radii.end());
    return radii[radii.size() / 2];
}

// Builds the tree over spheres[0, n).  Returns false (out unusable) for an empty scene.
inline bool build(const GpuSphere* s, uint32_t n, Tree& out) {
    out = Tree();
    if (n == 0) return false;
    const double lim = 1073741824.0;   // 2^30, the magnitude window of the bound
    std::vector<uint32_t> bounded, always;
    for (uint32_t i = 0; i < n; ++i) {
        const bool in_window = std::fabs((double)s[i].position.x) <= lim && std::fabs((double)s[i].position.y) <= lim &&
                               std::fabs((double)s[i].position.z) <= lim && std::fabs((double)s[i].radius) <= lim;
        if (in_window) bounded.push_back(i);   // false for NaN
        else always.push_back(i);
    }
    if (!bounded.empty()) {
        // spheres larger than a typical leaf would inflate every group above them: pull the largest out
        const double typical = typical_leaf_radius(s, bounded);
        std::vector<uint32_t> big;
        for (uint32_t i : bounded)
            if (std::fabs((double)s[i].radius) > typical) big.push_back(i);
        std::sort(big.begin(), big.end(), [&](uint32_t a, uint32_t b) {
            const double ra = std::fabs((double)s[a].radius), rb = std::fabs((double)s[b].radius);
            return ra != rb ? ra > rb : a < b;
        });
        if (big.size() + always.size() > kMaxAlways) big.resize(kMaxAlways > always.size() ? kMaxAlways - always.size() : 0);
        if (!big.empty()) {
            std::vector<uint8_t> flag(n, 0);
            for (uint32_t i : big) { flag[i] = 1; always.push_back(i); }
            std::vector<uint32_t> rest;
            for (uint32_t i : bounded)
                if (!flag[i]) rest.push_back(i);
            bounded.swap(rest);
        }
    }
    std::sort(always.begin(), always.end());

    Builder b{s, bounded, {}};
    Builder::Node root;   // its (at most two) children are descended into by every ray
    if (!bounded.empty()) {
        root.kids.push_back(b.make_subtree(0, bounded.size()));
        root.kid_rec.push_back({0.0f, 0.0f, 0.0f, kAlways});
        root.kid_r.push_back(0.0f);
    }
    if (!always.empty()) {
        root.kids.push_back(b.make_always(always));
        root.kid_rec.push_back({0.0f, 0.0f, 0.0f, kAlways});
        root.kid_r.push_back(0.0f);
    }
    b.nodes.push_back(std::move(root));
    const int root_i = (int)b.nodes.size() - 1;

    // number the nodes: internal ones breadth first from the root (node 0), then the leaves
    std::vector<int> order, number(b.nodes.size(), -1);
    order.push_back(root_i);
    for (size_t q = 0; q < order.size(); ++q)
        for (int k : b.nodes[order[q]].kids)
            if (!b.nodes[k].leaf) order.push_back(k);
    const uint32_t n_internal = (uint32_t)order.size();
    for (size_t q = 0; q < n_internal; ++q)
        for (int k : b.nodes[order[q]].kids)
            if (b.nodes[k].leaf) order.push_back(k);
    for (size_t q = 0; q < order.size(); ++q) number[order[q]] = (int)q;
    out.n_nodes = (uint32_t)order.size();
    out.first_leaf = n_internal;
    out.depth = b.depth_of(root_i);
    out.rec.assign((size_t)out.n_nodes * kStride * 4, 0.0f);
    for (size_t j = 0; j < (size_t)out.n_nodes * kStride; ++j) out.rec[j * 4 + 3] = kNever;
    out.link.assign((size_t)out.n_nodes * kWide, kNone);
    out.node_r.assign(out.n_nodes, 0.0f);
    out.child_r.assign((size_t)out.n_nodes * kWide, 0.0f);
    for (size_t q = 0; q < order.size(); ++q) {
        const Builder::Node& nd = b.nodes[order[q]];
        float* rec = &out.rec[q * kStride * 4];
        if (nd.leaf) {
            for (size_t k = 0; k < nd.spheres.size(); ++k) {
                const GpuSphere& sp = s[nd.spheres[k]];
                rec[k * 4 + 0] = sp.position.x; rec[k * 4 + 1] = sp.position.y; rec[k * 4 + 2] = sp.position.z;
                rec[k * 4 + 3] = rtc::broad_record_k(sp);
                out.link[q * kWide + k] = nd.spheres[k];
            }
        } else {
            for (size_t k = 0; k < nd.kids.size(); ++k) {
                memcpy(rec + k * 4, nd.kid_rec[k].data(), 16);
                out.link[q * kWide + k] = (uint32_t)number[nd.kids[k]];
                out.child_r[q * kWide + k] = nd.kid_r[k];
            }
            out.node_r[q] = nd.rmax;
        }
    }
    // entry list: descend level by level while the frontier stays short
    struct Entry { uint32_t node; std::array<float, 4> rec; float r; uint32_t slot; };
    std::vector<Entry> frontier;
    frontier.push_back({0u, {0.0f, 0.0f, 0.0f, kAlways}, 0.0f, kNone});
    for (;;) {
        std::vector<Entry> next;
        bool any_internal = false;
        for (const Entry& e : frontier) {
            if (e.node >= out.first_leaf) {
                next.push_back(e);
                continue;
            }
            any_internal = true;
            for (int k = 0; k < kWide; ++k) {
                const uint32_t child = out.link[(size_t)e.node * kWide + k];
                if (child == kNone) continue;
                Entry c;
                c.node = child;
                memcpy(c.rec.data(), &out.rec[((size_t)e.node * kStride + k) * 4], 16);
                c.r = c.rec[3] == kAlways ? 0.0f : out.node_r[e.node];
                c.slot = e.node * kStride + (uint32_t)k;
                next.push_back(c);
            }
        }
        if (!any_internal || next.size() > kMaxEntry) break;
        frontier.swap(next);
    }
    out.n_entry = (uint32_t)frontier.size();
    out.n_entry_pad = (out.n_entry + 3) / 4 * 4;
    out.entry_rec.assign((size_t)out.n_entry_pad * 4, 0.0f);
    out.entry_node.assign(out.n_entry_pad, kNone);
    out.entry_slot.assign(out.n_entry_pad, kNone);
    for (uint32_t j = 0; j < out.n_entry_pad; ++j) out.entry_rec[(size_t)j * 4 + 3] = kNever;
    for (uint32_t j = 0; j < out.n_entry; ++j) {
        memcpy(&out.entry_rec[(size_t)j * 4], frontier[j].rec.data(), 16);
        out.entry_node[j] = frontier[j].node;
        out.entry_slot[j] = frontier[j].slot;
        out.entry_r = std::max(out.entry_r, frontier[j].r);
    }
    return true;
}

}  // namespace rtt
