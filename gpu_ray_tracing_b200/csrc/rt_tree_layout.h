// rt_tree_layout.h — byte layout of the tree image (rt_tree.h) in device memory.  ONE definition, shared by the host
// packer (rtt::Tree::pack), the launch code and the kernels that walk the image, so the offsets cannot drift apart.
//
//   [rec        n_nodes * kStride float4   child records (C.xyz, k), 9-slot stride against bank conflicts
//   |link       n_nodes * kWide   u32      child node, or sphere index under a leaf node; 0xFFFFFFFF = empty
//   |node_r     n_nodes           f32      largest bounding radius among the node's children
//   |entry_rec  n_entry_pad       float4   the entry list's records ...
//   |entry_node n_entry_pad       u32      ... and the nodes they stand for
//   |child_r    n_nodes * kWide   f32      bounding radius per child slot (primary pass)       <- bounce_bytes ends here
//   |entry_slot n_entry_pad       u32]     child slot an entry's record came from (primary pass)
// every part padded to 16 bytes.
#pragma once
#include <stddef.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define RTT_HD __host__ __device__ __forceinline__
#else
#define RTT_HD inline
#endif

namespace rtt {

constexpr int kWide = 8;                // children per node
constexpr int kStride = kWide + 1;      // record slots per node in the image: the odd stride spreads the nodes over the banks

struct ImageLayout {
    size_t off_link, off_r, off_entry_rec, off_entry_node, off_child_r, off_entry_slot, bytes;
    uint32_t n_entry_pad;
};

RTT_HD size_t pad16(size_t b) { return (b + 15) / 16 * 16; }

RTT_HD ImageLayout image_layout(uint32_t n_nodes, uint32_t n_entry) {
    ImageLayout l;
    l.n_entry_pad = (n_entry + 3u) / 4u * 4u;
    l.off_link = (size_t)n_nodes * kStride * 16;
    l.off_r = l.off_link + pad16((size_t)n_nodes * kWide * 4);
    l.off_entry_rec = l.off_r + pad16((size_t)n_nodes * 4);
    l.off_entry_node = l.off_entry_rec + (size_t)l.n_entry_pad * 16;
    l.off_child_r = l.off_entry_node + pad16((size_t)l.n_entry_pad * 4);
    l.off_entry_slot = l.off_child_r + pad16((size_t)n_nodes * kWide * 4);
    l.bytes = l.off_entry_slot + pad16((size_t)l.n_entry_pad * 4);
    return l;
}

}  // namespace rtt
