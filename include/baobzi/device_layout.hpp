// device_layout.hpp -- the device-side ABI of libbaobzi.so: the node record and the by-value tree descriptor that
// every kernel of the library -- and any user kernel built on include/baobzi_device.cuh -- reads.  Plain C++ (no CUDA
// types), shipped under include/baobzi/ so that the public headers are self-contained.  The arrays behind the
// pointers are described in baobzi_b200/csrc/device_tree.hpp (reference: the persistent state of baobzi::Function,
// include/baobzi_template.hpp:539-560, flattened by build_cache 763-778).
#pragma once

#include <cstdint>

namespace bz {

constexpr uint32_t kLeafMark = 0xFFFFFFFFu; // NodeRec::child0 of a leaf
constexpr uint32_t kNotLeaf = 0xFFFFFFFFu;

struct alignas(64) NodeRec {
    double center[3];   // Box::center (template.hpp:93); unused dims are 0
    uint32_t child0;    // GLOBAL node index of child 0 (children are child0 .. child0 + 2^dim - 1), kLeafMark for a leaf
    uint32_t coeff16;   // leaf: offset of the coefficient block in units of 16 bytes (2 doubles)
    double inv_half[3]; // Box::inv_half_length (template.hpp:94)
    uint32_t out_dim;   // Node::output_dim (template.hpp:210)
    uint32_t leaf_rank; // dense leaf number, kNotLeaf for internal nodes
};
static_assert(sizeof(NodeRec) == 64, "NodeRec must be one 64-byte record");

// Passed to kernels by value (lives in the constant bank of the launch).
struct TreeView {
    double lower_left[3], upper_right[3], inv_bin_size[3];
    int n_sub[3];
    int dim, order;
    uint32_t n_bins, n_nodes, n_leaves, out_dim;
    const uint32_t *bin_root;
    const uint32_t *bin_leaf; // [n_bins] dense leaf rank when the cell's root is itself a leaf, else kNotLeaf
    const NodeRec *nodes;
    const double *coeffs;
    const uint32_t *leaf_node;
    const uint32_t *bin_first_leaf; // [n_bins + 1] dense rank of the first leaf of each grid cell's subtree
};

static_assert(sizeof(TreeView) == 160, "TreeView is part of the device ABI (include/baobzi_device.cuh, Baobzi.VIEW_BYTES)");

} // namespace bz
