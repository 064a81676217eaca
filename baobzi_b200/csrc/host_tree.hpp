// host_tree.hpp -- host-side image of a fitted baobzi function.
//
// This is the hand-over point between "how the tree came to be" (restore from a .baobzi file,
// or the host fitter behind baobzi_init) and "how it is evaluated" (device_tree.cu flattens it
// into the device arrays the sm_100a kernels read).
//
// It carries exactly the persistent state of the reference's baobzi::Function
// (include/baobzi_template.hpp:539-560, serialised by MSGPACK_DEFINE_MAP at 1081-1082), with the
// vector-of-subtrees-of-nodes collapsed into flat arrays in the reference's global node order
// (build_cache, template.hpp:763-778: subtree-major, breadth-first inside a subtree).
#pragma once

#include <cstdint>
#include <string>
#include <vector>

namespace bz {

constexpr uint32_t kNoChild = 0xFFFFFFFFu;            // Node::first_child_idx of a leaf (template.hpp:209)
constexpr uint64_t kNoCoeffs = 0xFFFFFFFFFFFFFFFFull; // Node::coeff_offset of an internal node (template.hpp:208)
constexpr int kFormatVersion = 5;                     // include/baobzi/header.h:4
constexpr uint32_t kMaxOutputDim = 1u << 20;          // sanity bound on output_dim_ read from a file

struct HostTree {
    int dim = 0, order = 0;
    uint32_t output_dim = 1;
    bool split_multi_eval = true; // persisted, meaningless on the GPU
    double tol = 0.0;

    // domain box (Function::box_), stored as centre + INVERSE half length like the file does
    double box_center[3] = {0, 0, 0}, box_inv_half_length[3] = {0, 0, 0};
    double lower_left[3] = {0, 0, 0}, upper_right[3] = {0, 0, 0}, inv_bin_size[3] = {0, 0, 0};
    int n_subtrees[3] = {1, 1, 1};

    // subtree k owns nodes [subtree_begin[k], subtree_begin[k+1]); size n_bins + 1
    std::vector<uint32_t> subtree_begin;
    // per node, global order; geometry is dim-strided (node*dim + j)
    std::vector<double> center, inv_half_length;
    std::vector<uint32_t> first_child; // subtree-LOCAL index of child 0, kNoChild for leaves
    std::vector<uint64_t> coeff_offset; // in doubles into coeffs, kNoCoeffs for internal nodes
    std::vector<uint32_t> node_output_dim;
    std::vector<double> coeffs;

    // fit statistics (not persisted by the reference either; zero after restore)
    int base_depth = 0, max_subtree_depth = -1;
    uint64_t n_func_evals = 0;
    uint32_t fit_ms = 0;

    int64_t n_bins() const { return (int64_t)subtree_begin.size() - 1; }
    int64_t n_nodes() const { return (int64_t)first_child.size(); }
    int64_t n_leaves() const;
    int64_t block_doubles() const; // order^dim
    bool is_leaf(int64_t node) const { return coeff_offset[node] != kNoCoeffs; }

    // structural sanity: child ranges inside their subtree, coefficient blocks inside coeffs, ...
    // Returns an empty string when fine, else a description.
    std::string validate() const;
};

inline bool supported_dim_order(int dim, int order) { // reference make_source.sh:11-13
    return dim >= 1 && dim <= 3 && order >= 6 && order <= 16 && order % 2 == 0;
}

// .baobzi v5.  read: false + message on any malformed input / wrong version.
bool read_baobzi_file(const char *path, HostTree &out, std::string &err);
bool read_baobzi_mem(const uint8_t *data, size_t len, HostTree &out, std::string &err);
bool write_baobzi_file(const char *path, const HostTree &t, std::string &err);
void write_baobzi_mem(const HostTree &t, std::vector<uint8_t> &out);

} // namespace bz
