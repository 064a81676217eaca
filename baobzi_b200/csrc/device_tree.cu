// device_tree.cu -- flatten a HostTree into the device arrays of device_tree.hpp and upload them.
//
// Replaces, for the GPU, the reference's build_cache (include/baobzi_template.hpp:763-778): the
// vector<FunctionTree> / vector<Node> / node_pointers_ indirections become one array of 64-byte node
// records addressed by the reference's own global node index, with child links rewritten from
// subtree-local (template.hpp:431, 489) to global so a descent step is a single dependent load.
// The stored centre / inverse-half-length doubles are copied verbatim (never recomputed) so that the
// strict `x > centre` comparisons and the affine map give the reference's bits.
#include "device_tree.hpp"
#include "host_tree.hpp"

#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <vector>

namespace bz {

namespace {
size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

constexpr size_t kGuardBytes = 4096;
constexpr unsigned char kGuardByte = 0xA5;
struct GuardedBuffer {
    char *base;
    size_t bytes; // payload size, rounded up to 256
    int device;
};
std::mutex g_guard_mu;
std::map<void *, GuardedBuffer> g_guarded; // payload pointer -> allocation
} // namespace

cudaError_t bz_malloc(void **p, size_t bytes) {
    static const bool guard = std::getenv("BAOBZI_B200_GUARD") && std::atoi(std::getenv("BAOBZI_B200_GUARD")) != 0;
    if (!guard)
        return cudaMalloc(p, bytes);
    const size_t pay = align_up(bytes ? bytes : 1, 256);
    char *base = nullptr;
    cudaError_t e = cudaMalloc((void **)&base, pay + 2 * kGuardBytes);
    if (e != cudaSuccess)
        return e;
    cudaMemset(base, kGuardByte, kGuardBytes);
    cudaMemset(base + kGuardBytes + bytes, kGuardByte, pay - bytes + kGuardBytes); // the slack behind the payload counts as guard
    e = cudaDeviceSynchronize();
    int dev = 0;
    cudaGetDevice(&dev);
    *p = base + kGuardBytes;
    std::lock_guard<std::mutex> lock(g_guard_mu);
    g_guarded[*p] = GuardedBuffer{base, bytes, dev};
    return e;
}

cudaError_t bz_free(void *p) {
    if (!p)
        return cudaSuccess;
    {
        std::lock_guard<std::mutex> lock(g_guard_mu);
        auto it = g_guarded.find(p);
        if (it != g_guarded.end()) {
            char *base = it->second.base;
            g_guarded.erase(it);
            return cudaFree(base);
        }
    }
    return cudaFree(p);
}

int guard_check(int64_t *bad_bytes, int64_t *buffers) {
    std::lock_guard<std::mutex> lock(g_guard_mu);
    int64_t bad = 0;
    int prev = 0;
    cudaGetDevice(&prev);
    cudaError_t e = cudaSuccess;
    std::vector<unsigned char> host;
    for (auto &kv : g_guarded) {
        const GuardedBuffer &g = kv.second;
        cudaSetDevice(g.device);
        e = cudaDeviceSynchronize();
        if (e != cudaSuccess)
            break;
        const size_t pay = align_up(g.bytes ? g.bytes : 1, 256), tail = pay - g.bytes + kGuardBytes;
        host.resize(kGuardBytes + tail);
        e = cudaMemcpy(host.data(), g.base, kGuardBytes, cudaMemcpyDeviceToHost);
        if (e == cudaSuccess)
            e = cudaMemcpy(host.data() + kGuardBytes, g.base + kGuardBytes + g.bytes, tail, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess)
            break;
        for (unsigned char c : host)
            bad += c != kGuardByte;
    }
    cudaSetDevice(prev);
    if (bad_bytes)
        *bad_bytes = bad;
    if (buffers)
        *buffers = (int64_t)g_guarded.size();
    return (int)e;
}

int upload_tree(const HostTree &h, DeviceTree &out) {
    const int dim = h.dim;
    const size_t nn = (size_t)h.n_nodes(), nb = (size_t)h.n_bins();

    // coefficient store: keep the reference's offsets when every block starts on a 16-byte boundary
    // (always the case for trees the reference writes: blocks are multiples of an even ORDER); otherwise
    // re-pack leaf by leaf.
    // ... and ascend with the node order (true for files the reference writes: subtrees are fitted cell by cell), so
    // that neighbouring leaves own neighbouring blocks -- the locality the bucketed path relies on
    bool aligned = true;
    uint64_t prev_off = 0;
    for (size_t g = 0; g < nn; ++g)
        if (h.is_leaf((int64_t)g)) {
            if ((h.coeff_offset[g] & 1) || h.coeff_offset[g] < prev_off)
                aligned = false;
            prev_off = h.coeff_offset[g];
        }
    std::vector<double> packed;
    std::vector<uint64_t> new_off;
    const std::vector<double> *coeffs = &h.coeffs;
    if (!aligned) {
        const size_t blk = (size_t)h.block_doubles();
        new_off.assign(nn, kNoCoeffs);
        for (size_t g = 0; g < nn; ++g) {
            if (!h.is_leaf((int64_t)g))
                continue;
            size_t len = blk;
            if (dim == 1)
                len = (size_t)h.order * (h.node_output_dim[g] ? h.node_output_dim[g] : 1);
            if (h.coeff_offset[g] + len > h.coeffs.size())
                len = h.coeffs.size() - h.coeff_offset[g];
            new_off[g] = packed.size();
            packed.insert(packed.end(), h.coeffs.begin() + h.coeff_offset[g], h.coeffs.begin() + h.coeff_offset[g] + len);
            if (packed.size() & 1)
                packed.push_back(0.0);
        }
        coeffs = &packed;
    }
    if (coeffs->size() / 2 >= 0xFFFFFFFFull)
        return (int)cudaErrorInvalidValue; // > 64 GB of coefficients: not addressable by NodeRec::coeff16

    std::vector<NodeRec> recs(nn);
    std::vector<uint32_t> leaf_node;
    std::vector<uint32_t> bin_root(nb), bin_leaf(nb, kNotLeaf), bin_first_leaf(nb + 1, 0);
    for (size_t s = 0; s < nb; ++s) {
        const uint32_t b = h.subtree_begin[s], e = h.subtree_begin[s + 1];
        bin_root[s] = b;
        bin_first_leaf[s] = (uint32_t)leaf_node.size();
        for (uint32_t g = b; g < e; ++g) {
            NodeRec &r = recs[g];
            std::memset(&r, 0, sizeof r);
            for (int j = 0; j < dim; ++j) {
                r.center[j] = h.center[(size_t)g * dim + j];
                r.inv_half[j] = h.inv_half_length[(size_t)g * dim + j];
            }
            r.out_dim = h.node_output_dim[g];
            if (h.is_leaf(g)) {
                r.child0 = kLeafMark;
                const uint64_t off = aligned ? h.coeff_offset[g] : new_off[g];
                r.coeff16 = (uint32_t)(off / 2);
                r.leaf_rank = (uint32_t)leaf_node.size();
                if (g == b)
                    bin_leaf[s] = r.leaf_rank; // undivided grid cell: the lookup ends at the bin table
                leaf_node.push_back(g);
            } else {
                r.child0 = b + h.first_child[g]; // subtree-local -> global
                r.coeff16 = 0;
                r.leaf_rank = kNotLeaf;
            }
        }
    }

    bin_first_leaf[nb] = (uint32_t)leaf_node.size();
    // one allocation, 256-byte aligned sections: nodes | bin_root | bin_leaf | leaf_node | coeffs (+ 32 KB of zero
    // slack so that vector loads near the end of the last block never leave the allocation)
    const size_t off_nodes = 0;
    const size_t off_bins = align_up(off_nodes + nn * sizeof(NodeRec), 256);
    const size_t off_binl = align_up(off_bins + nb * sizeof(uint32_t), 256);
    const size_t off_binf = align_up(off_binl + nb * sizeof(uint32_t), 256);
    const size_t off_leaf = align_up(off_binf + (nb + 1) * sizeof(uint32_t), 256);
    const size_t off_coef = align_up(off_leaf + leaf_node.size() * sizeof(uint32_t), 256);
    const size_t total = align_up(off_coef + coeffs->size() * sizeof(double) + 32768, 256);

    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess)
        return (int)e;
    void *blob = nullptr;
    e = cudaMalloc(&blob, total);
    if (e != cudaSuccess)
        return (int)e;
    char *base = static_cast<char *>(blob);
    e = cudaMemset(blob, 0, total);
    if (e == cudaSuccess)
        e = cudaMemcpy(base + off_nodes, recs.data(), nn * sizeof(NodeRec), cudaMemcpyHostToDevice);
    if (e == cudaSuccess)
        e = cudaMemcpy(base + off_bins, bin_root.data(), nb * sizeof(uint32_t), cudaMemcpyHostToDevice);
    if (e == cudaSuccess)
        e = cudaMemcpy(base + off_binl, bin_leaf.data(), nb * sizeof(uint32_t), cudaMemcpyHostToDevice);
    if (e == cudaSuccess)
        e = cudaMemcpy(base + off_binf, bin_first_leaf.data(), (nb + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice);
    if (e == cudaSuccess && !leaf_node.empty())
        e = cudaMemcpy(base + off_leaf, leaf_node.data(), leaf_node.size() * sizeof(uint32_t), cudaMemcpyHostToDevice);
    if (e == cudaSuccess && !coeffs->empty())
        e = cudaMemcpy(base + off_coef, coeffs->data(), coeffs->size() * sizeof(double), cudaMemcpyHostToDevice);
    // the copies above go through the legacy default stream (a pageable-source cudaMemcpy may return before its DMA
    // has landed); the kernels that read the tree run on arbitrary, possibly non-blocking streams
    if (e == cudaSuccess)
        e = cudaStreamSynchronize(nullptr);
    if (e != cudaSuccess) {
        cudaFree(blob);
        return (int)e;
    }

    out = DeviceTree();
    out.device = dev;
    out.blob = blob;
    out.blob_bytes = total;
    cudaDeviceGetAttribute(&out.sm_count, cudaDevAttrMultiProcessorCount, dev);
    TreeView &v = out.view;
    for (int j = 0; j < 3; ++j) {
        v.lower_left[j] = j < dim ? h.lower_left[j] : 0.0;
        v.upper_right[j] = j < dim ? h.upper_right[j] : 0.0;
        v.inv_bin_size[j] = j < dim ? h.inv_bin_size[j] : 0.0;
        v.n_sub[j] = j < dim ? h.n_subtrees[j] : 1;
    }
    v.dim = dim;
    v.order = h.order;
    v.n_bins = (uint32_t)nb;
    v.n_nodes = (uint32_t)nn;
    v.n_leaves = (uint32_t)leaf_node.size();
    v.out_dim = h.output_dim;
    v.nodes = reinterpret_cast<const NodeRec *>(base + off_nodes);
    v.bin_root = reinterpret_cast<const uint32_t *>(base + off_bins);
    v.bin_leaf = reinterpret_cast<const uint32_t *>(base + off_binl);
    v.bin_first_leaf = reinterpret_cast<const uint32_t *>(base + off_binf);
    v.leaf_node = reinterpret_cast<const uint32_t *>(base + off_leaf);
    v.coeffs = reinterpret_cast<const double *>(base + off_coef);
    if (dim == 1 && leaf_node.size() <= 16) { // few-leaf 1D kernel: coefficients travel as launch parameters
        out.leaf_coef_1d.resize(leaf_node.size() * (size_t)h.order);
        for (size_t l = 0; l < leaf_node.size(); ++l)
            for (int k = 0; k < h.order; ++k)
                out.leaf_coef_1d[l * h.order + k] = (*coeffs)[(size_t)recs[leaf_node[l]].coeff16 * 2 + k];
    }
    return 0;
}

void free_tree(DeviceTree &dt) {
    if (dt.blob) {
        int cur = 0;
        cudaGetDevice(&cur);
        if (cur != dt.device)
            cudaSetDevice(dt.device);
        cudaFree(dt.blob);
        if (cur != dt.device)
            cudaSetDevice(cur);
    }
    dt = DeviceTree();
}

} // namespace bz
