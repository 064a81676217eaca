// msgpack_io.cpp -- from-scratch reader/writer for the version-5 .baobzi container.
//
// A .baobzi file is two concatenated msgpack objects (reference Function::save,
// include/baobzi_template.hpp:914-919):
//   1. [dim, order, version]                               (include/baobzi/header.h:10-18)
//   2. a 10-key map, keys in this order (template.hpp:1081-1082):
//        box_, subtrees_, n_subtrees_, tol_, lower_left_, upper_right_, inv_bin_size_, coeffs_,
//        split_multi_eval_, output_dim_
//      Box          = [center, inv_half_length]                        (template.hpp:105)
//      FunctionTree = [nodes_]                                         (template.hpp:519)
//      Node         = [box_, first_child_idx, coeff_offset, output_dim] (template.hpp:382)
//      Eigen vector = ["__eigen__", rows, cols, v0, v1, ...] column-major
//                                                     (include/baobzi/eigen_matrix_plugin.h:30-42)
// The writer reproduces msgpack-c 3.3's encoding choices (smallest integer form, float64 for every
// double, fix/16/32 containers) so that a file read here and written back is byte-identical and the
// unmodified reference restores our files.
#include "host_tree.hpp"

#include <cstdio>
#include <cstring>

namespace bz {

// ------------------------------------------------------------------------------------------ reader
namespace {

class Reader {
  public:
    Reader(const uint8_t *d, size_t n) : p_(d), end_(d + n) {}
    bool ok() const { return !bad_; }
    void fail() { bad_ = true; }
    // bytes left: every msgpack object takes at least one, so a declared length above this is a lie and must not
    // reach resize() (a 4-byte length field could otherwise ask for 32 GB before the first payload byte is seen)
    size_t remaining() const { return (size_t)(end_ - p_); }
    uint64_t bounded_len(uint64_t n) {
        if (n > remaining()) {
            bad_ = true;
            return 0;
        }
        return n;
    }

    uint64_t be(int nbytes) {
        if ((size_t)(end_ - p_) < (size_t)nbytes) {
            bad_ = true;
            p_ = end_;
            return 0;
        }
        uint64_t v = 0;
        while (nbytes--)
            v = (v << 8) | *p_++;
        return v;
    }
    uint8_t tag() {
        if (p_ >= end_) {
            bad_ = true;
            return 0xc1;
        }
        return *p_++;
    }
    uint64_t array_len() {
        uint8_t t = tag();
        if ((t & 0xf0) == 0x90)
            return t & 0x0f;
        if (t == 0xdc)
            return be(2);
        if (t == 0xdd)
            return be(4);
        bad_ = true;
        return 0;
    }
    uint64_t map_len() {
        uint8_t t = tag();
        if ((t & 0xf0) == 0x80)
            return t & 0x0f;
        if (t == 0xde)
            return be(2);
        if (t == 0xdf)
            return be(4);
        bad_ = true;
        return 0;
    }
    std::string str() {
        uint8_t t = tag();
        uint64_t n = 0;
        if ((t & 0xe0) == 0xa0)
            n = t & 0x1f;
        else if (t == 0xd9)
            n = be(1);
        else if (t == 0xda)
            n = be(2);
        else if (t == 0xdb)
            n = be(4);
        else
            bad_ = true;
        if (bad_ || (uint64_t)(end_ - p_) < n) {
            bad_ = true;
            return std::string();
        }
        std::string s((const char *)p_, (size_t)n);
        p_ += n;
        return s;
    }
    // integers (any width / signedness) -- returned as the two's complement 64-bit pattern
    uint64_t integer() {
        uint8_t t = tag();
        if (t <= 0x7f)
            return t;
        if (t >= 0xe0)
            return (uint64_t)(int64_t)(int8_t)t;
        switch (t) {
        case 0xcc: return be(1);
        case 0xcd: return be(2);
        case 0xce: return be(4);
        case 0xcf: return be(8);
        case 0xd0: return (uint64_t)(int64_t)(int8_t)be(1);
        case 0xd1: return (uint64_t)(int64_t)(int16_t)be(2);
        case 0xd2: return (uint64_t)(int64_t)(int32_t)be(4);
        case 0xd3: return be(8);
        case 0xc2: return 0;
        case 0xc3: return 1;
        default: bad_ = true; return 0;
        }
    }
    // anything msgpack-c would convert to double
    double real() {
        if (p_ >= end_) {
            bad_ = true;
            return 0;
        }
        uint8_t t = *p_;
        if (t == 0xcb) {
            ++p_;
            uint64_t b = be(8);
            double d;
            std::memcpy(&d, &b, 8);
            return d;
        }
        if (t == 0xca) {
            ++p_;
            uint32_t b = (uint32_t)be(4);
            float f;
            std::memcpy(&f, &b, 4);
            return f;
        }
        if (t >= 0xd0 && t <= 0xd3)
            return (double)(int64_t)integer();
        if (t >= 0xe0)
            return (double)(int64_t)integer();
        return (double)integer();
    }
    bool boolean() {
        uint8_t t = tag();
        if (t == 0xc3)
            return true;
        if (t == 0xc2)
            return false;
        bad_ = true;
        return false;
    }
    void eigen(double *out, int n) {
        uint64_t len = array_len();
        std::string magic = str();
        uint64_t rows = integer(), cols = integer();
        if (magic != "__eigen__" || rows * cols != (uint64_t)n || len != (uint64_t)n + 3) {
            bad_ = true;
            return;
        }
        for (int i = 0; i < n && !bad_; ++i)
            out[i] = real();
    }
    void box(double *c, double *ihl, int dim) {
        if (array_len() != 2)
            bad_ = true;
        eigen(c, dim);
        eigen(ihl, dim);
    }

  private:
    const uint8_t *p_, *end_;
    bool bad_ = false;
};

} // namespace

bool read_baobzi_mem(const uint8_t *data, size_t len, HostTree &t, std::string &err) {
    Reader r(data, len);
    if (r.array_len() != 3) {
        err = "not a baobzi file (bad header)";
        return false;
    }
    int dim = (int)r.integer(), order = (int)r.integer(), version = (int)r.integer();
    if (!r.ok()) {
        err = "not a baobzi file (truncated header)";
        return false;
    }
    if (version != kFormatVersion) { // reference baobzi.cpp:94-97 (returns NULL silently)
        err = "unsupported .baobzi version " + std::to_string(version);
        return false;
    }
    if (!supported_dim_order(dim, order)) { // reference baobzi.cpp:102-108
        err = "Unable to initialize Baobzi function with variables (DIM, ORDER): (" + std::to_string(dim) + ", " +
              std::to_string(order) + ")";
        return false;
    }
    t = HostTree();
    t.dim = dim;
    t.order = order;
    bool seen_subtrees = false, seen_grid = false;
    uint64_t nkeys = r.map_len();
    for (uint64_t k = 0; k < nkeys && r.ok(); ++k) {
        std::string key = r.str();
        if (key == "box_") {
            r.box(t.box_center, t.box_inv_half_length, dim);
        } else if (key == "subtrees_") {
            seen_subtrees = true;
            uint64_t nsub = r.bounded_len(r.array_len());
            t.subtree_begin.assign(1, 0u);
            t.subtree_begin.reserve(nsub + 1);
            for (uint64_t s = 0; s < nsub && r.ok(); ++s) {
                if (r.array_len() != 1)
                    r.fail();
                uint64_t nn = r.bounded_len(r.array_len());
                size_t base = t.first_child.size();
                t.center.resize((base + nn) * dim);
                t.inv_half_length.resize((base + nn) * dim);
                t.first_child.resize(base + nn);
                t.coeff_offset.resize(base + nn);
                t.node_output_dim.resize(base + nn);
                for (uint64_t i = 0; i < nn && r.ok(); ++i) {
                    size_t g = base + i;
                    if (r.array_len() != 4)
                        r.fail();
                    r.box(&t.center[g * dim], &t.inv_half_length[g * dim], dim);
                    t.first_child[g] = (uint32_t)r.integer();
                    t.coeff_offset[g] = r.integer();
                    t.node_output_dim[g] = (uint32_t)r.integer();
                }
                t.subtree_begin.push_back((uint32_t)(base + nn));
            }
        } else if (key == "n_subtrees_") {
            double v[3];
            r.eigen(v, dim);
            for (int j = 0; j < dim; ++j)
                t.n_subtrees[j] = (int)v[j];
            seen_grid = true;
        } else if (key == "tol_") {
            t.tol = r.real();
        } else if (key == "lower_left_") {
            r.eigen(t.lower_left, dim);
        } else if (key == "upper_right_") {
            r.eigen(t.upper_right, dim);
        } else if (key == "inv_bin_size_") {
            r.eigen(t.inv_bin_size, dim);
        } else if (key == "coeffs_") {
            uint64_t n = r.bounded_len(r.array_len());
            t.coeffs.resize(n);
            for (uint64_t i = 0; i < n && r.ok(); ++i)
                t.coeffs[i] = r.real();
        } else if (key == "split_multi_eval_") {
            t.split_multi_eval = r.boolean();
        } else if (key == "output_dim_") {
            t.output_dim = (uint32_t)r.integer();
        } else {
            r.fail();
        }
    }
    if (!r.ok() || !seen_subtrees || !seen_grid) {
        err = "malformed .baobzi body";
        return false;
    }
    err = t.validate();
    return err.empty();
}

bool read_baobzi_file(const char *path, HostTree &out, std::string &err) {
    FILE *f = std::fopen(path, "rb");
    if (!f) {
        err = std::string("cannot open '") + path + "'";
        return false;
    }
    std::vector<uint8_t> buf;
    std::fseek(f, 0, SEEK_END);
    long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    buf.resize(n > 0 ? (size_t)n : 0);
    size_t got = buf.empty() ? 0 : std::fread(buf.data(), 1, buf.size(), f);
    std::fclose(f);
    if (got != buf.size()) {
        err = std::string("short read on '") + path + "'";
        return false;
    }
    return read_baobzi_mem(buf.data(), buf.size(), out, err);
}

// ------------------------------------------------------------------------------------------ writer
namespace {

class Writer {
  public:
    explicit Writer(std::vector<uint8_t> &o) : o_(o) {}
    void be(uint64_t v, int nbytes) {
        for (int s = (nbytes - 1) * 8; s >= 0; s -= 8)
            o_.push_back((uint8_t)(v >> s));
    }
    void uint(uint64_t v) { // msgpack-c pack_imp_uint*: smallest unsigned form
        if (v < 0x80)
            o_.push_back((uint8_t)v);
        else if (v < 0x100) {
            o_.push_back(0xcc);
            be(v, 1);
        } else if (v < 0x10000) {
            o_.push_back(0xcd);
            be(v, 2);
        } else if (v < 0x100000000ull) {
            o_.push_back(0xce);
            be(v, 4);
        } else {
            o_.push_back(0xcf);
            be(v, 8);
        }
    }
    void sint(int64_t v) { // msgpack-c pack_imp_int*: non-negative values use the unsigned forms
        if (v >= 0)
            return uint((uint64_t)v);
        if (v >= -32)
            o_.push_back((uint8_t)(int8_t)v);
        else if (v >= -128) {
            o_.push_back(0xd0);
            be((uint64_t)v, 1);
        } else if (v >= -32768) {
            o_.push_back(0xd1);
            be((uint64_t)v, 2);
        } else if (v >= -2147483648ll) {
            o_.push_back(0xd2);
            be((uint64_t)v, 4);
        } else {
            o_.push_back(0xd3);
            be((uint64_t)v, 8);
        }
    }
    void real(double d) {
        uint64_t b;
        std::memcpy(&b, &d, 8);
        o_.push_back(0xcb);
        be(b, 8);
    }
    void boolean(bool b) { o_.push_back(b ? 0xc3 : 0xc2); }
    void array(uint64_t n) {
        if (n < 16)
            o_.push_back((uint8_t)(0x90 | n));
        else if (n < 0x10000) {
            o_.push_back(0xdc);
            be(n, 2);
        } else {
            o_.push_back(0xdd);
            be(n, 4);
        }
    }
    void map(uint64_t n) {
        if (n < 16)
            o_.push_back((uint8_t)(0x80 | n));
        else if (n < 0x10000) {
            o_.push_back(0xde);
            be(n, 2);
        } else {
            o_.push_back(0xdf);
            be(n, 4);
        }
    }
    void str(const char *s) {
        size_t n = std::strlen(s);
        if (n < 32)
            o_.push_back((uint8_t)(0xa0 | n));
        else {
            o_.push_back(0xd9);
            be(n, 1);
        }
        o_.insert(o_.end(), s, s + n);
    }
    void eigen_real(const double *v, int n) {
        array((uint64_t)n + 3);
        str("__eigen__");
        uint((uint64_t)n);
        uint(1);
        for (int i = 0; i < n; ++i)
            real(v[i]);
    }
    void eigen_int(const int *v, int n) {
        array((uint64_t)n + 3);
        str("__eigen__");
        uint((uint64_t)n);
        uint(1);
        for (int i = 0; i < n; ++i)
            sint(v[i]);
    }
    void box(const double *c, const double *ihl, int dim) {
        array(2);
        eigen_real(c, dim);
        eigen_real(ihl, dim);
    }

  private:
    std::vector<uint8_t> &o_;
};

} // namespace

void write_baobzi_mem(const HostTree &t, std::vector<uint8_t> &out) {
    out.clear();
    out.reserve(64 + t.coeffs.size() * 9 + (size_t)t.n_nodes() * (size_t)(40 + 40 * t.dim));
    Writer w(out);
    const int dim = t.dim;
    w.array(3);
    w.sint(t.dim);
    w.sint(t.order);
    w.sint(kFormatVersion);

    w.map(10);
    w.str("box_");
    w.box(t.box_center, t.box_inv_half_length, dim);
    w.str("subtrees_");
    w.array((uint64_t)t.n_bins());
    for (int64_t s = 0; s < t.n_bins(); ++s) {
        w.array(1);
        uint32_t b = t.subtree_begin[s], e = t.subtree_begin[s + 1];
        w.array(e - b);
        for (uint32_t g = b; g < e; ++g) {
            w.array(4);
            w.box(&t.center[(size_t)g * dim], &t.inv_half_length[(size_t)g * dim], dim);
            w.uint(t.first_child[g]);
            w.uint(t.coeff_offset[g]);
            w.uint(t.node_output_dim[g]);
        }
    }
    w.str("n_subtrees_");
    w.eigen_int(t.n_subtrees, dim);
    w.str("tol_");
    w.real(t.tol);
    w.str("lower_left_");
    w.eigen_real(t.lower_left, dim);
    w.str("upper_right_");
    w.eigen_real(t.upper_right, dim);
    w.str("inv_bin_size_");
    w.eigen_real(t.inv_bin_size, dim);
    w.str("coeffs_");
    w.array(t.coeffs.size());
    for (double c : t.coeffs)
        w.real(c);
    w.str("split_multi_eval_");
    w.boolean(t.split_multi_eval);
    w.str("output_dim_");
    w.uint(t.output_dim);
}

bool write_baobzi_file(const char *path, const HostTree &t, std::string &err) {
    std::vector<uint8_t> buf;
    write_baobzi_mem(t, buf);
    FILE *f = std::fopen(path, "wb");
    if (!f) {
        err = std::string("cannot open '") + path + "' for writing";
        return false;
    }
    size_t put = std::fwrite(buf.data(), 1, buf.size(), f);
    bool ok = std::fclose(f) == 0 && put == buf.size();
    if (!ok)
        err = std::string("short write on '") + path + "'";
    return ok;
}

// ------------------------------------------------------------------------------------------ HostTree
int64_t HostTree::n_leaves() const {
    int64_t n = 0;
    for (uint64_t o : coeff_offset)
        n += (o != kNoCoeffs);
    return n;
}

int64_t HostTree::block_doubles() const {
    int64_t b = 1;
    for (int j = 0; j < dim; ++j)
        b *= order;
    return b;
}

std::string HostTree::validate() const {
    if (!supported_dim_order(dim, order))
        return "unsupported (dim, order)";
    int64_t bins = 1;
    for (int j = 0; j < dim; ++j) {
        if (n_subtrees[j] < 1)
            return "n_subtrees_ must be positive";
        bins *= n_subtrees[j];
    }
    if (bins != n_bins())
        return "subtree count does not match n_subtrees_";
    const size_t nn = (size_t)n_nodes();
    if (nn == 0 || center.size() != nn * dim || inv_half_length.size() != nn * dim || coeff_offset.size() != nn ||
        node_output_dim.size() != nn)
        return "inconsistent node arrays";
    if (nn >= 0xFFFFFFFFull)
        return "too many nodes";
    if (output_dim > kMaxOutputDim)
        return "output_dim out of range";
    if (coeffs.size() >= (uint64_t(1) << 35)) // the device layout addresses blocks in 16-byte units with 32 bits
        return "coefficient array too large";
    const int nchild = 1 << dim;
    // the scalar path reads output_dim blocks of ORDER doubles (template.hpp:373-374); the batched path
    // one block of order^dim
    for (int64_t s = 0; s < n_bins(); ++s) {
        uint32_t b = subtree_begin[s], e = subtree_begin[s + 1];
        if (e <= b)
            return "empty subtree";
        for (uint32_t g = b; g < e; ++g) {
            if (is_leaf(g)) {
                // 1D: the scalar / all-components entries read node_output_dim blocks of ORDER doubles
                // (template.hpp:373-374); otherwise one block of order^dim
                if (node_output_dim[g] > kMaxOutputDim)
                    return "node output_dim out of range";
                const uint64_t od = node_output_dim[g] < 1 ? 1 : node_output_dim[g];
                const uint64_t need = dim == 1 ? (uint64_t)order * od : (uint64_t)block_doubles();
                if (coeff_offset[g] > coeffs.size() || need > coeffs.size() - coeff_offset[g]) // no wrap-around
                    return "leaf coefficient block out of range";
            } else {
                // children follow their parent in the breadth-first order of a subtree (template.hpp:406-457):
                // a link at or before the node itself would make the descent loop for ever
                uint64_t c0 = first_child[g];
                if (first_child[g] == kNoChild || c0 + nchild > (uint64_t)(e - b) || c0 <= (uint64_t)(g - b))
                    return "child index out of range";
            }
        }
    }
    return std::string();
}

} // namespace bz
