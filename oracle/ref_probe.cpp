// TEST INFRASTRUCTURE ONLY -- not part of the product; only tests/ and the golden-vector
// generator may load the library built from this file.
//
// Leaf-index probe.  The reference's C ABI has no way to ask "which node did this point
// land in", so this translation unit (our code) instantiates the reference's own header
// template baobzi::Function<DIM,ORDER,0> from /root/reference/include/baobzi_template.hpp,
// restores a .baobzi file exactly the way the reference does
// (src/baobzi.cpp:75-116 + include/baobzi/macros.h:45-51) and calls
// Function::get_global_node_index (baobzi_template.hpp:846-849) for every in-domain point.
// Index space: all nodes (internal + leaf), subtree-major, BFS within a subtree.
// Out-of-domain points (baobzi_template.hpp:872-873) report -1.
//
// Built by oracle/Makefile into oracle/_ref/libbaobzi_refidx.so.
#include "baobzi_template.hpp"

#include <cstdint>
#include <fstream>
#include <sstream>
#include <string>

namespace {

template <int DIM, int ORDER>
int probe(const msgpack::object &obj, const double *x, int64_t n, int64_t *node_idx, int64_t *coeff_off) {
    using F = baobzi::Function<DIM, ORDER, 0>;
    F *f = new F();
    *f = obj.as<F>();
    f->build_cache();
    for (int64_t i = 0; i < n; ++i) {
        typename F::VecDimD xi(x + DIM * i);
        if ((xi.array() < f->lower_left_.array()).any() || (xi.array() >= f->upper_right_.array()).any()) {
            node_idx[i] = -1;
            if (coeff_off)
                coeff_off[i] = -1;
            continue;
        }
        std::size_t g = f->get_global_node_index(xi);
        node_idx[i] = (int64_t)g;
        if (coeff_off)
            coeff_off[i] = (int64_t)f->node_pointers_[g]->coeff_offset;
    }
    delete f;
    return 0;
}

} // namespace

#define CASE(D, O)                                                                                                     \
    case ((D) << 8 | (O)):                                                                                             \
        return probe<D, O>(obj, x, n, node_idx, coeff_off);

extern "C" int refidx_node_indices(const char *filename, const double *x, int64_t n, int64_t *node_idx,
                                   int64_t *coeff_off) {
    try {
        std::ostringstream buf;
        std::ifstream input(filename);
        buf << input.rdbuf();
        std::string data = buf.str();
        std::size_t offset = 0;
        msgpack::object_handle oh_header;
        msgpack::unpack(oh_header, data.data(), data.size(), offset);
        baobzi_header_t header = oh_header.get().as<baobzi_header_t>();
        msgpack::object_handle oh;
        msgpack::unpack(oh, data.data(), data.size(), offset);
        msgpack::object obj = oh.get();
        if (header.version != BAOBZI_HEADER_VERSION)
            return -2;
        switch (header.dim << 8 | header.order) {
            CASE(1, 6) CASE(1, 8) CASE(1, 10) CASE(1, 12) CASE(1, 14) CASE(1, 16)
            CASE(2, 6) CASE(2, 8) CASE(2, 10) CASE(2, 12) CASE(2, 14) CASE(2, 16)
            CASE(3, 6) CASE(3, 8) CASE(3, 10) CASE(3, 12) CASE(3, 14) CASE(3, 16)
        default:
            return -3;
        }
    } catch (std::exception &e) {
        return -1;
    }
}
