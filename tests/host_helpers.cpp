// Test helper: exposes the product's host-only pieces (reader/writer, fitter) without CUDA so the
// CPU test suite can exercise them.  Not part of the shipped library.
#include "../include/baobzi.h"
#include "fit.hpp"
#include "host_tree.hpp"

#include <cstdio>
#include <cstring>

extern "C" {

// read a .baobzi file with the product reader and write it back with the product writer
int bzt_roundtrip(const char *in, const char *out) {
    bz::HostTree t;
    std::string err;
    if (!bz::read_baobzi_file(in, t, err))
        return 1;
    if (!bz::write_baobzi_file(out, t, err))
        return 2;
    return 0;
}

// 0 = readable; otherwise nonzero and the message copied to msg
int bzt_read_status(const char *in, char *msg, int cap) {
    bz::HostTree t;
    std::string err;
    if (bz::read_baobzi_file(in, t, err))
        return 0;
    if (msg && cap > 0) {
        std::strncpy(msg, err.c_str(), (size_t)cap - 1);
        msg[cap - 1] = 0;
    }
    return 1;
}

// product fitter -> .baobzi file
int bzt_fit_save(const baobzi_input_t *input, const double *center, const double *half_length, const char *out) {
    bz::HostTree t;
    std::string err;
    if (!bz::fit_tree(*input, center, half_length, t, err)) {
        std::fprintf(stderr, "%s\n", err.c_str());
        return 1;
    }
    return bz::write_baobzi_file(out, t, err) ? 0 : 2;
}
}
