/* TEST INFRASTRUCTURE ONLY -- see baobzi_oracle.h.  Plain C11 restatement of the reference's
 * restore + batched evaluation path.  Compile with -ffp-contract=off: every fused multiply-add
 * below is an explicit fma() so the rounding sequence is exactly what is written. */
#include "baobzi_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------
 * Minimal msgpack pull-reader, enough for the version-5 .baobzi layout:
 * written by msgpack::pack in Function::save (template.hpp:914-919) with the adaptors at
 * template.hpp:29-62 (header), 105 (Box), 382 (Node), 519 (FunctionTree), 1081-1082 (Function map)
 * and include/baobzi/eigen_matrix_plugin.h:30-42 (Eigen vectors as ["__eigen__", rows, cols, v...]).
 * ------------------------------------------------------------------------------------------ */
typedef struct {
    const unsigned char *p, *end;
    int err;
} mp_t;

static uint64_t mp_be(mp_t *m, int n) {
    uint64_t v = 0;
    if (m->end - m->p < n) {
        m->err = 1;
        m->p = m->end;
        return 0;
    }
    for (int i = 0; i < n; ++i)
        v = (v << 8) | *m->p++;
    return v;
}

static int mp_tag(mp_t *m) {
    if (m->p >= m->end) {
        m->err = 1;
        return 0xc1; /* never-used tag */
    }
    return *m->p++;
}

static int64_t mp_array(mp_t *m) {
    int t = mp_tag(m);
    if ((t & 0xf0) == 0x90)
        return t & 0x0f;
    if (t == 0xdc)
        return (int64_t)mp_be(m, 2);
    if (t == 0xdd)
        return (int64_t)mp_be(m, 4);
    m->err = 1;
    return 0;
}

static int64_t mp_map(mp_t *m) {
    int t = mp_tag(m);
    if ((t & 0xf0) == 0x80)
        return t & 0x0f;
    if (t == 0xde)
        return (int64_t)mp_be(m, 2);
    if (t == 0xdf)
        return (int64_t)mp_be(m, 4);
    m->err = 1;
    return 0;
}

/* any msgpack number as double (msgpack-c's convert<double> accepts ints and both float widths) */
static double mp_number(mp_t *m, int *is_int, uint64_t *as_u64) {
    int t = mp_tag(m);
    uint64_t u;
    if (is_int)
        *is_int = 1;
    if (t <= 0x7f) {
        if (as_u64)
            *as_u64 = (uint64_t)t;
        return (double)t;
    }
    if (t >= 0xe0) {
        if (as_u64)
            *as_u64 = (uint64_t)(int64_t)(int8_t)t;
        return (double)(int8_t)t;
    }
    switch (t) {
    case 0xcc: u = mp_be(m, 1); if (as_u64) *as_u64 = u; return (double)u;
    case 0xcd: u = mp_be(m, 2); if (as_u64) *as_u64 = u; return (double)u;
    case 0xce: u = mp_be(m, 4); if (as_u64) *as_u64 = u; return (double)u;
    case 0xcf: u = mp_be(m, 8); if (as_u64) *as_u64 = u; return (double)u;
    case 0xd0: u = mp_be(m, 1); if (as_u64) *as_u64 = (uint64_t)(int64_t)(int8_t)u; return (double)(int8_t)u;
    case 0xd1: u = mp_be(m, 2); if (as_u64) *as_u64 = (uint64_t)(int64_t)(int16_t)u; return (double)(int16_t)u;
    case 0xd2: u = mp_be(m, 4); if (as_u64) *as_u64 = (uint64_t)(int64_t)(int32_t)u; return (double)(int32_t)u;
    case 0xd3: u = mp_be(m, 8); if (as_u64) *as_u64 = u; return (double)(int64_t)u;
    case 0xca: {
        uint32_t b = (uint32_t)mp_be(m, 4);
        float f;
        memcpy(&f, &b, 4);
        if (is_int) *is_int = 0;
        if (as_u64) *as_u64 = 0;
        return (double)f;
    }
    case 0xcb: {
        uint64_t b = mp_be(m, 8);
        double d;
        memcpy(&d, &b, 8);
        if (is_int) *is_int = 0;
        if (as_u64) *as_u64 = 0;
        return d;
    }
    case 0xc2: if (as_u64) *as_u64 = 0; return 0.0;
    case 0xc3: if (as_u64) *as_u64 = 1; return 1.0;
    default: m->err = 1; return 0.0;
    }
}

static double mp_double(mp_t *m) { return mp_number(m, NULL, NULL); }

static uint64_t mp_uint(mp_t *m) {
    uint64_t u = 0;
    int is_int = 0;
    mp_number(m, &is_int, &u);
    if (!is_int)
        m->err = 1;
    return u;
}

/* read a str into buf (NUL-terminated, truncated to cap-1); returns length */
static int64_t mp_str(mp_t *m, char *buf, int cap) {
    int t = mp_tag(m);
    int64_t n;
    if ((t & 0xe0) == 0xa0)
        n = t & 0x1f;
    else if (t == 0xd9)
        n = (int64_t)mp_be(m, 1);
    else if (t == 0xda)
        n = (int64_t)mp_be(m, 2);
    else if (t == 0xdb)
        n = (int64_t)mp_be(m, 4);
    else {
        m->err = 1;
        return 0;
    }
    if (m->end - m->p < n) {
        m->err = 1;
        return 0;
    }
    int64_t c = n < cap - 1 ? n : cap - 1;
    memcpy(buf, m->p, (size_t)c);
    buf[c] = 0;
    m->p += n;
    return n;
}

/* ["__eigen__", rows, cols, v0, ...] column-major (eigen_matrix_plugin.h:30-42); expects rows*cols == n */
static void mp_eigen(mp_t *m, double *out, int n) {
    char tag[16];
    int64_t len = mp_array(m);
    mp_str(m, tag, sizeof tag);
    uint64_t rows = mp_uint(m), cols = mp_uint(m);
    if (strcmp(tag, "__eigen__") != 0 || (int64_t)(rows * cols) != n || len != 3 + n) {
        m->err = 1;
        return;
    }
    for (int i = 0; i < n; ++i)
        out[i] = mp_double(m);
}

/* Box = [center, inv_half_length] (template.hpp:105) */
static void mp_box(mp_t *m, double *center, double *inv_hl, int dim) {
    if (mp_array(m) != 2)
        m->err = 1;
    mp_eigen(m, center, dim);
    mp_eigen(m, inv_hl, dim);
}

/* ------------------------------------------------------------------------------------------
 * restore: src/baobzi.cpp:75-116 + macros.h:45-51 + build_cache (template.hpp:763-778)
 * ------------------------------------------------------------------------------------------ */
static int supported(int dim, int order) { /* make_source.sh:11-13 */
    return dim >= 1 && dim <= 3 && order >= 6 && order <= 16 && (order % 2) == 0;
}

static void *grow(void *p, int64_t *cap, int64_t need, size_t elem) {
    if (need <= *cap)
        return p;
    int64_t c = *cap ? *cap : 1024;
    while (c < need)
        c *= 2;
    *cap = c;
    return realloc(p, (size_t)c * elem);
}

void oracle_free(oracle_tree_t *t) {
    if (!t)
        return;
    free(t->subtree_node_offsets);
    free(t->center);
    free(t->inv_half_length);
    free(t->first_child_idx);
    free(t->coeff_offset);
    free(t->node_output_dim);
    free(t->coeffs);
    free(t);
}

oracle_tree_t *oracle_restore_mem(const unsigned char *buf, uint64_t len) {
    mp_t m = {buf, buf + len, 0};
    /* header [dim, order, version] (header.h:10-18, template.hpp:34-58) */
    if (mp_array(&m) != 3)
        return NULL;
    int dim = (int)mp_uint(&m), order = (int)mp_uint(&m), version = (int)mp_uint(&m);
    if (m.err || version != 5 /* header.h:4, baobzi.cpp:94-97 */ || !supported(dim, order))
        return NULL;

    oracle_tree_t *t = (oracle_tree_t *)calloc(1, sizeof *t);
    t->dim = dim;
    t->order = order;
    t->output_dim = 1;
    t->split_multi_eval = 1;
    int64_t cap_c = 0, cap_h = 0, cap_f = 0, cap_o = 0, cap_d = 0;
    int have_subtrees = 0, have_nsub = 0;

    int64_t nkeys = mp_map(&m);
    for (int64_t k = 0; k < nkeys && !m.err; ++k) {
        char key[32];
        mp_str(&m, key, sizeof key);
        if (!strcmp(key, "box_")) {
            mp_box(&m, t->box_center, t->box_inv_half_length, dim);
        } else if (!strcmp(key, "subtrees_")) {
            have_subtrees = 1;
            t->n_bins = mp_array(&m);
            t->subtree_node_offsets = (int32_t *)malloc(sizeof(int32_t) * (size_t)(t->n_bins + 1));
            int64_t nn = 0;
            for (int64_t s = 0; s < t->n_bins && !m.err; ++s) {
                t->subtree_node_offsets[s] = (int32_t)nn;
                if (mp_array(&m) != 1) /* FunctionTree = [nodes_] (template.hpp:519) */
                    m.err = 1;
                int64_t cnt = mp_array(&m);
                t->center = (double *)grow(t->center, &cap_c, (nn + cnt) * dim, sizeof(double));
                t->inv_half_length = (double *)grow(t->inv_half_length, &cap_h, (nn + cnt) * dim, sizeof(double));
                t->first_child_idx = (uint32_t *)grow(t->first_child_idx, &cap_f, nn + cnt, sizeof(uint32_t));
                t->coeff_offset = (uint64_t *)grow(t->coeff_offset, &cap_o, nn + cnt, sizeof(uint64_t));
                t->node_output_dim = (uint32_t *)grow(t->node_output_dim, &cap_d, nn + cnt, sizeof(uint32_t));
                for (int64_t i = 0; i < cnt && !m.err; ++i, ++nn) {
                    /* Node = [box_, first_child_idx, coeff_offset, output_dim] (template.hpp:382) */
                    if (mp_array(&m) != 4)
                        m.err = 1;
                    mp_box(&m, t->center + nn * dim, t->inv_half_length + nn * dim, dim);
                    t->first_child_idx[nn] = (uint32_t)mp_uint(&m);
                    t->coeff_offset[nn] = mp_uint(&m);
                    t->node_output_dim[nn] = (uint32_t)mp_uint(&m);
                    if (t->coeff_offset[nn] != UINT64_MAX)
                        t->n_leaves++;
                }
            }
            t->n_nodes = nn;
            if (t->subtree_node_offsets)
                t->subtree_node_offsets[t->n_bins] = (int32_t)nn;
        } else if (!strcmp(key, "n_subtrees_")) {
            double v[3];
            mp_eigen(&m, v, dim);
            for (int j = 0; j < dim; ++j)
                t->n_subtrees[j] = (int)v[j];
            have_nsub = 1;
        } else if (!strcmp(key, "tol_")) {
            t->tol = mp_double(&m);
        } else if (!strcmp(key, "lower_left_")) {
            mp_eigen(&m, t->lower_left, dim);
        } else if (!strcmp(key, "upper_right_")) {
            mp_eigen(&m, t->upper_right, dim);
        } else if (!strcmp(key, "inv_bin_size_")) {
            mp_eigen(&m, t->inv_bin_size, dim);
        } else if (!strcmp(key, "coeffs_")) {
            t->n_coeffs = mp_array(&m);
            t->coeffs = (double *)malloc(sizeof(double) * (size_t)(t->n_coeffs ? t->n_coeffs : 1));
            for (int64_t i = 0; i < t->n_coeffs && !m.err; ++i)
                t->coeffs[i] = mp_double(&m);
        } else if (!strcmp(key, "split_multi_eval_")) {
            t->split_multi_eval = (int)mp_double(&m);
        } else if (!strcmp(key, "output_dim_")) {
            t->output_dim = (int)mp_uint(&m);
        } else {
            m.err = 1; /* MSGPACK_DEFINE_MAP has exactly these 10 keys (template.hpp:1081-1082) */
        }
    }
    int64_t prod = 1;
    for (int j = 0; j < dim; ++j)
        prod *= t->n_subtrees[j];
    if (m.err || !have_subtrees || !have_nsub || prod != t->n_bins || t->n_nodes == 0) {
        oracle_free(t);
        return NULL;
    }
    return t;
}

oracle_tree_t *oracle_restore(const char *filename) {
    FILE *f = fopen(filename, "rb");
    if (!f)
        return NULL;
    fseek(f, 0, SEEK_END);
    long len = ftell(f);
    fseek(f, 0, SEEK_SET);
    unsigned char *buf = (unsigned char *)malloc((size_t)(len > 0 ? len : 1));
    size_t got = fread(buf, 1, (size_t)len, f);
    fclose(f);
    oracle_tree_t *t = got == (size_t)len ? oracle_restore_mem(buf, (uint64_t)len) : NULL;
    free(buf);
    return t;
}

/* ------------------------------------------------------------------------------------------
 * lookup: bounds test, grid-cell floor transform, subtree descent
 * ------------------------------------------------------------------------------------------ */

/* template.hpp:872-873 (= 833): semi-open domain, out if any x<ll or any x>=ur.  NaN passes. */
static int out_of_domain(const oracle_tree_t *t, const double *x) {
    for (int j = 0; j < t->dim; ++j)
        if (x[j] < t->lower_left[j] || x[j] >= t->upper_right[j])
            return 1;
    return 0;
}

/* get_linear_bin, template.hpp:799-820: bin_j = (int)((x_j - lower_left_j) * inv_bin_size_j), x fastest.
 * The reference does not clamp (a point one ulp under upper_right can round to bin == n -> UB read);
 * we clamp to n-1, which is what the product does too. */
static int64_t linear_bin(const oracle_tree_t *t, const double *x) {
    int64_t lin = 0, stride = 1;
    for (int j = 0; j < t->dim; ++j) {
        const double x_bin = x[j] - t->lower_left[j];
        int b = (int)(x_bin * t->inv_bin_size[j]);
        if (b > t->n_subtrees[j] - 1)
            b = t->n_subtrees[j] - 1;
        if (b < 0)
            b = 0;
        lin += stride * b;
        stride *= t->n_subtrees[j];
    }
    return lin;
}

/* FunctionTree::get_node_index (template.hpp:482-493) + get_global_node_index (846-849):
 * child bit j = x[j] > center[j]; next = first_child_idx + bits, until is_leaf() (220). */
static int64_t global_node_index(const oracle_tree_t *t, const double *x) {
    const int64_t base = t->subtree_node_offsets[linear_bin(t, x)];
    int64_t cur = base;
    while (t->coeff_offset[cur] == UINT64_MAX) {
        uint32_t child = 0;
        for (int j = 0; j < t->dim; ++j)
            child |= (uint32_t)(x[j] > t->center[cur * t->dim + j]) << j;
        cur = base + t->first_child_idx[cur] + child;
    }
    return cur;
}

int64_t oracle_node_index(const oracle_tree_t *t, const double *x) {
    return out_of_domain(t, x) ? -1 : global_node_index(t, x);
}

void oracle_node_index_multi(const oracle_tree_t *t, const double *x, int64_t n, int64_t *idx) {
    for (int64_t i = 0; i < n; ++i)
        idx[i] = oracle_node_index(t, x + i * t->dim);
}

/* ------------------------------------------------------------------------------------------
 * cheb_eval (template.hpp:146-194).  FMA(a,b,c) is a*b+c with one rounding when arith ==
 * ORACLE_ARITH_FMA (what g++ -mfma -ffp-contract=fast makes of the reference's expressions,
 * established against the compiled reference) and two roundings otherwise.
 * ------------------------------------------------------------------------------------------ */
static inline int fused(int arith) { return arith == ORACLE_ARITH_AVX2 || arith == ORACLE_ARITH_AVX512; }
static inline double FMA(int arith, double a, double b, double c) { return fused(arith) ? fma(a, b, c) : a * b + c; }

/* template.hpp:146-161 Clenshaw, coefficients highest degree first */
static double cheb_eval_1d(int order, double x, const double *c, int arith) {
    const double x2 = 2 * x;
    double c0 = c[0];
    double c1 = c[1];
    for (int i = 2; i < order; ++i) {
        double tmp = c1;
        c1 = c[i] - c0;
        c0 = FMA(arith, c0, x2, tmp);
    }
    return FMA(arith, c0, x, c1);
}

/* T_0=1, T_1=x, T_n = (2x) T_{n-1} - T_{n-2}  (template.hpp:169-172, 182-187) */
static void cheb_T(int order, double x, double *T, int arith) {
    const double x2 = 2 * x;
    T[0] = 1.0;
    T[1] = x;
    for (int j = 2; j < order; ++j)
        T[j] = FMA(arith, x2, T[j - 1], -T[j - 2]);
}

/* y = C * v for a column-major order x order block: Eigen's GEMV (order >= 8) and its lazy
 * coefficient product (order 6) both accumulate each row sequentially over the columns
 * (extern/eigen/Eigen/src/Core/products/GeneralMatrixVector.h, ProductEvaluators.h). */
static void gemv_colmajor(int order, const double *C, const double *v, double *y, int arith) {
    for (int r = 0; r < order; ++r) {
        double acc = C[r] * v[0];
        for (int k = 1; k < order; ++k)
            acc = FMA(arith, C[r + k * order], v[k], acc);
        y[r] = acc;
    }
}

/* Eigen's completely unrolled reduction tree (Redux.h redux_novec_unroller / redux_vec_unroller):
 * the range is halved recursively, products at the leaves.  Element m of the range is
 * a[(start+m)*stride + lane] -- stride 1 for the scalar reduction of the 2D path, stride P (packet
 * width) for lane `lane` of the vectorised reduction of the 3D path.
 * FMA flavour (read off the -mavx2 -mfma object code of the compiled reference): wherever a sum has a
 * bare product as an operand GCC fuses it, the SECOND operand when both are products:
 *   e0 + e1        -> fma(a1, b1, a0*b0)
 *   e0 + (subtree) -> fma(a0, b0, subtree)                                                    */
static double dot_tree(const double *a, const double *b, int start, int len, int stride, int lane, int arith) {
#define EL(v, m) (v)[(m) * stride + lane]
    if (len == 1)
        return EL(a, start) * EL(b, start);
    const int half = len / 2;
    if (fused(arith) && len == 2)
        return fma(EL(a, start + 1), EL(b, start + 1), EL(a, start) * EL(b, start));
    if (fused(arith) && half == 1) /* len == 3: e0 + (e1 + e2) */
        return fma(EL(a, start), EL(b, start), dot_tree(a, b, start + 1, len - 1, stride, lane, arith));
#undef EL
    return dot_tree(a, b, start, half, stride, lane, arith) +
           dot_tree(a, b, start + half, len - half, stride, lane, arith);
}

/* template.hpp:163-177.  Tx is a strided row of a 2 x ORDER matrix, so Eigen's dot() here is the
 * scalar (non-vectorised) tree in every instruction-set build. */
static double cheb_eval_2d(int order, const double *x, const double *C, int arith) {
    double Tx[16], Ty[16], tmp[16];
    cheb_T(order, x[0], Tx, arith);
    cheb_T(order, x[1], Ty, arith);
    gemv_colmajor(order, C, Ty, tmp, arith);
    return dot_tree(Tx, tmp, 0, order, 1, 0, arith);
}

/* Vectorised dot of two contiguous ORDER-vectors (3D path): per-lane tree over the ORDER/P packets,
 * then Eigen's horizontal add (predux): Packet2d a0+a1; Packet4d (lo128+hi128) then the pair. */
static double dot_packets(int order, const double *a, const double *b, int P, int arith) {
    double acc[8];
    for (int l = 0; l < P; ++l)
        acc[l] = dot_tree(a, b, 0, order / P, P, l, arith);
    if (P == 8) /* AVX512 predux<Packet8d>: lo256+hi256, then the Packet4d pattern on pairs */
        return ((acc[0] + acc[4]) + (acc[1] + acc[5])) + ((acc[2] + acc[6]) + (acc[3] + acc[7]));
    if (P == 4)
        return (acc[0] + acc[2]) + (acc[1] + acc[3]);
    return acc[0] + acc[1];
}

/* template.hpp:179-194 */
static double cheb_eval_3d(int order, const double *x, const double *C, int arith) {
    double T[3][16], tmp[16];
    for (int i = 0; i < 3; ++i)
        cheb_T(order, x[i], T[i], arith);
    double res = 0.0;
    /* packet width Eigen picks for a fixed-size ORDER-vector of doubles (find_best_packet: the widest
     * available packet that divides ORDER): SSE2 Packet2d in the generic build, Packet4d with AVX/AVX2,
     * Packet8d with AVX512 */
    int P = 2;
    if (arith != ORACLE_ARITH_GENERIC && order % 4 == 0)
        P = 4;
    if (arith == ORACLE_ARITH_AVX512 && order % 8 == 0)
        P = 8;
    for (int i = 0; i < order; ++i) {
        gemv_colmajor(order, C + i * order * order, T[2], tmp, arith);
        double d = dot_packets(order, T[1], tmp, P, arith);
        res = FMA(arith, T[0][i], d, res);
    }
    return res;
}

double oracle_cheb_eval(int dim, int order, const double *xinterp, const double *coeffs, int arith) {
    if (dim == 1)
        return cheb_eval_1d(order, xinterp[0], coeffs, arith);
    if (dim == 2)
        return cheb_eval_2d(order, xinterp, coeffs, arith);
    return cheb_eval_3d(order, xinterp, coeffs, arith);
}

/* ------------------------------------------------------------------------------------------
 * Node::eval (template.hpp:366-375) and the batched / scalar drivers (866-889, 851-860)
 * ------------------------------------------------------------------------------------------ */
static void node_xinterp(const oracle_tree_t *t, int64_t node, const double *x, double *xi) {
    for (int j = 0; j < t->dim; ++j) /* (x - center) * inv_half_length, template.hpp:367 */
        xi[j] = (x[j] - t->center[node * t->dim + j]) * t->inv_half_length[node * t->dim + j];
}

void oracle_eval_multi(const oracle_tree_t *t, const double *x, double *res, int64_t n, int arith) {
    for (int64_t i = 0; i < n; ++i) {
        const double *xp = x + i * t->dim;
        if (out_of_domain(t, xp)) {
            res[i] = NAN; /* template.hpp:880-881 */
            continue;
        }
        int64_t node = global_node_index(t, xp);
        double xi[3];
        node_xinterp(t, node, xp, xi);
        res[i] = oracle_cheb_eval(t->dim, t->order, xi, t->coeffs + t->coeff_offset[node], arith);
    }
}

void oracle_eval(const oracle_tree_t *t, const double *x, double *y, int arith) {
    if (out_of_domain(t, x)) {
        for (int k = 0; k < t->output_dim; ++k)
            y[k] = NAN; /* template.hpp:852-855 (then UB in the reference; we stop here) */
        return;
    }
    int64_t node = global_node_index(t, x);
    double xi[3];
    node_xinterp(t, node, x, xi);
    for (uint32_t k = 0; k < t->node_output_dim[node]; ++k) /* template.hpp:373-374: block k at +k*ORDER */
        y[k] = oracle_cheb_eval(t->dim, t->order, xi, t->coeffs + t->coeff_offset[node] + k * t->order, arith);
}
