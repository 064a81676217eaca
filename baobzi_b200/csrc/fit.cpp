// fit.cpp -- host fitter (see fit.hpp).  Runtime-order implementation of the reference's adaptive
// fit; no Eigen, one small dense LU.  Geometry (box centres, inverse half lengths, grid constants) is
// produced with the same floating-point expressions as the reference so that trees agree with the
// reference's wherever the accept/refine decisions agree.
#include "fit.hpp"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <deque>
#include <vector>

namespace bz {
namespace {

constexpr int kMaxOrder = 16;

struct Cheb {
    int n = 0;
    double cosn[kMaxOrder];            // nodes on [-1,1], ascending (reference template.hpp:639-640)
    double lu[kMaxOrder][kMaxOrder];   // LU of A(j,i) = T_i(cosn[j]) (transposed Vandermonde, 607-622, 641)
    int piv[kMaxOrder];

    explicit Cheb(int order) : n(order) {
        for (int i = 0; i < n; ++i)
            cosn[n - i - 1] = std::cos(M_PI * (i + 0.5) / n);
        for (int j = 0; j < n; ++j) {
            lu[j][0] = 1.0;
            lu[j][1] = cosn[j];
            for (int i = 2; i < n; ++i)
                lu[j][i] = 2.0 * lu[j][i - 1] * cosn[j] - lu[j][i - 2];
        }
        // partial-pivot LU, in place
        for (int k = 0; k < n; ++k) {
            int p = k;
            for (int r = k + 1; r < n; ++r)
                if (std::fabs(lu[r][k]) > std::fabs(lu[p][k]))
                    p = r;
            piv[k] = p;
            if (p != k)
                for (int c = 0; c < n; ++c)
                    std::swap(lu[k][c], lu[p][c]);
            for (int r = k + 1; r < n; ++r) {
                lu[r][k] /= lu[k][k];
                for (int c = k + 1; c < n; ++c)
                    lu[r][c] -= lu[r][k] * lu[k][c];
            }
        }
    }

    // solve A c = f for one right-hand side with stride `s` (in place)
    void solve(double *f, int s) const {
        for (int k = 0; k < n; ++k)
            if (piv[k] != k)
                std::swap(f[k * s], f[piv[k] * s]);
        for (int r = 1; r < n; ++r) {
            double acc = f[r * s];
            for (int c = 0; c < r; ++c)
                acc -= lu[r][c] * f[c * s];
            f[r * s] = acc;
        }
        for (int r = n - 1; r >= 0; --r) {
            double acc = f[r * s];
            for (int c = r + 1; c < n; ++c)
                acc -= lu[r][c] * f[c * s];
            f[r * s] = acc / lu[r][r];
        }
    }

    void nodes(double lb, double ub, double *out) const { // template.hpp:628
        for (int i = 0; i < n; ++i)
            out[i] = 0.5 * ((lb + ub) + (ub - lb) * cosn[i]);
    }
};

struct BoxD { // Box (template.hpp:89-106): centre + inverse half length
    double c[3], ih[3];
};

BoxD make_box(const double *c, const double *hl, int dim) {
    BoxD b{};
    for (int j = 0; j < dim; ++j) {
        b.c[j] = c[j];
        b.ih[j] = 1.0 / hl[j];
    }
    return b;
}

// host-side cheb_eval for the 3D acceptance test (any consistent order of operations will do here)
double eval3(int n, const double *C, const double *u) {
    double T[3][kMaxOrder];
    for (int d = 0; d < 3; ++d) {
        T[d][0] = 1.0;
        T[d][1] = u[d];
        for (int j = 2; j < n; ++j)
            T[d][j] = 2.0 * u[d] * T[d][j - 1] - T[d][j - 2];
    }
    double res = 0.0;
    for (int i = 0; i < n; ++i) {
        double di = 0.0;
        for (int j = 0; j < n; ++j) {
            double acc = 0.0;
            for (int k = 0; k < n; ++k)
                acc += C[i * n * n + j + k * n] * T[2][k];
            di += T[1][j] * acc;
        }
        res += T[0][i] * di;
    }
    return res;
}

class Fitter {
  public:
    Fitter(const baobzi_input_t &in) : in_(in), dim_(in.dim), n_(in.order), od_(in.output_dim < 1 ? 1 : in.output_dim), ch_(in.order) {}

    // Node::fit (template.hpp:226-360).  true => leaf, coefficients in `out` (reference layout).
    bool fit_node(const BoxD &box, std::vector<double> &out) {
        double hl[3], lb[3], ub[3];
        for (int j = 0; j < dim_; ++j) {
            hl[j] = 1.0 / box.ih[j];
            lb[j] = box.c[j] - hl[j];
            ub[j] = box.c[j] + hl[j];
        }
        const int n = n_;
        if (dim_ == 1) {
            double xs[kMaxOrder];
            ch_.nodes(lb[0], ub[0], xs);
            vals_.assign((size_t)od_ * n, 0.0);
            for (int i = 0; i < n; ++i)
                in_.func(&xs[i], &vals_[(size_t)i * od_], in_.data);
            out.assign((size_t)n * od_, 0.0);
            for (int k = 0; k < od_; ++k) {
                double c[kMaxOrder];
                for (int i = 0; i < n; ++i)
                    c[i] = vals_[(size_t)i * od_ + k];
                ch_.solve(c, 1);
                // standard_error, vector case (template.hpp:116-120)
                double tail = std::max(std::fabs(c[n - 2]), std::fabs(c[n - 1]));
                double scale = std::max(1.0, std::fabs(c[0]));
                if (err_of(tail, scale) > in_.tol)
                    return false;
                for (int i = 0; i < n; ++i) // stored highest degree first (template.hpp:249-250)
                    out[(size_t)i + (size_t)n * k] = c[n - i - 1];
            }
            return true;
        }
        if (dim_ == 2) {
            double xs[kMaxOrder], ys[kMaxOrder];
            ch_.nodes(lb[0], ub[0], xs);
            ch_.nodes(lb[1], ub[1], ys);
            out.assign((size_t)n * n, 0.0); // column-major F(i,j) at i + j*n
            double tmp[kMaxOrder > 1 ? kMaxOrder : 1];
            (void)tmp;
            std::vector<double> y(od_);
            for (int i = 0; i < n; ++i)
                for (int j = 0; j < n; ++j) {
                    const double p[2] = {xs[i], ys[j]};
                    in_.func(p, y.data(), in_.data);
                    out[(size_t)i + (size_t)j * n] = y[0];
                }
            for (int j = 0; j < n; ++j) // along x for every column (template.hpp:289)
                ch_.solve(&out[(size_t)j * n], 1);
            for (int i = 0; i < n; ++i) // along y for every row (template.hpp:290)
                ch_.solve(&out[i], n);
            // standard_error, matrix case: anti-diagonal (template.hpp:121-128)
            double tail = 0.0;
            for (int i = 0; i < n; ++i)
                tail = std::max(tail, std::fabs(out[(size_t)i + (size_t)(n - i - 1) * n]));
            double scale = std::max(1.0, std::fabs(out[n - 1]));
            scale = std::max(scale, std::fabs(out[(size_t)(n - 1) * n]));
            return err_of(tail, scale) <= in_.tol;
        }
        // dim 3 (template.hpp:302-359)
        double xs[kMaxOrder], ys[kMaxOrder], zs[kMaxOrder];
        ch_.nodes(lb[0], ub[0], xs);
        ch_.nodes(lb[1], ub[1], ys);
        ch_.nodes(lb[2], ub[2], zs);
        const size_t n2 = (size_t)n * n;
        F_.assign(n2 * n, 0.0); // F(i,j,k) at i + j*n + k*n^2
        std::vector<double> y(od_);
        for (int i = 0; i < n; ++i)
            for (int j = 0; j < n; ++j)
                for (int k = 0; k < n; ++k) {
                    const double p[3] = {xs[i], ys[j], zs[k]};
                    in_.func(p, y.data(), in_.data);
                    F_[(size_t)i + (size_t)j * n + (size_t)k * n2] = y[0];
                }
        for (int k = 0; k < n; ++k)
            for (int j = 0; j < n; ++j) // along x
                ch_.solve(&F_[(size_t)j * n + (size_t)k * n2], 1);
        for (int k = 0; k < n; ++k)
            for (int i = 0; i < n; ++i) // along y
                ch_.solve(&F_[(size_t)i + (size_t)k * n2], n);
        for (int j = 0; j < n; ++j)
            for (int i = 0; i < n; ++i) // along z
                ch_.solve(&F_[(size_t)i + (size_t)j * n], (int)n2);
        // stored as ORDER slabs over x, each a column-major (y, z) matrix (template.hpp:331-335)
        out.assign(n2 * n, 0.0);
        for (int i = 0; i < n; ++i)
            for (int j = 0; j < n; ++j)
                for (int k = 0; k < n; ++k)
                    out[(size_t)i * n2 + j + (size_t)k * n] = F_[(size_t)i + (size_t)j * n + (size_t)k * n2];
        // resampling acceptance test on a uniform grid (template.hpp:339-356); relative error always
        for (int i = 0; i < n; ++i)
            for (int j = 0; j < n; ++j)
                for (int k = 0; k < n; ++k) {
                    const double g[3] = {(double)i, (double)j, (double)k};
                    double p[3], u[3];
                    for (int d = 0; d < 3; ++d) {
                        p[d] = (box.c[d] - hl[d]) + 2.0 * g[d] * hl[d] / n;
                        u[d] = (p[d] - box.c[d]) * box.ih[d];
                    }
                    const double test = eval3(n, out.data(), u);
                    in_.func(p, y.data(), in_.data);
                    const double rel = std::fabs((y[0] - test) / y[0]);
                    if (std::fabs(y[0]) > 1e-16 && rel > in_.tol)
                        return false;
                }
        extra_evals_ += (uint64_t)n2 * n;
        return true;
    }

    uint64_t evals_per_node() const {
        uint64_t e = 1;
        for (int j = 0; j < dim_; ++j)
            e *= (uint64_t)n_;
        return e;
    }

    const baobzi_input_t &in_;
    int dim_, n_, od_;
    Cheb ch_;
    std::vector<double> vals_, F_;
    uint64_t extra_evals_ = 0;

  private:
    double err_of(double tail, double scale) const {
        return in_.tol_type == BAOBZI_TOL_RELATIVE ? tail / scale : tail;
    }
};

void push_children(std::deque<BoxD> &q, const BoxD &parent, const double *half_width, int dim) {
    for (unsigned child = 0; child < (1u << dim); ++child) { // template.hpp:435-446, 680-694
        double c[3];
        for (int j = 0; j < dim; ++j)
            c[j] = parent.c[j] + (((child >> j) & 1) ? half_width[j] : -half_width[j]);
        q.push_back(make_box(c, half_width, dim));
    }
}

} // namespace

bool fit_tree(const baobzi_input_t &input, const double *center, const double *half_length, HostTree &t,
              std::string &err) {
    const auto t0 = std::chrono::steady_clock::now();
    const int dim = input.dim;
    Fitter fit(input);
    t = HostTree();
    t.dim = dim;
    t.order = input.order;
    t.output_dim = (uint32_t)(input.output_dim < 1 ? 1 : input.output_dim);
    t.split_multi_eval = input.split_multi_eval != 0;
    t.tol = input.tol;
    const BoxD domain = make_box(center, half_length, dim);
    for (int j = 0; j < dim; ++j) {
        t.box_center[j] = domain.c[j];
        t.box_inv_half_length[j] = domain.ih[j];
    }

    // ---- top level: whole-level breadth-first refinement until a level is not all parents (650-732)
    double lmin = half_length[0];
    for (int j = 1; j < dim; ++j)
        lmin = std::min(lmin, half_length[j]);
    for (int j = 0; j < dim; ++j)
        t.n_subtrees[j] = (int)(half_length[j] / lmin);
    std::deque<BoxD> q, maybe_q;
    q.push_back(domain);
    double hw[3];
    for (int j = 0; j < dim; ++j)
        hw[j] = half_length[j] * 0.5;
    std::vector<double> scratch;
    uint64_t root_evals = 0;
    int base_depth = 0;
    while (!q.empty()) {
        const size_t n_next = q.size();
        double leaf_fraction = 0.0;
        for (size_t i = 0; i < n_next; ++i) {
            BoxD box = q.front();
            q.pop_front();
            const bool leaf = fit.fit_node(box, scratch);
            if (!leaf || base_depth < input.min_depth) {
                push_children(q, box, hw, dim);
            } else {
                leaf_fraction += 1.0;
                push_children(maybe_q, box, hw, dim);
            }
        }
        root_evals += n_next * fit.evals_per_node();
        leaf_fraction /= (double)n_next;
        if (leaf_fraction < input.minimum_leaf_fraction) {
            while (!maybe_q.empty()) {
                q.push_back(maybe_q.front());
                maybe_q.pop_front();
            }
        }
        for (int j = 0; j < dim; ++j)
            hw[j] *= 0.5;
        if (((uint64_t)1 << (dim * (base_depth + 1))) == q.size()) {
            for (int j = 0; j < dim; ++j)
                t.n_subtrees[j] *= 2;
            base_depth++;
            if (base_depth > input.max_depth) {
                err = "Baobzi fit error: tree depth exceeded max allowed input depth";
                return false;
            }
        } else {
            break;
        }
    }

    // ---- grid geometry (734-741)
    double hl[3], bin_size[3];
    int64_t n_bins = 1;
    for (int j = 0; j < dim; ++j) {
        hl[j] = 1.0 / domain.ih[j];
        bin_size[j] = 2.0 * hl[j] / t.n_subtrees[j];
        t.inv_bin_size[j] = 0.5 * t.n_subtrees[j] / hl[j];
        t.lower_left[j] = domain.c[j] - hl[j];
        t.upper_right[j] = domain.c[j] + hl[j];
        n_bins *= t.n_subtrees[j];
    }

    // ---- one subtree per grid cell, x fastest (747-754 -> FunctionTree ctor 406-457)
    const int sub_max_depth = input.max_depth - base_depth;
    int deepest = 0;
    t.subtree_begin.assign(1, 0u);
    std::vector<double> coeffs_node;
    for (int64_t bin = 0; bin < n_bins; ++bin) {
        int b[3] = {0, 0, 0};
        if (dim == 1)
            b[0] = (int)bin;
        else if (dim == 2) {
            b[0] = (int)(bin % t.n_subtrees[0]);
            b[1] = (int)(bin / t.n_subtrees[0]);
        } else {
            b[0] = (int)(bin % t.n_subtrees[0]);
            b[1] = (int)((bin / t.n_subtrees[0]) % t.n_subtrees[1]);
            b[2] = (int)(bin / ((int64_t)t.n_subtrees[0] * t.n_subtrees[1]));
        }
        double pc[3], phl[3];
        for (int j = 0; j < dim; ++j) {
            pc[j] = ((double)b[j] + 0.5) * bin_size[j] + t.lower_left[j];
            phl[j] = 0.5 * bin_size[j];
        }
        const BoxD root = make_box(pc, phl, dim);
        std::deque<BoxD> sq;
        sq.push_back(root);
        double shw[3];
        for (int j = 0; j < dim; ++j)
            shw[j] = (1.0 / root.ih[j]) * 0.5;
        uint32_t next_child = 1;
        int depth = 0;
        const size_t base = t.first_child.size();
        while (!sq.empty()) {
            const size_t n_next = sq.size();
            for (size_t i = 0; i < n_next; ++i) {
                BoxD box = sq.front();
                sq.pop_front();
                for (int j = 0; j < dim; ++j) {
                    t.center.push_back(box.c[j]);
                    t.inv_half_length.push_back(box.ih[j]);
                }
                t.node_output_dim.push_back(t.output_dim);
                if (fit.fit_node(box, coeffs_node)) {
                    t.first_child.push_back(kNoChild);
                    t.coeff_offset.push_back(t.coeffs.size());
                    t.coeffs.insert(t.coeffs.end(), coeffs_node.begin(), coeffs_node.end());
                } else {
                    t.first_child.push_back(next_child);
                    t.coeff_offset.push_back(kNoCoeffs);
                    next_child += 1u << dim;
                    push_children(sq, box, shw, dim);
                }
            }
            if (!sq.empty())
                depth++;
            if (depth > sub_max_depth) {
                err = "Baobzi fit error: tree depth exceeded max allowed input depth";
                return false;
            }
            for (int j = 0; j < dim; ++j)
                shw[j] *= 0.5;
        }
        deepest = std::max(deepest, depth);
        (void)base;
        t.subtree_begin.push_back((uint32_t)t.first_child.size());
    }

    t.base_depth = base_depth;
    t.max_subtree_depth = deepest;
    t.n_func_evals = (uint64_t)t.n_nodes() * fit.evals_per_node() + root_evals; // as the reference reports it (600)
    t.fit_ms = (uint32_t)std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - t0).count();
    err = t.validate();
    return err.empty();
}

} // namespace bz
