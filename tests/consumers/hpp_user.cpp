// A user of the reference's C++ RAII wrapper (reference include/baobzi.hpp:10-44, compiled UNMODIFIED from where it
// lies; this file is ours).  The wrapper only calls the seven C functions, so it must work against libbaobzi.so
// (B200) unchanged: fit through the wrapper's init constructor, evaluate one point and a batch, save, restore through
// the wrapper's file constructor, and require the restored object to give the same bits.
#include <baobzi.hpp>

#include <cmath>
#include <cstdio>
#include <vector>

static void f2(const double *x, double *y, const void *) { *y = std::exp(std::cos(3.0 * x[0]) * std::sin(2.0 * x[1])); }

int main() {
    baobzi_input_t in = baobzi_input_default;
    in.dim = 2, in.order = 8, in.tol = 1e-9, in.func = f2;
    const double c[2] = {0.5, -0.25}, hl[2] = {1.0, 0.5};
    baobzi::Baobzi b(&in, c, hl);
    const int n = 4097;
    std::vector<double> x(2 * n), y(n), y2(n);
    for (int i = 0; i < n; ++i) {
        x[2 * i] = c[0] + hl[0] * (2.0 * ((i * 37) % n) / n - 1.0);
        x[2 * i + 1] = c[1] + hl[1] * (2.0 * ((i * 101) % n) / n - 1.0);
    }
    b(x.data(), y.data(), n);
    int bad = 0;
    for (int i = 0; i < n; ++i) {
        double want, one;
        f2(&x[2 * i], &want, nullptr);
        b(&x[2 * i], &one);
        bad += !(std::fabs(y[i] - want) <= 1e-8 * std::fabs(want)) || one != y[i]; // scalar twin == batched, bit for bit
    }
    b.save("hpp_user.baobzi");
    baobzi::Baobzi r(std::string("hpp_user.baobzi"));
    r(x.data(), y2.data(), n);
    for (int i = 0; i < n; ++i)
        bad += y2[i] != y[i];
    std::printf(bad ? "FAIL %d\n" : "OK %d\n", bad ? bad : n);
    return bad != 0;
}
