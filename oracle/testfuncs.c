/* TEST INFRASTRUCTURE ONLY.
 *
 * The benchmark / test callbacks of the reference, restated as a tiny C library so that
 * tests/, the golden generator and bench.py can hand a C function pointer to
 * baobzi_init (the reference's, through oracle/_ref, or ours).  Definitions follow
 *   examples/C/baobzi_timing.c:19-41   (testfun_1d, testfun_1d2, testfun_2d, testfun_2d_2, testfun_3d)
 *   tests/test_c.cpp:6-25              (testfun_3d1; the others coincide with timing.c's)
 * plus the oscillatory function of SURVEY.md section 8d (config 4): 2 + sin(Kx)cos(Ky).
 * `data` points at one double (scale factor or K) where the reference's does.
 */
#include <math.h>

void testfun_1d(const double *x, double *res, const void *data) {
    const double scale_factor = *(const double *)data;
    *res = scale_factor * log(x[0]);
}

void testfun_1d2(const double *x, double *res, const void *data) {
    const double scale_factor = *(const double *)data;
    res[0] = scale_factor * log(x[0]);
    res[1] = sin(x[0]);
}

void testfun_2d(const double *x, double *res, const void *data) {
    const double scale_factor = *(const double *)data;
    *res = scale_factor * exp(cos(5.0 * x[0]) * sin(5.0 * x[1]));
}

void testfun_2d_2(const double *x, double *res, const void *data) {
    (void)data;
    *res = exp(x[0] + 2 * sin(x[1])) * (x[0] * x[0] + log(2 + x[1]));
}

void testfun_3d(const double *x, double *res, const void *data) {
    (void)data;
    *res = exp(x[0] + 2 * sin(x[1])) * (x[0] * x[0] + log(2 + x[1] * x[2]));
}

void testfun_3d1(const double *x, double *res, const void *data) {
    const double scale_factor = *(const double *)data;
    *res = scale_factor * exp(cos(5.0 * x[0]) * sin(5.0 * x[1]) * cos(4.0 * x[2]));
}

void testfun_2d_osc(const double *x, double *res, const void *data) {
    const double K = *(const double *)data;
    *res = 2.0 + sin(K * x[0]) * cos(K * x[1]);
}

/* narrow gaussian of tests/test_template.cpp:163-181 style (forces refinement) */
void testfun_1d_gauss(const double *x, double *res, const void *data) {
    const double w = *(const double *)data;
    *res = 1.0 + exp(-x[0] * x[0] / (w * w));
}
