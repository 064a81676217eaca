/* benchfuncs.c -- the benchmark functions of BASELINE.json's configurations as C callbacks
 * (baobzi_input_func_t), so bench.py can fit the benchmark trees through baobzi_init without a Python
 * callback in the loop.  Definitions follow the reference's timing program (examples/C/baobzi_timing.c:19-41)
 * and SURVEY.md section 8d (config 4).  Built into baobzi_b200/lib/libbenchfuncs.so. */
#include <math.h>

void bench_fun_1d(const double *x, double *y, const void *data) { /* timing.c:19-22 */
    *y = *(const double *)data * log(x[0]);
}

void bench_fun_2d(const double *x, double *y, const void *data) { /* timing.c:30-33 */
    *y = *(const double *)data * exp(cos(5.0 * x[0]) * sin(5.0 * x[1]));
}

void bench_fun_3d(const double *x, double *y, const void *data) { /* timing.c:39-41 */
    (void)data;
    *y = exp(x[0] + 2 * sin(x[1])) * (x[0] * x[0] + log(2 + x[1] * x[2]));
}

void bench_fun_2d_osc(const double *x, double *y, const void *data) { /* 2 + sin(Kx) cos(Ky), bounded away from 0 */
    const double K = *(const double *)data;
    *y = 2.0 + sin(K * x[0]) * cos(K * x[1]);
}
