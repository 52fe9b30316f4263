// Host check of the table-driven exp used by k_expect (csrc/kernels.cu, exp_tab): same constants, same operation order,
// compared with libm exp over [-700, 1].  gcc -O2 -ffp-contract=off tools/exp_tab_check.c -lm && ./a.out  ->  max rel err 2.2e-16
#include <math.h>
#include <stdio.h>
#include <stdint.h>
#include <string.h>
#include <stdlib.h>
static double T[64];
static double exp_tab(double x) {
    const double INV = 0x1.71547652b82fep+6, C1 = 0x1.62e42fe000000p-7, C2 = 0x1.f473de6af278fp-36, MAGIC = 6755399441055744.0;
    double t = fma(x, INV, MAGIC);
    int64_t tb; memcpy(&tb, &t, 8);
    int n = (int) (uint32_t) tb;
    double tn = t - MAGIC;
    double r = fma(tn, -C1, x);
    r = fma(tn, -C2, r);
    double q = fma(r, 1.0 / 120, 1.0 / 24);
    q = fma(q, r, 1.0 / 6);
    q = fma(q, r, 0.5);
    q = fma(q, r, 1.0);
    double p = q * r;
    int j = n & 63, k = n >> 6;
    double res = fma(T[j], p, T[j]);
    int64_t rb; memcpy(&rb, &res, 8);
    rb += (int64_t) k << 52;
    memcpy(&res, &rb, 8);
    return x < -700.0 ? 0.0 : res;
}
int main() {
    for (int j = 0; j < 64; j++) T[j] = exp2(j / 64.0);
    double maxrel = 0; srand(1);
    for (long i = 0; i < 20000000; i++) {
        double x = -700.0 * ((double) rand() / RAND_MAX) * ((i & 1) ? 1.0 : 0.02) + ((i % 7 == 0) ? 0.5 : 0.0);
        double a = exp_tab(x), b = exp(x);
        double rel = fabs(a - b) / b;
        if (rel > maxrel) { maxrel = rel; }
    }
    printf("max rel err %.3e (ulp ~ %.2f)\n", maxrel, maxrel / 1.11e-16);
    printf("%g %g %g\n", exp_tab(-701), exp_tab(-INFINITY), exp_tab(0.0));
    return 0;
}
