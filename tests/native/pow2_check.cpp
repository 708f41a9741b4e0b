// Host cross-check of nclone_b200/csrc/pow2_exact.cuh against libm pow(x, 2.0).
// usage: pow2_check <n_samples> <seed> ; prints mismatches and fast-path statistics.
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cstdint>
#include "../../nclone_b200/csrc/pow2_exact.cuh"

static inline uint64_t splitmix(uint64_t& s) { uint64_t z = (s += 0x9e3779b97f4a7c15ULL); z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL; z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL; return z ^ (z >> 31); }

int main(int argc, char** argv) {
    long n = argc > 1 ? atol(argv[1]) : 10000000;
    uint64_t s = argc > 2 ? strtoull(argv[2], 0, 10) : 1;
    long bad_table = 0, bad_exact = 0, differ = 0, slow = 0;
    double worst = 0.0;   // largest distance to the rounding midpoint, in ulp, among pow != x*x cases
    volatile double two = 2.0;
    for (long it = 0; it < n; it++) {
        uint64_t u = splitmix(s);
        double x;
        int mode = it & 3;
        if (mode == 0) { // random mantissa, exponent in [-30, 12] (positions, speeds, normals)
            int e = (int)(splitmix(s) % 43) - 30;
            x = ldexp(1.0 + (double)(u >> 12) / 4503599627370496.0, e);
        } else if (mode == 1) { // uniform in [-1100, 1100]
            x = ((double)(u >> 11) / 9007199254740992.0 * 2200.0) - 1100.0;
        } else if (mode == 2) { // uniform in [-12, 12]
            x = ((double)(u >> 11) / 9007199254740992.0 * 24.0) - 12.0;
        } else { // near 1 / unit-vector components
            x = ((double)(u >> 11) / 9007199254740992.0 * 2.0) - 1.0;
        }
        if (u & 1) x = -x;
        double ref = pow(x, two);
        double got = pow2_exact(x);
        if (got != ref) { if (bad_exact < 10) printf("pow2_exact mismatch x=%a ref=%a got=%a\n", x, ref, got); bad_exact++; }
        double ax = fabs(x);
        if (ax >= 0x1p-100 && ax <= 0x1p100 && ax != 1.0) {
            double tp = pow2_table_path(x);
            if (tp != ref) { if (bad_table < 10) printf("table path mismatch x=%a ref=%a got=%a\n", x, ref, tp); bad_table++; }
        }
        double hi = x * x;
        if (ref != hi) {
            differ++;
            double lo = fma(x, x, -hi);
            int e; frexp(hi, &e);
            double ulp = ldexp(1.0, e - 53);
            double margin = (0.5 * ulp - fabs(lo)) / ulp;
            if (margin > worst) worst = margin;
        }
    }
    // exactly representable squares, including every x = +-2^k (x*x a power of two: the fast path's lo == 0 case)
    long exact_n = 0;
    for (int k = -100; k <= 100; k++)
        for (int m = 1; m <= 4096; m++) {
            double x = ldexp((double)m, k);
            for (int sg = 0; sg < 2; sg++, x = -x) {
                double ref = pow(x, two), got = pow2_exact(x);
                exact_n++;
                if (got != ref || ref != x * x) { if (bad_exact < 10) printf("exact-square mismatch x=%a ref=%a got=%a\n", x, ref, got); bad_exact++; }
            }
        }
    printf("exact squares checked: %ld\n", exact_n);
    printf("n=%ld pow!=x*x: %ld (%.3g)  table-path mismatches: %ld  pow2_exact mismatches: %ld  max midpoint margin among differing: %.6g ulp\n",
           n, differ, (double)differ / n, bad_table, bad_exact, worst);
    return (bad_table || bad_exact) ? 1 : 0;
}
