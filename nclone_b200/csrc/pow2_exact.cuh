// pow2_exact.cuh -- bit-exact emulation of glibc 2.39's pow(x, 2.0), the value CPython returns
// for `x ** 2` on the reference's platform (x86-64, ifunc variant __pow_fma).
//
// Algorithm: glibc sysdeps/ieee754/dbl-64/e_pow.c (ARM optimized-routines): log(x) as a
// double-double from a 128-entry table + degree-7 polynomial, exp(2*log x) from a 128-entry
// 2^(i/128) table + degree-5 polynomial.  The sequence of roundings below (which products are
// fused) is the one in this image's libm.so.6 (__pow_fma, disassembled); constants come from the
// same binary (tools/extract_pow_tables.py -> pow2_tables.h).  tests/test_pow2.py checks the
// host build of this file against libm pow on >1e9 inputs, and the device build against the host's.
//
// Fast path: hi = RN(x*x) and the exact residual lo = fma(x, x, -hi).  glibc bounds its pow at
// 0.509 ulp + 1.3*2^-68*|y log x| (e_pow.c header; here < 0.51 ulp), i.e. pow(x,2) == hi unless x^2
// lies within 0.01 ulp of a rounding midpoint; measured over 2e9 samples the largest such distance
// with pow != hi was 0.0087 ulp.  The table path is taken only inside 1/64 ulp of a midpoint
// (3 % of inputs), everything else returns hi.
//
// Compiles as CUDA device code and as host C++ (for the libm cross-check).
#pragma once
#include <stdint.h>
#ifdef __CUDACC__
#define POW2_FN __device__ __forceinline__
#define POW2_SLOW __device__ __noinline__
#define POW2_TABLE_QUALIFIER __device__ const
#define POW2_FMA(a, b, c) __fma_rn((a), (b), (c))
#define POW2_AS_U64(x) ((uint64_t)__double_as_longlong(x))
#define POW2_AS_F64(u) __longlong_as_double((long long)(u))
#define POW2_LD(p) __ldg(p)
#else
#include <math.h>
#include <string.h>
#define POW2_FN static inline
#define POW2_SLOW static
#define POW2_TABLE_QUALIFIER static const
#define POW2_FMA(a, b, c) fma((a), (b), (c))
static inline uint64_t pow2_as_u64(double x) { uint64_t u; memcpy(&u, &x, 8); return u; }
static inline double pow2_as_f64(uint64_t u) { double x; memcpy(&x, &u, 8); return x; }
#define POW2_AS_U64(x) pow2_as_u64(x)
#define POW2_AS_F64(u) pow2_as_f64(u)
#define POW2_LD(p) (*(p))
#endif
#include "pow2_tables.h"

#define POW2_K(name) POW2_AS_F64(POW2_##name)

// pow(|x|, 2.0) main path; valid for 2^-100 <= |x| <= 2^100, |x| != 1
POW2_SLOW double pow2_table_path(double x) {
    const uint64_t ix = POW2_AS_U64(x) & 0x7fffffffffffffffULL;      // y is an even integer: the sign is dropped
    // ---- log_inline
    const uint64_t tmp = ix - 0x3fe6955500000000ULL;
    const int i = (int)((tmp >> 45) & 127);
    const int64_t k = (int64_t)tmp >> 52;
    const uint64_t iz = ix - (tmp & 0xfff0000000000000ULL);
    const double z = POW2_AS_F64(iz), kd = (double)k;
    const double invc = POW2_AS_F64(POW2_LD(pow2_log_tab + 3 * i));
    const double logc = POW2_AS_F64(POW2_LD(pow2_log_tab + 3 * i + 1));
    const double logctail = POW2_AS_F64(POW2_LD(pow2_log_tab + 3 * i + 2));
    const double r = POW2_FMA(z, invc, -1.0);
    const double t1 = POW2_FMA(kd, POW2_K(LN2HI), logc);
    const double t2 = t1 + r;
    const double lo1 = POW2_FMA(kd, POW2_K(LN2LO), logctail);
    const double lo2 = t1 - t2 + r;
    const double ar = POW2_K(A0) * r, ar2 = r * ar, ar3 = r * ar2;
    const double hi = t2 + ar2;
    const double lo3 = POW2_FMA(ar, r, -ar2);
    const double lo4 = t2 - hi + ar2;
    double q = POW2_FMA(r, POW2_K(A6), POW2_K(A5));
    q = POW2_FMA(q, ar2, POW2_FMA(r, POW2_K(A4), POW2_K(A3)));
    q = POW2_FMA(ar2, q, POW2_FMA(r, POW2_K(A2), POW2_K(A1)));
    const double lo = POW2_FMA(ar3, q, lo1 + lo2 + lo3 + lo4);
    const double y = hi + lo;
    const double tail = hi - y + lo;
    // ---- pow: ehi + elo = 2 * (y + tail)
    const double ehi = 2.0 * y;
    const double elo = POW2_FMA(2.0, tail, POW2_FMA(y, 2.0, -ehi));
    // ---- exp_inline
    const double zz = POW2_FMA(ehi, POW2_K(INVLN2N), POW2_K(SHIFT));
    const uint64_t ki = POW2_AS_U64(zz);
    const double kd2 = zz - POW2_K(SHIFT);
    double rr = POW2_FMA(kd2, POW2_K(NEGLN2LON), POW2_FMA(kd2, POW2_K(NEGLN2HIN), ehi));
    rr = elo + rr;
    const int idx = 2 * (int)(ki & 127);
    const double etail = POW2_AS_F64(POW2_LD(pow2_exp_tab + idx));
    const uint64_t sbits = POW2_LD(pow2_exp_tab + idx + 1) + (ki << 45);
    const double r2 = rr * rr;
    const double p1 = POW2_FMA(rr, POW2_K(C3), POW2_K(C2));
    const double p0 = rr + etail;
    const double p2 = POW2_FMA(rr, POW2_K(C5), POW2_K(C4));
    const double t = POW2_FMA(p1, r2, p0);
    const double r4 = r2 * r2;
    const double tmpv = POW2_FMA(p2, r4, t);
    const double scale = POW2_AS_F64(sbits);
    return POW2_FMA(tmpv, scale, scale);
}

// Python `x ** 2` for a float x.
#ifdef __CUDACC__
// Fast-path test in one FMA: RN(hi + lo * 32/31) == hi  <=>  |lo| < 31/64 ulp(hi) (up to ties, which only
// widen the table path's share).  Just below a power of two the spacing halves, so the test fails there
// for lo < 0 at |lo| >= 31/128 ulp -- conservative, the table path is exact anyway; lo == 0 (an exactly
// representable square, incl. x = +-2^k and 0) always passes.  inf/nan give lo = nan and fail the test.
POW2_FN bool pow2_fast_ok(double hi, double lo) { return __fma_rn(lo, 0x1.0842108421084p+0, hi) == hi; }
POW2_FN double pow2_exact(double x) {
    const double hi = x * x;
    const double lo = __fma_rn(x, x, -hi);                     // exact: x*x == hi + lo
    if (pow2_fast_ok(hi, lo)) return hi;
    // outside [2^-200, 2^200] (incl. inf, nan, squares that underflow) the table path's preconditions do not hold: x*x
    if ((unsigned)((__double2hiint(hi) & 0x7ff00000) - 0x33700000) > (unsigned)(0x4c800000 - 0x33700000)) return hi;
    return pow2_table_path(x);
}
#else
POW2_FN double pow2_exact(double x) {
    const double hi = x * x;
    const double lo = POW2_FMA(x, x, -hi);                    // exact: x*x == hi + lo
    const uint64_t hb = POW2_AS_U64(hi);
    const uint64_t eb = hb & 0x7ff0000000000000ULL;
    if (eb - 0x3370000000000000ULL > 0x4c80000000000000ULL - 0x3370000000000000ULL) return hi;
    const double half_ulp = POW2_AS_F64(eb - 0x0350000000000000ULL);          // 2^(e-53)
    const double thr = half_ulp * 0.96875;
    const double alo = lo < 0 ? -lo : lo;
    if (alo < thr && ((hb & 0x000fffffffffffffULL) != 0 || lo == 0)) return hi;
    return pow2_table_path(x);
}
#endif

#ifdef __CUDACC__
// Out of line on purpose: the step kernel is bound by instruction fetch, and ~60 call sites sharing one hot copy
// (fast path included) beat 60 inlined copies.  Same for division and square root below (geom.cuh).
#ifdef NPP_OUTLINE_SQ
__device__ __noinline__ double SQ(double x) { return pow2_exact(x); }
#else
DEV double SQ(double x) { return pow2_exact(x); }
#endif
// x*x, and `inexact` set when that may differ from pow(x, 2.0) (x^2 within 1/64 ulp of a rounding
// midpoint, ~3 % of inputs).  With the flag clear the product IS the reference's value, so callers can
// run the literal reference expression on cheap squares and only redo flagged cases on the table path.
DEV double sq_flag(double x, bool& inexact) {
    const double hi = x * x;
    const double lo = __fma_rn(x, x, -hi);
    if (!pow2_fast_ok(hi, lo) && (unsigned)((__double2hiint(hi) & 0x7ff00000) - 0x33700000) <= (unsigned)(0x4c800000 - 0x33700000)) inexact = true;
    return hi;
}
#endif
