// pow2.cuh -- SQ(x): the reference's `x ** 2`.
//
// CPython evaluates float ** 2 as libm pow(x, 2.0) (floatobject.c:float_pow), which on glibc
// differs from the correctly rounded x*x by one ulp for ~1e-3 of inputs.
// NPP_SQ_MODE 0: x*x.  9 of the 130 bc_replays differ from the reference with it (death-ball speeds, 1 ulp).
// NPP_SQ_MODE 1: exact emulation of glibc's pow(x, 2.0) -- see pow2_exact.cuh.
#pragma once
#ifndef NPP_SQ_MODE
#define NPP_SQ_MODE 1
#endif
#if NPP_SQ_MODE == 0
DEV double SQ(double x) { return x * x; }
#else
#include "pow2_exact.cuh"
#endif
