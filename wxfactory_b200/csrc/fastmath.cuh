// Table-driven double-precision exp / log for the trace formation of the Euler kernels.
//
// The reference extrapolates rho and rho*theta to the element faces in log space (wx_factory/rhs/rhs_dfr.py:50-71):
// every node costs two logs and every trace point three exps, and with the CUDA library versions (~45 / ~30
// instructions each) the trace kernels are FP64-issue bound.  These versions do the usual argument reduction against a
// small table kept in shared memory (log: 128 x (1/c, log c); exp: 64 x 2^(j/64)) followed by a short polynomial,
// error <= 1.4 ulp on the ranges the kernels use (checked against libm in tests/test_fastmath.py).  All of them are
// branch-free (no special-case paths), so the several independent logs / exps / reciprocals a thread needs interleave.
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define WXB_HD __host__ __device__ __forceinline__
#define WXB_FM __device__ __forceinline__  // the math functions read __constant__ coefficients: device-only under nvcc
#else
#define WXB_HD inline
#define WXB_FM inline                      // plain C++ build (tests/test_fastmath.py): the same code on the host
#endif
#include <cmath>
#include <cstring>

#define WXB_LOGTAB_BITS 7
#define WXB_LOGTAB_N (1 << WXB_LOGTAB_BITS)
#define WXB_EXPTAB_BITS 6
#define WXB_EXPTAB_N (1 << WXB_EXPTAB_BITS)
// doubles in the combined table: [128] x (invc, logc_hi, logc_lo, 0), then [64] 2^(j/64)
#define WXB_LOGTAB_STRIDE 4
#define WXB_EXPTAB_OFF (WXB_LOGTAB_STRIDE * WXB_LOGTAB_N)
#define WXB_FASTTAB_DOUBLES (WXB_EXPTAB_OFF + WXB_EXPTAB_N)

namespace wxbfm {

WXB_HD uint64_t bits_of(double x) {
#ifdef __CUDA_ARCH__
    return (uint64_t)__double_as_longlong(x);
#else
    uint64_t u;
    memcpy(&u, &x, 8);
    return u;
#endif
}
WXB_HD double double_of(uint64_t u) {
#ifdef __CUDA_ARCH__
    return __longlong_as_double((long long)u);
#else
    double x;
    memcpy(&x, &u, 8);
    return x;
#endif
}

constexpr uint64_t LOG_OFF = 0x3fe6000000000000ULL;  // z = x * 2^-k lands in [0.6875, 1.375)

// Polynomial coefficients live in constant memory on the device: the FP64 instructions then take them as constant-bank
// operands instead of building every 64-bit literal with two moves.
#ifdef __CUDACC__
#define WXB_FM_CONST static __constant__ double
#else
#define WXB_FM_CONST static const double
#endif
WXB_FM_CONST FM_C[16] = {
    0x1.62e42fefa3800p-1,   // 0  ln2 hi (11 trailing zero bits: k*ln2_hi is exact)
    0x1.ef35793c76730p-45,  // 1  ln2 lo
    -1.0 / 6.0, 0.2, -0.25, 1.0 / 3.0, -0.5,  // 2..6  log1p(r) = r + r^2 (-1/2 + r (1/3 + r (-1/4 + r (1/5 - r/6))))
    0x1.71547652b82fep6,    // 7  64 / ln2
    -0x1.62e42fefa0000p-7,  // 8  -(ln2/64) hi
    -0x1.cf79abc9e3b3ap-46, // 9  -(ln2/64) lo
    1.0 / 120.0, 1.0 / 24.0, 1.0 / 6.0, 0.5,  // 10..13  e^r - 1 = r + r^2 (1/2 + r (1/6 + r (1/24 + r/120)))
    0x1.8p52,               // 14 rounding shift
    0.0};

// Natural logarithm; tab = the (invc, logc) pairs.  BRANCH-FREE: x <= 0, subnormal, inf or NaN gives NaN (a state that
// has blown up must still trip the callers' non-finite check, wx_factory/simulation/simulation.py:399-408; the value
// itself is of no interest then).  No branch means independent logs / exps of one thread interleave.
// accuracy A/B switches (scripts/gpu_r2_variants.sh): route one function family to the CUDA library version
#ifndef WXB_FM_LIB_LOG
#define WXB_FM_LIB_LOG 0
#endif
#ifndef WXB_FM_LIB_EXP
#define WXB_FM_LIB_EXP 0
#endif
#ifndef WXB_FM_LIB_DIV
#define WXB_FM_LIB_DIV 0
#endif

WXB_FM double fast_log(double x, const double* __restrict__ tab) {
    if (WXB_FM_LIB_LOG) return log(x);
    const uint64_t ix = bits_of(x);
    const uint32_t top = (uint32_t)(ix >> 32);
    const bool bad = top - 0x00100000u >= 0x7fe00000u;
    const uint64_t tmp = ix - LOG_OFF;
    const int i = (int)((tmp >> (52 - WXB_LOGTAB_BITS)) & (WXB_LOGTAB_N - 1));
    const int k = (int)((int64_t)tmp >> 52);
    const double z = double_of(ix - (tmp & 0xfff0000000000000ULL));
    // log c in two pieces: logc_hi is a multiple of 2^-43 like k*ln2_hi, so their sum w is EXACT (the result is then
    // good to ~0.6 ulp also where k*ln2 dominates, e.g. log(rho theta) ~ 5.7 -- the hydrostatic / acoustic balance of
    // the interface fluxes amplifies the rounding of exactly these logs)
    const double invc = tab[WXB_LOGTAB_STRIDE * i], logc = tab[WXB_LOGTAB_STRIDE * i + 1], logc_lo = tab[WXB_LOGTAB_STRIDE * i + 2];
    const double r = fma(z, invc, -1.0);  // |r| < 0.004
    const double kd = (double)k;
    const double w = fma(kd, FM_C[0], logc);
    const double r2 = r * r;
    double p = fma(r, FM_C[2], FM_C[3]);
    p = fma(r, p, FM_C[4]);
    p = fma(r, p, FM_C[5]);
    p = fma(r, p, FM_C[6]);
    const double hi = w + r;
    const double lo = (w - hi) + r;  // what the sum w + r lost
    const double y = fma(r2, p, fma(kd, FM_C[1], lo + logc_lo)) + hi;
    return bad ? double_of(0x7ff8000000000000ULL) : y;
}

// e^x; tab = the 64 values 2^(j/64).  BRANCH-FREE and unguarded: valid for |x| < 700 (the kernels feed it extrapolated
// logs of physical densities); NaN and +-inf give NaN.
WXB_FM double fast_exp(double x, const double* __restrict__ tab) {
    if (WXB_FM_LIB_EXP) return exp(x);
    double kd = fma(x, FM_C[7], FM_C[14]);  // x * 64/ln2, rounded to an integer in the low mantissa bits
    const uint64_t ki = bits_of(kd);
    kd -= FM_C[14];
    double r = fma(kd, FM_C[8], x);
    r = fma(kd, FM_C[9], r);
    const uint64_t sbits = bits_of(tab[ki & (WXB_EXPTAB_N - 1)]) + ((ki >> WXB_EXPTAB_BITS) << 52);
    const double s = double_of(sbits);
    double p = fma(r, FM_C[10], FM_C[11]);
    p = fma(r, p, FM_C[12]);
    p = fma(r, p, FM_C[13]);
    p = fma(r * r, p, r);
    return fma(s, p, s);
}

// 1/x and sqrt(x) without the library's special-case branches (one MUFU seed + Newton steps; within 1-2 ulp for
// normal positive x, NaN for 0 / inf / NaN where the library would give inf / 0)
WXB_FM double fast_rcp(double x) {
    if (WXB_FM_LIB_DIV) return 1.0 / x;
#ifdef __CUDA_ARCH__
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    double e = fma(-x, r, 1.0);
    r = fma(r, e, r);
    e = fma(-x, r, 1.0);
    return fma(r, e, r);
#else
    return 1.0 / x;
#endif
}
WXB_FM double fast_sqrt(double x) {
    if (WXB_FM_LIB_DIV) return sqrt(x);
#ifdef __CUDA_ARCH__
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    const double h = 0.5 * x;
    y = y * fma(-(h * y), y, 1.5);
    y = y * fma(-(h * y), y, 1.5);
    double s = x * y;
    const double e = fma(-s, s, x);
    return fma(e, 0.5 * y, s);
#else
    return sqrt(x);
#endif
}

// host: fill the combined table (long double arithmetic; each entry correctly rounded to double)
inline void fill_tables(double* tab) {
    for (int i = 0; i < WXB_LOGTAB_N; ++i) {
        const double lo = double_of(LOG_OFF + ((uint64_t)i << (52 - WXB_LOGTAB_BITS)));
        const double hi = double_of(LOG_OFF + ((uint64_t)(i + 1) << (52 - WXB_LOGTAB_BITS)));
        const long double c = 0.5L * ((long double)lo + (long double)hi);
        const double invc = (double)(1.0L / c);
        const long double logc = -logl((long double)invc);  // log of the c that invc really stands for
        const double hi43 = (double)(roundl(logc * 0x1p43L) * 0x1p-43L);
        tab[WXB_LOGTAB_STRIDE * i] = invc;
        tab[WXB_LOGTAB_STRIDE * i + 1] = hi43;
        tab[WXB_LOGTAB_STRIDE * i + 2] = (double)(logc - (long double)hi43);
        tab[WXB_LOGTAB_STRIDE * i + 3] = 0.0;
    }
    for (int j = 0; j < WXB_EXPTAB_N; ++j) tab[WXB_EXPTAB_OFF + j] = (double)exp2l((long double)j / WXB_EXPTAB_N);
}

}  // namespace wxbfm
