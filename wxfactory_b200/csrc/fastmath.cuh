// Table-driven double-precision exp / log for the trace formation of the Euler kernels.
//
// The reference extrapolates rho and rho*theta to the element faces in log space (wx_factory/rhs/rhs_dfr.py:50-71):
// every node costs two logs and every trace point three exps, and with the CUDA library versions (~45 / ~30
// instructions each) the trace kernels are FP64-issue bound.  These versions do the usual argument reduction against a
// small table kept in shared memory (log: 128 x (1/c, log c); exp: 64 x 2^(j/64)) followed by a short polynomial,
// ~16 / ~14 instructions, error <= 1 ulp on the ranges the kernels use (checked against libm in
// tests/test_fastmath.py); anything outside the fast range (x <= 0, subnormal, inf, NaN; |x| >= 700) goes to the
// library function, so a blown-up state still produces the NaN the callers' check looks for
// (wx_factory/simulation/simulation.py:399-408).
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define WXB_HD __host__ __device__ __forceinline__
#else
#define WXB_HD inline
#endif
#include <cmath>
#include <cstring>

#define WXB_LOGTAB_BITS 7
#define WXB_LOGTAB_N (1 << WXB_LOGTAB_BITS)
#define WXB_EXPTAB_BITS 6
#define WXB_EXPTAB_N (1 << WXB_EXPTAB_BITS)
// doubles in the combined table: [2*128] (invc, logc) pairs, then [64] 2^(j/64)
#define WXB_FASTTAB_DOUBLES (2 * WXB_LOGTAB_N + WXB_EXPTAB_N)

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

// natural logarithm; tab = the (invc, logc) pairs
WXB_HD double fast_log(double x, const double* __restrict__ tab) {
    const uint64_t ix = bits_of(x);
    const uint32_t top = (uint32_t)(ix >> 32);
    if (top - 0x00100000u >= 0x7fe00000u) return log(x);  // x <= 0, subnormal, inf, NaN
    const uint64_t tmp = ix - LOG_OFF;
    const int i = (int)((tmp >> (52 - WXB_LOGTAB_BITS)) & (WXB_LOGTAB_N - 1));
    const int k = (int)((int64_t)tmp >> 52);
    const double z = double_of(ix - (tmp & 0xfff0000000000000ULL));
    const double invc = tab[2 * i], logc = tab[2 * i + 1];
    const double r = fma(z, invc, -1.0);  // |r| < 0.004
    const double kd = (double)k;
    const double w = fma(kd, 0x1.62e42fefa3800p-1, logc);  // k*ln2_hi (exact: 11 trailing zero bits) + log c
    // log1p(r) = r - r^2/2 + r^3/3 - r^4/4 + r^5/5 - r^6/6 (next term < 2.4e-18)
    const double r2 = r * r;
    double p = fma(r, -1.0 / 6.0, 0.2);
    p = fma(r, p, -0.25);
    p = fma(r, p, 1.0 / 3.0);
    p = fma(r, p, -0.5);
    const double hi = w + r;
    const double lo = (w - hi) + r;  // what the sum w + r lost
    return fma(r2, p, fma(kd, 0x1.ef35793c76730p-45, lo)) + hi;
}

// e^x; tab = the 64 values 2^(j/64)
WXB_HD double fast_exp(double x, const double* __restrict__ tab) {
    if (!(fabs(x) < 700.0)) return exp(x);
    const double shift = 0x1.8p52;
    double kd = fma(x, 0x1.71547652b82fep6, shift);  // x * 64/ln2, rounded to an integer in the low mantissa bits
    const uint64_t ki = bits_of(kd);
    kd -= shift;
    double r = fma(kd, -0x1.62e42fefa0000p-7, x);    // ln2/64 hi
    r = fma(kd, -0x1.cf79abc9e3b3ap-46, r);          // ln2/64 lo
    const uint64_t sbits = bits_of(tab[ki & (WXB_EXPTAB_N - 1)]) + ((ki >> WXB_EXPTAB_BITS) << 52);
    const double s = double_of(sbits);
    // e^r - 1 = r + r^2/2 + r^3/6 + r^4/24 + r^5/120, |r| <= ln2/128 (next term < 3.6e-17 relative)
    double p = fma(r, 1.0 / 120.0, 1.0 / 24.0);
    p = fma(r, p, 1.0 / 6.0);
    p = fma(r, p, 0.5);
    p = fma(r * r, p, r);
    return fma(s, p, s);
}

// host: fill the combined table (long double arithmetic; each entry correctly rounded to double)
inline void fill_tables(double* tab) {
    for (int i = 0; i < WXB_LOGTAB_N; ++i) {
        const double lo = double_of(LOG_OFF + ((uint64_t)i << (52 - WXB_LOGTAB_BITS)));
        const double hi = double_of(LOG_OFF + ((uint64_t)(i + 1) << (52 - WXB_LOGTAB_BITS)));
        const long double c = 0.5L * ((long double)lo + (long double)hi);
        const double invc = (double)(1.0L / c);
        tab[2 * i] = invc;
        tab[2 * i + 1] = (double)(-logl((long double)invc));  // log of the c that invc really stands for
    }
    for (int j = 0; j < WXB_EXPTAB_N; ++j) tab[2 * WXB_LOGTAB_N + j] = (double)exp2l((long double)j / WXB_EXPTAB_N);
}

}  // namespace wxbfm
