// Scalar FP64 helpers for the hot kernels: exp and reciprocal with their constants as direct constant-bank operands.
//
// Why not ::exp: libdevice's exp materialises its 11 polynomial coefficients as 64-bit immediates, which on sm_100a become
// two UMOV each -- 22 extra issue slots per call.  The GERG kernel calls exp ~140 times per state point and was issue-bound
// on exactly that (profiles/r01a: UMOV 18 % of all warp instructions).  fast_exp keeps the coefficients in __constant__
// memory, i.e. as direct c[bank][offset] operands of the DFMAs, and has no slow path and no branches.
//
// Accuracy: Cody-Waite reduction with a two-part ln2 (FMA), degree-11 near-minimax polynomial with c0 = c1 = 1
// (tools/fit_exp_poly.py: 1.7e-17 relative with exact coefficients), result within 1 ulp of exp(x) over [-708, 709].
// Semantics at the edges: x < -708 returns exp(-708) ~ 3e-308 instead of a denormal/zero (a relative "error" that only matters
// below 1e-307, far under the absolute floor of the parity rule); x > 709 returns +inf; NaN propagates.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

#include "tq_jet.cuh"

namespace tq {

TQ_HD double bits_add_exponent(double p, int n) {
#if defined(__CUDA_ARCH__)
    const int hi = __double2hiint(p) + (n << 20);
    return __hiloint2double(hi, __double2loint(p));
#else
    uint64_t u;
    std::memcpy(&u, &p, 8);
    u += (uint64_t)(int64_t)n << 52;
    std::memcpy(&p, &u, 8);
    return p;
#endif
}
TQ_HD int bits_lo32(double t) {
#if defined(__CUDA_ARCH__)
    return __double2loint(t);
#else
    uint64_t u;
    std::memcpy(&u, &t, 8);
    return (int)(uint32_t)u;
#endif
}

// exp coefficients live in constant memory on the device so that they are direct operands of the DFMAs (a literal would be
// rebuilt from two 32-bit immediates before every use)
#define TQ_EXP_COEFFS                                                                                                   \
    { 0x1.71547652b82fep+0, 0x1.62e42fefa39efp-1, 0x1.abc9e3b39803fp-56, 0x1.af38a9b0ec855p-26, 0x1.289185613a3d6p-22,  \
      0x1.71de0dae63bb3p-19, 0x1.a019b90d2ae7ap-16, 0x1.a01a01a7c41d5p-13, 0x1.6c16c1788bd90p-10, 0x1.11111111109b3p-7, \
      0x1.5555555553d63p-5, 0x1.5555555555556p-3, 0x1.0000000000001p-1, 0x1.8p+52, -708.0, 709.0 }
#if defined(__CUDACC__)
static __constant__ double kExpC_dev[16] = TQ_EXP_COEFFS;
#endif
static const double kExpC_host[16] = TQ_EXP_COEFFS;

// The coefficients as a value type: a kernel loads it ONCE into registers (ExpK::load) and passes it down, so the
// polynomial's operands are registers instead of ~9 constant loads per call.
struct ExpK {
    double k[16];
    // opaque_zero: any run-time value that is 0 but that the compiler cannot prove to be 0.  With a compile-time-known
    // address ptxas re-materialises the constant loads before every use (9 loads per exp); an address it cannot fold keeps
    // the 13 coefficients resident in (uniform) registers for the whole kernel.
    TQ_HD static ExpK load(long long opaque_zero = 0) {
        ExpK r;
#if defined(__CUDA_ARCH__)
        const size_t base = __cvta_generic_to_constant(kExpC_dev) + (size_t)opaque_zero;
#pragma unroll
        for (int i = 0; i < 16; ++i) asm volatile("ld.const.f64 %0, [%1];" : "=d"(r.k[i]) : "l"(base + 8 * i));
#else
        for (int i = 0; i < 16; ++i) r.k[i] = kExpC_host[i];
#endif
        return r;
    }
};

TQ_HD double fast_exp(double x, const ExpK& K) {
    const double xc = (x < -708.0) ? -708.0 : x;
    const double t = fma(xc, K.k[0], 0x1.8p+52);  // x/ln2 rounded to nearest integer, in the low mantissa bits
    const double nf = t - 0x1.8p+52;
    double r = fma(-nf, K.k[1], xc);
    r = fma(-nf, K.k[2], r);
    double p = K.k[3];
#pragma unroll
    for (int k = 4; k <= 12; ++k) p = fma(p, r, K.k[k]);
    p = fma(p, r, 1.0);
    p = fma(p, r, 1.0);
    p = bits_add_exponent(p, bits_lo32(t));
    return (x > 709.0) ? INFINITY : p;
}
TQ_HD double fast_exp(double x) { return fast_exp(x, ExpK::load()); }

// 1/a for a normal, finite, non-zero a (the callers guard zero): hardware seed + two Newton steps, <= 1 ulp
TQ_HD double fast_rcp(double a) {
#if defined(__CUDA_ARCH__)
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(a));
    double e = fma(-a, y, 1.0);
    y = fma(y, e, y);
    e = fma(-a, y, 1.0);
    y = fma(y, e, y);
    return y;
#else
    return 1.0 / a;
#endif
}

}  // namespace tq
