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
      0x1.5555555553d63p-5, 0x1.5555555555556p-3, 0x1.0000000000001p-1, 0x1.8p+52, -708.0, 709.0,                      \
      /* [16..26]: 2^r = 1 + c1 r + .. + c11 r^11, |r| <= 1/2 (same fit, tools/fit_exp_poly.py) */                    \
      0x1.62e42fefa39efp-1, 0x1.ebfbdff82c590p-3, 0x1.c6b08d704a0c0p-5, 0x1.3b2ab6fba385bp-7, 0x1.5d87fe78a5dc3p-10,  \
      0x1.4309130379f8bp-13, 0x1.ffcbfc61f8673p-17, 0x1.62bfe45ac08ccp-20, 0x1.b525087314718p-24,                      \
      0x1.e5ea03c5ae3ccp-28, 0x1.e9bbe9e45e5a6p-32, 0.0, 0.0, 0.0, 0.0, 0.0 }
#if defined(__CUDACC__)
static __constant__ double kExpC_dev[32] = TQ_EXP_COEFFS;
#endif
static const double kExpC_host[32] = TQ_EXP_COEFFS;

// exp(x), coefficients read from constant memory at the point of use (fine for the few calls per state point that need the
// full argument range; the hot 2^y below keeps its coefficients in registers instead)
TQ_HD double fast_exp(double x) {
#if defined(__CUDA_ARCH__)
    const double* K = kExpC_dev;
#else
    const double* K = kExpC_host;
#endif
    const double xc = (x < -708.0) ? -708.0 : x;
    const double t = fma(xc, K[0], 0x1.8p+52);  // x/ln2 rounded to nearest integer, in the low mantissa bits
    const double nf = t - 0x1.8p+52;
    double r = fma(-nf, K[1], xc);
    r = fma(-nf, K[2], r);
    double p = K[3];
#pragma unroll
    for (int k = 4; k <= 12; ++k) p = fma(p, r, K[k]);
    p = fma(p, r, 1.0);
    p = fma(p, r, 1.0);
    p = bits_add_exponent(p, bits_lo32(t));
    return (x > 709.0) ? INFINITY : p;
}

// The 2^r coefficients as a value type: a kernel loads it ONCE (Exp2K::load) and passes it down, so that the polynomial's
// operands stay resident in (uniform) registers instead of being re-loaded from constant memory before every use.
struct Exp2K {
    double c[11];
    double magic;  // 1.5 * 2^52 in a REGISTER: as an immediate it cannot share a DFMA with a constant-bank operand (t), which costs
                   // two moves per call; c[10] is likewise wanted in a (vector) register so that the Horner chain can start with
                   // fma(c10, r, c9) without first copying c10 into the accumulator
    // opaque_zero: any run-time value that is 0 but that the compiler cannot prove to be 0.  With a compile-time-known
    // address ptxas re-materialises the constant loads before every use (~6 loads per call); an address it cannot fold
    // keeps the coefficients in registers for the whole kernel.  opaque_zero_thread: the same, but thread-dependent, so that the
    // two values loaded through it live in vector registers rather than uniform ones.
    TQ_HD static Exp2K load(long long opaque_zero = 0) {
        Exp2K r;
#if defined(__CUDA_ARCH__)
        const size_t base = __cvta_generic_to_constant(kExpC_dev + 16) + (size_t)opaque_zero;
#pragma unroll
        for (int i = 0; i < 11; ++i) asm volatile("ld.const.f64 %0, [%1];" : "=d"(r.c[i]) : "l"(base + 8 * i));
        asm volatile("ld.const.f64 %0, [%1];" : "=d"(r.magic) : "l"(base - 8 * 3));
#else
        for (int i = 0; i < 11; ++i) r.c[i] = kExpC_host[16 + i];
        r.magic = kExpC_host[13];
#endif
        return r;
    }
    // re-load c[10] and magic through a THREAD-dependent zero offset (see above): call once per thread
    TQ_HD void load_thread(long long opaque_zero_thread) {
#if defined(__CUDA_ARCH__)
        const size_t tb = __cvta_generic_to_constant(kExpC_dev) + (size_t)opaque_zero_thread;
        asm volatile("ld.const.f64 %0, [%1];" : "=d"(c[10]) : "l"(tb + 8 * 26));
        asm volatile("ld.const.f64 %0, [%1];" : "=d"(magic) : "l"(tb + 8 * 13));
#else
        (void)opaque_zero_thread;
#endif
    }
};

// 2^y for |y| < 1000 (NOT checked: the caller bounds its argument).  r = y - rint(y) is exact, so the whole error is the
// polynomial's (1.7e-17) plus Horner rounding: <= 1 ulp.  14 FP64 instructions, no branches.
TQ_HD double fast_exp2_bounded(double y, const Exp2K& K) {
    const double t = y + K.magic;
    const double r = y - (t - K.magic);
    double p = fma(K.c[10], r, K.c[9]);
#pragma unroll
    for (int k = 8; k >= 0; --k) p = fma(p, r, K.c[k]);
    p = fma(p, r, 1.0);
    return bits_add_exponent(p, bits_lo32(t));
}
// 2^(a*b): the product is fused into the rounding step (one DFMA for a*b + magic), the reduced argument r = a*b - n is then
// exact to one rounding of a*b (fma(a, b, -n))
TQ_HD double fast_exp2_prod(double a, double b, const Exp2K& K) {
    const double t = fma(a, b, K.magic);
    const double r = fma(a, b, -(t - K.magic));
    double p = fma(K.c[10], r, K.c[9]);
#pragma unroll
    for (int k = 8; k >= 0; --k) p = fma(p, r, K.c[k]);
    p = fma(p, r, 1.0);
    return bits_add_exponent(p, bits_lo32(t));
}

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
