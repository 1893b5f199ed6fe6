// Per-model specialised kernels for the GERG-2004/2008 (and JSON multifluid) residual Helmholtz energy: when a model is built, its
// term list is turned into STRAIGHT-LINE sm_100a code -- one CUDA C source per (model, derivative shape), compiled with NVRTC to a
// cubin and loaded into the process (tq_jit.cu).  This replaces the term-program interpreter of tq_gerg.cuh on the hot path:
// ncu r01k showed 9 981 instructions per state point of which only 4 103 were FP64 -- the rest (LDCU 879, LDC 663, IMAD 800,
// BRA 634, UIADD3 576) was the interpreter walking the record stream.  In the generated code every coefficient is a constant-bank
// operand of the DFMA that uses it, every loop is gone, zero binomials are never multiplied, and (for the density-only tables,
// NT = 0) tau^t is evaluated once per DISTINCT exponent t instead of once per (density group, t).
//
// Mathematics: identical to tq_gerg.cuh (reference: /root/reference/include/teqp/models/GERG/GERG.hpp:103-122, 179-205, 257-274,
// 322-340, 1178-1191).  With the scaled perturbation delta (1 + s):
//     alphar = sum_shapes  E_h(s) * B_h(s),   E_h = Taylor coefficients of exp(h(delta (1+s))),
//     B_h[j] = sum_{groups g of shape h} binom(d_g, j) delta^d_g C_g,      C_g = sum_t tau^t sum_{terms (g,t)} n W_w(x)
#pragma once
#include <memory>
#include <string>
#include <tuple>
#include <vector>

#include "tq_gerg.cuh"

namespace tq {

struct GergTerm {
    int w;        // weight index: component i, or nc + q for departure pair q
    double n, t;
    int kind;     // GergGroupKind (density shape)
    int d, l;
    double c, eta, eps, beta, gamma;
    int tk = 0;   // GergTauKind (temperature shape), parameters tp1..tp3 (tq_gerg.cuh)
    double tp1 = 0, tp2 = 0, tp3 = 0;
};
using GergTauKey = std::tuple<double, int, double, double, double>;  // t, tk, tp1, tp2, tp3
inline GergTauKey tau_key(const GergTerm& r) { return GergTauKey{r.t, r.tk, r.tp1, r.tp2, r.tp3}; }

// everything the generator needs, in the reference's order (pairs lexicographic i<j)
struct GergSource {
    int nc = 0, ndep = 0;
    bool pure_shortcut = false;   // single-component multifluid: the mole fraction does not multiply the sum (multifluid.hpp:172-175)
    bool all_plain_tau = true;    // no term has a non-plain temperature shape (then temperature derivatives are binomial factors)
    bool has_nonanalytic = false; // the model has non-analytic terms: jet sweeps only
    std::vector<double> Tc, vc;
    std::vector<GergRedPair> pair;
    std::vector<int> dep_i, dep_j;
    std::vector<double> depF;
    std::vector<GergTerm> terms;
};

struct JitSource {
    std::string code;            // one translation unit: compiles under NVRTC (kernel `tqj_kernel`) and under g++ (function `tqj_host`)
    std::vector<double> consts;  // the constant pool the kernel receives as its __grid_constant__ parameter
    int rows_out = 0;            // (NT+1)*(ND+1)
    int nexp = 0, nfma_terms = 0, nshapes = 0, ngroups = 0;  // statistics
    bool w_in_smem = false;      // weights staged in shared memory (large mixtures) instead of registers
    int nw = 0;
};

// straight-line source for the Lambda^r table of shape (NT, ND).  NT == 0: tau^t once per distinct t, per-group accumulators in
// registers; NT > 0: groups one after the other, tau^t per (group, t), temperature orders by t (t-1) ... /n!
JitSource gerg_generate(const GergSource& src, int NT, int ND);

// What a corresponding-states term-program model is built from (GERG-2004/2008 tables, or the JSON-driven multifluid kind)
struct TermProgramInput {
    std::string family;
    std::vector<std::string> names;
    std::vector<double> Tc, vc;          // reducing temperature, reducing molar volume (1/rho_r)
    struct Pair {                        // pairs (i < j) in lexicographic order, betas oriented for (i, j)
        double betaT = 1, gammaT = 1, betaV = 1, gammaV = 1, F = 0;
        std::vector<GergTerm> terms;     // departure function terms (w is filled in by the builder)
    };
    std::vector<Pair> pairs;
    std::vector<GergTerm> pure;          // w = component index
    std::vector<GergNA> na;              // non-analytic terms (w = component index)
    double Rgas = 0;                     // constant gas constant ...
    std::vector<double> Ri;              // ... or per-fluid values, mole-fraction weighted when r_weighted
    bool r_weighted = false, limit_form = false, pure_shortcut = false;
    std::string meta = "{}";
};
std::unique_ptr<Model> make_term_program_model(const TermProgramInput& in);

// ---- NVRTC runtime (tq_jit.cu) ------------------------------------------------------------------------------------------
struct JitKernel;  // opaque: a loaded cubin + kernel handle
JitKernel* jit_compile(const std::string& code, const char* entry, std::string* log_out, std::vector<char>* cubin_out, bool load);
void jit_release(JitKernel*);
// launches entry(consts, npts, T, rho, x, x_stride, x_bcast, out) with one thread per point, 128 threads per CTA
void jit_launch_table(const JitKernel*, const std::vector<double>& consts, size_t smem_bytes, const DeviceBatch& in, double* out, cudaStream_t st);
int jit_kernel_regs(const JitKernel*);

}  // namespace tq
