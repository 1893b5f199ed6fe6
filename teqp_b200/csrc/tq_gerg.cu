// GERG-2004/2008 residual models: names -> term program (host) and the batched Lambda^r_nm kernel (device).
// See tq_gerg.cuh for the factorisation.  Reference factories: /root/reference/interface/CPP/model_factory_GERG.cpp:13-18,
// model classes /root/reference/include/teqp/models/GERG/GERG.hpp:902-937 (2004) and :1166-1202 (2008).
#include <algorithm>
#include <cmath>
#include <cctype>
#include <map>
#include <tuple>

#include <cstdlib>
#include <mutex>

#include "tq_common.hpp"
#include "tq_gerg.cuh"
#include "tq_gerg_iso.cuh"
#include "tq_gerg_iso_coop.cuh"
#include "tq_gerg_jit.hpp"

namespace tq {
void gergjet_launch_isoprops(const GergProgram&, const DeviceBatch&, const IsoPropsOut&, cudaStream_t);
void gergjet_launch_ATrhoXiXj(const GergProgram&, const DeviceBatch&, int, int, double*, cudaStream_t);
void gergjet_launch_ATrhoXiXjXk(const GergProgram&, const DeviceBatch&, int, int, int, double*, cudaStream_t);
void gergjet_launch_virial(const GergProgram&, int NT, int ND, const DeviceBatch&, double*, cudaStream_t);   // tq_ex_gergjet.cu
void gergjet_launch_sigma(const GergProgram&, const DeviceBatch&, const double*, double*, cudaStream_t);
void gergjet_launch_isochoric(const GergProgram&, const DeviceBatch&, const IsochoricOut&, cudaStream_t);
void gergtd_launch(const GergProgram&, int which, const DeviceBatch&, int i, int j, int k, double* out, cudaStream_t);   // tq_ex_gergtd.cu

namespace {

// ---------------------------------------------------------------------------------------------------------------------
// published coefficient tables (generated, see tools/dump_gerg_tables.cpp)
// ---------------------------------------------------------------------------------------------------------------------
struct GergFluid {
    const char* name;
    double rhoc_molm3, Tc_K, M_kgmol;
    std::vector<double> n, t, d, c, l;
};
struct GergPair {
    size_t i, j;
    double betaV, gammaV, betaT, gammaT, Fij;
    std::vector<double> n, t, d, eta, beta, gamma, epsilon;
};
#include "data/gerg_tables.inc"

constexpr int kBlock = kGergWS;

// ---------------------------------------------------------------------------------------------------------------------
// device
// ---------------------------------------------------------------------------------------------------------------------
template<int NT, int ND>
__global__ void __launch_bounds__(kBlock) gerg_ar_table_kernel(const __grid_constant__ GergProgram P, const DeviceBatch in,
                                                               double* __restrict__ out) {
    extern __shared__ double Wsh[];  // [nw][kBlock]; each thread touches only its own column -> no barriers needed
    double* W = Wsh + threadIdx.x;
    Exp2K K = Exp2K::load(in.npts >> 62);  // npts < 2^62: the shift is 0, but only at run time
    const long long npts = in.npts;
    const long long stride = (long long)gridDim.x * kBlock;
    for (long long k = (long long)blockIdx.x * kBlock + threadIdx.x; k < npts; k += stride) {
        const double T = __ldg(in.T + k);
        // the sign bit of T is 0 for every valid temperature, and thread-dependent
        K.load_thread((long long)(((unsigned)__double2hiint(T) >> 31) << 3));
        const double rho = __ldg(in.rho + k);
        for (int i = 0; i < P.nc; ++i) W[i * kBlock] = __ldg(in.x + (long long)i * in.x_stride + (in.x_bcast ? 0 : k));
        double A[NT + 1][ND + 1];
        gerg_point<NT, ND>(P, K, T, rho, W, A);
        double fn = 1.0;
#pragma unroll
        for (int n = 0; n <= NT; ++n) {
            if (n >= 2) fn *= double(n);
            double fm = 1.0;
#pragma unroll
            for (int m = 0; m <= ND; ++m) {
                if (m >= 2) fm *= double(m);
                out[(long long)(n * (ND + 1) + m) * npts + k] = (fn * fm) * A[n][m];
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// host: names -> program
// ---------------------------------------------------------------------------------------------------------------------
std::string lower(std::string s) {
    for (auto& ch : s) ch = (char)std::tolower((unsigned char)ch);
    return s;
}
double binom_real(double a, int m) {
    double r = 1.0;
    for (int k = 1; k <= m; ++k) r = r * (a - (k - 1)) / k;
    return r;
}

using RawTerm = GergTerm;  // tq_gerg_jit.hpp

class GergModel final : public Model {
public:
    GergProgram P{};
    GergSource src;  // the term list the per-model kernels are generated from (tq_gerg_jit.hpp)
    std::vector<std::string> names;
    double Rgas = 8.314472;  // GERG.hpp:1175 (2008), :906 (2004)
    std::vector<double> Ri;  // per-fluid gas constants (multifluid, mole-fraction weighted)
    std::string fam, meta_json = "{}";
    std::string meta() const override { return meta_json; }
    GergIsoDeviceTables iso_tab;   // tables of the lane-cooperative isochoric kernels (tq_gerg_iso_coop.cuh), built and uploaded at first use
    // TEQPB200_ISO_COOP_MINROWS: use the lane-cooperative form only for mixtures whose thread-per-point staging area has at least this many rows
    bool coop_pays() const {
        static const int min_rows = [] { const char* e = std::getenv("TEQPB200_ISO_COOP_MINROWS"); return e ? std::atoi(e) : 0; }();
        return gerg_iso_rows(P.nc, P.nw, true) >= min_rows;
    }
    bool taudelta_ok() const { return fam == "multifluid"; }   // the reference's GERG2008ResidualModel has no alphar_taudelta (derivs.hpp:137-140)
    int ncomp() const override { return P.nc; }
    const char* family() const override { return fam.c_str(); }
    double R(const double* x, int n) const override {
        if (!P.r_weighted) return Rgas;
        double r = 0;
        for (int i = 0; i < n && i < P.nc; ++i) r += x[i] * Ri[i];   // MoleFractionWeighted::get_R (multifluid_gas_constant.hpp)
        return r;
    }
    const void* params_blob(size_t* n) const override { *n = sizeof(P); return &P; }

    void launch_ATrhoXi_table(const DeviceBatch& in, double* out, cudaStream_t st) const override { gerg_launch_ATrhoXi(P, in, out, st); }
    void launch_ATrhoXiXj_table(const DeviceBatch& in, int i, int j, double* out, cudaStream_t st) const override { gergjet_launch_ATrhoXiXj(P, in, i, j, out, st); }
    void launch_ATrhoXiXjXk_table(const DeviceBatch& in, int i, int j, int k, double* out, cudaStream_t st) const override { gergjet_launch_ATrhoXiXjXk(P, in, i, j, k, out, st); }
    void launch_isoprops(const DeviceBatch& in, const IsoPropsOut& o, cudaStream_t st) const override {
        // analytic pass (tq_gerg_iso.cuh) where it applies; TEQPB200_GERG_ISOPROPS_JETS forces the jet sweeps (A/B and cross-check)
        static const bool jets = std::getenv("TEQPB200_GERG_ISOPROPS_JETS") != nullptr;
        if (!jets && src.all_plain_tau && !src.has_nonanalytic && !P.r_weighted && !P.pure_shortcut) {
            if (!coop_pays() || !gerg_coop_launch_isoprops(iso_tab, P, Rgas, in, o, st)) gerg_launch_isoprops(P, Rgas, in, o, st);
        }
        else gergjet_launch_isoprops(P, in, o, st);
    }
    // GERG200XReducing::get_Tr / get_rhor (GERG.hpp:179-205) on the host, same pair order and skip rule as the kernels
    bool reducing(const double* x, int n, double* Tr, double* rhor) const override {
        if (n != P.nc) throw InvalidArgument("mole fractions must be of size " + std::to_string(P.nc));
        double YT = 0, YV = 0;
        int q = 0;
        for (int i = 0; i < P.nc; ++i) {
            YT += x[i] * x[i] * P.Tc[i]; YV += x[i] * x[i] * P.vc[i];
            for (int j = i + 1; j < P.nc; ++j, ++q) {
                const GergRedPair& pr = P.pair[q];
                const double dT = pr.beta2T * x[i] + x[j], dV = pr.beta2V * x[i] + x[j];
                if (dT == 0.0 || dV == 0.0) continue;
                const double num = x[i] * x[j] * (x[i] + x[j]);
                YT += num / dT * pr.YT2; YV += num / dV * pr.YV2;
            }
        }
        *Tr = YT; *rhor = 1.0 / YV;
        return true;
    }
    void launch_virial_table(int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) const override { gergjet_launch_virial(P, NT, ND, in, out, st); }
    void launch_sigma_derivs(const DeviceBatch& in, const double* v, double* out, cudaStream_t st) const override { gergjet_launch_sigma(P, in, v, out, st); }
    void launch_isochoric(const DeviceBatch& in, const IsochoricOut& o, cudaStream_t st) const override {
        // the analytic kernel (tq_gerg_iso.cuh) knows tau^t temperature shapes and a constant R; JSON multifluid models with Gaussian /
        // GaoB / Lemmon2005 terms or the mole-fraction-weighted R go through the jet sweeps
        if (src.all_plain_tau && !src.has_nonanalytic && !P.r_weighted && !P.pure_shortcut) {
            if (!coop_pays() || !gerg_coop_launch_isochoric(iso_tab, P, Rgas, in, o, st)) gerg_launch_isochoric(P, Rgas, in, o, st);
        }
        else gergjet_launch_isochoric(P, in, o, st);
    }
    // Psir / gradient / Hessian AND the derived properties in one launch of the lane-cooperative kernel (false: not applicable, launch separately)
    bool launch_fused(const DeviceBatch& in, const IsochoricOut& o, const IsoPropsOut& ip, cudaStream_t st) const {
        static const bool jets = std::getenv("TEQPB200_GERG_ISOPROPS_JETS") != nullptr;
        if (jets || !src.all_plain_tau || src.has_nonanalytic || P.r_weighted || P.pure_shortcut || !coop_pays()) return false;
        if (o.lnphi || o.phi) return false;   // the Hessian instantiation serves Psir, grad, hess
        return gerg_coop_launch_fused(iso_tab, P, Rgas, in, o, ip, st);
    }
    void launch_Ataudelta_table(int which, const DeviceBatch& in, int i, int j, int k, double* out, cudaStream_t st) const override {
        if (!taudelta_ok()) throw NotImplementedError("Cannot take derivatives of a class that doesn't define the alphar_taudelta method");
        gergtd_launch(P, which, in, i, j, k, out, st);
    }
    template<int NT, int ND> void launch(const DeviceBatch& in, double* out, cudaStream_t st) const {
        const size_t smem = (size_t)P.nw * kBlock * sizeof(double);
        static_assert(kGergMaxW * kBlock * sizeof(double) <= 48 * 1024, "W staging must fit the default 48 KB dynamic shared memory");
        // one state point per thread, one CTA per 128 points: with ~11 000 instructions per point the CTA launch cost is noise,
        // and the hardware CTA scheduler balances the SMs (a persistent grid-stride grid left 30 % of the warp slots idle in
        // its tail: profiles/r01c)
        const int grid = grid_for(in.npts, kBlock, 1 << 30);
        gerg_ar_table_kernel<NT, ND><<<grid, kBlock, smem, st>>>(P, in, out);
        TQ_CUDA(cudaGetLastError());
    }
    // the compiled shapes; out is laid out for the REQUESTED (NT, ND), so a wider kernel writes through a scratch-free
    // remap only when shapes match exactly -- otherwise the caller (capi) asks for a compiled shape and slices.
    // ---- per-model specialised kernels (NVRTC), one per derivative shape, compiled at first use and kept for the model's lifetime ----
    struct Jit { JitSource s; JitKernel* k = nullptr; };
    mutable std::mutex jit_mtx;
    mutable std::map<std::pair<int, int>, Jit> jit;
    ~GergModel() override { for (auto& [key, j] : jit) jit_release(j.k); }
    static bool shape_ok(int NT, int ND) { return (NT == 0 && ND <= kMaxDerivD) || (ND == 0 && NT <= kMaxDerivT) || (NT <= 2 && ND <= 4); }
    const Jit& specialised(int NT, int ND) const {
        std::lock_guard<std::mutex> g(jit_mtx);
        Jit& j = jit[{NT, ND}];
        if (!j.k) {
            j.s = gerg_generate(src, NT, ND);
            j.k = jit_compile(j.s.code, "tqj_kernel", nullptr, nullptr, true);
        }
        return j;
    }
    bool ar_table_ready(int NT, int ND) const override {
        if ((NT > 0 && !src.all_plain_tau) || src.has_nonanalytic) return true;   // jet sweep: a compiled-in kernel
        std::lock_guard<std::mutex> g(jit_mtx);
        auto it = jit.find({NT, ND});
        return it != jit.end() && it->second.k != nullptr;
    }
    std::string jit_source(int NT, int ND, std::vector<double>* consts) const override {
        JitSource s = gerg_generate(src, NT, ND);
        if (consts) *consts = s.consts;
        return s.code;
    }
    std::vector<char> jit_cubin(int NT, int ND) const override {
        std::vector<char> cubin;
        jit_compile(gerg_generate(src, NT, ND).code, "tqj_kernel", nullptr, &cubin, false);
        return cubin;
    }
    void launch_ar_table(int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) const override {
        if (in.npts == 0) return;
        if (NT < 0 || ND < 0 || !shape_ok(NT, ND))
            throw InvalidArgument("derivative table (" + std::to_string(NT) + "," + std::to_string(ND) + ") is not a supported shape for " + fam);
        if ((NT > 0 && !src.all_plain_tau) || src.has_nonanalytic) {
            // temperature derivatives of Gaussian / GaoB / Lemmon2005 temperature shapes are not binomial factors: jet sweep
            gergjet_launch_virial(P, NT, ND, in, out, st);
            return;
        }
        static const bool interp = std::getenv("TEQPB200_GERG_INTERPRETER") != nullptr;  // A/B measurements only (profiles/)
        if (!interp || P.ntf > 0 || P.pure_shortcut) {
            const Jit& j = specialised(NT, ND);
            jit_launch_table(j.k, j.s.consts, j.s.w_in_smem ? (size_t)j.s.nw * 128 * sizeof(double) : 0, in, out, st);
            return;
        }
#define TQ_CASE(a, b) if (NT == a && ND == b) { launch<a, b>(in, out, st); return; }
        TQ_CASE(0, 0) TQ_CASE(0, 1) TQ_CASE(0, 2) TQ_CASE(0, 3) TQ_CASE(0, 4) TQ_CASE(0, 5) TQ_CASE(0, 6)
        TQ_CASE(1, 0) TQ_CASE(2, 0) TQ_CASE(3, 0) TQ_CASE(4, 0)
        TQ_CASE(1, 1) TQ_CASE(1, 2) TQ_CASE(2, 1) TQ_CASE(2, 2) TQ_CASE(1, 3) TQ_CASE(1, 4) TQ_CASE(2, 3) TQ_CASE(2, 4)
#undef TQ_CASE
        throw InvalidArgument("derivative table (" + std::to_string(NT) + "," + std::to_string(ND) + ") is not a compiled shape for " + fam);
    }
};

template<class F, class PR>
std::unique_ptr<Model> build_gerg(const char* fam, const std::vector<std::string>& names_in, const F* fluids, size_t nfluids, const PR* pairs, size_t npairs) {
    TermProgramInput in;
    in.family = fam;
    in.Rgas = 8.314472;   // GERG.hpp:1175 (2008), :906 (2004)
    const size_t N = names_in.size();
    if (N == 0) throw InvalidArgument("names must not be empty");
    std::vector<size_t> idx(N);
    for (size_t i = 0; i < N; ++i) {
        const std::string nm = lower(names_in[i]);
        size_t f = nfluids;
        for (size_t q = 0; q < nfluids; ++q) if (nm == fluids[q].name) { f = q; break; }
        if (f == nfluids) throw InvalidArgument("Unable to load pure info for " + names_in[i]);
        idx[i] = f;
        in.names.push_back(nm);
        in.Tc.push_back(fluids[f].Tc_K);
        in.vc.push_back(1.0 / fluids[f].rhoc_molm3);
    }
    auto find_pair = [&](size_t a, size_t b) -> const PR& {
        for (size_t q = 0; q < npairs; ++q) if (pairs[q].i == a && pairs[q].j == b) return pairs[q];
        throw InvalidArgument(std::string("Unable to obtain BIP for the pair {") + fluids[a].name + "," + fluids[b].name + "}");
    };
    // pure-fluid terms (GERG.hpp:117-119)
    for (size_t i = 0; i < N; ++i) {
        const F& fl = fluids[idx[i]];
        for (size_t k = 0; k < fl.n.size(); ++k) {
            RawTerm r{};
            r.w = (int)i; r.n = fl.n[k]; r.t = fl.t[k];
            r.d = (int)fl.d[k]; r.l = (int)fl.l[k]; r.c = fl.c[k];
            r.kind = (r.c == 0.0) ? GG_POLY : GG_EXP;
            if (r.kind == GG_POLY) r.l = 0;
            in.pure.push_back(r);
        }
    }
    // reducing pairs and departure terms
    for (size_t i = 0; i < N; ++i) {
        for (size_t j = i + 1; j < N; ++j) {
            if (idx[i] == idx[j]) throw InvalidArgument("duplicate component " + names_in[i]);
            const bool fwd = idx[i] < idx[j];
            const PR& pr = fwd ? find_pair(idx[i], idx[j]) : find_pair(idx[j], idx[i]);
            TermProgramInput::Pair q;
            // reversed order: reciprocal betas (GERG.hpp:757-763)
            q.betaT = fwd ? pr.betaT : 1 / pr.betaT; q.betaV = fwd ? pr.betaV : 1 / pr.betaV;
            q.gammaT = pr.gammaT; q.gammaV = pr.gammaV; q.F = pr.Fij;
            if (pr.Fij != 0.0) {
                for (size_t k = 0; k < pr.n.size(); ++k) {
                    RawTerm r{};
                    r.n = pr.n[k]; r.t = pr.t[k]; r.d = (int)pr.d[k];
                    r.eta = pr.eta[k]; r.eps = pr.epsilon[k]; r.beta = pr.beta[k]; r.gamma = pr.gamma[k];
                    if (r.eta == 0.0 && r.beta == 0.0) { r.kind = GG_POLY; r.eps = r.gamma = 0; }
                    else r.kind = GG_GAUSS;
                    q.terms.push_back(r);
                }
            }
            in.pairs.push_back(std::move(q));
        }
    }
    return make_term_program_model(in);
}

}  // namespace

// Compiles a term list into the device term program (groups x temperature shapes x terms) shared by every kernel of the family
std::unique_ptr<Model> make_term_program_model(const TermProgramInput& in) {
    auto m = std::make_unique<GergModel>();
    m->fam = in.family;
    m->names = in.names;
    m->meta_json = in.meta;
    GergProgram& P = m->P;
    const size_t N = in.names.size();
    if (N == 0) throw InvalidArgument("at least one component is needed");
    if (N > (size_t)kGergMaxC) throw InvalidArgument("too many components (the device term program holds " + std::to_string(kGergMaxC) + ")");
    if (in.pairs.size() != N * (N - 1) / 2) throw InvalidArgument("internal: pair list does not match the component count");
    P.nc = (int)N;
    m->Rgas = in.Rgas;
    P.Rgas = in.Rgas;
    P.r_weighted = in.r_weighted ? 1 : 0;
    P.limit_form = in.limit_form ? 1 : 0;
    P.pure_shortcut = (in.pure_shortcut && N == 1) ? 1 : 0;
    m->Ri = in.Ri;
    for (size_t i = 0; i < N; ++i) { P.Tc[i] = in.Tc[i]; P.vc[i] = in.vc[i]; if (in.r_weighted) P.Ri[i] = in.Ri.at(i); }
    if (in.na.size() > (size_t)kGergMaxNA) throw InvalidArgument("too many non-analytic terms for the device term program");
    P.nna = (int)in.na.size();
    for (size_t q = 0; q < in.na.size(); ++q) P.na[q] = in.na[q];

    std::vector<RawTerm> raw = in.pure;
    int pidx = 0, ndep = 0;
    for (size_t i = 0; i < N; ++i) {
        for (size_t j = i + 1; j < N; ++j, ++pidx) {
            const TermProgramInput::Pair& pr = in.pairs[pidx];
            const double Tci = P.Tc[i], Tcj = P.Tc[j], vci = P.vc[i], vcj = P.vc[j];
            const double YT = pr.betaT * pr.gammaT * std::sqrt(Tci * Tcj);
            const double s = std::cbrt(vci) + std::cbrt(vcj);
            const double YV = 1.0 / 8.0 * pr.betaV * pr.gammaV * (s * s * s);
            P.pair[pidx].beta2T = pr.betaT * pr.betaT; P.pair[pidx].YT2 = 2.0 * YT;
            P.pair[pidx].beta2V = pr.betaV * pr.betaV; P.pair[pidx].YV2 = 2.0 * YV;
            if (pr.F != 0.0 && !pr.terms.empty()) {
                if (ndep >= kGergMaxDep) throw InvalidArgument("too many departure pairs");
                P.dep_i[ndep] = (unsigned char)i; P.dep_j[ndep] = (unsigned char)j; P.depF[ndep] = pr.F;
                for (RawTerm r : pr.terms) { r.w = (int)N + ndep; raw.push_back(r); }
                ++ndep;
            }
        }
    }
    P.npair = pidx; P.ndep = ndep; P.nw = (int)N + ndep;
    {
        GergSource& S = m->src;
        S.nc = (int)N; S.ndep = ndep;
        S.pure_shortcut = P.pure_shortcut != 0;
        S.Tc.assign(P.Tc, P.Tc + N); S.vc.assign(P.vc, P.vc + N);
        S.pair.assign(P.pair, P.pair + pidx);
        for (int q = 0; q < ndep; ++q) { S.dep_i.push_back(P.dep_i[q]); S.dep_j.push_back(P.dep_j[q]); S.depF.push_back(P.depF[q]); }
        S.terms = raw;
        for (const auto& r : raw) if (r.tk != GT_PLAIN) S.all_plain_tau = false;
        S.has_nonanalytic = P.nna > 0;
    }

    // non-plain temperature shapes
    std::map<GergTauKey, int> tfidx;   // (ignoring t) -> 1 + index into P.tf
    auto tf_of = [&](const RawTerm& r) -> int {
        if (r.tk == GT_PLAIN) return 0;
        const GergTauKey k{0.0, r.tk, r.tp1, r.tp2, r.tp3};
        auto it = tfidx.find(k);
        if (it != tfidx.end()) return it->second;
        if (P.ntf >= kGergMaxTF) throw InvalidArgument("too many distinct temperature shapes (Gaussian / GaoB / Lemmon2005 terms) for the device term program");
        P.tf[P.ntf] = GergTauFn{r.tk, 0, r.tp1, r.tp2, r.tp3};
        tfidx[k] = ++P.ntf;
        return P.ntf;
    };

    // group by density shape, then by temperature shape
    // key = (exponent shape h, then d): groups that share h are consecutive, d ascending (running product of delta)
    using GKey = std::tuple<int, int, double, double, double, double, double, int>;
    std::map<GKey, std::map<GergTauKey, std::vector<const RawTerm*>>> groups;
    for (const auto& r : raw) groups[GKey{r.kind, r.l, r.c, r.eta, r.eps, r.beta, r.gamma, r.d}][tau_key(r)].push_back(&r);
    if (groups.size() > (size_t)kGergMaxGroup) throw InvalidArgument("too many distinct density groups (" + std::to_string(groups.size()) + ") for the device term program");
    int g = 0, tq = 0, term = 0, rec = 0;
    for (const auto& [key, bytau] : groups) {
        GergGroup& G = P.group[g];
        G.kind = std::get<0>(key); G.l = std::get<1>(key); G.d = std::get<7>(key);
        G.c = std::get<2>(key); G.eta = std::get<3>(key); G.eps = std::get<4>(key); G.beta = std::get<5>(key); G.gamma = std::get<6>(key);
        G.new_h = 1;
        G.dstep = G.d;
        if (g > 0) {
            const GergGroup& Q = P.group[g - 1];
            if (Q.kind == G.kind && Q.l == G.l && Q.c == G.c && Q.eta == G.eta && Q.eps == G.eps && Q.beta == G.beta && Q.gamma == G.gamma && Q.d <= G.d) { G.new_h = 0; G.dstep = G.d - Q.d; }
        }
        for (int mm = 1; mm <= kGergND; ++mm) { G.bd[mm - 1] = binom_real(G.d, mm); G.bl[mm - 1] = binom_real(G.l, mm); }
        G.ntq = (int)bytau.size();
        for (const auto& [tkey, terms] : bytau) {
            const double t = std::get<0>(tkey);
            const int rest = (int)terms.size() - 1, npair = (rest + 1) / 2;
            if (rec + 2 + 2 * npair > kGergMaxRec) throw InvalidArgument("too many terms for the device term program");
            if (!(std::fabs(t) <= kGergMaxAbsT)) throw InvalidArgument("temperature exponent out of the device term program's range");
            P.rec[rec] = GergRec{t, npair, terms[0]->w * kGergWS * 8};
            P.rec[rec + 1] = GergRec{terms[0]->n, tf_of(*terms[0]), 0};
            rec += 2;
            for (int k = 0; k < 2 * npair; ++k, ++rec)
                P.rec[rec] = (k < rest) ? GergRec{terms[k + 1]->n, terms[k + 1]->w * kGergWS * 8, 0} : GergRec{0.0, 0, 0};
            term += (int)terms.size();
            ++tq;
        }
        ++g;
    }
    P.ngroup = g; P.ntq = tq; P.nterm = term; P.nrec = rec;
    return m;
}

namespace {
}  // namespace

// tq_capi.cu (build_Psir_fgradHessian_and_properties_batch): true when the model is a GERG / multifluid term-program model and the fused launch ran
bool gerg_try_launch_fused(const Model& m, const DeviceBatch& in, const IsochoricOut& o, const IsoPropsOut& ip, cudaStream_t st) {
    const GergModel* gm = dynamic_cast<const GergModel*>(&m);
    return gm && gm->launch_fused(in, o, ip, st);
}

std::unique_ptr<Model> make_gerg_model(const std::string& kind, const Json& spec) {
    const std::vector<std::string> names = spec.at("names").as_strings();
    if (kind == "GERG2008resid")
        return build_gerg("GERG2008resid", names, gerg2008_fluids, sizeof(gerg2008_fluids) / sizeof(gerg2008_fluids[0]), gerg2008_pairs, sizeof(gerg2008_pairs) / sizeof(gerg2008_pairs[0]));
    if (kind == "GERG2004resid")
        return build_gerg("GERG2004resid", names, gerg2004_fluids, sizeof(gerg2004_fluids) / sizeof(gerg2004_fluids[0]), gerg2004_pairs, sizeof(gerg2004_pairs) / sizeof(gerg2004_pairs[0]));
    throw InvalidArgument("unknown GERG kind " + kind);
}

}  // namespace tq
