// SAFT-VR-Mie and genericSAFT (SAFT-VR-Mie + association): JSON -> device parameter blocks, and the generic jet-sweep launches.
// Host tables follow the reference's constructors: SAFTVRMieChainContributionTerms (/root/reference/include/teqp/models/saftvrmie.hpp:
// 106-237, 263-281), SAFTVRMieNonpolarfactory (:1117-1154), Association::factory / make_mapper / make_D
// (/root/reference/include/teqp/models/association/association.hpp:71-151, 185-262), DufalData::apply_mixing_rules
// (association_types.hpp:57-74), GenericSAFT (/root/reference/include/teqp/models/saft/genericsaft.hpp:20-62).
#include <algorithm>
#include <cmath>
#include <map>
#include <set>

#include "tq_assoc.cuh"
#include "tq_pcassoc.cuh"
#include <cstdlib>

#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_mie.cuh"

namespace tq {

// kernels: tq_gsaft_k{1..4}.cu (one translation unit per jet-shape group, compiled in parallel)
void gsaft_launch_ar_rho(const GsaftParams&, int ND, const DeviceBatch&, double*, cudaStream_t);
void gsaft_launch_ar_mixed(const GsaftParams&, int NT, int ND, const DeviceBatch&, double*, cudaStream_t);
void gsaft_launch_ATrhoXi(const GsaftParams&, const DeviceBatch&, double*, cudaStream_t);
void gsaft_launch_isochoric(const GsaftParams&, const DeviceBatch&, const IsochoricOut&, cudaStream_t);
void gsaft_launch_ATrhoXi_part0(const GsaftParams&, const DeviceBatch&, double*, cudaStream_t);   // tq_ax_gsaft_{0,1,2}.cu
void gsaft_launch_ATrhoXi_part1(const GsaftParams&, const DeviceBatch&, double*, cudaStream_t);
void gsaft_launch_ATrhoXi_part2(const GsaftParams&, const DeviceBatch&, double*, cudaStream_t);
void gsaft_launch_cross2(const GsaftParams&, const DeviceBatch&, int, int, int, double*, cudaStream_t);   // tq_xx_gsaft_{2,3}.cu (directional jets)
void gsaft_launch_cross3(const GsaftParams&, const DeviceBatch&, int, int, int, double*, cudaStream_t);
void gsaft_launch_sigma(const GsaftParams&, const DeviceBatch&, const double*, double*, cudaStream_t);   // tq_ex_gsaft.cu
void gsaft_launch_ip_part0(const GsaftParams&, const DeviceBatch&, const IsoPropsOut&, double*, const double*, cudaStream_t);   // tq_ip_gsaft_{L,C,X}.cu
void gsaft_launch_ip_part1(const GsaftParams&, const DeviceBatch&, const IsoPropsOut&, double*, const double*, cudaStream_t);
void gsaft_launch_ip_part2(const GsaftParams&, const DeviceBatch&, const IsoPropsOut&, double*, const double*, cudaStream_t);

void fill_pcsaft(PcsaftParams&, const Json&);   // tq_pcsaft.cu
// genericSAFT with a PC-SAFT non-polar term: tq_pcassoc_k{1,2,4}.cu, tq_ax_pcassoc_{0,1,2}.cu, tq_ip_pcassoc_{L,C,X}.cu, tq_xx_pcassoc_{2,3}.cu, tq_ex_pcassoc.cu
void pcassoc_launch_ar_rho(const PcAssocParams&, int ND, const DeviceBatch&, double*, cudaStream_t);
void pcassoc_launch_ar_mixed(const PcAssocParams&, int NT, int ND, const DeviceBatch&, double*, cudaStream_t);
void pcassoc_launch_isochoric(const PcAssocParams&, const DeviceBatch&, const IsochoricOut&, cudaStream_t);
void pcassoc_launch_ATrhoXi_part0(const PcAssocParams&, const DeviceBatch&, double*, cudaStream_t);
void pcassoc_launch_ATrhoXi_part1(const PcAssocParams&, const DeviceBatch&, double*, cudaStream_t);
void pcassoc_launch_ATrhoXi_part2(const PcAssocParams&, const DeviceBatch&, double*, cudaStream_t);
void pcassoc_launch_cross2(const PcAssocParams&, const DeviceBatch&, int, int, int, double*, cudaStream_t);
void pcassoc_launch_cross3(const PcAssocParams&, const DeviceBatch&, int, int, int, double*, cudaStream_t);
void pcassoc_launch_sigma(const PcAssocParams&, const DeviceBatch&, const double*, double*, cudaStream_t);
void pcassoc_launch_ip_part0(const PcAssocParams&, const DeviceBatch&, const IsoPropsOut&, double*, const double*, cudaStream_t);
void pcassoc_launch_ip_part1(const PcAssocParams&, const DeviceBatch&, const IsoPropsOut&, double*, const double*, cudaStream_t);
void pcassoc_launch_ip_part2(const PcAssocParams&, const DeviceBatch&, const IsoPropsOut&, double*, const double*, cudaStream_t);

namespace {

#include "data/dufal_tables.inc"

// Lafitte 2013, Table 3 (phi) and Eq. A18 matrix (A): published constants of the model
const double kPhi[7][7] = {{7.5365557, -359.44, 1550.9, -1.19932, -1911.28, 9236.9, 10},
                           {-37.60463, 1825.6, -5070.1, 9.063632, 21390.175, -129430, 10},
                           {71.745953, -3168.0, 6534.6, -17.9482, -51320.7, 357230, 0.57},
                           {-46.83552, 1884.2, -3288.7, 11.34027, 37064.54, -315530, -6.7},
                           {-2.467982, -0.82376, -2.7171, 20.52142, 1103.742, 1390.2, -8},
                           {-0.50272, -3.1935, 2.0883, -56.6377, -3264.61, -4518.2, 0},
                           {8.0956883, 3.7090, 0, 40.53683, 2556.181, 4241.6, 0}};
const double kA[4][4] = {{0.81096, 1.7888, -37.578, 92.284}, {1.0205, -19.341, 151.26, -463.50}, {-1.9057, 22.845, -228.14, 973.92}, {1.0885, -6.1962, 106.98, -677.64}};

struct LibEntry { const char* name; double m, sigma_m, eps, lr, la; };
const LibEntry kLibrary[] = {  // saftvrmie.hpp:50-54
    {"Methane", 1.0000, 3.7412e-10, 153.36, 12.650, 6}, {"Ethane", 1.4373, 3.7257e-10, 206.12, 12.400, 6}, {"Propane", 1.6845, 3.9056e-10, 239.89, 13.006, 6}};

class GsaftModel final : public Model {
public:
    GsaftParams p{};
    std::string fam;
    int ncomp() const override { return p.mie.nc; }
    const char* family() const override { return fam.c_str(); }
    double R(const double*, int) const override { return p.mie.R; }
    const void* params_blob(size_t* n) const override { *n = sizeof(p); return &p; }
    void launch_ar_table(int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) const override {
        generic_check_ar_shape(NT, ND);
        if (in.npts == 0) return;
        if (NT == 0) gsaft_launch_ar_rho(p, ND, in, out, st);
        else gsaft_launch_ar_mixed(p, NT, ND, in, out, st);
    }
    void launch_virial_table(int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) const override { launch_ar_table(NT, ND, in, out, st); }
    void launch_sigma_derivs(const DeviceBatch& in, const double* v, double* out, cudaStream_t st) const override { gsaft_launch_sigma(p, in, v, out, st); }
    void launch_ATrhoXiXj_table(const DeviceBatch& in, int i, int j, double* out, cudaStream_t st) const override { gsaft_launch_cross2(p, in, i, j, -1, out, st); }
    void launch_ATrhoXiXjXk_table(const DeviceBatch& in, int i, int j, int k, double* out, cudaStream_t st) const override { gsaft_launch_cross3(p, in, i, j, k, out, st); }
    void launch_ATrhoXi_table(const DeviceBatch& in, double* out, cudaStream_t st) const override {
        // three small-jet kernels (tq_generic_dir.cuh); TEQPB200_ATRHOXI_RECT runs the rectangular J3<2,2,3> sweep of round 1 (A/B)
        static const bool rect = std::getenv("TEQPB200_ATRHOXI_RECT") != nullptr;
        if (rect) { gsaft_launch_ATrhoXi(p, in, out, st); return; }
        gsaft_launch_ATrhoXi_part0(p, in, out, st);
        gsaft_launch_ATrhoXi_part1(p, in, out, st);
        gsaft_launch_ATrhoXi_part2(p, in, out, st);
    }
    void launch_isochoric(const DeviceBatch& in, const IsochoricOut& o, cudaStream_t st) const override { gsaft_launch_isochoric(p, in, o, st); }
    // derived isochoric properties: three single-shape kernels + the Hessian of the isochoric kernel (scratch on the stream)
    void launch_isoprops(const DeviceBatch& in, const IsoPropsOut& o, cudaStream_t st) const override {
        if (in.npts == 0) return;
        const int nc = p.mie.nc;
        double *Ls = nullptr, *H = nullptr;
        TQ_CUDA(cudaMallocAsync(&Ls, sizeof(double) * 6 * (size_t)in.npts, st));
        gsaft_launch_ip_part0(p, in, o, Ls, nullptr, st);
        const bool vec = o.dpdT || o.d2PsirdTdrhoi || o.dpdrhovec || o.pmv || o.chempotVLE || o.dchempotdT || o.dlnphi_dT || o.dlnphi_drho || o.dlnphi_dv;
        if (vec) gsaft_launch_ip_part1(p, in, o, Ls, nullptr, st);
        if (o.dlnphi_dx) {
            TQ_CUDA(cudaMallocAsync(&H, sizeof(double) * (size_t)nc * nc * (size_t)in.npts, st));
            IsochoricOut io; io.hess = H;
            gsaft_launch_isochoric(p, in, io, st);
            gsaft_launch_ip_part2(p, in, o, Ls, H, st);
            TQ_CUDA(cudaFreeAsync(H, st));
        }
        TQ_CUDA(cudaFreeAsync(Ls, st));
    }
};

// genericSAFT whose non-polar term is PC-SAFT (genericsaft.hpp:26-28): the same launch plan as GsaftModel over PcAssocDev
class PcAssocModel final : public Model {
public:
    PcAssocParams p{};
    int ncomp() const override { return p.pc.nc; }
    const char* family() const override { return "genericSAFT-PCSAFT"; }
    double R(const double*, int) const override { return p.pc.R; }
    const void* params_blob(size_t* n) const override { *n = sizeof(p); return &p; }
    void launch_ar_table(int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) const override {
        generic_check_ar_shape(NT, ND);
        if (in.npts == 0) return;
        if (NT == 0) pcassoc_launch_ar_rho(p, ND, in, out, st);
        else pcassoc_launch_ar_mixed(p, NT, ND, in, out, st);
    }
    void launch_virial_table(int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) const override { launch_ar_table(NT, ND, in, out, st); }
    void launch_sigma_derivs(const DeviceBatch& in, const double* v, double* out, cudaStream_t st) const override { pcassoc_launch_sigma(p, in, v, out, st); }
    void launch_ATrhoXiXj_table(const DeviceBatch& in, int i, int j, double* out, cudaStream_t st) const override { pcassoc_launch_cross2(p, in, i, j, -1, out, st); }
    void launch_ATrhoXiXjXk_table(const DeviceBatch& in, int i, int j, int k, double* out, cudaStream_t st) const override { pcassoc_launch_cross3(p, in, i, j, k, out, st); }
    void launch_ATrhoXi_table(const DeviceBatch& in, double* out, cudaStream_t st) const override {
        pcassoc_launch_ATrhoXi_part0(p, in, out, st);
        pcassoc_launch_ATrhoXi_part1(p, in, out, st);
        pcassoc_launch_ATrhoXi_part2(p, in, out, st);
    }
    void launch_isochoric(const DeviceBatch& in, const IsochoricOut& o, cudaStream_t st) const override { pcassoc_launch_isochoric(p, in, o, st); }
    void launch_isoprops(const DeviceBatch& in, const IsoPropsOut& o, cudaStream_t st) const override {
        if (in.npts == 0) return;
        const int nc = p.pc.nc;
        double *Ls = nullptr, *H = nullptr;
        TQ_CUDA(cudaMallocAsync(&Ls, sizeof(double) * 6 * (size_t)in.npts, st));
        pcassoc_launch_ip_part0(p, in, o, Ls, nullptr, st);
        const bool vec = o.dpdT || o.d2PsirdTdrhoi || o.dpdrhovec || o.pmv || o.chempotVLE || o.dchempotdT || o.dlnphi_dT || o.dlnphi_drho || o.dlnphi_dv;
        if (vec) pcassoc_launch_ip_part1(p, in, o, Ls, nullptr, st);
        if (o.dlnphi_dx) {
            TQ_CUDA(cudaMallocAsync(&H, sizeof(double) * (size_t)nc * nc * (size_t)in.npts, st));
            IsochoricOut io; io.hess = H;
            pcassoc_launch_isochoric(p, in, io, st);
            pcassoc_launch_ip_part2(p, in, o, Ls, H, st);
            TQ_CUDA(cudaFreeAsync(H, st));
        }
        TQ_CUDA(cudaFreeAsync(Ls, st));
    }
};

std::vector<std::vector<double>> square_matrix(const Json& j, size_t N, const char* what) {
    std::vector<std::vector<double>> k(N, std::vector<double>(N, 0.0));
    if (j.size() != N) throw InvalidArgument(std::string("Provided length of ") + what + " of " + std::to_string(N) + " does not match the dimension of the kmat of " + std::to_string(j.size()));
    for (size_t i = 0; i < N; ++i) {
        if (!j[i].is_array() || j[i].size() != N) throw InvalidArgument("provided matrix is not square");
        for (size_t c = 0; c < N; ++c) k[i][c] = j[i][c].as_double();
    }
    return k;
}

void fill_mie(MieParams& p, const Json& spec) {
    std::vector<double> m, sig_m, eps, lr, la;
    const char* what = "names";
    if (spec.contains("names")) {
        for (const std::string& nm : spec.at("names").as_strings()) {
            bool found = false;
            for (const auto& e : kLibrary)
                if (nm == e.name) { m.push_back(e.m); sig_m.push_back(e.sigma_m); eps.push_back(e.eps); lr.push_back(e.lr); la.push_back(e.la); found = true; }
            if (!found) throw InvalidArgument("Bad name:" + nm);
        }
    } else if (spec.contains("coeffs")) {
        what = "coeffs";
        const Json& jc = spec.at("coeffs");
        for (size_t i = 0; i < jc.size(); ++i) {
            const Json& c = jc[i];
            (void)c.at("name").as_string();
            m.push_back(c.at("m").as_double());
            sig_m.push_back(c.contains("sigma_m") ? c.at("sigma_m").as_double() : c.at("sigma_Angstrom").as_double() / 1e10);
            eps.push_back(c.at("epsilon_over_k").as_double());
            lr.push_back(c.at("lambda_r").as_double());
            la.push_back(c.at("lambda_a").as_double());
            (void)c.at("BibTeXKey");
            for (const char* key : {"(mu^*)^2", "(Q^*)^2", "Q_Cm2", "Q_DA", "mu_Cm", "mu_D"})
                if (c.contains(key)) throw NotImplementedError("dipolar/quadrupolar SAFT-VR-Mie contributions are not implemented on the device path");
        }
    } else {
        throw InvalidArgument("you must provide names or coeffs, but not both");
    }
    const size_t N = m.size();
    if (N == 0 || N > (size_t)kMieMaxC) throw InvalidArgument("SAFT-VR-Mie supports 1.." + std::to_string(kMieMaxC) + " components on the device path");
    std::vector<std::vector<double>> kmat(N, std::vector<double>(N, 0.0));
    if (spec.contains("kmat") && spec.at("kmat").is_array() && spec.at("kmat").size() > 0) kmat = square_matrix(spec.at("kmat"), N, what);
    p.nc = (int)N;
    p.R = 6.02214076e23 * 1.380649e-23;  // get_R_gas, /root/reference/include/teqp/constants.hpp:22-26
    p.phi73 = kPhi[3][6]; p.phi74 = kPhi[4][6];
    std::vector<double> sigA(N);
    for (size_t i = 0; i < N; ++i) {
        sigA[i] = sig_m[i] * 1e10;
        p.m[i] = m[i]; p.sigma[i] = sigA[i]; p.eps[i] = eps[i]; p.la[i] = la[i]; p.lr[i] = lr[i];
    }
    for (size_t i = 0; i < N; ++i) {
        for (size_t j = i; j < N; ++j) {
            MiePairData& q = p.pair[mie_pair_index((int)i, (int)j, (int)N)];
            q.lr = 3 + std::sqrt((lr[i] - 3) * (lr[j] - 3));
            q.la = 3 + std::sqrt((la[i] - 3) * (la[j] - 3));
            q.C = q.lr / (q.lr - q.la) * std::pow(q.lr / q.la, q.la / (q.lr - q.la));
            const double alpha = q.C * (1 / (q.la - 3) - 1 / (q.lr - 3));
            q.sigma = (sigA[i] + sigA[j]) / 2.0;
            q.sigma3 = q.sigma * q.sigma * q.sigma;
            // Lafitte's combining rule, Eq. A55 (the nonpolar factory passes no flags: saftvrmie.hpp:1117-1154)
            q.eps = (1.0 - kmat[i][j]) * std::sqrt(std::pow(sigA[i], 3) * std::pow(sigA[j], 3) * eps[i] * eps[j]) / std::pow(q.sigma, 3);
            const double lam[5] = {q.la, q.lr, 2.0 * q.la, 2.0 * q.lr, q.lr + q.la};
            for (int s = 0; s < 5; ++s) {
                const double l = lam[s];
                const double b[4] = {1, 1.0 / l, 1.0 / std::pow(l, 2), 1.0 / std::pow(l, 3)};
                for (int r = 0; r < 4; ++r) {
                    double acc = 0;
                    for (int c = 0; c < 4; ++c) acc += kA[r][c] * b[c];
                    q.c[s][r] = acc;
                }
            }
            for (int k = 1; k <= 6; ++k) {
                double num = 0, den = 0;
                for (int n = 0; n < 4; ++n) num += kPhi[n][k - 1] * std::pow(alpha, n);
                for (int n = 4; n < 7; ++n) den += kPhi[n][k - 1] * std::pow(alpha, n - 3);
                q.f[k - 1] = num / (1 + den);
            }
            q.g7 = kPhi[0][6] * (-std::tanh(kPhi[1][6] * (kPhi[2][6] - alpha)) + 1.0);
            q.pad_ = 0;
            if (i == j) { p.Cii[i] = q.C; p.epsii[i] = q.eps; }
        }
    }
}

void fill_assoc(AssocParams& a, const Json& j, size_t N) {
    const Json& ms = j.at("molecule_sites");
    if (ms.size() != N) throw InvalidArgument("Wrong size of molefracs; should be " + std::to_string(ms.size()));
    // options (association.hpp:39-44, 153-162)
    a.alpha = 0.5; a.rtol = 1e-12; a.atol = 1e-12; a.max_iters = 100;
    bool allow_explicit = true;
    std::vector<bool> self_mask;
    std::map<std::string, std::vector<std::string>> partners;
    std::set<std::string> types;
    for (size_t i = 0; i < N; ++i) for (const std::string& s : ms[i].as_strings()) types.insert(s);
    bool have_partners = false;
    if (j.contains("options")) {
        const Json& o = j.at("options");
        a.alpha = o.value("alpha", a.alpha); a.rtol = o.value("rtol", a.rtol); a.atol = o.value("atol", a.atol);
        a.max_iters = (int)o.value("max_iters", (double)a.max_iters);
        // association_types.hpp:20-28: "CS" (Carnahan-Starling, the default) or "KG"; the device term evaluates CS only
        const std::string rd = o.value("radial_dist", std::string("CS"));
        if (rd == "KG") throw NotImplementedError("association option radial_dist = KG is not implemented on the device path (CS is)");
        if (rd != "CS") throw InvalidArgument("bad radial_dist flag:" + rd);
        if (o.contains("self_association_mask")) {
            const Json& sm = o.at("self_association_mask");
            for (size_t i = 0; i < sm.size(); ++i) self_mask.push_back(sm[i].as_bool());
            if (self_mask.size() != N) throw InvalidArgument("self_association_mask is of the wrong size");
        }
        if (o.contains("interaction_partners")) {
            have_partners = true;
            const Json& ip = o.at("interaction_partners");
            for (const std::string& t : types)
                if (ip.contains(t)) partners[t] = ip.at(t).as_strings();
            for (const auto& kv : ip.obj) {
                if (!types.count(kv.first)) throw InvalidArgument("Site is invalid in interaction_partners: " + kv.first);
                for (const std::string& pt : kv.second.as_strings())
                    if (!types.count(pt)) throw InvalidArgument("Partner " + pt + " is invalid for site " + kv.first);
            }
        }
    }
    if (!have_partners)
        for (const std::string& s1 : types)
            for (const std::string& s2 : types)
                if (s1 != s2) partners[s1].push_back(s2);
    // site groups: per molecule, unique site names in std::set (alphabetical) order (association.hpp:86-111)
    std::vector<std::string> type_of;
    int S = 0;
    for (size_t i = 0; i < N; ++i) {
        const std::vector<std::string> sites = ms[i].as_strings();
        std::set<std::string> uniq(sites.begin(), sites.end());
        a.first_site[i] = S; a.nsites_of[i] = (int)uniq.size();
        for (const std::string& s : uniq) {
            if (S >= kAssocMaxS) throw InvalidArgument("too many association site groups for the device path");
            a.comp[S] = (int)i;
            a.count[S] = (double)std::count(sites.begin(), sites.end(), s);
            type_of.push_back(s);
            ++S;
        }
    }
    a.nsite = S;
    for (int I = 0; I < S; ++I) {
        for (int J = 0; J < S; ++J) {
            double d = 0;
            const bool masked = !self_mask.empty() && a.comp[I] == a.comp[J] && !self_mask[a.comp[I]];
            if (!masked) {
                bool ok = partners.empty();
                if (!ok) {
                    auto it = partners.find(type_of[I]);
                    if (it == partners.end()) throw std::out_of_range("map::at");
                    ok = std::find(it->second.begin(), it->second.end(), type_of[J]) != it->second.end();
                }
                if (ok) d = a.count[J];
            }
            a.D[I * kAssocMaxS + J] = d;
        }
    }
    a.explicit_ok = (allow_explicit && N == 1 && S == 2) ? 1 : 0;
    const double Rc = 8.31446261815324;  // constants::R_CODATA2017
    if (!j.contains("Delta_rule")) throw InvalidArgument("The Delta_rule has not been specified");
    const std::string rule = j.at("Delta_rule").as_string();
    auto vec = [&](const char* k) { std::vector<double> v = j.at(k).as_doubles(); if (v.size() != N) throw InvalidArgument(std::string("wrong length of ") + k); return v; };
    if (rule == "CR1") {
        a.rule = 0;
        const auto b = vec("b / m^3/mol"), beta = vec("beta"), e = vec("epsilon / J/mol");
        for (size_t i = 0; i < N; ++i) {
            a.b[i] = b[i];
            for (size_t c = 0; c < N; ++c) {
                a.pref[i * kAssocMaxC + c] = (b[i] + b[c]) / 2.0 * std::sqrt(beta[i] * beta[c]);
                a.eps1[i * kAssocMaxC + c] = (e[i] + e[c]) / 2.0 / Rc;
            }
        }
    } else if (rule == "Dufal") {
        a.rule = 1;
        const auto sig = vec("sigma / m");
        std::vector<double> e;
        if (j.contains("epsilon / J/mol")) e = vec("epsilon / J/mol");
        else if (j.contains("epsilon/kB / K")) { e = vec("epsilon/kB / K"); for (auto& v : e) v *= Rc; }
        else throw InvalidArgument("One of the epsilon variables must be provided");
        const auto lr = vec("lambda_r"), eHB = vec("epsilon_HB / J/mol"), KHB = vec("K_HB / m^3");
        const auto kmat = square_matrix(j.at("kmat"), N, "components");
        for (size_t i = 0; i < N; ++i) {
            for (size_t c = 0; c < N; ++c) {
                const size_t ic = i * kAssocMaxC + c;
                const double s3 = std::pow((sig[i] + sig[c]) / 2.0, 3) ;
                a.sigma3[ic] = ((sig[i] + sig[c]) / 2.0) * ((sig[i] + sig[c]) / 2.0) * ((sig[i] + sig[c]) / 2.0);
                (void)s3;
                const double ss = sig[i] * sig[c];
                a.epsk[ic] = (1 - kmat[i][c]) * std::sqrt(ss * ss * ss) / a.sigma3[ic] * std::sqrt(e[i] * e[c]) / Rc;
                a.eps1[ic] = std::sqrt(eHB[i] * eHB[c]) / Rc;
                const double kk = (std::cbrt(KHB[i]) + std::cbrt(KHB[c])) / 2.0;
                a.pref[ic] = kk * kk * kk;
                if (c >= i) {
                    const double lam = 3 + std::sqrt((lr[i] - 3) * (lr[c] - 3));
                    double* out = a.a[mie_pair_index((int)i, (int)c, (int)N)];
                    int q = 0;
                    for (int ii = 0; ii <= 10; ++ii)
                        for (int jj = 0; ii + jj <= 10; ++jj, ++q) {
                            double aij = 0.0, l2k = 1.0;
                            for (int k = 0; k <= 6; ++k) { aij += kDufalB[k][ii][jj] * l2k; l2k *= lam; }
                            out[q] = aij;
                        }
                }
            }
        }
    } else {
        throw InvalidArgument("The Delta_rule has not been specified");
    }
}

}  // namespace

std::unique_ptr<Model> make_saftvrmie_model(const Json& spec) {
    auto mdl = std::make_unique<GsaftModel>();
    mdl->fam = "SAFT-VR-Mie";
    fill_mie(mdl->p.mie, spec);
    mdl->p.has_assoc = 0;
    return mdl;
}

std::unique_ptr<Model> make_genericsaft_model(const Json& spec) {
    const Json& np = spec.at("nonpolar");
    const std::string kind = np.at("kind").as_string();
    if (kind == "PCSAFT" || kind == "PC-SAFT") {
        auto pm = std::make_unique<PcAssocModel>();
        fill_pcsaft(pm->p.pc, np.at("model"));
        if (pm->p.pc.nc > kAssocMaxC) throw InvalidArgument("genericSAFT supports at most " + std::to_string(kAssocMaxC) + " components on the device path");
        pm->p.has_assoc = 0;
        if (spec.contains("association")) {
            const Json& as = spec.at("association");
            const std::string ak = as.at("kind").as_string();
            if (ak != "canonical" && ak != "Dufal") throw InvalidArgument("Not valid association kind:" + ak);
            fill_assoc(pm->p.assoc, as.at("model"), (size_t)pm->p.pc.nc);
            pm->p.has_assoc = 1;
        }
        return pm;
    }
    auto mdl = std::make_unique<GsaftModel>();
    mdl->fam = "genericSAFT";
    if (kind == "SAFTVRMie" || kind == "SAFT-VR-Mie") fill_mie(mdl->p.mie, np.at("model"));
    else if (kind == "Johnson+Johnson" || kind == "softSAFT" || kind == "2CLJF" || kind == "2CLJ")
        throw NotImplementedError("genericSAFT nonpolar kind " + kind + " is not implemented on the device path (SAFT-VR-Mie is)");
    else throw InvalidArgument("Not valid nonpolar kind:" + kind);
    mdl->p.has_assoc = 0;
    if (spec.contains("association")) {
        const Json& as = spec.at("association");
        const std::string ak = as.at("kind").as_string();
        if (ak != "canonical" && ak != "Dufal") throw InvalidArgument("Not valid association kind:" + ak);
        fill_assoc(mdl->p.assoc, as.at("model"), (size_t)mdl->p.mie.nc);
        mdl->p.has_assoc = 1;
    }
    return mdl;
}

}  // namespace tq
