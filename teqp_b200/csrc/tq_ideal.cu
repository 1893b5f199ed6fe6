// Ideal-gas models: JSON -> flat term list, and the generic jet-sweep launches.
//   "IdealHelmholtz"                      /root/reference/include/teqp/ideal_eosterms.hpp:252-368 (PureIdealHelmholtz, IdealHelmholtz),
//                                         factory /root/reference/interface/CPP/model_factory_IdealHelmholtz.cpp:9-11
//   "GERG2004idealgas" "GERG2008idealgas" /root/reference/include/teqp/models/GERG/GERG.hpp:343-423 (GERG200XAlphaig), :939-961, :1204-1224,
//                                         factory /root/reference/interface/CPP/model_factory_GERG.cpp:19-24
// The coefficient tables in data/gerg_ideal_tables.inc are generated from the reference's accessors (tools/dump_gerg_ideal_tables.cpp).
#include <cmath>
#include <cstring>

#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_ideal.cuh"

namespace tq {
void ideal_launch_isoprops(const IdealParams&, const DeviceBatch&, const IsoPropsOut&, cudaStream_t);                  // tq_ip_ideal.cu
void ideal_launch_ATrhoXiXj(const IdealParams&, const DeviceBatch&, int, int, double*, cudaStream_t);                  // tq_xx_ideal.cu
void ideal_launch_ATrhoXiXjXk(const IdealParams&, const DeviceBatch&, int, int, int, double*, cudaStream_t);
void ideal_launch_sigma(const IdealParams&, const DeviceBatch&, const double*, double*, cudaStream_t);

namespace {

struct GergIdealFluid { const char* name; double Tc_K, rhoc_molm3, n0[8], theta0[8]; };
#include "data/gerg_ideal_tables.inc"

class IdealModel final : public Model {
public:
    IdealParams p{};
    std::string fam;
    int ncomp() const override { return p.nc; }
    const char* family() const override { return fam.c_str(); }
    double R(const double*, int) const override { return p.R; }
    const void* params_blob(size_t* n) const override { *n = sizeof(p); return &p; }
    void launch_ar_table(int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) const override {
        generic_launch_ar_table<IdealDev>(p, p.nc, NT, ND, in, out, st);
    }
    void launch_ATrhoXi_table(const DeviceBatch& in, double* out, cudaStream_t st) const override { generic_launch_ATrhoXi<IdealDev>(p, p.nc, in, out, st); }
    void launch_isochoric(const DeviceBatch& in, const IsochoricOut& o, cudaStream_t st) const override { generic_launch_isochoric<IdealDev>(p, p.nc, in, o, st); }
    void launch_isoprops(const DeviceBatch& in, const IsoPropsOut& o, cudaStream_t st) const override { ideal_launch_isoprops(p, in, o, st); }
    void launch_ATrhoXiXj_table(const DeviceBatch& in, int i, int j, double* out, cudaStream_t st) const override { ideal_launch_ATrhoXiXj(p, in, i, j, out, st); }
    void launch_ATrhoXiXjXk_table(const DeviceBatch& in, int i, int j, int k, double* out, cudaStream_t st) const override { ideal_launch_ATrhoXiXjXk(p, in, i, j, k, out, st); }
    void launch_sigma_derivs(const DeviceBatch& in, const double* v, double* out, cudaStream_t st) const override { ideal_launch_sigma(p, in, v, out, st); }
    // no virial table: alpha^ig has ln(rho), the rho -> 0 limit does not exist (the reference returns non-finite values there)
};

struct TermList {
    IdealParams& p;
    int n = 0;
    void add(int type, double a, double b = 0, double c = 0, double d = 0) {
        if (n >= kIdealMaxTerms) throw InvalidArgument("ideal-gas model has more than " + std::to_string(kIdealMaxTerms) + " terms in total (device-path limit)");
        IdealTerm& t = p.terms[n++];
        t.type = type; t.pad_ = 0; t.a = a; t.b = b; t.c = c; t.d = d;
    }
};

std::vector<double> same_length(const Json& term, const char* key, size_t n) {
    std::vector<double> v = term.at(key).as_doubles();
    if (v.size() != n) throw InvalidArgument(std::string("ideal-gas term arrays must have the same length: ") + key);
    return v;
}

// PureIdealHelmholtz::PureIdealHelmholtz, ideal_eosterms.hpp:255-303
void add_pure_terms(const Json& jpure, int i, TermList& L) {
    if (jpure.is_array() || !jpure.is_object())
        throw InvalidArgument("JSON data passed to PureIdealHelmholtz must be an object and contain the fields \"R\" and \"terms\"");
    (void)jpure.at("R").as_double();
    const Json& terms = jpure.at("terms");
    IdealParams& p = L.p;
    for (size_t k = 0; k < terms.size(); ++k) {
        const Json& term = terms[k];
        if (!term.is_object()) throw InvalidArgument("JSON data for pure fluid must be an object");
        const std::string type = term.at("type").as_string();
        if (type == "Constant") {
            L.add(IG_CONSTANT, term.at("a").as_double());
        } else if (type == "Lead") {   // ln(rho) + a_1 + a_2/T
            p.lnrho_coef[i] += 1.0;
            L.add(IG_CONSTANT, term.at("a_1").as_double());
            L.add(IG_INVT, term.at("a_2").as_double());
        } else if (type == "LogT") {
            L.add(IG_LOGT, term.at("a").as_double());
        } else if (type == "PowerT") {
            const std::vector<double> n = term.at("n").as_doubles(), t = same_length(term, "t", n.size());
            for (size_t q = 0; q < n.size(); ++q) L.add(IG_POWERT, n[q], t[q]);
        } else if (type == "PlanckEinstein") {
            const std::vector<double> n = term.at("n").as_doubles(), th = same_length(term, "theta", n.size());
            for (size_t q = 0; q < n.size(); ++q) L.add(IG_PLANCK, n[q], th[q]);
        } else if (type == "PlanckEinsteinGeneralized") {
            const std::vector<double> n = term.at("n").as_doubles(), c = same_length(term, "c", n.size()), d = same_length(term, "d", n.size()),
                                      th = same_length(term, "theta", n.size());
            for (size_t q = 0; q < n.size(); ++q) L.add(IG_PLANCKGEN, n[q], th[q], c[q], d[q]);
        } else if (type == "GERG2004Cosh") {
            const std::vector<double> n = term.at("n").as_doubles(), th = same_length(term, "theta", n.size());
            for (size_t q = 0; q < n.size(); ++q) L.add(IG_LNCOSH, n[q], th[q]);
        } else if (type == "GERG2004Sinh") {
            const std::vector<double> n = term.at("n").as_doubles(), th = same_length(term, "theta", n.size());
            for (size_t q = 0; q < n.size(); ++q) L.add(IG_LNSINH, n[q], th[q]);
        } else if (type == "Cp0Constant") {
            L.add(IG_CP0CONST, term.at("c").as_double(), term.at("T_0").as_double());
        } else if (type == "Cp0PowerT") {
            const double c = term.at("c").as_double(), t = term.at("t").as_double(), T0 = term.at("T_0").as_double();
            L.add(IG_CP0POWT, c, t, std::pow(T0, t + 1) / (t + 1), std::pow(T0, t) / t);
        } else {
            throw InvalidArgument("Don't understand this type: " + type);
        }
    }
}

// AlphaigCoeffs::recalc_integration_constants, GERG.hpp:49-74: n0[1], n0[2] such that h = s = 0 at 298.15 K, 101325 Pa (2x2 system)
void gerg_integration_constants(const GergIdealFluid& f, double T, double rho, double Rs, double& n1, double& n2) {
    const double* n0 = f.n0;
    const double* th = f.theta0;
    const double Tci = f.Tc_K, rhoci = f.rhoc_molm3;
    const double a00_0 = Rs, a00_1 = Rs * Tci / T;
    const double a00_2 = std::log(rho / rhoci) + Rs * (n0[3] * std::log(Tci / T)
        + ((n0[4] != 0) ? n0[4] * std::log(std::fabs(std::sinh(th[4] * Tci / T))) : 0)
        + ((n0[6] != 0) ? n0[6] * std::log(std::fabs(std::sinh(th[6] * Tci / T))) : 0)
        - ((n0[5] != 0) ? n0[5] * std::log(std::fabs(std::cosh(th[5] * Tci / T))) : 0)
        - ((n0[7] != 0) ? n0[7] * std::log(std::fabs(std::cosh(th[7] * Tci / T))) : 0));
    const double a10_0 = 0, a10_1 = Rs * Tci / T;
    const double a10_2 = Rs * (n0[3]
        + ((n0[4] != 0) ? n0[4] * th[4] * Tci / T / std::tanh(th[4] * Tci / T) : 0)
        + ((n0[6] != 0) ? n0[6] * th[6] * Tci / T / std::tanh(th[6] * Tci / T) : 0)
        - n0[5] * th[5] * Tci / T * std::tanh(th[5] * Tci / T) - n0[7] * th[7] * Tci / T * std::tanh(th[7] * Tci / T));
    // [a10_0, a10_1; a10_0 - a00_0, a10_1 - a00_1] [n1; n2] = [-1 - a10_2; -a10_2 + a00_2]
    const double A00 = a10_0, A01 = a10_1, A10 = a10_0 - a00_0, A11 = a10_1 - a00_1;
    const double b0 = -1 - a10_2, b1 = -a10_2 + a00_2;
    const double det = A00 * A11 - A01 * A10;
    n1 = (b0 * A11 - A01 * b1) / det;
    n2 = (A00 * b1 - A10 * b0) / det;
}

}  // namespace

std::unique_ptr<Model> make_ideal_helmholtz_model(const Json& jpures) {
    if (!jpures.is_array()) throw InvalidArgument("JSON data passed to IdealHelmholtz must be an array");
    const size_t N = jpures.size();
    if (N == 0 || N > (size_t)kIdealMaxC) throw InvalidArgument("IdealHelmholtz supports 1.." + std::to_string(kIdealMaxC) + " components on the device path");
    auto mdl = std::make_unique<IdealModel>();
    mdl->fam = "idealgas";
    IdealParams& p = mdl->p;
    std::memset(&p, 0, sizeof(p));
    p.nc = (int)N;
    p.gerg_rule = 0;
    p.R = 8.31446261815324;   // IdealHelmholtz::R, ideal_eosterms.hpp:363-366
    TermList L{p};
    for (size_t i = 0; i < N; ++i) {
        p.first[i] = L.n;
        p.rho_div[i] = 1.0;
        add_pure_terms(jpures[i], (int)i, L);
    }
    p.first[N] = L.n;
    return mdl;
}

std::unique_ptr<Model> make_gerg_idealgas_model(const std::string& kind, const Json& spec) {
    const bool is2008 = (kind == "GERG2008idealgas");
    const GergIdealFluid* tab = is2008 ? gerg2008_ideal : gerg2004_ideal;
    const size_t ntab = is2008 ? sizeof(gerg2008_ideal) / sizeof(gerg2008_ideal[0]) : sizeof(gerg2004_ideal) / sizeof(gerg2004_ideal[0]);
    const std::vector<std::string> names = spec.at("names").as_strings();
    const size_t N = names.size();
    if (N == 0 || N > (size_t)kIdealMaxC) throw InvalidArgument("GERG ideal-gas models support 1.." + std::to_string(kIdealMaxC) + " components");
    auto mdl = std::make_unique<IdealModel>();
    mdl->fam = "idealgas";   // not "GERG...": that prefix names the residual term-program family
    IdealParams& p = mdl->p;
    std::memset(&p, 0, sizeof(p));
    p.nc = (int)N;
    p.gerg_rule = 1;
    p.R = 8.314472;                           // GERG.hpp:946-948, 1210-1212
    const double Rstar = 8.314510, R = 8.314472;   // GERG.hpp:350-351
    const double Rs = Rstar / R;
    TermList L{p};
    for (size_t i = 0; i < N; ++i) {
        const GergIdealFluid* f = nullptr;
        for (size_t q = 0; q < ntab; ++q) if (names[i] == tab[q].name) f = &tab[q];
        if (!f) throw InvalidArgument("Not able to get ideal-gas coefficients for " + names[i]);
        const double T0 = 298.15, p0 = 101325, rho0 = p0 / (R * T0);   // GERG.hpp:378
        double n1, n2;
        gerg_integration_constants(*f, T0, rho0, Rs, n1, n2);
        p.first[i] = L.n;
        p.lnrho_coef[i] = 1.0;
        p.rho_div[i] = f->rhoc_molm3;
        const double Tc = f->Tc_K;
        // alphaig_pure, GERG.hpp:397-409: ln(rho/rhoc) + Rs (n1 + n2 Tc/T + n3 ln(Tc/T) + n4 ln|sinh(th4 Tc/T)| + n6 ln|sinh(th6 Tc/T)|
        //                                                     - n5 ln|cosh(th5 Tc/T)| - n7 ln|cosh(th7 Tc/T)|), a term present when its theta != 0
        L.add(IG_CONSTANT, Rs * n1);
        L.add(IG_INVT, Rs * (n2 * Tc));
        L.add(IG_LOGRATIO, Rs * f->n0[3], Tc);
        if (f->theta0[4] != 0) L.add(IG_LNSINH, Rs * f->n0[4], f->theta0[4] * Tc);
        if (f->theta0[6] != 0) L.add(IG_LNSINH, Rs * f->n0[6], f->theta0[6] * Tc);
        if (f->theta0[5] != 0) L.add(IG_LNCOSH, -(Rs * f->n0[5]), f->theta0[5] * Tc);
        if (f->theta0[7] != 0) L.add(IG_LNCOSH, -(Rs * f->n0[7]), f->theta0[7] * Tc);
    }
    p.first[N] = L.n;
    return mdl;
}

}  // namespace tq
