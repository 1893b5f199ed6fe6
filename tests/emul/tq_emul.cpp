// TEST HARNESS ONLY -- never shipped, never linked into libteqpb200.so, never used by bench.py or the Python package.
//
// Replays the per-state-point __host__ __device__ functions of the kernels (gerg_point, cubic_table_point, generic_*_point)
// in a plain host loop, on the very parameter block the product's JSON ingestion builds.  Purpose: the build container has no
// GPU, so this lets the CPU-only test tier (-m "not gpu") check the kernels' MATHEMATICS against the oracle before a GPU box is
// spent on them.  The GPU parity tests (-m gpu) go through the C ABI of libteqpb200.so and do not touch this file.
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../teqp_b200/csrc/tq_common.hpp"
#include "../../teqp_b200/csrc/tq_cubic.cuh"
#include "../../teqp_b200/csrc/tq_generic.cuh"
#include "../../teqp_b200/csrc/tq_generic_dir.cuh"
#include "../../teqp_b200/csrc/tq_gerg.cuh"
#include "../../teqp_b200/csrc/tq_gerg_iso.cuh"
#include "../../teqp_b200/csrc/tq_gerg_iso_coop.cuh"
#include "../../teqp_b200/csrc/tq_gerg_jet.cuh"
#include "../../teqp_b200/csrc/tq_pcsaft.cuh"
#include "../../teqp_b200/csrc/tq_assoc.cuh"
#include "../../teqp_b200/csrc/tq_ideal.cuh"
#include "../../teqp_b200/csrc/tq_qcpr.cuh"
#include "../../teqp_b200/csrc/tq_rkpr.cuh"
#include "../../teqp_b200/csrc/tq_apr.cuh"
#include "../../teqp_b200/csrc/tq_pcassoc.cuh"
#include "../../teqp_b200/csrc/tq_vle.cuh"

using namespace tq;

namespace {
void put(const std::string& m, char* e, int n) { if (e && n > 0) { std::strncpy(e, m.c_str(), n - 1); e[n - 1] = 0; } }

template<int NT, int ND>
void gerg_loop(const GergProgram& P, long long npts, const double* T, const double* rho, const double* x, long long xs, double* out) {
    const Exp2K K = Exp2K::load();
    std::vector<double> W((size_t)P.nw * kGergWS);  // same column stride as a CTA's staging area; column 0 is used
    for (long long k = 0; k < npts; ++k) {
        for (int i = 0; i < P.nc; ++i) W[(size_t)i * kGergWS] = x[(long long)i * xs + (xs ? k : 0)];
        double A[NT + 1][ND + 1];
        gerg_point<NT, ND>(P, K, T[k], rho[k], W.data(), A);
        double fn = 1;
        for (int n = 0; n <= NT; ++n) {
            if (n >= 2) fn *= n;
            double fm = 1;
            for (int m = 0; m <= ND; ++m) { if (m >= 2) fm *= m; out[(long long)(n * (ND + 1) + m) * npts + k] = fn * fm * A[n][m]; }
        }
    }
}

template<int NT>
void cubic_loop(const CubicParams& p, int ND, long long npts, const double* T, const double* rho, const double* x, long long xs, double* out) {
    for (long long k = 0; k < npts; ++k) {
        double xx[kCubicMaxC];
        for (int i = 0; i < p.nc; ++i) xx[i] = x[(long long)i * xs + (xs ? k : 0)];
        cubic_table_point<NT, kCubicMaxC>(p, T[k], rho[k], xx, ND, out + k, npts);
    }
}

template<class M, int CT, int CD>
void generic_loop(const typename M::Params& p, int nc, int NT, int ND, long long npts, const double* T, const double* rho, const double* x, long long xs, double* out) {
    for (long long k = 0; k < npts; ++k) {
        double xx[M::kMaxC];
        for (int i = 0; i < nc; ++i) xx[i] = x[(long long)i * xs + (xs ? k : 0)];
        generic_ar_point<M, M::kMaxC, CT, CD>(p, nc, T[k], rho[k], xx, NT, ND, out + k, npts);
    }
}
// virial tables: unit-seeded density jet at rho = 0 (what launch_virial_table runs), T-major [NT+1][ND+1]
template<class M, int CT, int CD>
void virial_loop(const typename M::Params& p, int nc, int NT, int ND, long long npts, const double* T, const double* x, long long xs, double* out) {
    for (long long k = 0; k < npts; ++k) {
        double xx[M::kMaxC];
        for (int i = 0; i < nc; ++i) xx[i] = x[(long long)i * xs + (xs ? k : 0)];
        generic_ar_point<M, M::kMaxC, CT, CD>(p, nc, T[k], 0.0, xx, NT, ND, out + k, npts, true);
    }
}
template<class M>
void virial_table(const typename M::Params& p, int nc, int NT, int ND, long long npts, const double* T, const double* x, long long xs, double* out) {
    if (NT == 0 && ND <= 4) virial_loop<M, 0, 4>(p, nc, NT, ND, npts, T, x, xs, out);
    else if (NT == 0) virial_loop<M, 0, 6>(p, nc, NT, ND, npts, T, x, xs, out);
    else if (NT <= 2 && ND <= 4) virial_loop<M, 2, 4>(p, nc, NT, ND, npts, T, x, xs, out);
    else throw InvalidArgument("shape not compiled in the emulation harness");
}
template<class M>
void sigma_loop(const typename M::Params& p, int nc, long long npts, const double* T, const double* rv, const double* v, double* out) {
    for (long long k = 0; k < npts; ++k) {
        double r[M::kMaxC], d[M::kMaxC];
        for (int i = 0; i < nc; ++i) { r[i] = rv[(long long)i * npts + k]; d[i] = v[(long long)i * npts + k]; }
        generic_sigma_point<M, M::kMaxC>(p, nc, T[k], r, d, out + k, npts);
    }
}
template<class M>
void generic_table(const typename M::Params& p, int nc, int NT, int ND, long long npts, const double* T, const double* rho, const double* x, long long xs, double* out) {
    if (NT == 0 && ND <= 4) generic_loop<M, 0, 4>(p, nc, NT, ND, npts, T, rho, x, xs, out);
    else if (NT == 0) generic_loop<M, 0, 6>(p, nc, NT, ND, npts, T, rho, x, xs, out);
    else if (ND == 0) generic_loop<M, 4, 0>(p, nc, NT, ND, npts, T, rho, x, xs, out);
    else if (NT <= 2 && ND <= 4) generic_loop<M, 2, 4>(p, nc, NT, ND, npts, T, rho, x, xs, out);
    else throw InvalidArgument("shape not compiled in the emulation harness");
}
template<class M>
void generic_atrhoxi(const typename M::Params& p, int nc, long long npts, const double* T, const double* rho, const double* x, double* out) {
    for (long long k = 0; k < npts; ++k) {
        double xx[M::kMaxC];
        for (int i = 0; i < nc; ++i) xx[i] = x[(long long)i * npts + k];
        generic_ATrhoXi_point<M, M::kMaxC>(p, nc, T[k], rho[k], xx, out + k, npts);
    }
}
template<class M>
void generic_iso(const typename M::Params& p, int nc, long long npts, const double* T, const double* rv, const IsochoricOut& o) {
    for (long long k = 0; k < npts; ++k) {
        double r[M::kMaxC];
        for (int i = 0; i < nc; ++i) r[i] = rv[(long long)i * npts + k];
        generic_isochoric_point<M, M::kMaxC>(p, nc, T[k], r, o, k, npts);
    }
}
// the product launches the lane-cooperative form (tq_gerg_iso_coop.cuh) unless TEQPB200_ISO_COOP_L=0; one lane plays all the lanes here
bool coop_off() { const char* e = std::getenv("TEQPB200_ISO_COOP_L"); return e && std::atoi(e) == 0; }
void gerg_iso_loop(const GergProgram& P, double Rgas, long long npts, const double* T, const double* rv, const IsochoricOut& o) {
    if (!coop_off()) {
        const std::vector<double> blob = gerg_iso_build_blob(P);
        const GergIsoHdr& h = *reinterpret_cast<const GergIsoHdr*>(blob.data());
        std::vector<double> W((size_t)gerg_coop_rows(h.nc, h.nw, h.nT, h.ngroup, true));
        for (long long k = 0; k < npts; ++k) {
            for (int i = 0; i < P.nc; ++i) W[(size_t)i] = rv[(long long)i * npts + k];
            if (o.hess) gerg_iso_coop_point<1, true>(blob.data(), Rgas, T[k], W.data(), 0, true, o, k, npts);
            else gerg_iso_coop_point<1, false>(blob.data(), Rgas, T[k], W.data(), 0, true, o, k, npts);
        }
        return;
    }
    constexpr int WS = 32;
    const Exp2K K = Exp2K::load();
    std::vector<double> W((size_t)gerg_iso_rows(P.nc, P.nw, true) * WS);
    for (long long k = 0; k < npts; ++k) {
        for (int i = 0; i < P.nc; ++i) W[(size_t)i * WS] = rv[(long long)i * npts + k];
        if (o.hess) gerg_iso_point<WS, true>(P, K, Rgas, T[k], W.data(), o, k, npts);
        else gerg_iso_point<WS, false>(P, K, Rgas, T[k], W.data(), o, k, npts);
    }
}
void gerg_ip_loop(const GergProgram& P, double Rgas, long long npts, const double* T, const double* rv, const IsoPropsOut& ip) {
    const IsochoricOut none{};
    if (!coop_off()) {
        const std::vector<double> blob = gerg_iso_build_blob(P);
        const GergIsoHdr& h = *reinterpret_cast<const GergIsoHdr*>(blob.data());
        std::vector<double> W((size_t)gerg_coop_rows(h.nc, h.nw, h.nT, h.ngroup, true));
        for (long long k = 0; k < npts; ++k) {
            for (int i = 0; i < P.nc; ++i) W[(size_t)i] = rv[(long long)i * npts + k];
            gerg_iso_coop_point<1, true>(blob.data(), Rgas, T[k], W.data(), 0, true, none, k, npts, &ip);
        }
        return;
    }
    constexpr int WS = 32;
    const Exp2K K = Exp2K::load();
    std::vector<double> W((size_t)gerg_iso_rows(P.nc, P.nw, true) * WS);
    for (long long k = 0; k < npts; ++k) {
        for (int i = 0; i < P.nc; ++i) W[(size_t)i * WS] = rv[(long long)i * npts + k];
        gerg_iso_point<WS, true>(P, K, Rgas, T[k], W.data(), none, k, npts, &ip);
    }
}
// cross-composition tables by directional jets (tq_generic_dir.cuh): PC-SAFT, SAFT-VR-Mie
template<class M>
void dir_xx(const typename M::Params& p, int nc, int which, int i, int j, int k, long long npts, const double* T, const double* rho, const double* x, double* out) {
    for (long long q = 0; q < npts; ++q) {
        double xx[M::kMaxC];
        for (int c = 0; c < nc; ++c) xx[c] = x[(long long)c * npts + q];
        if (which == 2) dir_ATrhoXiXj_point<M, M::kMaxC>(p, nc, T[q], rho[q], xx, i, j, out + q, npts);
        else dir_ATrhoXiXjXk_point<M, M::kMaxC>(p, nc, T[q], rho[q], xx, i, j, k, out + q, npts);
    }
}
// cross-composition tables: which = 2 -> (i, j) table [36][npts]; which = 3 -> (i, j, k) table [32][npts]
template<class M>
void generic_xx(const typename M::Params& p, int nc, int which, int i, int j, int k, long long npts, const double* T, const double* rho, const double* x, double* out) {
    for (long long q = 0; q < npts; ++q) {
        double xx[M::kMaxC];
        for (int c = 0; c < nc; ++c) xx[c] = x[(long long)c * npts + q];
        if (which == 2) generic_ATrhoXiXj_point<M, M::kMaxC, M::kCrossOrder>(p, nc, T[q], rho[q], xx, i, j, out + q, npts);
        else generic_ATrhoXiXjXk_point<M, M::kMaxC>(p, nc, T[q], rho[q], xx, i, j, k, out + q, npts);
    }
}
template<class M>
void generic_ip(const typename M::Params& p, int nc, long long npts, const double* T, const double* rv, const IsoPropsOut& o) {
    for (long long k = 0; k < npts; ++k) {
        double r[M::kMaxC];
        for (int i = 0; i < nc; ++i) r[i] = rv[(long long)i * npts + k];
        generic_isoprops_point<M, M::kMaxC>(p, nc, T[k], r, o, k, npts);
    }
}
}  // namespace

extern "C" {

// mode 0: Lambda_nm table for the kernel family the product would launch; mode 1: force the generic jet sweep (cubic only)
int emul_ar_table(const char* json, int mode, int NT, int ND, long long npts, const double* T, const double* rho, const double* x,
                  int ncomp, int broadcast, double* out, char* err, int errlen) {
    try {
        auto m = tq::make_model(Json::parse(json));
        if (ncomp != m->ncomp()) throw InvalidArgument("wrong number of components");
        size_t nb = 0;
        const void* blob = m->params_blob(&nb);
        const std::string fam = m->family();
        const long long xs = broadcast ? 0 : npts;
        if (fam == "multifluid") {   // JSON multifluid: the jet sweep (what the product runs where the generated kernel does not apply)
            const GergProgram& P = *static_cast<const GergProgram*>(blob);
            generic_table<GergJetDev>(P, P.nc, NT, ND, npts, T, rho, x, xs, out);
            return 0;
        }
        if (fam.rfind("GERG", 0) == 0) {
            const GergProgram& P = *static_cast<const GergProgram*>(blob);
#define C(a, b) if (NT == a && ND == b) { gerg_loop<a, b>(P, npts, T, rho, x, xs, out); return 0; }
            C(0, 0) C(0, 1) C(0, 2) C(0, 3) C(0, 4) C(0, 5) C(0, 6) C(1, 0) C(2, 0) C(3, 0) C(4, 0) C(1, 1) C(1, 2) C(2, 1) C(2, 2) C(2, 4)
#undef C
            throw InvalidArgument("shape not compiled in the emulation harness");
        } else if (fam == "cubic" || fam == "vdW") {
            const CubicParams& p = *static_cast<const CubicParams*>(blob);
            if (mode == 1) { generic_table<CubicDev>(p, p.nc, NT, ND, npts, T, rho, x, xs, out); return 0; }
            switch (NT) {
                case 0: cubic_loop<0>(p, ND, npts, T, rho, x, xs, out); break;
                case 1: cubic_loop<1>(p, ND, npts, T, rho, x, xs, out); break;
                case 2: cubic_loop<2>(p, ND, npts, T, rho, x, xs, out); break;
                case 3: cubic_loop<3>(p, ND, npts, T, rho, x, xs, out); break;
                default: cubic_loop<4>(p, ND, npts, T, rho, x, xs, out); break;
            }
            return 0;
        } else if (fam == "PCSAFT") {
            const PcsaftParams& p = *static_cast<const PcsaftParams*>(blob);
            generic_table<PcsaftDev>(p, p.nc, NT, ND, npts, T, rho, x, xs, out);
            return 0;
        } else if (fam == "SAFT-VR-Mie" || fam == "genericSAFT") {
            const GsaftParams& p = *static_cast<const GsaftParams*>(blob);
            generic_table<GsaftDev>(p, p.mie.nc, NT, ND, npts, T, rho, x, xs, out);
            return 0;
        }
        if (fam == "genericSAFT-PCSAFT") {
            const PcAssocParams& p = *static_cast<const PcAssocParams*>(blob);
            generic_table<PcAssocDev>(p, p.pc.nc, NT, ND, npts, T, rho, x, xs, out);
            return 0;
        }
        if (fam == "idealgas") {
            const IdealParams& p = *static_cast<const IdealParams*>(blob);
            generic_table<IdealDev>(p, p.nc, NT, ND, npts, T, rho, x, xs, out);
            return 0;
        }
        if (fam == "QCPR") {
            const QcprParams& p = *static_cast<const QcprParams*>(blob);
            generic_table<QcprDev>(p, p.nc, NT, ND, npts, T, rho, x, xs, out);
            return 0;
        }
        if (fam == "RKPR") {
            const RkprParams& p = *static_cast<const RkprParams*>(blob);
            generic_table<RkprDev>(p, p.nc, NT, ND, npts, T, rho, x, xs, out);
            return 0;
        }
        if (fam == "advancedPRaEres") {
            const AprParams& p = *static_cast<const AprParams*>(blob);
            generic_table<AprDev>(p, p.nc, NT, ND, npts, T, rho, x, xs, out);
            return 0;
        }
        throw InvalidArgument("family not wired in the emulation harness: " + fam);
    } catch (const std::exception& e) { put(e.what(), err, errlen); return 1; }
}

// what = 0: virial table [NT+1][ND+1][npts] at rho = 0 (x broadcast if broadcast != 0);  what = 1: sigma derivatives out[5][npts] (x = rhovec, v)
int emul_extras(const char* json, int what, int NT, int ND, long long npts, const double* T, const double* x, const double* v, int ncomp, int broadcast,
                double* out, char* err, int errlen) {
    try {
        auto m = tq::make_model(Json::parse(json));
        if (ncomp != m->ncomp()) throw InvalidArgument("wrong number of components");
        size_t nb = 0;
        const void* blob = m->params_blob(&nb);
        const std::string fam = m->family();
        const long long xs = broadcast ? 0 : npts;
#define TQ_FAM(M, P, NC) { const typename M::Params& p = *static_cast<const typename M::Params*>(blob); \
            if (what == 0) virial_table<M>(p, NC, NT, ND, npts, T, x, xs, out); else sigma_loop<M>(p, NC, npts, T, x, v, out); return 0; }
        if (fam.rfind("GERG", 0) == 0 || fam == "multifluid") TQ_FAM(GergJetDev, GergProgram, p.nc)
        if (fam == "cubic" || fam == "vdW") TQ_FAM(CubicDev, CubicParams, p.nc)
        if (fam == "PCSAFT") TQ_FAM(PcsaftDev, PcsaftParams, p.nc)
        if (fam == "SAFT-VR-Mie" || fam == "genericSAFT") TQ_FAM(GsaftDev, GsaftParams, p.mie.nc)
        if (fam == "idealgas" && what == 1) TQ_FAM(IdealDev, IdealParams, p.nc)
        if (fam == "QCPR" && what == 1) TQ_FAM(QcprDev, QcprParams, p.nc)
        if (fam == "RKPR") TQ_FAM(RkprDev, RkprParams, p.nc)
        if (fam == "advancedPRaEres") TQ_FAM(AprDev, AprParams, p.nc)
        if (fam == "genericSAFT-PCSAFT") TQ_FAM(PcAssocDev, PcAssocParams, p.pc.nc)
#undef TQ_FAM
        throw InvalidArgument("family not wired in the emulation harness: " + fam);
    } catch (const std::exception& e) { put(e.what(), err, errlen); return 1; }
}

// get_AtaudeltaXi table (which = 1; layout of emul_ATrhoXi_table), XiXj (2), XiXjXk (3): multifluid only; tau, delta given
int emul_Ataudelta(const char* json, int which, int i, int j, int k, long long npts, const double* tau, const double* delta, const double* x, int ncomp,
                   double* out, char* err, int errlen) {
    try {
        auto m = tq::make_model(Json::parse(json));
        if (ncomp != m->ncomp()) throw InvalidArgument("wrong number of components");
        if (std::string(m->family()) != "multifluid") throw NotImplementedError("Cannot take derivatives of a class that doesn't define the alphar_taudelta method");
        size_t nb = 0;
        const GergProgram& P = *static_cast<const GergProgram*>(m->params_blob(&nb));
        std::vector<double> Tinv((size_t)npts);
        for (long long q = 0; q < npts; ++q) Tinv[q] = 1.0 / tau[q];
        if (which == 1) generic_atrhoxi<GergTauDeltaDev>(P, ncomp, npts, Tinv.data(), delta, x, out);
        else generic_xx<GergTauDeltaDev>(P, ncomp, which, i, j, k, npts, Tinv.data(), delta, x, out);
        return 0;
    } catch (const std::exception& e) { put(e.what(), err, errlen); return 1; }
}

int emul_ATrhoXi_table(const char* json, long long npts, const double* T, const double* rho, const double* x, int ncomp, double* out,
                       char* err, int errlen) {
    try {
        auto m = tq::make_model(Json::parse(json));
        if (ncomp != m->ncomp()) throw InvalidArgument("wrong number of components");
        size_t nb = 0;
        const void* blob = m->params_blob(&nb);
        const std::string fam = m->family();
        if (fam == "cubic" || fam == "vdW") generic_atrhoxi<CubicDev>(*static_cast<const CubicParams*>(blob), ncomp, npts, T, rho, x, out);
        else if (fam == "PCSAFT") generic_atrhoxi<PcsaftDev>(*static_cast<const PcsaftParams*>(blob), ncomp, npts, T, rho, x, out);
        else if (fam == "idealgas") generic_atrhoxi<IdealDev>(*static_cast<const IdealParams*>(blob), ncomp, npts, T, rho, x, out);
        else if (fam == "QCPR") generic_atrhoxi<QcprDev>(*static_cast<const QcprParams*>(blob), ncomp, npts, T, rho, x, out);
        else if (fam == "RKPR") generic_atrhoxi<RkprDev>(*static_cast<const RkprParams*>(blob), ncomp, npts, T, rho, x, out);
        else if (fam == "advancedPRaEres") generic_atrhoxi<AprDev>(*static_cast<const AprParams*>(blob), ncomp, npts, T, rho, x, out);
        else if (fam.rfind("GERG", 0) == 0 || fam == "multifluid") generic_atrhoxi<GergJetDev>(*static_cast<const GergProgram*>(blob), ncomp, npts, T, rho, x, out);
        else if (fam == "SAFT-VR-Mie" || fam == "genericSAFT") {
            // the three small-jet parts the product launches for this family (GsaftModel::launch_ATrhoXi_table)
            const GsaftParams& p = *static_cast<const GsaftParams*>(blob);
            for (long long k = 0; k < npts; ++k) {
                double xx[GsaftDev::kMaxC];
                for (int i = 0; i < ncomp; ++i) xx[i] = x[(long long)i * npts + k];
                split_ATrhoXi_point<GsaftDev, GsaftDev::kMaxC, 1, 1>(p, ncomp, T[k], rho[k], xx, out + k, npts);
                split_ATrhoXi_point<GsaftDev, GsaftDev::kMaxC, 2, 0>(p, ncomp, T[k], rho[k], xx, out + k, npts);
                split_ATrhoXi_point<GsaftDev, GsaftDev::kMaxC, 0, 2>(p, ncomp, T[k], rho[k], xx, out + k, npts);
            }
        }
        else if (fam == "genericSAFT-PCSAFT") {
            // the three small-jet parts the product launches for this family (PcAssocModel::launch_ATrhoXi_table)
            const PcAssocParams& p = *static_cast<const PcAssocParams*>(blob);
            for (long long k = 0; k < npts; ++k) {
                double xx[PcAssocDev::kMaxC];
                for (int i = 0; i < ncomp; ++i) xx[i] = x[(long long)i * npts + k];
                split_ATrhoXi_point<PcAssocDev, PcAssocDev::kMaxC, 1, 1>(p, ncomp, T[k], rho[k], xx, out + k, npts);
                split_ATrhoXi_point<PcAssocDev, PcAssocDev::kMaxC, 2, 0>(p, ncomp, T[k], rho[k], xx, out + k, npts);
                split_ATrhoXi_point<PcAssocDev, PcAssocDev::kMaxC, 0, 2>(p, ncomp, T[k], rho[k], xx, out + k, npts);
            }
        }
        else throw InvalidArgument("family not wired in the emulation harness: " + fam);
        return 0;
    } catch (const std::exception& e) { put(e.what(), err, errlen); return 1; }
}

int emul_cross_table(const char* json, int which, int i, int j, int k, long long npts, const double* T, const double* rho, const double* x,
                     int ncomp, double* out, char* err, int errlen) {
    try {
        auto m = tq::make_model(Json::parse(json));
        if (ncomp != m->ncomp()) throw InvalidArgument("wrong number of components");
        size_t nb = 0;
        const void* blob = m->params_blob(&nb);
        const std::string fam = m->family();
        if (fam == "cubic" || fam == "vdW") generic_xx<CubicDev>(*static_cast<const CubicParams*>(blob), ncomp, which, i, j, k, npts, T, rho, x, out);
        else if (fam == "idealgas") generic_xx<IdealDev>(*static_cast<const IdealParams*>(blob), ncomp, which, i, j, k, npts, T, rho, x, out);
        else if (fam == "QCPR") generic_xx<QcprDev>(*static_cast<const QcprParams*>(blob), ncomp, which, i, j, k, npts, T, rho, x, out);
        else if (fam == "RKPR") generic_xx<RkprDev>(*static_cast<const RkprParams*>(blob), ncomp, which, i, j, k, npts, T, rho, x, out);
        else if (fam == "advancedPRaEres") generic_xx<AprDev>(*static_cast<const AprParams*>(blob), ncomp, which, i, j, k, npts, T, rho, x, out);
        else if (fam.rfind("GERG", 0) == 0 || fam == "multifluid") generic_xx<GergJetDev>(*static_cast<const GergProgram*>(blob), ncomp, which, i, j, k, npts, T, rho, x, out);
        else if (fam == "PCSAFT") dir_xx<PcsaftDev>(*static_cast<const PcsaftParams*>(blob), ncomp, which, i, j, k, npts, T, rho, x, out);
        else if (fam == "SAFT-VR-Mie" || fam == "genericSAFT") dir_xx<GsaftDev>(*static_cast<const GsaftParams*>(blob), ncomp, which, i, j, k, npts, T, rho, x, out);
        else if (fam == "genericSAFT-PCSAFT") dir_xx<PcAssocDev>(*static_cast<const PcAssocParams*>(blob), ncomp, which, i, j, k, npts, T, rho, x, out);
        else throw InvalidArgument("family not wired in the emulation harness: " + fam);
        return 0;
    } catch (const std::exception& e) { put(e.what(), err, errlen); return 1; }
}

// outs: the 12 pointers of IsoPropsOut in declaration order (null = skip)
int emul_isoprops(const char* json, long long npts, const double* T, const double* rhovec, int ncomp, double** outs, char* err, int errlen) {
    try {
        auto m = tq::make_model(Json::parse(json));
        if (ncomp != m->ncomp()) throw InvalidArgument("wrong number of components");
        size_t nb = 0;
        const void* blob = m->params_blob(&nb);
        const std::string fam = m->family();
        IsoPropsOut o;
        o.pr = outs[0]; o.splus = outs[1]; o.dpdT = outs[2]; o.d2PsirdTdrhoi = outs[3]; o.dpdrhovec = outs[4]; o.pmv = outs[5];
        o.chempotVLE = outs[6]; o.dchempotdT = outs[7]; o.dlnphi_dT = outs[8]; o.dlnphi_drho = outs[9]; o.dlnphi_dv = outs[10]; o.dlnphi_dx = outs[11];
        if (fam == "cubic" || fam == "vdW") generic_ip<CubicDev>(*static_cast<const CubicParams*>(blob), ncomp, npts, T, rhovec, o);
        else if (fam == "idealgas") generic_ip<IdealDev>(*static_cast<const IdealParams*>(blob), ncomp, npts, T, rhovec, o);
        else if (fam == "QCPR") generic_ip<QcprDev>(*static_cast<const QcprParams*>(blob), ncomp, npts, T, rhovec, o);
        else if (fam == "RKPR") generic_ip<RkprDev>(*static_cast<const RkprParams*>(blob), ncomp, npts, T, rhovec, o);
        else if (fam == "advancedPRaEres") generic_ip<AprDev>(*static_cast<const AprParams*>(blob), ncomp, npts, T, rhovec, o);
        else if (fam == "PCSAFT") generic_ip<PcsaftDev>(*static_cast<const PcsaftParams*>(blob), ncomp, npts, T, rhovec, o);
        else if (fam == "SAFT-VR-Mie" || fam == "genericSAFT") {
            // the three single-shape parts the product launches for this family (GsaftModel::launch_isoprops)
            const GsaftParams& p = *static_cast<const GsaftParams*>(blob);
            std::vector<double> Ls((size_t)6 * npts), H;
            if (o.dlnphi_dx) { H.resize((size_t)ncomp * ncomp * npts); IsochoricOut io; io.hess = H.data(); generic_iso<GsaftDev>(p, ncomp, npts, T, rhovec, io); }
            for (long long k = 0; k < npts; ++k) {
                double r[GsaftDev::kMaxC];
                for (int i = 0; i < ncomp; ++i) r[i] = rhovec[(long long)i * npts + k];
                generic_ip_L_point<GsaftDev, GsaftDev::kMaxC>(p, ncomp, T[k], r, o, Ls.data(), k, npts);
                generic_ip_C_point<GsaftDev, GsaftDev::kMaxC>(p, ncomp, T[k], r, o, Ls.data(), k, npts);
                if (o.dlnphi_dx) generic_ip_X_point<GsaftDev, GsaftDev::kMaxC>(p, ncomp, T[k], r, o, Ls.data(), H.data(), k, npts);
            }
        }
        else if (fam == "genericSAFT-PCSAFT") {
            // the three single-shape parts the product launches for this family (PcAssocModel::launch_isoprops)
            const PcAssocParams& p = *static_cast<const PcAssocParams*>(blob);
            std::vector<double> Ls((size_t)6 * npts), H;
            if (o.dlnphi_dx) { H.resize((size_t)ncomp * ncomp * npts); IsochoricOut io; io.hess = H.data(); generic_iso<PcAssocDev>(p, ncomp, npts, T, rhovec, io); }
            for (long long k = 0; k < npts; ++k) {
                double r[PcAssocDev::kMaxC];
                for (int i = 0; i < ncomp; ++i) r[i] = rhovec[(long long)i * npts + k];
                generic_ip_L_point<PcAssocDev, PcAssocDev::kMaxC>(p, ncomp, T[k], r, o, Ls.data(), k, npts);
                generic_ip_C_point<PcAssocDev, PcAssocDev::kMaxC>(p, ncomp, T[k], r, o, Ls.data(), k, npts);
                if (o.dlnphi_dx) generic_ip_X_point<PcAssocDev, PcAssocDev::kMaxC>(p, ncomp, T[k], r, o, Ls.data(), H.data(), k, npts);
            }
        }
        else if (fam.rfind("GERG", 0) == 0 || fam == "multifluid") {
            const GergProgram& P = *static_cast<const GergProgram*>(blob);
            // same dispatch as GergModel::launch_isoprops
            if (P.ntf == 0 && P.nna == 0 && !P.r_weighted && !P.pure_shortcut && !std::getenv("TEQPB200_GERG_ISOPROPS_JETS")) gerg_ip_loop(P, P.Rgas, npts, T, rhovec, o);
            else generic_ip<GergJetDev>(P, ncomp, npts, T, rhovec, o);
        }
        else throw InvalidArgument("family not wired in the emulation harness: " + fam);
        return 0;
    } catch (const std::exception& e) { put(e.what(), err, errlen); return 1; }
}

int emul_isochoric(const char* json, long long npts, const double* T, const double* rhovec, int ncomp, double* f, double* g, double* H,
                   double* lnphi, double* phi, char* err, int errlen) {
    try {
        auto m = tq::make_model(Json::parse(json));
        if (ncomp != m->ncomp()) throw InvalidArgument("wrong number of components");
        size_t nb = 0;
        const void* blob = m->params_blob(&nb);
        const std::string fam = m->family();
        IsochoricOut o; o.Psir = f; o.grad = g; o.hess = H; o.lnphi = lnphi; o.phi = phi;
        if (fam == "cubic" || fam == "vdW") generic_iso<CubicDev>(*static_cast<const CubicParams*>(blob), ncomp, npts, T, rhovec, o);
        else if (fam == "idealgas") generic_iso<IdealDev>(*static_cast<const IdealParams*>(blob), ncomp, npts, T, rhovec, o);
        else if (fam == "QCPR") generic_iso<QcprDev>(*static_cast<const QcprParams*>(blob), ncomp, npts, T, rhovec, o);
        else if (fam == "RKPR") generic_iso<RkprDev>(*static_cast<const RkprParams*>(blob), ncomp, npts, T, rhovec, o);
        else if (fam == "advancedPRaEres") generic_iso<AprDev>(*static_cast<const AprParams*>(blob), ncomp, npts, T, rhovec, o);
        else if (fam == "PCSAFT") generic_iso<PcsaftDev>(*static_cast<const PcsaftParams*>(blob), ncomp, npts, T, rhovec, o);
        else if (fam == "SAFT-VR-Mie" || fam == "genericSAFT") generic_iso<GsaftDev>(*static_cast<const GsaftParams*>(blob), ncomp, npts, T, rhovec, o);
        else if (fam == "genericSAFT-PCSAFT") generic_iso<PcAssocDev>(*static_cast<const PcAssocParams*>(blob), ncomp, npts, T, rhovec, o);
        else if (fam.rfind("GERG", 0) == 0 || fam == "multifluid") {
            const GergProgram& P = *static_cast<const GergProgram*>(blob);
            // same dispatch as GergModel::launch_isochoric: analytic kernel for tau^t shapes and a constant R, jet sweeps otherwise
            if (P.ntf == 0 && P.nna == 0 && !P.r_weighted && !P.pure_shortcut) gerg_iso_loop(P, P.Rgas, npts, T, rhovec, o);
            else generic_iso<GergJetDev>(P, ncomp, npts, T, rhovec, o);
        }
        else throw InvalidArgument("family not wired in the emulation harness: " + fam);
        return 0;
    } catch (const std::exception& e) { put(e.what(), err, errlen); return 1; }
}

// one Newton update of the batched pure-fluid saturation solver (pure_vle_update_kernel, tq_vle.cuh): A[3][2n], rho2[2n], done[n]
int emul_pure_vle_update(long long n, const double* A, double* rho2, double* done) {
    const long long n2 = 2 * n;
    for (long long k = 0; k < n; ++k) {
        if (done[k] != 0.0) continue;
        double rL = rho2[k], rV = rho2[n + k];
        if (pure_vle_newton_point(A[k], A[n2 + k], A[2 * n2 + k], A[n + k], A[n2 + n + k], A[2 * n2 + n + k], rL, rV)) done[k] = 1.0;
        else { rho2[k] = rL; rho2[n + k] = rV; }
    }
    return 0;
}

// one Newton update of the batched binary-mixture solver (mix_vle_update_kernel, tq_vle.cuh): f[2n], g[2][2n], H[4][2n], rv2[2][2n], xspec[2][n], code[n]
int emul_mix_vle_update(long long n, double R, const double* T, const double* xspec, const double* f, const double* g, const double* H, double atol, double reltol,
                        double axtol, double relxtol, int last_iter, double* rv2, double* code) {
    const long long n2 = 2 * n;
    for (long long k = 0; k < n; ++k) {
        if (code[k] != 0.0) continue;
        double x[4] = {rv2[k], rv2[n2 + k], rv2[n + k], rv2[n2 + n + k]};
        const double gL[2] = {g[k], g[n2 + k]}, gV[2] = {g[n + k], g[n2 + n + k]};
        double HL[4], HV[4];
        for (int q = 0; q < 4; ++q) { HL[q] = H[(long long)q * n2 + k]; HV[q] = H[(long long)q * n2 + n + k]; }
        const double xs[2] = {xspec[k], xspec[n + k]};
        double c = 0.0;
        mix_vle_newton_point(R * T[k], f[k], gL, HL, f[n + k], gV, HV, xs, atol, reltol, axtol, relxtol, last_iter != 0, x, c);
        if (c != 5.0) { rv2[k] = x[0]; rv2[n2 + k] = x[1]; rv2[n + k] = x[2]; rv2[n2 + n + k] = x[3]; }
        code[k] = c;
    }
    return 0;
}

// one Newton update of mix_VLE_Tp_batch (mix_vle_tp_update_kernel, tq_vle.cuh)
int emul_mix_vle_tp_update(long long n, double R, const double* T, const double* p, const double* f, const double* g, const double* H, double* rv2) {
    const long long n2 = 2 * n;
    for (long long k = 0; k < n; ++k) {
        double x[4] = {rv2[k], rv2[n2 + k], rv2[n + k], rv2[n2 + n + k]};
        const double gL[2] = {g[k], g[n2 + k]}, gV[2] = {g[n + k], g[n2 + n + k]};
        double HL[4], HV[4];
        for (int q = 0; q < 4; ++q) { HL[q] = H[(long long)q * n2 + k]; HV[q] = H[(long long)q * n2 + n + k]; }
        mix_vle_tp_newton_point(R * T[k], p[k], f[k], gL, HL, f[n + k], gV, HV, x);
        rv2[k] = x[0]; rv2[n2 + k] = x[1]; rv2[n + k] = x[2]; rv2[n2 + n + k] = x[3];
    }
    return 0;
}

}  // extern "C"
