// PC-SAFT: JSON -> parameter block (mirrors PCSAFTfactory, /root/reference/include/teqp/models/pcsaft.hpp:455-513, and the
// built-in three-fluid library :51-55) and the generic jet-sweep launches.
#include <cmath>

#include "tq_common.hpp"
#include "tq_generic.cuh"
#include "tq_pcsaft.cuh"

namespace tq {
void pcsaft_launch_isoprops(const PcsaftParams&, const DeviceBatch&, const IsoPropsOut&, cudaStream_t);

// the kernels live in tq_pcsaft_k{1..4}.cu (one translation unit per jet-shape group, compiled in parallel)
void pcsaft_launch_ar_rho(const PcsaftParams&, int ND, const DeviceBatch&, double*, cudaStream_t);
void pcsaft_launch_ar_mixed(const PcsaftParams&, int NT, int ND, const DeviceBatch&, double*, cudaStream_t);
void pcsaft_launch_ATrhoXi(const PcsaftParams&, const DeviceBatch&, double*, cudaStream_t);
void pcsaft_launch_isochoric(const PcsaftParams&, const DeviceBatch&, const IsochoricOut&, cudaStream_t);
void pcsaft_launch_cross2(const PcsaftParams&, const DeviceBatch&, int, int, int, double*, cudaStream_t);   // tq_xx_pcsaft_{2,3}.cu (directional jets)
void pcsaft_launch_cross3(const PcsaftParams&, const DeviceBatch&, int, int, int, double*, cudaStream_t);
void pcsaft_launch_sigma(const PcsaftParams&, const DeviceBatch&, const double*, double*, cudaStream_t);   // tq_ex_pcsaft.cu

namespace {

// universal model constants, published tables (see header comment of tq_pcsaft.cuh)
const double GS2001_a[3][7] = {
    {0.9105631445, 0.6361281449, 2.6861347891, -26.547362491, 97.759208784, -159.59154087, 91.297774084},
    {-0.3084016918, 0.1860531159, -2.5030047259, 21.419793629, -65.255885330, 83.318680481, -33.746922930},
    {-0.0906148351, 0.4527842806, 0.5962700728, -1.7241829131, -4.1302112531, 13.776631870, -8.6728470368}};
const double GS2001_b[3][7] = {
    {0.7240946941, 2.2382791861, -4.0025849485, -21.003576815, 26.855641363, 206.55133841, -355.60235612},
    {-0.5755498075, 0.6995095521, 3.8925673390, -17.215471648, 192.67226447, -161.82646165, -165.20769346},
    {0.0976883116, -0.2557574982, -9.1558561530, 20.642075974, -38.804430052, 93.626774077, -29.666905585}};
const double L2012_a[3][7] = {
    {0.836215101666, 2.201683842453, -11.25210310939, 37.841836899902, -68.035304263396, 69.952369867326, -42.828905226651},
    {-0.411727190935913, 2.37426400571265, -21.0620603419144, 105.65718855671, -298.665894225644, 468.695983731173, -316.673589664169},
    {0.0319867672916212, -2.75137756155503, 25.7581175334397, -103.082737163044, 239.569856365622, -320.622085430506, 186.50494276364}};
const double L2012_b[3][7] = {
    {0.627209841336118, 4.02517816132384, -22.6660554011051, 70.0153445172765, -129.548046066679, 150.197401680241, -86.9664928046989},
    {-0.622507280536237, 1.86478114256654, -22.5660800172653, 153.069895818654, -321.710866534244, 453.083735030445, -233.232187026907},
    {-0.0303061275320169, 2.59554209415371, 11.5803588817289, -145.915305288352, 354.901071174401, -151.675970200232, -282.837925415568}};
const double L2014_a[3][7] = {
    {0.961597, 0.414449, 0.689253, -7.43899, 31.8755, -54.8833, 27.3613},
    {-0.333416, 0.358440, -0.219088, 0.285586, 5.88256, -22.4931, 21.2109},
    {-0.0632483, 0.287134, -0.105309, 7.16607, -37.5008, 68.8002, -40.3177}};
const double L2014_b[3][7] = {
    {0.548398, 2.07176, -1.84013, -29.9683, 160.445, -206.106, 51.6201},
    {-0.277226, 0.0621105, 4.29640, -39.9453, 214.911, -232.563, 60.9734},
    {-0.150684, -0.507877, 1.24227, -48.7052, 63.1155, 205.380, -262.437}};

struct LibEntry { const char* name; double m, sigma, eps; };
const LibEntry kLibrary[] = {  // Gross & Sadowski 2001, Table 2 (pcsaft.hpp:53-55)
    {"Methane", 1.0000, 3.7039, 150.03}, {"Ethane", 1.6069, 3.5206, 191.42}, {"Propane", 2.0020, 3.6184, 208.11}};

class PcsaftModel final : public Model {
public:
    PcsaftParams p{};
    int ncomp() const override { return p.nc; }
    const char* family() const override { return "PCSAFT"; }
    double R(const double*, int) const override { return p.R; }
    const void* params_blob(size_t* n) const override { *n = sizeof(p); return &p; }
    void launch_ar_table(int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) const override {
        generic_check_ar_shape(NT, ND);
        if (in.npts == 0) return;
        if (NT == 0) pcsaft_launch_ar_rho(p, ND, in, out, st);
        else pcsaft_launch_ar_mixed(p, NT, ND, in, out, st);
    }
    void launch_virial_table(int NT, int ND, const DeviceBatch& in, double* out, cudaStream_t st) const override { launch_ar_table(NT, ND, in, out, st); }
    void launch_sigma_derivs(const DeviceBatch& in, const double* v, double* out, cudaStream_t st) const override { pcsaft_launch_sigma(p, in, v, out, st); }
    void launch_ATrhoXiXj_table(const DeviceBatch& in, int i, int j, double* out, cudaStream_t st) const override { pcsaft_launch_cross2(p, in, i, j, -1, out, st); }
    void launch_ATrhoXiXjXk_table(const DeviceBatch& in, int i, int j, int k, double* out, cudaStream_t st) const override { pcsaft_launch_cross3(p, in, i, j, k, out, st); }
    void launch_ATrhoXi_table(const DeviceBatch& in, double* out, cudaStream_t st) const override { pcsaft_launch_ATrhoXi(p, in, out, st); }
    void launch_isoprops(const DeviceBatch& in, const IsoPropsOut& o, cudaStream_t st) const override { pcsaft_launch_isoprops(p, in, o, st); }
    void launch_isochoric(const DeviceBatch& in, const IsochoricOut& o, cudaStream_t st) const override { pcsaft_launch_isochoric(p, in, o, st); }
};

}  // namespace

// PCSAFTfactory (pcsaft.hpp:455-513) into the parameter block; also used by genericSAFT with a PC-SAFT non-polar term (tq_gsaft.cu)
void fill_pcsaft(PcsaftParams& p, const Json& spec) {
    const double (*a)[7] = GS2001_a;
    const double (*b)[7] = GS2001_b;
    if (spec.contains("ab")) {
        const std::string src = spec.at("ab").as_string();
        if (src == "Liang-IECR-2012") { a = L2012_a; b = L2012_b; }
        else if (src == "Liang-IECR-2014") { a = L2014_a; b = L2014_b; }
        else throw InvalidArgument("Don't know what to do with this source for a&b: " + src);
    }
    for (int r = 0; r < 3; ++r) for (int k = 0; k < 7; ++k) { p.a[r][k] = a[r][k]; p.b[r][k] = b[r][k]; }
    p.R = 6.02214076e23 * 1.380649e-23;  // get_R_gas, /root/reference/include/teqp/constants.hpp:22-26

    std::vector<double> m, sig, eps;
    if (spec.contains("names")) {
        for (const std::string& nm : spec.at("names").as_strings()) {
            bool found = false;
            for (const auto& e : kLibrary) if (nm == e.name) { m.push_back(e.m); sig.push_back(e.sigma); eps.push_back(e.eps); found = true; }
            if (!found) throw InvalidArgument("Bad name:" + nm);
        }
    } else if (spec.contains("coeffs")) {
        const Json& jc = spec.at("coeffs");
        for (size_t i = 0; i < jc.size(); ++i) {
            const Json& c = jc[i];
            (void)c.at("name").as_string();
            m.push_back(c.at("m").as_double());
            sig.push_back(c.at("sigma_Angstrom").as_double());
            eps.push_back(c.at("epsilon_over_k").as_double());
            (void)c.at("BibTeXKey");
            double polar = 0;
            if (c.contains("(mu^*)^2") && c.contains("nmu")) polar += std::fabs(c.at("(mu^*)^2").as_double() * c.at("nmu").as_double());
            if (c.contains("(Q^*)^2") && c.contains("nQ")) polar += std::fabs(c.at("(Q^*)^2").as_double() * c.at("nQ").as_double());
            if (polar != 0) throw NotImplementedError("dipolar/quadrupolar PC-SAFT contributions are not implemented on the device path");
        }
    } else {
        throw InvalidArgument("you must provide names or coeffs, but not both");
    }
    const size_t N = m.size();
    if (N == 0 || N > (size_t)kPcsaftMaxC) throw InvalidArgument("PC-SAFT supports 1.." + std::to_string(kPcsaftMaxC) + " components on the device path");
    p.nc = (int)N;
    for (size_t i = 0; i < N; ++i) { p.m[i] = m[i]; p.sigma[i] = sig[i]; p.eps[i] = eps[i]; }
    for (size_t i = 0; i < N * N; ++i) p.kmat[i] = 0.0;
    if (spec.contains("kmat") && spec.at("kmat").is_array() && spec.at("kmat").size() > 0) {
        const Json& jk = spec.at("kmat");
        if (jk.size() != N) throw InvalidArgument("Provided length of names of " + std::to_string(N) + " does not match the dimension of the kmat of " + std::to_string(jk.size()));
        for (size_t i = 0; i < N; ++i) {
            if (!jk[i].is_array() || jk[i].size() != N) throw InvalidArgument("provided matrix is not square");
            for (size_t j = 0; j < N; ++j) p.kmat[i * N + j] = jk[i][j].as_double();
        }
    }
}

std::unique_ptr<Model> make_pcsaft_model(const Json& spec) {
    auto mdl = std::make_unique<PcsaftModel>();
    fill_pcsaft(mdl->p, spec);
    return mdl;
}

}  // namespace tq
