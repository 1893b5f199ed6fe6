// TEST INFRASTRUCTURE ONLY -- the reference-held known-answer values of the SAFT families, evaluated with the reference's OWN classes the
// way its own tests do, so that tests/test_oracle_pins.py can assert them against the numbers stored in the reference's test-suite:
//   src/tests/catch_test_SAFTVRMie.cxx:57-80   SAFTVRMieChainContributionTerms::get_core_calcs (a1kB, a2kB2, a3kB3, alphar_mono, alphar_chain; Clapeyron.jl)
//   src/tests/catch_test_association.cxx:28-52 Association::get_Delta / successive_substitution, ethanol + water (CR1, CS radial distribution)
//   src/tests/catch_test_SAFTgeneric.cxx:160-180 Dufal water: non-bonded site fraction X_A at 290 K (Clapeyron.jl)
// These exercise the model arithmetic through the AD shim's plain-double path and the association solver, independently of any kernel.
#include "teqp/models/saftvrmie.hpp"
#include "teqp/models/association/association.hpp"
#include "teqp/models/saft/genericsaft.hpp"

using namespace teqp;

extern "C" {

// out[5] = a1kB, a2kB2, a3kB3, alphar_mono, alphar_chain
int ref_pin_saftvrmie_core(double* out) {
    try {
        auto m = (Eigen::ArrayXd(1) << 1.4373).finished();
        auto eps = (Eigen::ArrayXd(1) << 206.12).finished();
        auto sigma = (Eigen::ArrayXd(1) << 3.7257e-10).finished();
        auto lambda_r = (Eigen::ArrayXd(1) << 12.4).finished();
        auto lambda_a = (Eigen::ArrayXd(1) << 6.0).finished();
        Eigen::ArrayXXd kmat = Eigen::ArrayXXd::Zero(1, 1);
        SAFTVRMie::SAFTVRMieChainContributionTerms terms(m, eps, sigma, lambda_r, lambda_a, kmat);
        auto z = (Eigen::ArrayXd(1) << 1.0).finished();
        auto core = terms.get_core_calcs(300.0, 12000.0, z);
        out[0] = core.a1kB; out[1] = core.a2kB2; out[2] = core.a3kB3; out[3] = core.alphar_mono; out[4] = core.alphar_chain;
        return 0;
    } catch (...) { return 1; }
}

// out[5] = Delta(0,0), Delta(0,3), Delta(3,3), X_A[0], X_A[3]
int ref_pin_association(double* out) {
    try {
        auto b_m3mol = (Eigen::ArrayXd(2) << 0.0491 / 1e3, 0.0145 / 1e3).finished();
        auto beta = (Eigen::ArrayXd(2) << 8e-3, 69.2e-3).finished();
        auto epsilon_Jmol = (Eigen::ArrayXd(2) << 215.00 * 100, 166.55 * 100).finished();
        std::vector<std::vector<std::string>> molecule_sites = {{"e", "H"}, {"e", "e", "H", "H"}};
        association::AssociationOptions opt;
        opt.radial_dist = association::radial_dists::CS;
        opt.max_iters = 1000;
        opt.interaction_partners = {{"e", {"H"}}, {"H", {"e"}}};
        association::Association a(b_m3mol, beta, epsilon_Jmol, molecule_sites, opt);
        auto molefracs = (Eigen::ArrayXd(2) << 0.3, 0.7).finished();
        auto Ngroups = a.mapper.to_siteid.size();
        Eigen::ArrayXd X_init = Eigen::ArrayXd::Ones(Ngroups);
        double v = 3.0680691201961814e-5, T = 303.15;
        auto Delta = a.get_Delta(T, 1 / v, molefracs);
        Eigen::ArrayXd X_A = a.successive_substitution(T, 1 / v, molefracs, X_init);
        out[0] = Delta(0, 0); out[1] = Delta(0, 3); out[2] = Delta(3, 3); out[3] = X_A[0]; out[4] = X_A[3];
        return 0;
    } catch (...) { return 1; }
}

// out[1] = X_A[0] of Dufal-2015 water at 290 K, 1.011 g/cm3
int ref_pin_dufal_water(double* out) {
    try {
        auto spec = nlohmann::json::parse(R"({
          "nonpolar": {"kind": "SAFT-VR-Mie", "model": {"coeffs": [{"name": "Water", "BibTeXKey": "Dufal-2015", "m": 1.0, "sigma_Angstrom": 3.0555,
                        "epsilon_over_k": 418.00, "lambda_r": 35.823, "lambda_a": 6.0}]}},
          "association": {"kind": "Dufal", "model": {"sigma / m": [3.0555e-10], "epsilon / J/mol": [3475.445374388054], "lambda_r": [35.823],
                        "epsilon_HB / J/mol": [13303.140189045183], "K_HB / m^3": [496.66e-30], "kmat": [[0.0]], "Delta_rule": "Dufal",
                        "molecule_sites": [["e","e","H","H"]]}}})");
        saft::genericsaft::GenericSAFT gen(spec);
        const auto& a = std::get<association::Association>(gen.association.value());
        auto z = (Eigen::ArrayXd(1) << 1.0).finished();
        double rhomolar = 1.011 * 100 * 100 * 100 / 1000 / 0.018015;
        Eigen::ArrayXd X_init(2); X_init = 1.0;
        auto X = a.successive_substitution(290.0, rhomolar, z, X_init);
        out[0] = X[0];
        return 0;
    } catch (...) { return 1; }
}

}  // extern "C"
