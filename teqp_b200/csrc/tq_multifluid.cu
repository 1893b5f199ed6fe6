// JSON-driven multi-fluid models ("kind": "multifluid"): host-side ingestion of the reference's fluid files into the device term program.
//
// Reference: /root/reference/include/teqp/models/multifluid.hpp
//   multifluidfactory :1095-1160 (spec keys: components, root, BIP, departure, flags), build_multifluid_model :1074-1086,
//   make_pure_components_JSON :1020-1072, _build_multifluid_model :948-995 (per-fluid gas constants, identifier selection, F, BIPs),
//   get_EOS_terms :530-819 (term types and how Power / Lemmon2005 blocks are split), build_departure_function :236-477,
//   multifluid_reducing.hpp:11-87 get_BIPdep (hash -> name -> CAS passes, "estimate" flag), :90-113 (xi/zeta form), :134-148 get_Tcvc,
//   multifluid_eosterms.hpp:13-243 (the term formulas).
// The fluid data themselves (teqp/fluiddata/dev/fluids/*.json, dev/mixtures/*.json) are NOT part of this repository: they are read from
// the caller-supplied "root" (or given inline in the spec), exactly as the reference does.
//
// Every term  n tau^t delta^d exp(h_delta(delta) + h_tau(tau))  of the supported types factorises into a density shape and a temperature shape
// (tq_gerg.cuh), so a JSON model compiles into the same term program as GERG-2008 and runs on the same kernels (tq_gerg_codegen.cu for the
// density tables, tq_gerg_jet.cuh for everything that needs jets).  Not factorisable, answered with NotImplemented: ResidualHelmholtzNonAnalytic
// (CO2, water), Chebyshev2D departure functions, ResidualHelmholtzGenericCubic / PCSAFTGrossSadowski2001 pseudo-terms, REFPROP FLD / HMX.BNC input.
#include <algorithm>
#include <cctype>
#include <cmath>
#include <fstream>
#include <set>
#include <sstream>

#include "tq_common.hpp"
#include "tq_gerg_jit.hpp"

namespace tq {

namespace {

bool is_file(const std::string& p) {
    if (p.empty() || p.size() > 4096) return false;
    std::ifstream f(p);
    return f.good() && f.peek() != std::ifstream::traits_type::eof();
}
Json load_json_file(const std::string& path) {
    std::ifstream f(path);
    if (!f) throw InvalidArgument("Path to be loaded does not exist: " + path);
    std::stringstream ss;
    ss << f.rdbuf();
    try { return Json::parse(ss.str()); }
    catch (const std::exception&) { throw InvalidArgument("File at " + path + " is not valid JSON"); }
}
// multilevel_JSON_load (json_tools.hpp:82-125): nothing -> default path; object / non-empty array -> itself; string -> file or JSON text
Json multilevel_load(const Json& j, const std::string& default_path) {
    if (j.is_null() || (j.is_array() && j.size() == 0) || (j.is_string() && j.as_string().empty())) {
        if (default_path.empty()) throw InvalidArgument("default path was not provided, and cannot load this thing");
        return load_json_file(default_path);
    }
    if (j.is_object() || j.is_array()) return j;
    if (j.is_string()) {
        const std::string& s = j.as_string();
        if (s.rfind("PATH::", 0) == 0) return load_json_file(s.substr(6));
        if (is_file(s)) return load_json_file(s);
        return Json::parse(s);
    }
    throw InvalidArgument("Unable to load the argument to multilevel_JSON_load");
}
std::string upper(std::string s) { for (auto& c : s) c = (char)std::toupper((unsigned char)c); return s; }

std::vector<double> arr_or_zero(const Json& term, const char* key, size_t N) {
    if (term.contains(key) && term.at(key).size() > 0) return term.at(key).as_doubles();
    return std::vector<double>(N, 0.0);
}
void same_length(const Json& term, std::initializer_list<const char*> keys, const char* what) {
    std::set<size_t> L;
    for (auto k : keys) L.insert(term.at(k).size());
    if (L.size() != 1) throw InvalidArgument(std::string("Lengths are not all identical in ") + what + " term");
}
int as_int_exponent(double v, const char* what) {
    if (v != std::floor(v) || v < 0 || v > 64) throw InvalidArgument(std::string("Non-integer entry in ") + what + " found");
    return (int)v;
}

// n tau^t delta^d exp(-c delta^l): JustPowerEOSTerm / PowerEOSTerm / ExponentialEOSTerm (multifluid_eosterms.hpp:13-103)
GergTerm power_term(double n, double t, double d, double c, double l) {
    GergTerm r{};
    r.n = n; r.t = t; r.d = as_int_exponent(d, "d");
    r.l = as_int_exponent(l, "l"); r.c = c;
    if (r.l == 0 && c != 0.0) { r.n = n * std::exp(-c); r.c = 0.0; }   // exp(-c delta^0) is a constant factor
    r.kind = (r.c == 0.0) ? GG_POLY : GG_EXP;
    if (r.kind == GG_POLY) { r.l = 0; r.c = 0; }
    return r;
}

// The terms of one "alphar" array of a fluid file (get_EOS_terms, multifluid.hpp:530-819)
void add_pure_terms(const Json& alphar, int w, std::vector<GergTerm>& out, std::vector<GergNA>& na) {
    if (!alphar.is_array() || alphar.size() == 0) throw InvalidArgument("alphar array cannot be empty");
    for (size_t q = 0; q < alphar.size(); ++q) {
        const Json& term = alphar[q];
        const std::string type = term.at("type").as_string();
        const size_t first = out.size();
        if (type == "ResidualHelmholtzPower") {
            const size_t N = term.at("n").size();
            const auto n = arr_or_zero(term, "n", N), t = arr_or_zero(term, "t", N), d = arr_or_zero(term, "d", N), l = arr_or_zero(term, "l", N);
            for (size_t k = 0; k < N; ++k) out.push_back(power_term(n[k], t[k], d[k], l[k] > 0 ? 1.0 : 0.0, l[k]));   // c_i = 1 if l_i > 0 (:583-588)
        } else if (type == "ResidualHelmholtzExponential") {
            same_length(term, {"n", "t", "d", "g", "l"}, "exponential");
            const auto n = term.at("n").as_doubles(), t = term.at("t").as_doubles(), d = term.at("d").as_doubles(), g = term.at("g").as_doubles(), l = term.at("l").as_doubles();
            for (size_t k = 0; k < n.size(); ++k) {
                // exp(-g delta^l) with l cast to int (ExponentialEOSTerm :83-103); l = 0 folds exp(-g) into n
                const int li = (int)l[k];
                out.push_back(power_term(n[k], t[k], d[k], g[k], (double)li));
            }
        } else if (type == "ResidualHelmholtzGaussian") {
            same_length(term, {"n", "t", "d", "eta", "beta", "gamma", "epsilon"}, "Gaussian");
            const auto n = term.at("n").as_doubles(), t = term.at("t").as_doubles(), d = term.at("d").as_doubles(), eta = term.at("eta").as_doubles(),
                       beta = term.at("beta").as_doubles(), gamma = term.at("gamma").as_doubles(), eps = term.at("epsilon").as_doubles();
            for (size_t k = 0; k < n.size(); ++k) {
                // n tau^t delta^d exp(-eta (delta - eps)^2 - beta (tau - gamma)^2)   (GaussianEOSTerm :139-165)
                GergTerm r{};
                r.n = n[k]; r.t = t[k]; r.d = as_int_exponent(d[k], "d");
                r.kind = GG_GAUSS; r.eta = eta[k]; r.eps = eps[k]; r.beta = 0.0; r.gamma = 0.0;
                if (beta[k] != 0.0) { r.tk = GT_GAUSS; r.tp1 = beta[k]; r.tp2 = gamma[k]; }
                out.push_back(r);
            }
        } else if (type == "ResidualHelmholtzGaoB") {
            same_length(term, {"n", "t", "d", "eta", "beta", "gamma", "epsilon", "b"}, "GaoB");
            const auto n = term.at("n").as_doubles(), t = term.at("t").as_doubles(), d = term.at("d").as_doubles(), eta = term.at("eta").as_doubles(),
                       beta = term.at("beta").as_doubles(), gamma = term.at("gamma").as_doubles(), eps = term.at("epsilon").as_doubles(), b = term.at("b").as_doubles();
            for (size_t k = 0; k < n.size(); ++k) {
                // n tau^t delta^d exp(eta_json (delta - eps)^2 + 1 / (beta (tau - gamma)^2 + b)): the factory flips the sign of eta (:763) and the
                // term subtracts it again (GaoBEOSTerm :226-235), i.e. the density exponent is -(-eta_json) (delta - eps)^2
                GergTerm r{};
                r.n = n[k]; r.t = t[k]; r.d = as_int_exponent(d[k], "d");
                r.kind = GG_GAUSS; r.eta = -eta[k]; r.eps = eps[k];
                r.tk = GT_GAOB; r.tp1 = beta[k]; r.tp2 = gamma[k]; r.tp3 = b[k];
                out.push_back(r);
            }
        } else if (type == "ResidualHelmholtzLemmon2005") {
            same_length(term, {"n", "t", "d", "m", "l"}, "Lemmon2005");
            const auto n = term.at("n").as_doubles(), t = term.at("t").as_doubles(), d = term.at("d").as_doubles(), mm = term.at("m").as_doubles(), l = term.at("l").as_doubles();
            for (size_t k = 0; k < n.size(); ++k) {
                // n tau^t delta^d exp(-delta^l - tau^m); m = 0 and/or l = 0 rows are plain power / exponential terms (:650-693)
                GergTerm r = power_term(n[k], t[k], d[k], l[k] > 0 ? 1.0 : 0.0, l[k]);
                if (mm[k] > 0 && l[k] > 0) { r.tk = GT_POWEXP; r.tp1 = 1.0; r.tp2 = mm[k]; }
                else if (mm[k] != 0) throw InvalidArgument("Term subdivision failed in Lemmon2005 term");
                out.push_back(r);
            }
        } else if (type == "ResidualHelmholtzDoubleExponential") {
            same_length(term, {"n", "t", "d", "ld", "gd", "lt", "gt"}, "double exponential");
            const auto n = term.at("n").as_doubles(), t = term.at("t").as_doubles(), d = term.at("d").as_doubles(), ld = term.at("ld").as_doubles(),
                       gd = term.at("gd").as_doubles(), lt = term.at("lt").as_doubles(), gt = term.at("gt").as_doubles();
            for (size_t k = 0; k < n.size(); ++k) {
                // n tau^t delta^d exp(-gd delta^ld - gt tau^lt)   (DoubleExponentialEOSTerm :108-133)
                GergTerm r = power_term(n[k], t[k], d[k], gd[k], (double)(int)ld[k]);
                if (gt[k] != 0.0) { r.tk = GT_POWEXP; r.tp1 = gt[k]; r.tp2 = lt[k]; }
                out.push_back(r);
            }
        } else if (type == "ResidualHelmholtzNonAnalytic") {
            same_length(term, {"n", "A", "B", "C", "D", "a", "b", "beta"}, "nonanalytic");
            const auto n = term.at("n").as_doubles(), A = term.at("A").as_doubles(), B = term.at("B").as_doubles(), Cc = term.at("C").as_doubles(),
                       D = term.at("D").as_doubles(), a = term.at("a").as_doubles(), b = term.at("b").as_doubles(), beta = term.at("beta").as_doubles();
            for (size_t k = 0; k < n.size(); ++k) na.push_back(GergNA{n[k], A[k], B[k], Cc[k], D[k], a[k], b[k], beta[k], w, 0});
        } else if (type == "ResidualHelmholtzGenericCubic" || type == "ResidualHelmholtzPCSAFTGrossSadowski2001") {
            throw NotImplementedError("multifluid pseudo-term type " + type + " is not implemented in libteqpb200");
        } else {
            throw InvalidArgument("Bad term type: " + type);
        }
        for (size_t k = first; k < out.size(); ++k) out[k].w = w;
    }
}

// Departure function terms (build_departure_function, multifluid.hpp:236-477)
std::vector<GergTerm> departure_terms(const Json& j) {
    std::vector<GergTerm> out;
    const std::string type = j.at("type").as_string();
    auto gauss_block = [&](bool gerg_form) {
        same_length(j, {"n", "t", "d", "eta", "beta", "gamma", "epsilon"}, gerg_form ? "GERG" : "Gaussian+Exponential");
        const int Npower = (int)j.at("Npower").as_int();
        const auto n = j.at("n").as_doubles(), t = j.at("t").as_doubles(), d = j.at("d").as_doubles(), eta = j.at("eta").as_doubles(),
                   beta = j.at("beta").as_doubles(), gamma = j.at("gamma").as_doubles(), eps = j.at("epsilon").as_doubles();
        const auto l = j.contains("l") ? j.at("l").as_doubles() : std::vector<double>(n.size(), 0.0);
        for (int k = 0; k < (int)n.size(); ++k) {
            if (k < Npower) { out.push_back(power_term(n[k], t[k], d[k], l[k] > 0 ? 1.0 : 0.0, l[k])); continue; }
            GergTerm r{};
            r.n = n[k]; r.t = t[k]; r.d = as_int_exponent(d[k], "d"); r.kind = GG_GAUSS; r.eta = eta[k]; r.eps = eps[k];
            if (gerg_form) { r.beta = beta[k]; r.gamma = gamma[k]; }                          // exp(-eta (delta-eps)^2 - beta (delta - gamma))   GERG2004EOSTerm :171-193
            else if (beta[k] != 0.0) { r.tk = GT_GAUSS; r.tp1 = beta[k]; r.tp2 = gamma[k]; }  // exp(-eta (delta-eps)^2 - beta (tau - gamma)^2)   GaussianEOSTerm
            if (r.eta == 0.0 && r.beta == 0.0 && r.tk == GT_PLAIN) { r.kind = GG_POLY; r.eps = r.gamma = 0; }
            out.push_back(r);
        }
    };
    if (type == "Exponential") {
        const size_t N = j.at("n").size();
        if (N == 0) return out;
        const auto n = arr_or_zero(j, "n", N), t = arr_or_zero(j, "t", N), d = arr_or_zero(j, "d", N);
        std::vector<double> l(N, 0.0), c(N, 0.0);
        if (j.contains("l") && j.at("l").size() > 0) {
            l = j.at("l").as_doubles();
            if (l.size() != N) throw InvalidArgument("Lengths are not all identical in exponential term");
            if (j.contains("c")) c = arr_or_zero(j, "c", N);
            else for (size_t k = 0; k < N; ++k) c[k] = l[k] > 0 ? 1.0 : 0.0;
        }
        for (size_t k = 0; k < N; ++k) out.push_back(power_term(n[k], t[k], d[k], c[k], l[k]));
    } else if (type == "GERG-2004" || type == "GERG-2008") {
        gauss_block(true);
    } else if (type == "Gaussian+Exponential") {
        gauss_block(false);
    } else if (type == "DoubleExponential") {
        same_length(j, {"n", "t", "d", "ld", "gd", "lt", "gt"}, "double exponential");
        const auto n = j.at("n").as_doubles(), t = j.at("t").as_doubles(), d = j.at("d").as_doubles(), ld = j.at("ld").as_doubles(), gd = j.at("gd").as_doubles(),
                   lt = j.at("lt").as_doubles(), gt = j.at("gt").as_doubles();
        for (size_t k = 0; k < n.size(); ++k) {
            GergTerm r = power_term(n[k], t[k], d[k], gd[k], (double)(int)ld[k]);
            if (gt[k] != 0.0) { r.tk = GT_POWEXP; r.tp1 = gt[k]; r.tp2 = lt[k]; }
            out.push_back(r);
        }
    } else if (type == "none") {
    } else if (type == "Chebyshev2D") {
        throw NotImplementedError("Chebyshev2D departure functions are not implemented in libteqpb200");
    } else {
        throw InvalidArgument("Bad departure term type: " + type + ". Options are {Exponential,GERG-2004,GERG-2008,Gaussian+Exponential,none,DoubleExponential,Chebyshev2D}");
    }
    return out;
}

// get_BIPdep (multifluid_reducing.hpp:11-87): the collection entry of a pair and whether the order is swapped
struct BipHit { const Json* el = nullptr; bool swap = false; bool estimated = false; };
BipHit find_bip(const Json& collection, const std::string& id0, const std::string& id1, const Json& flags) {
    const bool have_flags = flags.is_object();
    if (have_flags && flags.contains("force-estimate") && flags.contains("estimate")) {   // :15-28
        if (flags.at("estimate").as_string() != "Lorentz-Berthelot") throw InvalidArgument("estimation scheme is not understood:" + flags.at("estimate").as_string());
        BipHit h; h.estimated = true; return h;
    }
    const std::string u0 = upper(id0), u1 = upper(id1);
    // 0-th pass: hashes (entries without them are skipped); first pass: names (both case-insensitive)
    for (int pass = 0; pass < 2; ++pass) {
        const char* k1 = pass == 0 ? "hash1" : "Name1";
        const char* k2 = pass == 0 ? "hash2" : "Name2";
        for (size_t q = 0; q < collection.size(); ++q) {
            const Json& el = collection[q];
            if (pass == 0 && !el.contains(k1)) continue;
            const std::string n1 = upper(el.at(k1).as_string()), n2 = upper(el.at(k2).as_string());
            if (u0 == n1 && u1 == n2) return BipHit{&el, false, false};
            if (u0 == n2 && u1 == n1) return BipHit{&el, true, false};
        }
    }
    for (size_t q = 0; q < collection.size(); ++q) {
        const Json& el = collection[q];
        const std::string& c1 = el.at("CAS1").as_string(); const std::string& c2 = el.at("CAS2").as_string();
        if (id0 == c1 && id1 == c2) return BipHit{&el, false, false};
        if (id0 == c2 && id1 == c1) return BipHit{&el, true, false};
    }
    if (have_flags && flags.contains("estimate")) {
        if (flags.at("estimate").as_string() != "Lorentz-Berthelot") throw InvalidArgument("estimation scheme is not understood:" + flags.at("estimate").as_string());
        BipHit h; h.estimated = true; return h;
    }
    throw InvalidArgument("Can't match the binary pair for: " + id0 + "/" + id1);
}

const Json& find_departure(const Json& depcollection, const std::string& name) {
    for (size_t q = 0; q < depcollection.size(); ++q) if (depcollection[q].at("Name").as_string() == name) return depcollection[q];
    throw InvalidArgument("Bad departure function name: " + name);
}

}  // namespace

std::unique_ptr<Model> make_multifluid_model(const Json& spec) {
    if (spec.contains("HMX.BNC")) throw NotImplementedError("REFPROP HMX.BNC / FLD input is not implemented in libteqpb200");
    const Json flags = spec.contains("flags") ? spec.at("flags") : Json();
    const std::string root = spec.contains("root") ? spec.at("root").as_string() : std::string();
    const Json& components = spec.at("components");
    if (!components.is_array() || components.size() == 0) throw InvalidArgument("Must be an array");
    const size_t N = components.size();

    // ---- pure fluids (make_pure_components_JSON, multifluid.hpp:1020-1072; names are looked up under root/dev/fluids) ------------------
    std::vector<Json> pure;
    for (size_t i = 0; i < N; ++i) {
        const Json& c = components[i];
        if (c.is_object()) { pure.push_back(c); continue; }
        const std::string s = c.as_string();
        if (s.rfind("PATH::", 0) == 0) { pure.push_back(load_json_file(s.substr(6))); continue; }
        if (s.rfind("FLDPATH::", 0) == 0 || s.rfind("FLD::", 0) == 0) throw NotImplementedError("REFPROP FLD input is not implemented in libteqpb200");
        if (is_file(s)) { pure.push_back(load_json_file(s)); continue; }
        const std::string cand = root + "/dev/fluids/" + s + ".json";
        if (!root.empty() && is_file(cand)) { pure.push_back(load_json_file(cand)); continue; }
        if (!s.empty() && s[0] == '{') { pure.push_back(Json::parse(s)); continue; }
        throw InvalidArgument("Could not load any of the candidates:" + s + " (fluid files are looked up by file name under <root>/dev/fluids; the alias map of the reference is not built)");
    }

    TermProgramInput in;
    in.family = "multifluid";
    in.limit_form = true;
    in.pure_shortcut = true;
    std::vector<std::string> idCAS, idName, idRP, idHash;
    for (size_t i = 0; i < N; ++i) {
        const Json& eos = pure[i].at("EOS")[0];
        const Json& red = eos.at("STATES").at("reducing");                       // get_Tcvc, multifluid_reducing.hpp:134-148
        in.Tc.push_back(red.at("T").as_double());
        in.vc.push_back(1.0 / red.at("rhomolar").as_double());
        in.Ri.push_back(eos.at("gas_constant").as_double());
        add_pure_terms(eos.at("alphar"), (int)i, in.pure, in.na);
        const Json& info = pure[i].at("INFO");
        idName.push_back(info.at("NAME").as_string());
        idCAS.push_back(info.at("CAS").as_string());
        idRP.push_back(info.at("REFPROP_NAME").as_string());
        if (info.contains("HASH")) idHash.push_back(info.at("HASH").as_string());
        in.names.push_back(info.at("NAME").as_string());
    }
    in.r_weighted = true;                                                         // MoleFractionWeighted (multifluid.hpp:973)
    if (flags.is_object() && flags.contains("Rmodel") && flags.at("Rmodel").as_string() == "CODATA") { in.r_weighted = false; in.Rgas = 8.31446261815324; }

    // ---- mixture part ---------------------------------------------------------------------------------------------------------------------
    if (N > 1) {
        const Json BIP = multilevel_load(spec.contains("BIP") ? spec.at("BIP") : Json(), root.empty() ? std::string() : root + "/dev/mixtures/mixture_binary_pairs.json");
        Json dep;
        if (spec.contains("departure")) dep = multilevel_load(spec.at("departure"), root.empty() ? std::string() : root + "/dev/mixtures/mixture_departure_functions.json");
        else if (!root.empty() && is_file(root + "/dev/mixtures/mixture_departure_functions.json")) dep = load_json_file(root + "/dev/mixtures/mixture_departure_functions.json");
        // select_identifier (multifluid.hpp:870-896): the first identifier kind, in the std::map order CAS < Name < REFPROP < hash, that matches every pair
        std::vector<std::pair<std::string, const std::vector<std::string>*>> kinds = {{"CAS", &idCAS}, {"Name", &idName}, {"REFPROP", &idRP}};
        if (idHash.size() == N) kinds.push_back({"hash", &idHash});
        const std::vector<std::string>* ids = nullptr;
        for (auto& [key, v] : kinds) {
            try {
                for (size_t i = 0; i < N; ++i) for (size_t j = i + 1; j < N; ++j) find_bip(BIP, (*v)[i], (*v)[j], flags);
                ids = v;
                break;
            } catch (const std::exception&) {}
        }
        if (!ids) throw InvalidArgument("Unable to match any of the identifier options: CAS,Name,REFPROP");
        for (size_t i = 0; i < N; ++i) {
            for (size_t j = i + 1; j < N; ++j) {
                const BipHit h = find_bip(BIP, (*ids)[i], (*ids)[j], flags);
                TermProgramInput::Pair p;
                if (!h.estimated) {
                    const Json& el = *h.el;
                    if (el.contains("betaT") && el.contains("gammaT") && el.contains("betaV") && el.contains("gammaV")) {
                        p.betaT = el.at("betaT").as_double(); p.gammaT = el.at("gammaT").as_double();
                        p.betaV = el.at("betaV").as_double(); p.gammaV = el.at("gammaV").as_double();
                        if (h.swap) { p.betaT = 1.0 / p.betaT; p.betaV = 1.0 / p.betaV; }   // backwards order of components (:97-100)
                    } else if (el.contains("xi") && el.contains("zeta")) {                    // :102-108
                        const double xi = el.at("xi").as_double(), zeta = el.at("zeta").as_double();
                        const double Tc0 = in.Tc[i], Tc1 = in.Tc[j], vc0 = in.vc[i], vc1 = in.vc[j];
                        p.gammaT = 0.5 * (Tc0 + Tc1 + xi) / (2 * std::sqrt(Tc0 * Tc1));
                        p.gammaV = 4.0 * (vc0 + vc1 + zeta) / (0.25 * std::pow(1 / std::pow(1 / vc0, 1.0 / 3.0) + 1 / std::pow(1 / vc1, 1.0 / 3.0), 3));
                        p.betaT = p.betaV = 1.0;
                    } else {
                        throw InvalidArgument("Could not understand what to do with this binary model specification: " + el.dump());
                    }
                    p.F = el.at("F").as_double();
                    const std::string fn = el.contains("function") ? el.at("function").as_string() : std::string();
                    if (!fn.empty()) {
                        if (!dep.is_array() || dep.size() == 0) throw InvalidArgument("No departure functions were loaded, unable to select requested function: " + fn);
                        p.terms = departure_terms(find_departure(dep, fn));
                    }
                }
                in.pairs.push_back(std::move(p));
            }
        }
    }
    in.meta = "{\"kind\": \"multifluid\", \"components\": " + std::to_string(N) + "}";
    return make_term_program_model(in);
}

}  // namespace tq
