// TEST INFRASTRUCTURE ONLY -- link-time stand-ins for the non-pure virtual members of teqp::cppinterface::AbstractModel that the
// reference implements in /root/reference/interface/CPP/AbstractModel_implementation.cpp:14-77.  That TU cannot be compiled
// here (it pulls in include/teqp/algorithms/critical_tracing.hpp:15 -> boost/numeric/odeint, a missing blob); these members
// are the phase-equilibrium / critical-curve *callers* of the hot path, never used by the oracle.
#include "teqp/cpp/teqpcpp.hpp"
#include "teqp/exceptions.hpp"
namespace teqp { namespace cppinterface {
#define NOPE throw teqp::NotImplementedError("algorithm layer is not part of the oracle build")
std::tuple<EArrayd, EArrayd> AbstractModel::get_drhovecdp_Tsat(const double, const REArrayd&, const REArrayd&) const { NOPE; }
std::tuple<EArrayd, EArrayd> AbstractModel::get_drhovecdT_psat(const double, const REArrayd&, const REArrayd&) const { NOPE; }
double AbstractModel::get_dpsat_dTsat_isopleth(const double, const REArrayd&, const REArrayd&) const { NOPE; }
nlohmann::json AbstractModel::trace_VLE_isotherm_binary(const double, const EArrayd&, const EArrayd&, const std::optional<TVLEOptions>&) const { NOPE; }
nlohmann::json AbstractModel::trace_VLE_isobar_binary(const double, const double, const EArrayd&, const EArrayd&, const std::optional<PVLEOptions>&) const { NOPE; }
std::tuple<VLE_return_code, EArrayd, EArrayd> AbstractModel::mix_VLE_Tx(const double, const REArrayd&, const REArrayd&, const REArrayd&, const double, const double, const double, const double, const int) const { NOPE; }
MixVLEReturn AbstractModel::mix_VLE_Tp(const double, const double, const REArrayd&, const REArrayd&, const std::optional<MixVLETpFlags>&) const { NOPE; }
std::tuple<VLE_return_code, double, EArrayd, EArrayd> AbstractModel::mixture_VLE_px(const double, const REArrayd&, const double, const REArrayd&, const REArrayd&, const std::optional<MixVLEpxFlags>&) const { NOPE; }
nlohmann::json AbstractModel::trace_critical_arclength_binary(const double, const EArrayd&, const std::optional<std::string>&, const std::optional<TCABOptions>&) const { NOPE; }
EArrayd AbstractModel::get_drhovec_dT_crit(const double, const REArrayd&) const { NOPE; }
double AbstractModel::get_dp_dT_crit(const double, const REArrayd&) const { NOPE; }
EArray2 AbstractModel::get_criticality_conditions(const double, const REArrayd&) const { NOPE; }
EigenData AbstractModel::eigen_problem(const double, const REArrayd&, const std::optional<REArrayd>&) const { NOPE; }
double AbstractModel::get_minimum_eigenvalue_Psi_Hessian(const double, const REArrayd&) const { NOPE; }
}}
