// Oracle shim (test infrastructure): stands in for <boost/multiprecision/cpp_bin_float.hpp>, which the reference's
// type-erasure layer needs only so that get_Ar01ep/get_Ar02ep/get_Ar03ep compile
// (/root/reference/include/teqp/cpp/deriv_adapter.hpp:148-171).  The boost bundle is a missing blob
// (/root/reference/.MISSING_LARGE_BLOBS:1).  `number<cpp_bin_float<N>>` is mapped onto `long double`; the *ep entry points are
// therefore NOT extended precision in the oracle build and are never called by the tests.
#pragma once
#include <Eigen/Core>
namespace boost { namespace multiprecision {
template<unsigned Digits> struct cpp_bin_float {};
#if defined(SHIM_EP_DOUBLE)
// (the multifluid term container's std::visit needs one return type for every alternative; GenericCubicTerm / PCSAFTGrossSadowski2001Term
// return double for a long double density, so that TU maps the stand-in onto double)
template<typename Backend> using number = double;
#else
template<typename Backend> using number = long double;
#endif
}}
namespace Eigen {
template<typename Op> struct ScalarBinaryOpTraits<double, long double, Op> { typedef long double ReturnType; };
template<typename Op> struct ScalarBinaryOpTraits<long double, double, Op> { typedef long double ReturnType; };
}
