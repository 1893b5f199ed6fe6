// Oracle build prelude (test infrastructure), force-included (-include) ahead of every reference translation unit:
//  * <iostream>: /root/reference/include/teqp/types.hpp:133 uses std::cout without including it;
//  * the boost::multiprecision stand-in and teqp/finite_derivs.hpp, which deriv_adapter.hpp:148-171 uses outside the
//    TEQP_MULTIPRECISION_ENABLED guard that would have included them (deriv_adapter.hpp:7-12).
#pragma once
#include <iostream>
#include <complex>
#include "boost/multiprecision/cpp_bin_float.hpp"
#include "teqp/finite_derivs.hpp"
