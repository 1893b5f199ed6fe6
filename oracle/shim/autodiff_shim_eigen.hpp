// TEST INFRASTRUCTURE ONLY -- Eigen glue for the oracle's AD shim (see autodiff_shim.hpp for provenance).
// Provides what /root/reference/include/teqp/derivs.hpp:1031-1124 uses: ArrayXdual, ArrayXdual2nd,
// gradient(f, wrt(v), at(v)) and hessian(f, wrt(v), at(v), u, g), plus NumTraits/ScalarBinaryOpTraits so that
// Eigen array expressions mixing double with AD scalars compile.
#pragma once
#include "autodiff_shim.hpp"
#include <Eigen/Dense>

namespace Eigen {
template<typename T, typename G> struct NumTraits<autodiff::detail::Dual<T,G>> : NumTraits<double> {
    typedef autodiff::detail::Dual<T,G> Real; typedef Real NonInteger; typedef Real Nested; typedef double Literal;
    enum { IsComplex = 0, IsInteger = 0, IsSigned = 1, RequireInitialization = 1, ReadCost = 2, AddCost = 4, MulCost = 8 };
};
template<size_t N, typename T> struct NumTraits<autodiff::detail::Real<N,T>> : NumTraits<double> {
    typedef autodiff::detail::Real<N,T> Real; typedef Real NonInteger; typedef Real Nested; typedef double Literal;
    enum { IsComplex = 0, IsInteger = 0, IsSigned = 1, RequireInitialization = 1, ReadCost = 2, AddCost = 4, MulCost = 8 };
};
#define SHIM_BINOP_TRAITS(HEAD, AD) \
HEAD struct ScalarBinaryOpTraits<AD, double, Op> { typedef AD ReturnType; }; \
HEAD struct ScalarBinaryOpTraits<double, AD, Op> { typedef AD ReturnType; }; \
HEAD struct ScalarBinaryOpTraits<AD, int, Op> { typedef AD ReturnType; }; \
HEAD struct ScalarBinaryOpTraits<int, AD, Op> { typedef AD ReturnType; };
#define SHIM_BINOP_TRAITS_Q(HEAD, AD) \
HEAD struct ScalarBinaryOpTraits<AD, __float128, Op> { typedef AD ReturnType; }; \
HEAD struct ScalarBinaryOpTraits<__float128, AD, Op> { typedef AD ReturnType; };
#define SHIM_COMMA ,
SHIM_BINOP_TRAITS(template<typename T SHIM_COMMA typename G SHIM_COMMA typename Op>, autodiff::detail::Dual<T SHIM_COMMA G>)
SHIM_BINOP_TRAITS(template<size_t N SHIM_COMMA typename T SHIM_COMMA typename Op>, autodiff::detail::Real<N SHIM_COMMA T>)
#if defined(SHIM_QUAD)
// __float128 as an Eigen scalar and as the innermost coefficient type of Real / Dual (oracle/ref_truth.cpp only)
template<> struct NumTraits<__float128> : GenericNumTraits<__float128> {
    typedef __float128 Real; typedef __float128 NonInteger; typedef __float128 Nested; typedef __float128 Literal;
    enum { IsComplex = 0, IsInteger = 0, IsSigned = 1, RequireInitialization = 0, ReadCost = 2, AddCost = 8, MulCost = 16 };
    static inline __float128 epsilon() { return 1.92592994438723585305597794258492732e-34Q; }
    static inline __float128 dummy_precision() { return 1e-30Q; }
    static inline __float128 highest() { return 1.18973149535723176508575932662800702e4932Q; }
    static inline __float128 lowest() { return -1.18973149535723176508575932662800702e4932Q; }
    static inline int digits10() { return 33; }
    static inline int digits() { return 113; }
};
template<typename Op> struct ScalarBinaryOpTraits<__float128, double, Op> { typedef __float128 ReturnType; };
template<typename Op> struct ScalarBinaryOpTraits<double, __float128, Op> { typedef __float128 ReturnType; };
template<typename Op> struct ScalarBinaryOpTraits<__float128, int, Op> { typedef __float128 ReturnType; };
template<typename Op> struct ScalarBinaryOpTraits<int, __float128, Op> { typedef __float128 ReturnType; };
SHIM_BINOP_TRAITS_Q(template<typename T SHIM_COMMA typename G SHIM_COMMA typename Op>, autodiff::detail::Dual<T SHIM_COMMA G>)
SHIM_BINOP_TRAITS_Q(template<size_t N SHIM_COMMA typename T SHIM_COMMA typename Op>, autodiff::detail::Real<N SHIM_COMMA T>)
#endif
#undef SHIM_COMMA
#undef SHIM_BINOP_TRAITS
#undef SHIM_BINOP_TRAITS_Q
}

namespace autodiff {
using ArrayXreal = Eigen::Array<real, Eigen::Dynamic, 1>;
using ArrayXdual = Eigen::Array<dual, Eigen::Dynamic, 1>;
using ArrayXdual1st = ArrayXdual;
using ArrayXdual2nd = Eigen::Array<dual2nd, Eigen::Dynamic, 1>;
using VectorXdual = Eigen::Matrix<dual, Eigen::Dynamic, 1>;
using VectorXdual2nd = Eigen::Matrix<dual2nd, Eigen::Dynamic, 1>;

namespace detail {

// gradient(f, wrt(x), at(x)) with x an Eigen vector of first-order (or higher) duals: one sweep per entry
template<typename Fun, typename Vec, typename Arg>
Eigen::VectorXd gradient(const Fun& f, const Wrt<Vec>& w, const At<Arg>& a) {
    auto& x = std::get<0>(w.args);
    const Eigen::Index n = x.size();
    Eigen::VectorXd g(n);
    for (Eigen::Index i = 0; i < n; ++i) {
        seed<1>(x[i], 1.0);
        auto u = std::apply(f, a.args);
        seed<1>(x[i], 0.0);
        g[i] = derivative<1>(u);
    }
    return g;
}

// hessian(f, wrt(x), at(x), u, g): fills the function value u, gradient g, returns the dense Hessian
template<typename Fun, typename Vec, typename Arg, typename U, typename GVec>
Eigen::MatrixXd hessian(const Fun& f, const Wrt<Vec>& w, const At<Arg>& a, U& u, GVec& g) {
    auto& x = std::get<0>(w.args);
    const Eigen::Index n = x.size();
    Eigen::MatrixXd H(n, n);
    g.resize(n);
    for (Eigen::Index i = 0; i < n; ++i) {
        for (Eigen::Index j = i; j < n; ++j) {
            seed<1>(x[i], 1.0);
            seed<2>(x[j], 1.0);
            u = std::apply(f, a.args);
            seed<1>(x[i], 0.0);
            seed<2>(x[j], 0.0);
            H(i, j) = H(j, i) = derivative<2>(u);
            if (j == i) g[i] = derivative<1>(u);
        }
    }
    return H;
}
} // namespace detail
using detail::gradient;
using detail::hessian;
} // namespace autodiff
