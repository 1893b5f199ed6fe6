// Oracle shim (test infrastructure): stands in for <autodiff/forward/dual/eigen.hpp>.
#pragma once
#include "../../../autodiff_shim_eigen.hpp"
