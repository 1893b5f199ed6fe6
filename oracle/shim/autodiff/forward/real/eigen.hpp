// Oracle shim (test infrastructure): stands in for <autodiff/forward/real/eigen.hpp>.
#pragma once
#include "../../../autodiff_shim_eigen.hpp"
