// Oracle shim (test infrastructure): stands in for <autodiff/forward/real.hpp> of autodiff @ cb2d6e8b (absent from /root/reference).
#pragma once
#include "../../autodiff_shim.hpp"
