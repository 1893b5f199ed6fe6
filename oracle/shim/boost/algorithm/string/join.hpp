// Oracle shim (test infrastructure): boost::algorithm::join as used at /root/reference/include/teqp/models/multifluid.hpp:27,482.
#pragma once
#include <string>
namespace boost { namespace algorithm {
template<typename Seq> std::string join(const Seq& parts, const std::string& sep) {
    std::string out; bool first = true;
    for (const auto& p : parts) { if (!first) out += sep; out += p; first = false; }
    return out;
}
}}
