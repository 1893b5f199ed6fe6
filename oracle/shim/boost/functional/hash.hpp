// Oracle shim (test infrastructure): the one boost facility used by /root/reference/include/teqp/models/GERG/GERG.hpp:17,598,1008
// (boost::hash of a pair of strings as an unordered_map hasher).  The boost bundle is a missing blob (.MISSING_LARGE_BLOBS:1).
#pragma once
#include <functional>
#include <string>
#include <utility>
namespace boost {
template<typename K> struct hash;
template<> struct hash<std::pair<std::string, std::string>> {
    std::size_t operator()(const std::pair<std::string, std::string>& p) const noexcept {
        std::size_t h1 = std::hash<std::string>{}(p.first), h2 = std::hash<std::string>{}(p.second);
        return h1 ^ (h2 + 0x9e3779b97f4a7c15ULL + (h1 << 6) + (h1 >> 2));
    }
};
}
