// Minimal stand-in for <boost/functional/hash.hpp> (boost is not installed in this image).
// TEST INFRASTRUCTURE ONLY: lets the UNMODIFIED /root/reference/src/pairwise.cpp compile into
// oracle/_ref/pairwise_ref.  Only boost::hash<std::pair<uint32_t,uint32_t>> is used by the
// reference (src/pairwise.cpp:25); the hash only decides hash-map iteration order, i.e. the TSV
// row order, which the reference leaves unspecified.  Written from the published boost
// hash_combine recipe, not copied from boost.
#pragma once
#include <cstddef>
#include <functional>
#include <utility>

namespace boost {

template <class T>
inline void hash_combine(std::size_t& seed, const T& v) {
    seed ^= std::hash<T>()(v) + 0x9e3779b9 + (seed << 6) + (seed >> 2);
}

template <class T>
struct hash {
    std::size_t operator()(const T& v) const { return std::hash<T>()(v); }
};

template <class A, class B>
struct hash<std::pair<A, B>> {
    std::size_t operator()(const std::pair<A, B>& p) const {
        std::size_t seed = 0;
        hash_combine(seed, p.first);
        hash_combine(seed, p.second);
        return seed;
    }
};

}  // namespace boost
