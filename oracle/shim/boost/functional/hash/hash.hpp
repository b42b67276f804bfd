// Build shim (test infrastructure): boost::hash_range as used at the
// reference's common/utils.cpp:14 (off the query path; hash_combine recipe).
#pragma once
#include <cstddef>
#include <functional>
namespace boost {
template <class It>
inline std::size_t hash_range(It first, It last) {
    std::size_t seed = 0;
    for (; first != last; ++first) {
        std::hash<typename std::decay<decltype(*first)>::type> h;
        seed ^= h(*first) + 0x9e3779b9 + (seed << 6) + (seed >> 2);
    }
    return seed;
}
}
