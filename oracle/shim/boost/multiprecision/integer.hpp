// Build shim (test infrastructure): the one boost::multiprecision symbol the
// reference's common/utils.cpp:277 uses, so the oracle builds without boost.
#pragma once
#include <cstdint>
namespace boost { namespace multiprecision {
inline unsigned msb(uint64_t a) { return 63u - static_cast<unsigned>(__builtin_clzll(a)); }
} }
