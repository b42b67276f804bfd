// AVX-512 form of the rrr_vector<63> block walk (ggq_format.hpp: Binomial63::decode_batch) — host code of the LOAD path,
// compiled by the host compiler alone (no CUDA front end in the way of <immintrin.h>), for this one function only
// (target attribute: the rest of the library stays baseline x86-64); chosen at run time when the CPU has AVX-512F.
//
// 16 blocks are walked together, 8 per zmm register set: per position one 8-lane gather of C(62 - i, k), an unsigned
// compare into a mask, and masked subtract / decrement / or.  Measured 53 ns per block against ~125 ns for the
// eight-lane scalar walk and 300-550 ns for the one-block loop (dense blocks, one Sapphire-Rapids-class core).
#include <cstdint>

#if defined(__x86_64__) && defined(__GNUC__)
#include <immintrin.h>

extern "C" int ggq_rrr_simd_available() {
    static const int ok = __builtin_cpu_supports("avx512f") ? 1 : 0;
    return ok;
}

// walk: Binomial63::walk (walk[n][0] = 2^63).  k / nr: 16 lanes (unused lanes: k = 0, nr = 0).  Returns 1, or 0 when some
// lane did not end at k == 0 and nr == 0 (a corrupt block).
extern "C" __attribute__((target("avx512f"))) int ggq_rrr_walk16(const uint64_t (*walk)[64], const unsigned *k_in,
                                                                const uint64_t *nr_in, uint64_t *bits_out) {
    __m512i k0 = _mm512_cvtepu32_epi64(_mm256_loadu_si256(reinterpret_cast<const __m256i *>(k_in)));
    __m512i k1 = _mm512_cvtepu32_epi64(_mm256_loadu_si256(reinterpret_cast<const __m256i *>(k_in + 8)));
    __m512i n0 = _mm512_loadu_si512(nr_in), n1 = _mm512_loadu_si512(nr_in + 8);
    __m512i b0 = _mm512_setzero_si512(), b1 = b0;
    const __m512i minus1 = _mm512_set1_epi64(-1);
    for (unsigned i = 0; i < 63; ++i) {
        const void *row = walk[62 - i];
        const __m512i bit = _mm512_set1_epi64(static_cast<long long>(uint64_t(1) << i));
        const __m512i c0 = _mm512_i64gather_epi64(k0, row, 8), c1 = _mm512_i64gather_epi64(k1, row, 8);
        const __mmask8 t0 = _mm512_cmpge_epu64_mask(n0, c0), t1 = _mm512_cmpge_epu64_mask(n1, c1);   // the bit is set
        n0 = _mm512_mask_sub_epi64(n0, t0, n0, c0);
        n1 = _mm512_mask_sub_epi64(n1, t1, n1, c1);
        k0 = _mm512_mask_add_epi64(k0, t0, k0, minus1);
        k1 = _mm512_mask_add_epi64(k1, t1, k1, minus1);
        b0 = _mm512_mask_or_epi64(b0, t0, b0, bit);
        b1 = _mm512_mask_or_epi64(b1, t1, b1, bit);
    }
    _mm512_storeu_si512(bits_out, b0);
    _mm512_storeu_si512(bits_out + 8, b1);
    const __m512i rest = _mm512_or_si512(_mm512_or_si512(k0, n0), _mm512_or_si512(k1, n1));
    return _mm512_reduce_or_epi64(rest) == 0 ? 1 : 0;
}

#else

extern "C" int ggq_rrr_simd_available() { return 0; }
extern "C" int ggq_rrr_walk16(const uint64_t (*)[64], const unsigned *, const uint64_t *, uint64_t *) { return 0; }

#endif
