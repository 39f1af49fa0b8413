// edsbwt_pack_patterns: pattern bytes -> the 2-bit words edsbwt_search_into_packed takes (layout: include/edsbwt.h,
// internal.cuh PackedPats).  Host-only, compiled by g++ (no device code): this is the one place where the product does
// arithmetic on the CPU, and it is text ingestion, the counterpart of the reference's std::getline loop
// (EDSBWTsearch.cpp:207-241) -- done once per pattern set, not per search.
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#if defined(__x86_64__)
#include <immintrin.h>
#endif

#include "../../include/edsbwt.h"

namespace {

// bit j of x -> bit 2j
inline uint64_t spread32(uint64_t x) {
    x = (x | (x << 16)) & 0x0000FFFF0000FFFFull;
    x = (x | (x << 8)) & 0x00FF00FF00FF00FFull;
    x = (x | (x << 4)) & 0x0F0F0F0F0F0F0F0Full;
    x = (x | (x << 2)) & 0x3333333333333333ull;
    x = (x | (x << 1)) & 0x5555555555555555ull;
    return x;
}

struct Alphabet { uint8_t code[256]; uint8_t letter[4]; int nletters; };

// n <= 32 characters at p -> codes, character j at bits 2j; false when a byte is not a letter
inline bool pack_scalar(const uint8_t* p, uint32_t n, const Alphabet& al, uint64_t& w) {
    uint64_t v = 0; uint32_t bad = 0;
    for (uint32_t j = 0; j < n; ++j) { const uint32_t c = al.code[p[j]]; bad |= c; v |= (uint64_t)(c & 3u) << (2 * j); }
    w = v;
    return bad < 4;
}

#if defined(__x86_64__)
__attribute__((target("avx2"))) inline bool pack32_avx2(const uint8_t* p, const Alphabet& al, uint64_t& w) {
    const __m256i v = _mm256_loadu_si256((const __m256i*)p);
    const __m256i e0 = _mm256_cmpeq_epi8(v, _mm256_set1_epi8((char)al.letter[0]));
    const __m256i e1 = al.nletters > 1 ? _mm256_cmpeq_epi8(v, _mm256_set1_epi8((char)al.letter[1])) : _mm256_setzero_si256();
    const __m256i e2 = al.nletters > 2 ? _mm256_cmpeq_epi8(v, _mm256_set1_epi8((char)al.letter[2])) : _mm256_setzero_si256();
    const __m256i e3 = al.nletters > 3 ? _mm256_cmpeq_epi8(v, _mm256_set1_epi8((char)al.letter[3])) : _mm256_setzero_si256();
    const uint32_t b0 = (uint32_t)_mm256_movemask_epi8(_mm256_or_si256(e1, e3));       // low bit of the code
    const uint32_t b1 = (uint32_t)_mm256_movemask_epi8(_mm256_or_si256(e2, e3));       // high bit
    const uint32_t ok = (uint32_t)_mm256_movemask_epi8(_mm256_or_si256(_mm256_or_si256(e0, e1), _mm256_or_si256(e2, e3)));
    w = spread32(b0) | (spread32(b1) << 1);
    return ok == 0xffffffffu;
}
#endif

void pack_range(const uint8_t* bytes, uint32_t pat_len, uint32_t lo_pat, uint32_t hi_pat, const Alphabet& al, bool avx2,
                uint64_t* words, uint32_t* invalid, uint32_t* n_bad) {
    const uint32_t stride = (pat_len + 31) / 32;
    uint32_t nbad = 0;
    for (uint32_t i = lo_pat; i < hi_pat; ++i) {
        const uint8_t* p = bytes + (uint64_t)i * pat_len;
        bool ok = true;
        for (uint32_t hi = pat_len, w = 0; hi > 0; ++w) {
            const uint32_t lo = hi > 32 ? hi - 32 : 0, n = hi - lo;
            uint64_t cw;
#if defined(__x86_64__)
            if (n == 32 && avx2) ok &= pack32_avx2(p + lo, al, cw);
            else
#endif
                ok &= pack_scalar(p + lo, n, al, cw);
            words[(uint64_t)i * stride + w] = cw << (2 * (32 - n));
            hi = lo;
        }
        if (!ok) { ++nbad; __atomic_fetch_or(&invalid[i >> 5], 1u << (i & 31), __ATOMIC_RELAXED); }
    }
    *n_bad = nbad;
}

}  // namespace

extern "C" int edsbwt_pack_patterns_alphabet(const uint8_t* alphabet, uint32_t sigma, const uint8_t* pat_bytes, uint32_t pat_len, uint32_t npat,
                                             int nthreads, uint64_t* words, uint32_t* invalid, uint32_t* n_invalid) {
    if (!alphabet || sigma < 2 || sigma > 5 || !pat_len || (npat && (!pat_bytes || !words || !invalid))) return EDSBWT_E_ARG;
    Alphabet al; memset(al.code, 255, sizeof al.code); memset(al.letter, 0, sizeof al.letter);
    al.nletters = (int)sigma - 1;
    for (uint32_t a = 1; a < sigma; ++a) { al.code[alphabet[a]] = (uint8_t)(a - 1); al.letter[a - 1] = alphabet[a]; }
    memset(invalid, 0, (size_t)((npat + 31) / 32) * 4);
    bool avx2 = false;
#if defined(__x86_64__)
    avx2 = __builtin_cpu_supports("avx2");
#endif
    if (nthreads < 1) nthreads = 1;
    if ((uint32_t)nthreads > npat / 4096 + 1) nthreads = (int)(npat / 4096 + 1);
    std::vector<uint32_t> bad((size_t)nthreads, 0);
    if (nthreads == 1) pack_range(pat_bytes, pat_len, 0, npat, al, avx2, words, invalid, &bad[0]);
    else {
        std::vector<std::thread> th;
        // cut at multiples of 32 patterns: two threads never touch the same word of the invalid bitmap
        const uint32_t per = (uint32_t)((((uint64_t)npat + nthreads - 1) / nthreads + 31) & ~31ull);
        for (int t = 0; t < nthreads; ++t) {
            const uint64_t lo = (uint64_t)t * per, hi = lo + per < npat ? lo + per : npat;
            if (lo >= npat) break;
            th.emplace_back(pack_range, pat_bytes, pat_len, (uint32_t)lo, (uint32_t)hi, std::cref(al), avx2, words, invalid, &bad[t]);
        }
        for (auto& x : th) x.join();
    }
    if (n_invalid) { uint32_t s = 0; for (uint32_t b : bad) s += b; *n_invalid = s; }
    return EDSBWT_OK;
}
