// hostscan_avx2.cpp -- the two byte scans of the read indexer (hostpipe.cpp) with 256-bit compares.  This file alone
// is compiled with -mavx2; hostpipe.cpp calls it only when the CPU reports AVX2 (scalar memchr loops otherwise).
#include <immintrin.h>
#include <stddef.h>
#include <stdint.h>

extern "C" size_t ccbh_count_newlines_avx2(const uint8_t* p, size_t n)
{
    const __m256i nl = _mm256_set1_epi8('\n');
    size_t c = 0, i = 0;
    for (; i + 32 <= n; i += 32)
        c += (size_t)__builtin_popcount((unsigned)_mm256_movemask_epi8(_mm256_cmpeq_epi8(_mm256_loadu_si256((const __m256i*)(p + i)), nl)));
    for (; i < n; i++) c += p[i] == '\n';
    return c;
}

// offsets (relative to p) of every '\n' in p[0, n), ascending; out must hold n entries; returns how many
extern "C" size_t ccbh_newline_offsets_avx2(const uint8_t* p, size_t n, uint32_t* out)
{
    const __m256i nl = _mm256_set1_epi8('\n');
    size_t c = 0, i = 0;
    for (; i + 32 <= n; i += 32) {
        unsigned m = (unsigned)_mm256_movemask_epi8(_mm256_cmpeq_epi8(_mm256_loadu_si256((const __m256i*)(p + i)), nl));
        while (m) { out[c++] = (uint32_t)(i + (size_t)__builtin_ctz(m)); m &= m - 1; }
    }
    for (; i < n; i++) if (p[i] == '\n') out[c++] = (uint32_t)i;
    return c;
}

// does p[0, n) hold a record marker ('>' or '@') that stands alone on its line ("\n>\n", "\n@\n")?  Such a record's name
// scan runs into the next line (src/CuCLARK_hh.hh:1370): only the sequential indexer reproduces that.
extern "C" int ccbh_has_bare_marker_avx2(const uint8_t* p, size_t n)
{
    const __m256i nl = _mm256_set1_epi8('\n'), a = _mm256_set1_epi8('@'), g = _mm256_set1_epi8('>');
    uint64_t carry_nl = 0, carry_mk = 0;                 // the last two positions of the previous block, in bits 0..1
    size_t i = 0;
    for (; i + 32 <= n; i += 32) {
        const __m256i v = _mm256_loadu_si256((const __m256i*)(p + i));
        const uint64_t mn = (uint32_t)_mm256_movemask_epi8(_mm256_cmpeq_epi8(v, nl));
        const uint64_t mm = (uint32_t)_mm256_movemask_epi8(_mm256_or_si256(_mm256_cmpeq_epi8(v, a), _mm256_cmpeq_epi8(v, g)));
        const uint64_t NL = (mn << 2) | carry_nl, MK = (mm << 2) | carry_mk;     // bit j = position i - 2 + j
        if (NL & (MK >> 1) & (NL >> 2)) return 1;
        carry_nl = mn >> 30; carry_mk = mm >> 30;
    }
    for (size_t j = i > 2 ? i - 2 : 0; j + 2 < n; j++)
        if (p[j] == '\n' && (p[j + 1] == '@' || p[j + 1] == '>') && p[j + 2] == '\n') return 1;
    return 0;
}
