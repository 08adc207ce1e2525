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

// 2-bit packing of whole groups of 32 bases (hostpipe.cpp pack_reads, src/CuCLARK_hh.hh:1660-1697): p[0, n) -> four u16
// containers per group, eight bases per container, first base in the top bits, A=3 C=2 G=1 T/U=0, either case.  Stops in
// front of the first group that holds anything but a base (groups of 32, then at most one of 16 and one of 8); returns the
// number of characters consumed (a multiple of 8) -- the caller's byte path deals with what follows.
extern "C" size_t ccbh_pack_bases_avx2(const uint8_t* p, size_t n, uint16_t* out)
{
    // by low nibble of the character: which lower-case letter it must be, and its code
    const __m256i want = _mm256_setr_epi8(0, 'a', 0, 'c', 't', 'u', 0, 'g', 0, 0, 0, 0, 0, 0, 0, 0,
                                          0, 'a', 0, 'c', 't', 'u', 0, 'g', 0, 0, 0, 0, 0, 0, 0, 0);
    const __m256i code = _mm256_setr_epi8(0, 3, 0, 2, 0, 0, 0, 1, 0, 0, 0, 0, 0, 0, 0, 0,
                                          0, 3, 0, 2, 0, 0, 0, 1, 0, 0, 0, 0, 0, 0, 0, 0);
    const __m256i low4 = _mm256_set1_epi8(0x0F), lower = _mm256_set1_epi8(0x20);
    const __m256i w41 = _mm256_set1_epi16(0x0104);                 // bytes (4, 1): first base of a pair in the high bits
    const __m256i w161 = _mm256_set1_epi32(0x00010010);            // words (16, 1)
    // low bytes of the four 32-bit lanes of each half -> two little-endian u16 (first four bases in the high byte)
    const __m256i gather = _mm256_setr_epi8(4, 0, 12, 8, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1,
                                            4, 0, 12, 8, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1);
    size_t i = 0;
    for (; i + 32 <= n; i += 32) {
        const __m256i v = _mm256_loadu_si256((const __m256i*)(p + i));
        const __m256i nib = _mm256_and_si256(v, low4);
        // a character with the top bit set keeps it through the OR and never equals a letter
        const __m256i ok = _mm256_cmpeq_epi8(_mm256_or_si256(v, lower), _mm256_shuffle_epi8(want, nib));
        if (_mm256_movemask_epi8(ok) != -1) break;
        const __m256i c = _mm256_shuffle_epi8(code, nib);
        const __m256i q = _mm256_madd_epi16(_mm256_maddubs_epi16(c, w41), w161);      // one byte per four bases, in 32-bit lanes
        const __m256i g = _mm256_shuffle_epi8(q, gather);
        uint32_t lo = (uint32_t)_mm256_cvtsi256_si32(g), hi = (uint32_t)_mm256_extract_epi32(g, 4);
        __builtin_memcpy(out, &lo, 4); __builtin_memcpy(out + 2, &hi, 4);
        out += 4;
    }
    // one more group of 16, then of 8 (a 150-base read ends with 16 + 6)
    for (int w = 16; w >= 8; w >>= 1) {
        if (i + (size_t)w > n) continue;
        __m128i v;
        if (w == 16) v = _mm_loadu_si128((const __m128i*)(p + i));
        else         v = _mm_loadl_epi64((const __m128i*)(p + i));
        const __m128i nib = _mm_and_si128(v, _mm256_castsi256_si128(low4));
        const __m128i ok = _mm_cmpeq_epi8(_mm_or_si128(v, _mm256_castsi256_si128(lower)), _mm_shuffle_epi8(_mm256_castsi256_si128(want), nib));
        if ((_mm_movemask_epi8(ok) & (w == 16 ? 0xFFFF : 0xFF)) != (w == 16 ? 0xFFFF : 0xFF)) break;
        const __m128i c = _mm_shuffle_epi8(_mm256_castsi256_si128(code), nib);
        const __m128i q = _mm_madd_epi16(_mm_maddubs_epi16(c, _mm256_castsi256_si128(w41)), _mm256_castsi256_si128(w161));
        const uint32_t g = (uint32_t)_mm_cvtsi128_si32(_mm_shuffle_epi8(q, _mm256_castsi256_si128(gather)));
        if (w == 16) { __builtin_memcpy(out, &g, 4); out += 2; }
        else         { const uint16_t h = (uint16_t)g; __builtin_memcpy(out, &h, 2); out += 1; }
        i += (size_t)w;
    }
    return i;
}
