// kmer.cuh -- k-mer arithmetic shared by host and device code.
//
// Encoding follows the reference so that canonical values are identical to the
// ones stored in its database files:
//   * x-encoding A=3 C=2 G=1 T/U=0, first base most significant
//     (reference src/CuCLARK_hh.hh:294-297, src/CuClarkDB.cu:1063-1083)
//   * reverse complement R(x) and canonical c = min(x, R(x))
//     (reference src/CuClarkDB.cu:1195-1206)
// Everything after the canonical value (hash, bucket, stored key) is this
// implementation's own table layout, see DESIGN.md "Table".
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define CCB_HD __host__ __device__ __forceinline__
#else
#define CCB_HD inline
#endif

namespace ccb {

// Labels 0..0xFFFD are targets.
constexpr uint64_t kEmpty      = ~0ull;     // never a canonical k-mer (poly-A is not canonical; k<32 values are < 2^2k)
constexpr uint32_t kLabelTomb  = 0xFFFEu;   // k-mer seen under >1 label (not discriminative)
constexpr uint32_t kLabelNone  = 0xFFFFu;
constexpr int      kSectorKeys = 3;         // 3 x (8 B key + 2 B label) + 2 B meta = one 32-byte sector
constexpr int      kMaxHops    = 24;        // lines a probe sequence may visit (4 sectors each)
constexpr int      kWin        = 12;        // minimizer window (m-mers per k-mer) when k >= kWin + 11

CCB_HD int code_of(unsigned char c)
{   // >=0: x-encoding code; -1: anything else
    switch (c) {
    case 'A': case 'a': return 3;
    case 'C': case 'c': return 2;
    case 'G': case 'g': return 1;
    case 'T': case 't': case 'U': case 'u': return 0;
    default: return -1;
    }
}

CCB_HD uint64_t kmask(int k) { return k >= 32 ? ~0ull : ((1ull << (2 * k)) - 1); }

CCB_HD uint64_t revcomp(uint64_t x, int k)
{
    uint64_t r = x;
    r = ((r >> 2)  & 0x3333333333333333ull) | ((r & 0x3333333333333333ull) << 2);
    r = ((r >> 4)  & 0x0F0F0F0F0F0F0F0Full) | ((r & 0x0F0F0F0F0F0F0F0Full) << 4);
    r = ((r >> 8)  & 0x00FF00FF00FF00FFull) | ((r & 0x00FF00FF00FF00FFull) << 8);
    r = ((r >> 16) & 0x0000FFFF0000FFFFull) | ((r & 0x0000FFFF0000FFFFull) << 16);
    r = (r >> 32) | (r << 32);
    return (~r) >> (64 - 2 * k);
}

CCB_HD uint64_t canonical(uint64_t x, int k)
{
    uint64_t r = revcomp(x, k);
    return x < r ? x : r;
}

// ---- minimizer addressing (DESIGN.md "Table v2") ---------------------------------
// m-mer length: k >= 23 -> m = k - 11 (window of 12 m-mers); shorter k-mers are their own minimizer.
CCB_HD int mmer_len(int k) { return k >= kWin + 11 ? k - (kWin - 1) : k; }

// order/identity hash of a canonical m-mer: odd multiply mod 2^64 (injective), compared as u64
CCB_HD uint64_t phi64(uint64_t cm) { return cm * 0x9E3779B97F4A7C15ull; }

// min over the m-mers of a k-mer of phi64(canonical m-mer); identical for a k-mer and its reverse complement
CCB_HD uint64_t minimizer_phi(uint64_t x, int k, int m)
{
    const uint64_t mm = kmask(m);
    uint64_t best = ~0ull;
    for (int j = 0; j + m <= k; j++) {
        uint64_t f = (x >> (2 * (k - m - j))) & mm;
        uint64_t r = revcomp(f, m);
        uint64_t p = phi64(f < r ? f : r);
        best = p < best ? p : best;
    }
    return best;
}

CCB_HD uint32_t mulhi32(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a * b) >> 32); }

// 128-byte line of the table that is home to every k-mer whose minimizer hashes to `phi`
CCB_HD uint32_t line_of(uint64_t phi, uint32_t n_lines)
{
    uint64_t z = (phi ^ (phi >> 29)) * 0xBF58476D1CE4E5B9ull;     // the minimum is concentrated near 0: remix
    return mulhi32((uint32_t)(z >> 32), n_lines);
}

// h-th line (h >= 1) of the probe sequence of the k-mers whose home line is `home`
CCB_HD uint32_t hop_line(uint32_t home, int h, uint32_t n_lines)
{
    uint64_t z = (((uint64_t)home << 8) | (uint32_t)h) * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 32)) * 0xBF58476D1CE4E5B9ull;
    return mulhi32((uint32_t)(z >> 32), n_lines);
}

// sector (0..3) of the home line where the probe sequence of canonical k-mer c starts
CCB_HD uint32_t sector_of(uint64_t c)
{
    return ((uint32_t)c * 0x9E3779B1u + (uint32_t)(c >> 32) * 0x85EBCA77u) >> 30;
}

struct HashParams {
    uint32_t n_lines;  // home lines (4 sectors each)
    int      k, m;     // k-mer and minimizer length
    uint64_t kmsk;     // 2^(2k)-1
};

}  // namespace ccb
