// hostpipe.cpp -- see hostpipe.hpp
#include "hostpipe.hpp"
#include "../../include/cuclark_b200.h"

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <map>
#include <mutex>
#include <stdio.h>
#include <string.h>
#include <thread>
#include <chrono>
#include <memory>
#include <stdlib.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#ifndef _GNU_SOURCE
#define _GNU_SOURCE
#endif

namespace ccbh {

// ---------------------------------------------------------------------------
// A1: indexing
// ---------------------------------------------------------------------------
namespace {

inline bool is_sep(uint8_t c) { return c == ' ' || c == '\t' || c == '\n'; }   // src/CuCLARK_hh.hh:302

struct Rec { uint64_t ns, ne, ss, se, len; };

// The reference scans a name with a pre-increment (src/CuCLARK_hh.hh:1370, 1499): the first
// byte after the marker is never tested, and one byte past the end may be read (it sees the
// zero padding of the mapping, which is not a separator).
inline size_t name_end(const uint8_t* m, size_t nb, size_t i)
{
    while (i < nb) {
        ++i;
        if (i < nb && is_sep(m[i])) break;
    }
    return i;
}

inline size_t past_newline(const uint8_t* m, size_t nb, size_t i)
{   // while (i < nb && m[i++] != '\n') {}
    if (i >= nb) return i;
    const void* p = memchr(m + i, '\n', nb - i);
    return p ? (size_t)((const uint8_t*)p - m) + 1 : nb;
}

inline size_t to_newline(const uint8_t* m, size_t nb, size_t i)
{   // while (i < nb && m[i] != '\n') i++;
    if (i >= nb) return i;
    const void* p = memchr(m + i, '\n', nb - i);
    return p ? (size_t)((const uint8_t*)p - m) : nb;
}

// FASTA record whose '>' is at p.  Returns the position of the next record's '>' (or >= nb).
inline size_t fasta_record(const uint8_t* m, size_t nb, size_t p, Rec& r)
{
    size_t i = p + 1;
    r.ns = i;
    i = name_end(m, nb, i);
    r.ne = i;
    i = past_newline(m, nb, i);
    r.ss = r.se = i;
    uint64_t lines = 0;
    while (i < nb && m[i] != '>') {            // a '>' only ends the record at the start of a line
        i = to_newline(m, nb, i);
        lines++;
        r.se = i++;
    }
    r.len = (r.se - r.ss + 1) - lines;         // bytes minus newlines (:1388-1389); 1 for an empty record
    return i;
}

// FASTQ record starting at p (four lines, no validation, :1496-1524).  Returns the next record's start.
inline size_t fastq_record(const uint8_t* m, size_t nb, size_t p, Rec& r)
{
    size_t i = p + 1;
    r.ns = i;
    i = name_end(m, nb, i);
    r.ne = i;
    i = past_newline(m, nb, i);
    r.ss = i;
    i = to_newline(m, nb, i);
    r.se = i++;
    r.len = r.se - r.ss;
    i = past_newline(m, nb, i);
    i = past_newline(m, nb, i);
    return i;
}

inline void push(ReadSet& o, const Rec& r)
{
    o.name_s.push_back(r.ns); o.name_e.push_back(r.ne);
    o.seq_s.push_back(r.ss); o.seq_e.push_back(r.se); o.len.push_back(r.len);
}

void index_sequential(const uint8_t* m, size_t nb, bool fasta, ReadSet& o)
{
    Rec r;
    size_t p = 0;
    if (fasta) {
        for (;;) {
            p = fasta_record(m, nb, p, r);
            push(o, r);
            if (p >= nb) break;
        }
    } else {
        for (;;) {
            p = fastq_record(m, nb, p, r);
            push(o, r);
            if (p + 1 >= nb) break;                // (++i) >= nb, :1521
        }
    }
}

// ---- newline events, 256 bits at a time when the CPU can (hostscan_avx2.cpp), memchr otherwise ------------------
extern "C" size_t ccbh_count_newlines_avx2(const uint8_t* p, size_t n);
extern "C" size_t ccbh_newline_offsets_avx2(const uint8_t* p, size_t n, uint32_t* out);
extern "C" int ccbh_has_bare_marker_avx2(const uint8_t* p, size_t n);
extern "C" size_t ccbh_pack_bases_avx2(const uint8_t* p, size_t n, uint16_t* out);

const bool kAvx2 = __builtin_cpu_supports("avx2") && !getenv("CCB_NO_AVX2");

size_t count_newlines(const uint8_t* p, size_t n)
{
    if (kAvx2) return ccbh_count_newlines_avx2(p, n);
    size_t c = 0;
    const uint8_t* e = p + n;
    while (p < e) { const void* q = memchr(p, '\n', (size_t)(e - p)); if (!q) break; c++; p = (const uint8_t*)q + 1; }
    return c;
}

// positions of the newlines from `pos` on, one window of text at a time
struct NewlineStream {
    static constexpr size_t kWin = 8192;
    const uint8_t* m; size_t nb, pos;
    uint32_t off[kWin]; size_t n = 0, at = 0, base = 0;
    NewlineStream(const uint8_t* m_, size_t nb_, size_t from) : m(m_), nb(nb_), pos(from) {}
    inline bool next(size_t& q)
    {
        while (at == n) {
            if (pos >= nb) return false;
            const size_t len = std::min(kWin, nb - pos);
            base = pos; at = 0;
            if (kAvx2) n = ccbh_newline_offsets_avx2(m + pos, len, off);
            else {
                n = 0;
                const uint8_t* p = m + pos; const uint8_t* e = p + len;
                while (p < e) { const void* x = memchr(p, '\n', (size_t)(e - p)); if (!x) break; off[n++] = (uint32_t)((const uint8_t*)x - (m + pos)); p = (const uint8_t*)x + 1; }
            }
            pos += len;
        }
        q = base + off[at++];
        return true;
    }
};

// first separator after ns, looking from ns + 1 on (name_end), when a newline is known to lie at q >= ns + 1
inline size_t name_end_before(const uint8_t* m, size_t ns, size_t q)
{
    size_t i = ns + 1;
    while (i < q && m[i] != ' ' && m[i] != '\t') i++;
    return i;
}

// FASTQ records whose first byte lies in (b, e] at a line whose index is a multiple of four (`line` = index of the line
// that holds byte b); the record at byte 0 belongs to the range that starts there.  Same records, same fields as
// fastq_record() -- the newlines just come from one pass instead of four memchr calls per record.
// returns false for a record whose marker stands alone on its line (only the sequential scan reproduces it)
bool index_range_fastq(const uint8_t* m, size_t nb, size_t b, size_t e, uint64_t line, bool first, ReadSet& o)
{
    Rec r{};
    NewlineStream nl(m, nb, b);
    bool in_rec = false; int stage = 0; size_t rec_at = 0;
    auto start = [&](size_t p) -> bool {           // a record begins at p
        if (m[p + 1] == '\n') return false;
        in_rec = true; stage = 0; rec_at = p; r.ns = p + 1;
        return true;
    };
    if (first) { if (nb > 1 && !start(0)) return false; if (nb <= 1) { fastq_record(m, nb, 0, r); push(o, r); return true; } }
    size_t q;
    for (;;) {
        if (!nl.next(q)) {                          // end of the text inside a record: the plain parser knows every corner
            if (in_rec) { fastq_record(m, nb, rec_at, r); push(o, r); }
            return true;
        }
        line++;
        if (in_rec) {
            stage++;
            if (stage == 1) { r.ne = name_end_before(m, r.ns, q); r.ss = q + 1; }
            else if (stage == 2) { r.se = q; r.len = q - r.ss; }
            else if (stage == 4) { push(o, r); in_rec = false; }
        }
        if (!in_rec) {
            if (q >= e) return true;
            const size_t p = q + 1;
            if ((line & 3) == 0 && p + 1 < nb && !start(p)) return false;
        }
    }
}

// FASTA records whose '>' starts a line inside [b, e)
bool index_range_fasta(const uint8_t* m, size_t nb, size_t b, size_t e, ReadSet& o)
{
    Rec r{};
    bool in_rec = false, header = false; size_t rec_at = 0; uint64_t lines = 0;
    auto start = [&](size_t p) -> bool {
        if (p + 1 >= nb || m[p + 1] == '\n') return false;
        in_rec = true; header = true; rec_at = p; r.ns = p + 1; lines = 0;
        return true;
    };
    size_t from = b;
    if (b == 0 || m[b - 1] == '\n') { if (m[b] == '>' && !start(b)) return false; }
    NewlineStream nl(m, nb, from);
    size_t q;
    for (;;) {
        if (!nl.next(q)) {
            if (in_rec) { fasta_record(m, nb, rec_at, r); push(o, r); }
            return true;
        }
        const size_t c = q + 1;                     // the next line starts here
        if (in_rec) {
            if (header) { r.ne = name_end_before(m, r.ns, q); r.ss = r.se = c; header = false; }
            else { lines++; r.se = q; }
            if (c >= nb || m[c] == '>') {           // the record ends: at the end of the text, or where a line starts with '>'
                r.len = (r.se - r.ss + 1) - lines;
                push(o, r);
                in_rec = false;
            }
        }
        if (!in_rec) {
            if (c >= e || c >= nb) return true;
            if (m[c] == '>' && !start(c)) return false;
        }
    }
}

}  // namespace

int index_reads(const uint8_t* m, size_t nb, int threads, ReadSet& out)
{
    out.clear();
    out.map = m; out.nb = nb;
    if (nb == 0 || (m[0] != '>' && m[0] != '@')) return -1;
    out.fasta = m[0] == '>';
    const bool fasta = out.fasta;
    int T = threads < 1 ? 1 : threads;
    if (nb < (size_t)(1 << 20)) T = 1;
    if (T == 1) { index_sequential(m, nb, fasta, out); return 0; }

    // Parallel over byte ranges.  Record starts are found locally ('>' at the start of a line;
    // every fourth line for FASTQ), each record is then parsed by the same code as the
    // sequential path.  The only input for which local detection and the sequential scan can
    // disagree is a record whose marker is directly followed by a newline (the name scan then
    // runs into the next line, see name_end): such files are re-indexed sequentially.
    std::vector<size_t> cut((size_t)T + 1);
    for (int t = 0; t <= T; t++) cut[(size_t)t] = nb / (size_t)T * (size_t)t;
    cut[(size_t)T] = nb;
    std::vector<uint64_t> newlines((size_t)T, 0);
    std::vector<ReadSet> part((size_t)T);
    std::atomic<bool> odd(false);
    if (!fasta) {
        std::vector<std::thread> th;
        for (int t = 0; t < T; t++) th.emplace_back([&, t] {
            const uint64_t c = count_newlines(m + cut[(size_t)t], cut[(size_t)t + 1] - cut[(size_t)t]);
            newlines[(size_t)t] = c;
        });
        for (auto& x : th) x.join();
    }
    std::vector<uint64_t> line0((size_t)T, 0);      // index of the line that contains byte cut[t]
    for (int t = 1; t < T; t++) line0[(size_t)t] = line0[(size_t)t - 1] + newlines[(size_t)t - 1];
    {
        std::vector<std::thread> th;
        for (int t = 0; t < T; t++) th.emplace_back([&, t] {
            ReadSet& o = part[(size_t)t];
            const size_t b = cut[(size_t)t], e = cut[(size_t)t + 1];
            Rec r;
            {   // room for the records of this range up front (FASTQ: a quarter of its lines; FASTA: a guess), so the
                // five vectors are not copied over and over while they grow; untouched reservation costs nothing
                const size_t guess = (fasta ? (e - b) / 100 : (size_t)newlines[(size_t)t] / 4) + 16;
                o.name_s.reserve(guess); o.name_e.reserve(guess); o.seq_s.reserve(guess); o.seq_e.reserve(guess); o.len.reserve(guess);
            }
            if (fasta) { if (!index_range_fasta(m, nb, b, e, o)) odd = true; }
            else if (!index_range_fastq(m, nb, b, e, line0[(size_t)t], t == 0, o)) odd = true;
            (void)r;
        });
        for (auto& x : th) x.join();
    }
    if (odd) { out.clear(); index_sequential(m, nb, fasta, out); return 0; }
    // concatenate: sizes first, then every thread copies its own part (which also spreads the first touch)
    std::vector<size_t> at((size_t)T + 1, 0);
    for (int t = 0; t < T; t++) at[(size_t)t + 1] = at[(size_t)t] + part[(size_t)t].n();
    const size_t total = at[(size_t)T];
    out.name_s.resize(total); out.name_e.resize(total); out.seq_s.resize(total); out.seq_e.resize(total); out.len.resize(total);
    {
        std::vector<std::thread> th;
        for (int t = 0; t < T; t++) th.emplace_back([&, t] {
            const ReadSet& p = part[(size_t)t];
            const size_t o = at[(size_t)t], n = p.n();
            if (!n) return;
            memcpy(out.name_s.data() + o, p.name_s.data(), n * 8); memcpy(out.name_e.data() + o, p.name_e.data(), n * 8);
            memcpy(out.seq_s.data() + o, p.seq_s.data(), n * 8);   memcpy(out.seq_e.data() + o, p.seq_e.data(), n * 8);
            memcpy(out.len.data() + o, p.len.data(), n * 8);
            part[(size_t)t].clear();
        });
        for (auto& x : th) x.join();
    }
    return 0;
}

// ---------------------------------------------------------------------------
// A2: 2-bit packing
// ---------------------------------------------------------------------------
namespace {

constexpr uint32_t kMaxPart = 65528;     // multiple of 8 below 2^16: longer ACGT runs are split

struct CodeTable {
    uint8_t t[256];
    CodeTable()
    {   // 0..3: code (A=3 C=2 G=1 T/U=0, src/CuCLARK_hh.hh:294-297); 4: newline (skipped, :1673); 5: ends a part
        for (int i = 0; i < 256; i++) t[i] = 5;
        t[(int)'\n'] = 4;
        t[(int)'A'] = t[(int)'a'] = 3; t[(int)'C'] = t[(int)'c'] = 2; t[(int)'G'] = t[(int)'g'] = 1;
        t[(int)'T'] = t[(int)'t'] = t[(int)'U'] = t[(int)'u'] = 0;
    }
};
const CodeTable kCodes;

// two characters at once: (code << 2 | code) of a pair of bases, 0xFF when either is not a base
struct PairTable {
    uint8_t t[65536];
    PairTable()
    {
        for (int a = 0; a < 256; a++)
            for (int b = 0; b < 256; b++) {
                const uint8_t ca = kCodes.t[a], cb = kCodes.t[b];
                t[a | (b << 8)] = (ca < 4 && cb < 4) ? (uint8_t)((ca << 2) | cb) : (uint8_t)0xFF;     // little endian: first char in the low byte
            }
    }
};
const PairTable kPairs;

struct PartWriter {
    uint16_t* c; size_t cc, cap; int k;
    size_t hdr = 0; uint32_t plen = 0; uint32_t acc = 0, cnt = 0; bool open = false; uint64_t hist = 0;
    bool overflow = false;

    inline void put(uint32_t code)
    {
        acc = (acc << 2) | code; hist = (hist << 2) | code; plen++;
        if (++cnt == 8) { if (cc < cap) c[cc++] = (uint16_t)acc; else overflow = true; acc = 0; cnt = 0; }
    }
    inline void begin() { hdr = cc; if (cc < cap) cc++; else overflow = true; plen = 0; acc = 0; cnt = 0; open = true; }
    inline void end()
    {   // keep the part iff it holds at least one k-mer (:1646-1650, :1700-1703)
        if (!open) return;
        if (cnt) { if (cc < cap) c[cc++] = (uint16_t)(acc << (2 * (8 - cnt))); else overflow = true; }
        if (plen < (uint32_t)k) cc = hdr; else if (hdr < cap) c[hdr] = (uint16_t)plen;
        open = false; acc = 0; cnt = 0;
    }
    inline void base(uint32_t code)
    {
        if (!open) begin();
        else if (plen == kMaxPart) {               // split: next part restarts k-1 bases back
            const uint64_t h = hist;
            end(); begin();
            for (int j = k - 2; j >= 0; j--) put((uint32_t)(h >> (2 * j)) & 3u);
        }
        put(code);
    }
};

}  // namespace

size_t pack_bound(const ReadSet& r, size_t first, size_t count, int k)
{
    size_t tot = 0;
    for (size_t i = first; i < first + count; i++) {
        if (r.len[i] < (uint64_t)k) continue;
        size_t b = (size_t)(r.seq_e[i] - r.seq_s[i]);
        // bytes/8 containers, two more u16 per kept part (header + rounding); a part is at least k bases and one
        // separator long -- the same worst case as spec_containers()
        tot += b / 8 + 2 * (b / (size_t)(k + 1) + 1) + 4 + (b / kMaxPart) * 8;
    }
    return tot + 8;
}

namespace {
// one read's sequence bytes [p, e) through the writer (the body of pack_reads; also driven record by record by the
// streaming pipeline, while the bytes the indexer has just walked are still in the first-level cache)
inline void pack_one(PartWriter& w, const uint8_t* p, const uint8_t* e)
{
    while (p < e) {
        // fastest path (AVX2): the writer sits on a container boundary and whole groups of 32 / 16 / 8 bases follow
        if (kAvx2 && w.cnt == 0 && p + 32 <= e) {
            size_t room = (size_t)(e - p);
            const size_t left = kMaxPart - (w.open ? w.plen : 0u);                  // bases the current part may still take
            if (room > left) room = left;
            const size_t slots = w.cap - w.cc - (w.open ? 0u : 1u);                 // (a new part's header comes first)
            if (w.cc + 1 < w.cap && room / 8 > slots) room = slots * 8;
            if (room >= 32 && w.cc + 5 <= w.cap) {
                const size_t at = w.open ? w.cc : w.cc + 1;
                const size_t done = ccbh_pack_bases_avx2(p, room & ~(size_t)7, w.c + at);
                if (done) {
                    if (!w.open) w.begin();
                    w.cc += done / 8; w.plen += (uint32_t)done;
                    const uint16_t* l = w.c + w.cc - 4;                             // (done >= 32: at least four containers)
                    w.hist = ((uint64_t)l[0] << 48) | ((uint64_t)l[1] << 32) | ((uint64_t)l[2] << 16) | l[3];
                    p += done;
                    continue;
                }
            }
        }
        // fast path: the writer sits on a container boundary and the next eight characters are bases
        if (w.cnt == 0 && p + 8 <= e && w.cc < w.cap && (!w.open || w.plen + 8 <= kMaxPart)) {
            uint16_t q[4];
            memcpy(q, p, 8);
            const uint32_t v0 = kPairs.t[q[0]], v1 = kPairs.t[q[1]], v2 = kPairs.t[q[2]], v3 = kPairs.t[q[3]];
            if (!((v0 | v1 | v2 | v3) & 0x80u)) {
                if (!w.open) { w.begin(); if (w.overflow) break; }
                const uint32_t c16 = (v0 << 12) | (v1 << 8) | (v2 << 4) | v3;
                if (w.cc < w.cap) w.c[w.cc++] = (uint16_t)c16; else w.overflow = true;
                w.hist = (w.hist << 16) | c16; w.plen += 8;
                p += 8;
                continue;
            }
        }
        const uint32_t code = kCodes.t[*p++];
        if (code < 4) w.base(code);
        else if (code == 5) w.end();
    }
    w.end();
}
}  // namespace

int64_t pack_reads(const ReadSet& r, size_t first, size_t count, int k,
                   uint32_t* offsets, uint16_t* cont, size_t cap)
{
    PartWriter w{cont, 0, cap, k};
    const uint8_t* m = r.map;
    for (size_t j = 0; j < count; j++) {
        const size_t i = first + j;
        offsets[j] = (uint32_t)w.cc;
        if (r.len[i] < (uint64_t)k) continue;       // :1632, reads without a k-mer store nothing
        pack_one(w, m + r.seq_s[i], m + r.seq_e[i]);
        if (w.overflow) return -1;
        if (w.cc > 0xFFFFFFFFull) return -2;        // :1709
    }
    offsets[count] = (uint32_t)w.cc;
    return (int64_t)w.cc;
}

// ---------------------------------------------------------------------------
// A10: paired-end join (src/file.cc:205-268), in memory
// ---------------------------------------------------------------------------
namespace {

struct LineReader {
    FILE* f; char* buf = nullptr; size_t cap = 0; ssize_t n = 0;
    explicit LineReader(FILE* f_) : f(f_) {}
    ~LineReader() { free(buf); if (f) fclose(f); }
    bool next()
    {   // getLineFromFile: strips one trailing '\n' only
        n = getline(&buf, &cap, f);
        if (n < 0) { n = 0; return false; }
        if (n > 0 && buf[n - 1] == '\n') buf[--n] = 0;
        return true;
    }
};

// first token of a header line split at ' ', '/', '\t' and the '@' marker itself
inline void pair_id(const char* s, size_t n, const char*& b, size_t& len)
{
    size_t i = 0;
    auto sep = [](char c) { return c == ' ' || c == '/' || c == '\t' || c == '@'; };
    while (i < n && sep(s[i])) i++;
    b = s + i;
    size_t j = i;
    while (j < n && !sep(s[j])) j++;
    len = j - i;
}

}  // namespace

int join_pairs(const char* file1, const char* file2, TextBuf& out, std::string& err, int threads)
{
    // small files, and whatever cannot be mapped: line by line
    size_t min_bytes = (size_t)1 << 20;
    if (const char* e = getenv("CCB_JOIN_MIN_BYTES")) min_bytes = (size_t)atoll(e);                                         // (tests)
    struct Map {
        int fd = -1; void* p = MAP_FAILED; size_t n = 0;
        explicit Map(const char* f)
        {
            struct stat sb;
            fd = open(f, O_RDONLY);
            if (fd != -1 && fstat(fd, &sb) == 0 && S_ISREG(sb.st_mode) && sb.st_size > 0) { n = (size_t)sb.st_size; p = mmap(nullptr, n, PROT_READ, MAP_PRIVATE, fd, 0); }
        }
        ~Map() { if (p != MAP_FAILED) munmap(p, n); if (fd != -1) close(fd); }
        bool ok() const { return p != MAP_FAILED; }
    };
    if (!getenv("CCB_JOIN_SEQUENTIAL")) {
        Map a(file1), b(file2);
        if (a.ok() && b.ok() && a.n >= min_bytes && b.n >= min_bytes)
            return join_pairs_texts((const uint8_t*)a.p, a.n, (const uint8_t*)b.p, b.n, out, err, threads);
    }
    FILE* fa = fopen(file1, "r"); FILE* fb = fopen(file2, "r");
    if (!fa || !fb) { err = std::string("Failed to find/read ") + (!fa ? file1 : file2); if (fa) fclose(fa); if (fb) fclose(fb); return 1; }
    return join_pairs_streams(fa, fb, out, err);
}

int join_pairs_texts(const uint8_t* a, size_t na, const uint8_t* b, size_t nb, TextBuf& out, std::string& err, int threads)
{
    auto sequential = [&]() -> int {
        FILE* fa = fmemopen((void*)a, na, "r"); FILE* fb = fmemopen((void*)b, nb, "r");
        if (!fa || !fb) { err = "Failed to find/read the paired-end reads"; if (fa) fclose(fa); if (fb) fclose(fb); return 1; }
        return join_pairs_streams(fa, fb, out, err);
    };
    if (na == 0 || nb == 0 || getenv("CCB_JOIN_SEQUENTIAL")) return sequential();
    int T = threads > 0 ? threads : (int)std::thread::hardware_concurrency();
    if (T < 1) T = 1;
    if (T > 64) T = 64;
    // lines of both texts, counted slice by slice
    struct Side { const uint8_t* m; size_t n; std::vector<size_t> cut; std::vector<uint64_t> before; uint64_t lines = 0; };
    Side S[2] = {{a, na, {}, {}}, {b, nb, {}, {}}};
    for (Side& s : S) {
        s.cut.resize((size_t)T + 1); s.before.assign((size_t)T + 1, 0);
        for (int t = 0; t <= T; t++) s.cut[(size_t)t] = s.n / (size_t)T * (size_t)t;
        s.cut[(size_t)T] = s.n;
    }
    {
        std::vector<std::thread> th;
        for (int t = 0; t < T; t++) th.emplace_back([&, t] {
            for (Side& s : S) s.before[(size_t)t + 1] = count_newlines(s.m + s.cut[(size_t)t], s.cut[(size_t)t + 1] - s.cut[(size_t)t]);
        });
        for (auto& x : th) x.join();
    }
    for (Side& s : S) {
        for (int t = 0; t < T; t++) s.before[(size_t)t + 1] += s.before[(size_t)t];      // newlines in front of slice t + 1
        s.lines = s.before[(size_t)T] + (s.m[s.n - 1] != '\n' ? 1u : 0u);
    }
    if (S[0].lines != S[1].lines || S[0].lines == 0 || (S[0].lines & 3)) return sequential();
    const uint64_t R = S[0].lines / 4;
    const int W = (int)std::min<uint64_t>((uint64_t)T, R);
    // where line number `line` starts: behind the line-th newline, which lies in the slice with before[t] < line <= before[t+1]
    auto line_start = [&](const Side& s, uint64_t line) -> size_t {
        if (line == 0) return 0;
        size_t t = (size_t)(std::lower_bound(s.before.begin(), s.before.end(), line) - s.before.begin()) - 1;
        NewlineStream nl(s.m, s.cut[t + 1], s.cut[t]);
        size_t q = 0;
        for (uint64_t i = s.before[t]; i < line; i++) nl.next(q);
        return q + 1;
    };
    std::vector<TextBuf> piece((size_t)W);
    std::atomic<int> odd(0), mismatch(0);
    {
        std::vector<std::thread> th;
        for (int w = 0; w < W; w++) th.emplace_back([&, w] {
            const uint64_t r0 = R * (uint64_t)w / (uint64_t)W, r1 = R * (uint64_t)(w + 1) / (uint64_t)W;
            size_t pa = line_start(S[0], 4 * r0), pb = line_start(S[1], 4 * r0);
            NewlineStream na_(a, na, pa), nb_(b, nb, pb);
            TextBuf& o = piece[(size_t)w];
            o.reserve((size_t)((double)(na + nb) / (double)W * 0.56) + 4096);
            auto next = [](NewlineStream& s, size_t n, size_t& q) { if (!s.next(q)) q = n; };   // the last line may lack its newline
            for (uint64_t r = r0; r < r1 && !odd && !mismatch; r++) {
                size_t ha, sa, xa, ya, hb, sb, xb, yb;
                next(na_, na, ha); next(na_, na, sa); next(na_, na, xa); next(na_, na, ya);
                next(nb_, nb, hb); next(nb_, nb, sb); next(nb_, nb, xb); next(nb_, nb, yb);
                // header lines [pa, ha), [pb, hb); sequences [ha + 1, sa), [hb + 1, sb)
                if (!(ha > pa && hb > pb && a[pa] == '@' && b[pb] == '@')) { odd = 1; break; }
                const char *ia, *ib; size_t la, lb;
                pair_id((const char*)a + pa, ha - pa, ia, la); pair_id((const char*)b + pb, hb - pb, ib, lb);
                if (la != lb || memcmp(ia, ib, la) != 0) { mismatch = 1; break; }
                const size_t qa = std::min(ha + 1, sa), qb = std::min(hb + 1, sb);
                const size_t need = 1 + la + 1 + (sa - qa) + 1 + (sb - qb) + 1, at = o.size();
                o.resize(at + need);
                char* d = o.data() + at;
                *d++ = '>'; memcpy(d, ia, la); d += la; *d++ = '\n';
                memcpy(d, a + qa, sa - qa); d += sa - qa; *d++ = 'N';
                memcpy(d, b + qb, sb - qb); d += sb - qb; *d++ = '\n';
                pa = std::min(ya + 1, na); pb = std::min(yb + 1, nb);
            }
        });
        for (auto& x : th) x.join();
    }
    if (mismatch) { out.clear(); err = "Error: read id does not match between files!"; return 1; }
    if (odd) return sequential();                        // a fourth line that is no header: the line-by-line join knows what the reference does
    std::vector<size_t> at((size_t)W + 1, 0);
    for (int w = 0; w < W; w++) at[(size_t)w + 1] = at[(size_t)w] + piece[(size_t)w].size();
    out.clear();
    out.resize(at[(size_t)W]);
    {
        std::vector<std::thread> th;
        for (int w = 0; w < W; w++) th.emplace_back([&, w] {
            if (!piece[(size_t)w].empty()) memcpy(out.data() + at[(size_t)w], piece[(size_t)w].data(), piece[(size_t)w].size());
            TextBuf().swap(piece[(size_t)w]);
        });
        for (auto& x : th) x.join();
    }
    return 0;
}

int join_pairs_streams(FILE* fa, FILE* fb, TextBuf& out, std::string& err)
{
    LineReader a(fa), b(fb);
    if (!a.f || !b.f) { err = "Failed to find/read the paired-end reads"; return 1; }
    if (!a.next() || !b.next() || a.buf[0] != b.buf[0]) { err = "Error: the files have different format!"; return 1; }
    if (a.buf[0] != '@') { err = "Error: paired-end reads must be FASTQ files!"; return 1; }
    rewind(a.f); rewind(b.f);
    out.clear();
    while (a.next() && b.next()) {
        if (!(a.n && b.n && a.buf[0] == '@' && b.buf[0] == '@')) continue;
        const char *ia, *ib; size_t la, lb;
        pair_id(a.buf, (size_t)a.n, ia, la); pair_id(b.buf, (size_t)b.n, ib, lb);
        if (la != lb || memcmp(ia, ib, la) != 0) { err = "Error: read id does not match between files!"; return 1; }
        out.push_back('>'); out.insert(out.end(), ia, ia + la); out.push_back('\n');
        if (!(a.next() && b.next())) { err = "Error: Found read without sequence"; return 1; }
        out.insert(out.end(), a.buf, a.buf + a.n); out.push_back('N');
        out.insert(out.end(), b.buf, b.buf + b.n); out.push_back('\n');
        if (a.next() && b.next()) { if (a.next() && b.next()) continue; }     // '+' and quality lines
    }
    return 0;
}

// ---------------------------------------------------------------------------
// A7: CSV
// ---------------------------------------------------------------------------
void csv_header(std::string& out, bool extended, const std::vector<std::string>& names)
{
    out += "Object_ID";
    if (extended) for (size_t t = 1; t < names.size(); t++) { out += ','; out += names[t]; }
    out += ",Gamma,Assignment,Score,Confidence\n";
}

namespace {
// printf("%g") of a ratio of two small integers is by far the most expensive part of a row and takes few
// distinct values in a batch: a direct-mapped memo keyed by the integers keeps the text exactly printf's.
struct FmtMemo {
    struct E { uint64_t key; uint8_t len; char s[39]; };
    std::vector<E> e;
    FmtMemo() : e(4096) { for (auto& x : e) x.key = ~0ull; }
    template <typename F> void put(std::string& out, uint64_t key, F&& make)
    {
        E& x = e[(size_t)((key * 0x9E3779B97F4A7C15ull) >> 52)];
        if (x.key != key) {
            char buf[96];
            const int m = make(buf, sizeof buf);
            if (m < 0 || m >= (int)sizeof x.s) { out.append(buf, (size_t)std::max(m, 0)); return; }
            x.key = key; x.len = (uint8_t)m; memcpy(x.s, buf, (size_t)m);
        }
        out.append(x.s, x.len);
    }
    // the same, written at p (which has room for sizeof E::s bytes); returns the length, -1 when the text did not fit the memo
    template <typename F> int put_at(char* p, uint64_t key, F&& make)
    {
        E& x = e[(size_t)((key * 0x9E3779B97F4A7C15ull) >> 52)];
        if (x.key != key) {
            char buf[96];
            const int m = make(buf, sizeof buf);
            if (m < 0 || m >= (int)sizeof x.s) return -1;
            x.key = key; x.len = (uint8_t)m; memcpy(x.s, buf, (size_t)m);
        }
        memcpy(p, x.s, sizeof x.s);                      // fixed-size copy (the caller advances by the length)
        return x.len;
    }
};
}  // namespace

void csv_rows(std::string& out, const ReadSet& r, size_t first, size_t count, int k, bool paired, bool extended,
              const std::vector<std::string>& names, const uint16_t* final5, const uint16_t* rows, int row_stride)
{
    char num[96];
    FmtMemo memo_gamma, memo_tail;
    const size_t n_targets = names.size() - 1;
    if (!extended) {
        // the common case, row by row into a block of raw memory: name (<= 39) + gamma (<= 38) + target + tail (<= 38)
        size_t max_name = 0;
        for (const std::string& s : names) max_name = std::max(max_name, s.size());
        const size_t row_max = 39 + 39 + max_name + 39 + 39;      // (+ the memo's fixed-size copies overhang by < 39)
        constexpr size_t kBlock = 256;
        std::vector<char> buf(kBlock * row_max + 64);
        size_t j = 0;
        while (j < count) {
            const size_t stop = std::min(count, j + kBlock);
            char* p = buf.data();
            for (; j < stop; j++) {
                const size_t i = first + j;
                const uint16_t* f = final5 + 5 * j;
                const unsigned total = f[0], ib = f[1], best = f[2], second = f[4];
                size_t nl = (size_t)(r.name_e[i] - r.name_s[i]);
                if (nl >= 40) nl = 39;
                const char* nm = (const char*)r.map + r.name_s[i];
                const void* z = memchr(nm, 0, nl);
                if (z) nl = (size_t)((const char*)z - nm);
                memcpy(p, nm, nl); p += nl;
                const uint32_t norm = paired ? (uint32_t)r.len[i] - 1u : (uint32_t)r.len[i];
                auto gamma_text = [&](char* b, size_t cap) {
                    const double gamma = (double)total / (((double)norm - (double)k) + 1.0);
                    return snprintf(b, cap, ",%g,", gamma);
                };
                int m = memo_gamma.put_at(p, ((uint64_t)total << 32) | norm, gamma_text);
                if (m < 0) m = gamma_text(p, 39);
                p += m;
                const std::string& tn = names[ib < names.size() ? ib : 0];
                memcpy(p, tn.data(), tn.size()); p += tn.size();
                auto tail_text = [&](char* b, size_t cap) {
                    double delta = (double)(best + second);
                    delta = (delta < 0.001) ? 0 : ((double)best) / delta;
                    return snprintf(b, cap, ",%u,%g\n", best, delta);
                };
                m = memo_tail.put_at(p, ((uint64_t)best << 16) | second, tail_text);
                if (m < 0) m = tail_text(p, 39);
                p += m;
            }
            out.append(buf.data(), (size_t)(p - buf.data()));
        }
        return;
    }
    for (size_t j = 0; j < count; j++) {
        const size_t i = first + j;
        const uint16_t* f = final5 + 5 * j;
        const unsigned total = f[0], ib = f[1], best = f[2], second = f[4];
        // name: at most OBJECTNAMEMAX-1 = 39 bytes, ended early by a NUL (strncpy, :2102-2105)
        size_t nl = (size_t)(r.name_e[i] - r.name_s[i]);
        if (nl >= 40) nl = 39;
        const char* nm = (const char*)r.map + r.name_s[i];
        const void* z = memchr(nm, 0, nl);
        if (z) nl = (size_t)((const char*)z - nm);
        out.append(nm, nl);
        if (extended) {
            const uint16_t* row = rows + (size_t)row_stride * j;
            unsigned w = 0;
            for (unsigned h = 0; h < row[0]; h++) {
                const unsigned ti = row[1 + 2 * h], sc = row[2 + 2 * h];
                for (; w < ti; w++) out += ",0";
                int m = snprintf(num, sizeof num, ",%d", (int)sc);
                out.append(num, (size_t)m); w++;
            }
            for (; w < n_targets; w++) out += ",0";
        }
        const uint32_t norm = paired ? (uint32_t)r.len[i] - 1u : (uint32_t)r.len[i];     // ITYPE is u32; NBN = 1 (:2107)
        memo_gamma.put(out, ((uint64_t)total << 32) | norm, [&](char* b, size_t cap) {
            const double gamma = (double)total / (((double)norm - (double)k) + 1.0);     // :2110
            return snprintf(b, cap, ",%g,", gamma);
        });
        out += names[ib < names.size() ? ib : 0];
        memo_tail.put(out, ((uint64_t)best << 16) | second, [&](char* b, size_t cap) {
            double delta = (double)(best + second);
            delta = (delta < 0.001) ? 0 : ((double)best) / delta;                          // :2111-2112
            return snprintf(b, cap, ",%u,%g\n", best, delta);
        });
    }
}

// ---------------------------------------------------------------------------
// pipeline: workers pack -> submit -> wait -> format; text is appended in batch order
// ---------------------------------------------------------------------------
namespace {

struct DevSlots {
    ccb_handle* h;
    std::mutex mu; std::condition_variable cv;
    std::vector<int> free_slots;
    std::mutex submit_mu;                          // queryBatch is serialised per device (omp critical(cudaQuery))
};

struct OrderedSink {
    FILE* f = nullptr; FILE* dump = nullptr;
    std::mutex mu;
    size_t next = 0;
    std::map<size_t, std::pair<std::string, std::vector<uint16_t>>> ready;
    bool io_error = false;
    void put(size_t b, std::string&& text, std::vector<uint16_t>&& raw)
    {
        std::lock_guard<std::mutex> g(mu);
        ready.emplace(b, std::make_pair(std::move(text), std::move(raw)));
        for (auto it = ready.find(next); it != ready.end(); it = ready.find(next)) {
            const std::string& s = it->second.first;
            if (!s.empty() && fwrite(s.data(), 1, s.size(), f) != s.size()) io_error = true;
            if (dump && !it->second.second.empty()) fwrite(it->second.second.data(), 2, it->second.second.size(), dump);
            ready.erase(it);
            next++;
        }
    }
};

}  // namespace

namespace {
constexpr size_t kMaxBatchReads = 1u << 17, kMaxBatchBytes = 32u << 20;
// the most a batch can need: bytes/8 containers, two more per part, a part at least k bases and a separator long
inline size_t spec_containers(int k) { return kMaxBatchBytes / 8 + 2 * (kMaxBatchBytes / (size_t)(k + 1) + kMaxBatchReads) + 64; }
}  // namespace

int reserve_slots(const std::vector<ccb_handle*>& devs, int k, bool extended)
{
    int rc = 0;
    for (ccb_handle* h : devs) { const int r = ccb_batch_alloc(h, (uint32_t)kMaxBatchReads, spec_containers(k), extended); if (r && !rc) rc = r; }
    return rc;
}

// ---------------------------------------------------------------------------
// Streaming pipeline for large inputs: the text is cut into chunks of ~40 MB; a worker indexes ITS chunk, packs it,
// submits it, formats its rows -- no global index, no serial stage in front of the GPU work.  What a chunk needs from the
// rest of the file is one number (FASTQ: the index of the line it starts in), found by a newline count over all chunks
// first (256-bit compares, memory speed).  Anything unusual -- a marker alone on its line, a chunk without a record
// start (a read longer than a chunk), a record that does not fit a slot -- and the whole file goes through the two-phase
// path below instead; both produce the same bytes.
// ---------------------------------------------------------------------------
namespace {
// ---- FASTQ in chunks cut at arbitrary bytes (the streaming pipeline; index_fastq_chunked for the tests) -------------
constexpr size_t kNoStart = ~(size_t)0;

// The records of chunk (b, e]: those whose first byte p satisfies b < p <= e; the record at byte 0 belongs to the chunk
// that starts there (at_origin).  The chunk's first record is found locally -- the first line in (b, e] that starts with
// '@' two lines above a line that starts with '+' -- and every fourth line from there is taken as the reference's
// positional parser does.  first = where the first record starts (kNoStart: none found), next = where the record after
// the last one taken starts (nb: the text ended inside a record).  emit(rec) per record; false = a marker alone on its
// line, or emit said stop.
template <typename Emit>
bool fastq_chunk_walk(const uint8_t* m, size_t nb, size_t b, size_t e, bool at_origin, size_t& first, size_t& next, Emit&& emit)
{
    size_t p = kNoStart;
    if (at_origin) p = 0;
    else {
        NewlineStream nl(m, nb, b);
        size_t q0, q1, q2;
        if (nl.next(q0) && nl.next(q1) && nl.next(q2)) {
            for (;;) {
                const size_t st = q0 + 1;             // a line starts here
                if (st > e) break;
                if (m[st] == '@' && q2 + 1 < nb && m[q2 + 1] == '+') { p = st; break; }
                q0 = q1; q1 = q2;
                if (!nl.next(q2)) break;
            }
        }
    }
    first = p; next = kNoStart;
    if (p == kNoStart) return true;
    if (p + 1 >= nb) { next = p; return true; }
    NewlineStream nl(m, nb, p);
    for (;;) {
        if (m[p + 1] == '\n') return false;           // a marker alone on its line: sequential indexer only
        size_t q1, q2, q3, q4;
        Rec r;
        if (!nl.next(q1) || !nl.next(q2) || !nl.next(q3) || !nl.next(q4)) {
            fastq_record(m, nb, p, r);                // the text ends inside the record: the plain parser knows every corner
            next = nb;
            return emit(r);
        }
        r.ns = p + 1; r.ne = name_end_before(m, r.ns, q1); r.ss = q1 + 1; r.se = q2; r.len = q2 - r.ss;
        if (!emit(r)) return false;
        p = q4 + 1;
        if (q4 >= e || p + 1 >= nb) { next = p; return true; }
    }
}

// the pieces are the sequential parse iff every chunk starts where the one before stopped (induction from byte 0)
bool fastq_chunks_consistent(const std::vector<size_t>& first, const std::vector<size_t>& next, size_t nb, size_t chunk)
{
    if (first.empty()) return true;
    size_t expect = next[0];
    for (size_t c = 1; c < first.size(); c++) {
        const size_t e = std::min(nb, c * chunk + chunk);
        if (first[c] == kNoStart) { if (!(expect > e || expect + 1 >= nb)) return false; }
        else if (first[c] != expect) return false;
        else expect = next[c];
    }
    return true;
}
}  // namespace

// the same walk without a device: every chunk indexed on its own, the pieces checked, the records in order
int index_fastq_chunked(const uint8_t* m, size_t nb, size_t chunk, ReadSet& out)
{
    out.clear();
    out.map = m; out.nb = nb; out.fasta = false;
    if (nb < 2 || m[0] != '@' || chunk == 0) return -1;
    if (m[1] == '\n') return 1;
    const size_t C = (nb + chunk - 1) / chunk;
    std::vector<size_t> first(C, kNoStart), next(C, kNoStart);
    for (size_t c = 0; c < C; c++) {
        const size_t b = c * chunk, e = std::min(nb, b + chunk);
        if (!fastq_chunk_walk(m, nb, b, e, c == 0, first[c], next[c], [&](const Rec& r) { push(out, r); return true; })) return 1;
    }
    return fastq_chunks_consistent(first, next, nb, chunk) ? 0 : 1;
}

namespace {
constexpr int kNotStreamable = -12345;
constexpr size_t kStreamMinBytes = (size_t)128 << 20;

int classify_stream(const std::vector<ccb_handle*>& devs, const uint8_t* m, size_t nb, const char* result_csv,
                    const std::vector<std::string>& names, const ClassifyOptions& opt, ClassifyStats& stats, std::string& err)
{
    auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t0 = now();
    const bool fasta = m[0] == '>';
    size_t chunk = fasta ? ((size_t)28 << 20) : ((size_t)40 << 20);            // <= kMaxBatchBytes of sequence, ~130 k short reads
    if (const char* e = getenv("CCB_STREAM_CHUNK")) { const long long v = atoll(e); if (v >= 1024) chunk = (size_t)v; }    // (tests)
    const size_t C = (nb + chunk - 1) / chunk;
    const int T = std::max(1, opt.threads);
    // slots first (pinning 17 MB x slots takes 0.1 s): on a thread of their own while the chunks are surveyed
    std::vector<int> alloc_rc(devs.size(), 0);
    std::thread alloc([&] {
        for (size_t d = 0; d < devs.size(); d++) alloc_rc[d] = ccb_batch_alloc(devs[d], (uint32_t)kMaxBatchReads, spec_containers(opt.k), opt.extended);
    });
    // FASTQ: no survey.  Every worker finds the first record of its chunk by itself (a line that starts with '@' two
    // lines above a line that starts with '+': in a well-formed file only a header is that) and takes every fourth line
    // from there, as the reference does; where a chunk's records end and where the next chunk's begin is compared once
    // all are done -- equal everywhere means the pieces are exactly the sequential parse (induction from byte 0), anything
    // else (quality lines that look like headers in front of sequences that start with '+', missing lines) gives the
    // file to the two-phase path.  The text is then walked ONCE: index, 2-bit packing and the copy of the names happen
    // record by record while the bytes are in the first-level cache.
    // FASTA: survey (markers alone on a line, chunks without a record start), then index and pack per chunk.
    std::vector<size_t> first_start(C, kNoStart), next_start(C, kNoStart);
    std::atomic<bool> unusual(false);
    if (fasta) {
        std::atomic<size_t> next(0);
        auto survey = [&] {
            for (size_t c = next++; c < C && !unusual; c = next++) {
                const size_t b = c * chunk, e = std::min(nb, b + chunk);
                const size_t lo = b ? b - 1 : 0, hi = std::min(nb, e + 1);          // patterns that straddle a chunk boundary
                if (kAvx2 ? ccbh_has_bare_marker_avx2(m + lo, hi - lo) : 1) {
                    if (kAvx2) { unusual = true; return; }
                    for (size_t j = lo; j + 2 < hi; j++)
                        if (m[j] == '\n' && (m[j + 1] == '@' || m[j + 1] == '>') && m[j + 2] == '\n') { unusual = true; return; }
                }
                if (c && e - b >= 2 && !memmem(m + b - 1, e - b + 1, "\n>", 2)) { unusual = true; return; }
            }
        };
        std::vector<std::thread> th;
        for (int t = 0; t < T; t++) th.emplace_back(survey);
        for (auto& x : th) x.join();
    }
    if (nb > 1 && m[1] == '\n') unusual = true;
    alloc.join();
    for (size_t d = 0; d < devs.size(); d++)
        if (alloc_rc[d]) { err = ccb_last_error(devs[d]); return alloc_rc[d]; }
    if (unusual) return kNotStreamable;
    stats.seconds_index = now() - t0;
    stats.seconds_alloc = 0;
    const size_t max_cont = spec_containers(opt.k);

    std::vector<std::unique_ptr<DevSlots>> ds;
    for (ccb_handle* h : devs) {
        ds.emplace_back(new DevSlots());
        ds.back()->h = h;
        for (int s = 0; s < ccb_batch_slots(h); s++) ds.back()->free_slots.push_back(s);
    }
    OrderedSink sink;
    sink.f = fopen(result_csv, "w");
    if (!sink.f) { err = std::string("Failed to create/open file result: ") + result_csv; return 1; }
    if (opt.dump_final5) sink.dump = fopen(opt.dump_final5, "wb");
    {
        std::string hdr; csv_header(hdr, opt.extended, names);
        fwrite(hdr.data(), 1, hdr.size(), sink.f);
    }
    std::atomic<size_t> next_chunk(0);
    std::atomic<uint64_t> n_reads(0), n_batches(0);
    std::atomic<int> failed(0);
    std::atomic<bool> give_up(false);
    std::mutex err_mu, stat_mu;
    auto fail = [&](int rc, const std::string& msg) { std::lock_guard<std::mutex> g(err_mu); if (!failed) { failed = rc ? rc : 1; err = msg; } };
    auto work = [&](int wid) {
        double c_idx = 0, c_pack = 0, c_wait = 0, c_fmt = 0, c_sink = 0;
        struct Fold { std::mutex& mu; ClassifyStats& st; double &i, &a, &b, &c, &d;
                      ~Fold() { std::lock_guard<std::mutex> g(mu); st.cpu_index += i; st.cpu_pack += a; st.cpu_wait += b; st.cpu_format += c; st.cpu_sink += d; } }
            fold{stat_mu, stats, c_idx, c_pack, c_wait, c_fmt, c_sink};
        ReadSet rs;
        rs.map = m; rs.nb = nb; rs.fasta = fasta;
        std::vector<char> name_buf;
        const std::vector<std::string>& names_of_targets = names;
        DevSlots& d = *ds[(size_t)wid % ds.size()];
        for (;;) {
            const size_t c = next_chunk.fetch_add(1);
            if (c >= C || failed || give_up) return;
            double t1 = now();
            const size_t b = c * chunk, e = std::min(nb, b + chunk);
            rs.clear();
            if (!fasta) {
                // ---- FASTQ: one walk over the chunk; a batch is packed while it is indexed ----
                std::string text; std::vector<uint16_t> raw;
                int slot = -1; uint32_t* off = nullptr; uint16_t *cont = nullptr, *fin = nullptr, *rows = nullptr;
                PartWriter w{nullptr, 0, 0, opt.k};
                size_t cnt = 0, bytes = 0, n_chunk = 0;
                std::vector<char>& names = name_buf;           // (shadows the target names on purpose: csv_rows gets names_of_targets)
                names.clear();
                bool bad = false;
                auto release = [&] { { std::lock_guard<std::mutex> g(d.mu); d.free_slots.push_back(slot); } d.cv.notify_one(); slot = -1; };
                auto flush = [&]() -> bool {                     // submit what the slot holds, format its rows
                    if (slot < 0) return true;
                    if (cnt == 0) { release(); return true; }
                    c_idx += now() - t1; t1 = now();
                    if (w.overflow || w.cc > 0xFFFFFFFFull) { release(); give_up = true; return false; }
                    off[cnt] = (uint32_t)w.cc;
                    int rc;
                    {
                        std::lock_guard<std::mutex> g(d.submit_mu);
                        rc = ccb_batch_submit(d.h, slot, (uint32_t)cnt, (uint64_t)w.cc, opt.extended ? CCB_EXTENDED : 0);
                    }
                    if (!rc) rc = ccb_batch_wait(d.h, slot);
                    if (rc) { fail(rc, ccb_last_error(d.h)); release(); return false; }
                    c_wait += now() - t1; t1 = now();
                    rs.map = (const uint8_t*)names.data();
                    csv_rows(text, rs, 0, cnt, opt.k, opt.paired, opt.extended, names_of_targets, fin, rows, opt.row_stride);
                    if (sink.dump) raw.insert(raw.end(), fin, fin + cnt * 5);
                    release();
                    n_batches++; n_chunk += cnt;
                    rs.clear(); names.clear(); cnt = 0; bytes = 0;
                    c_fmt += now() - t1; t1 = now();
                    return true;
                };
                auto emit = [&](const Rec& r) -> bool {          // one record: its name, its length, its bases into the slot
                    const size_t sb = (size_t)(r.se - r.ss);
                    if (sb > kMaxBatchBytes) { give_up = true; return false; }          // one read larger than a slot: two-phase path
                    if (slot >= 0 && (cnt == kMaxBatchReads || bytes + sb > kMaxBatchBytes) && !flush()) return false;
                    if (slot < 0) {
                        c_idx += now() - t1; t1 = now();
                        {
                            std::unique_lock<std::mutex> g(d.mu);
                            d.cv.wait(g, [&] { return !d.free_slots.empty(); });
                            slot = d.free_slots.back(); d.free_slots.pop_back();
                        }
                        ccb_batch_buffers(d.h, slot, &off, &cont, &fin, &rows);
                        w = PartWriter{cont, 0, max_cont, opt.k};
                        c_wait += now() - t1; t1 = now();
                        if (text.capacity() == 0) text.reserve((size_t)(e - b) / 6);
                    }
                    size_t nl = (size_t)(r.ne - r.ns);
                    if (nl > 40) nl = 40;                         // csv_rows prints at most 39 bytes of it
                    const size_t at = names.size();
                    names.insert(names.end(), (const char*)m + r.ns, (const char*)m + r.ns + nl);
                    rs.name_s.push_back(at); rs.name_e.push_back(at + nl); rs.len.push_back(r.len);
                    off[cnt] = (uint32_t)w.cc;
                    if (r.len >= (uint64_t)opt.k) pack_one(w, m + r.ss, m + r.se);
                    cnt++; bytes += sb;
                    return true;
                };
                if (!fastq_chunk_walk(m, nb, b, e, c == 0, first_start[c], next_start[c], emit)) bad = true;
                if (bad) { if (slot >= 0) release(); if (!failed) give_up = true; sink.put(c, "", {}); return; }
                if (!flush()) { sink.put(c, "", {}); return; }
                c_idx += now() - t1; t1 = now();
                n_reads += n_chunk;
                sink.put(c, std::move(text), std::move(raw));
                c_sink += now() - t1;
                continue;
            }
            const bool ok = index_range_fasta(m, nb, b, e, rs);
            c_idx += now() - t1;
            if (!ok) { give_up = true; sink.put(c, "", {}); return; }
            const size_t n = rs.n();
            std::string text; std::vector<uint16_t> raw;
            text.reserve(n * 48);
            size_t first = 0;
            while (first < n && !failed && !give_up) {
                size_t cnt = 0, bytes = 0;                      // a sub-batch within the slot's limits
                while (first + cnt < n && cnt < kMaxBatchReads) {
                    const size_t sb = (size_t)(rs.seq_e[first + cnt] - rs.seq_s[first + cnt]);
                    if (cnt && bytes + sb > kMaxBatchBytes) break;
                    bytes += sb; cnt++;
                }
                if (bytes > kMaxBatchBytes) { give_up = true; break; }          // one read larger than a slot: two-phase path
                int slot;
                {
                    std::unique_lock<std::mutex> g(d.mu);
                    d.cv.wait(g, [&] { return !d.free_slots.empty(); });
                    slot = d.free_slots.back(); d.free_slots.pop_back();
                }
                auto release = [&] { { std::lock_guard<std::mutex> g(d.mu); d.free_slots.push_back(slot); } d.cv.notify_one(); };
                uint32_t* off; uint16_t *cont, *fin, *rows;
                ccb_batch_buffers(d.h, slot, &off, &cont, &fin, &rows);
                t1 = now();
                const int64_t nc = pack_reads(rs, first, cnt, opt.k, off, cont, max_cont);
                c_pack += now() - t1; t1 = now();
                if (nc < 0) { release(); give_up = true; break; }
                int rc;
                {
                    std::lock_guard<std::mutex> g(d.submit_mu);
                    rc = ccb_batch_submit(d.h, slot, (uint32_t)cnt, (uint64_t)nc, opt.extended ? CCB_EXTENDED : 0);
                }
                if (!rc) rc = ccb_batch_wait(d.h, slot);
                if (rc) { fail(rc, ccb_last_error(d.h)); release(); break; }
                c_wait += now() - t1; t1 = now();
                csv_rows(text, rs, first, cnt, opt.k, opt.paired, opt.extended, names, fin, rows, opt.row_stride);
                if (sink.dump) raw.insert(raw.end(), fin, fin + cnt * 5);
                release();
                c_fmt += now() - t1;
                first += cnt; n_batches++;
            }
            n_reads += n;
            t1 = now();
            sink.put(c, std::move(text), std::move(raw));
            c_sink += now() - t1;
        }
    };
    const int workers = (int)std::max<size_t>(1, std::min<size_t>((size_t)T, C));
    std::vector<std::thread> th;
    for (int w = 0; w < workers; w++) th.emplace_back(work, w);
    for (auto& t : th) t.join();
    const bool io_err = sink.io_error || fclose(sink.f) != 0;
    if (sink.dump) fclose(sink.dump);
    if (failed) return failed;
    if (!fasta && !give_up && !fastq_chunks_consistent(first_start, next_start, nb, chunk)) give_up = true;
    if (give_up) return kNotStreamable;                  // nothing of the partial file is kept: the caller starts over
    stats.reads = n_reads; stats.batches = n_batches;
    stats.seconds_total = now() - t0;
    if (io_err) { err = std::string("Failed to write ") + result_csv; return 1; }
    return 0;
}
}  // namespace

int classify_text(const std::vector<ccb_handle*>& devs, const uint8_t* map, size_t nb, const char* result_csv,
                  const std::vector<std::string>& names, const ClassifyOptions& opt, ClassifyStats& stats, std::string& err)
{
    size_t stream_min = kStreamMinBytes;
    if (const char* e = getenv("CCB_STREAM_MIN_BYTES")) stream_min = (size_t)atoll(e);                                        // (tests)
    if (nb >= stream_min && nb > 2 && map && (map[0] == '>' || map[0] == '@') && !getenv("CCB_NO_STREAM")) {
        const int rc = classify_stream(devs, map, nb, result_csv, names, opt, stats, err);
        if (rc != kNotStreamable) return rc;
        stats = ClassifyStats();
    }
    auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t0 = now();
    // For a large input the pinned slots are allocated while the file is being indexed (0.1 s of cudaHostAlloc
    // next to 0.3 s of indexing per 10 M reads).  Their size is the most a batch can need -- bytes/8 containers,
    // two more per part, a part at least k bases and a separator long -- unless a single read is larger than a
    // batch; the exact figures are checked below and the slots re-made if they do not fit.
    const bool early = nb >= ((size_t)256 << 20);
    const size_t spec_reads = kMaxBatchReads, spec_cont = spec_containers(opt.k);
    std::vector<int> early_rc(devs.size(), 0);
    std::thread early_alloc;
    if (early) early_alloc = std::thread([&] {
        for (size_t d = 0; d < devs.size(); d++) early_rc[d] = ccb_batch_alloc(devs[d], (uint32_t)spec_reads, spec_cont, opt.extended);
    });
    ReadSet rs;
    const int irc = index_reads(map, nb, opt.threads, rs);
    if (irc != 0) { if (early) early_alloc.join(); err = "Failed to recognize the format of the file."; return 255; }
    stats.seconds_index = now() - t0;
    const size_t n = rs.n();
    stats.reads = n;

    // batches: bounded by sequence bytes and read count so that pinned slots stay modest and
    // there are enough of them in flight to hide packing behind the GPU and vice versa
    // (and small enough that every worker gets several: 10 M x 150 bp reads -> 77 batches)
    std::vector<size_t> bstart{0};
    bool oversize = false;                     // some read alone exceeds the batch byte limit
    {
        size_t bytes = 0, cnt = 0;
        for (size_t i = 0; i < n; i++) {
            size_t b = (size_t)(rs.seq_e[i] - rs.seq_s[i]);
            oversize = oversize || b > kMaxBatchBytes;
            if (cnt && (cnt >= kMaxBatchReads || bytes + b > kMaxBatchBytes)) { bstart.push_back(i); bytes = 0; cnt = 0; }
            bytes += b; cnt++;
        }
        bstart.push_back(n);
    }
    const size_t nbatches = bstart.size() - 1;
    stats.batches = nbatches;
    size_t max_reads = 1, max_cont = 1;
    // with slots sized for the largest batch the rule allows, the per-batch bound (one more pass over the index) is
    // only needed when a single read is larger than that
    const bool need_bound = !early || oversize;
    for (size_t b = 0; b < nbatches; b++) {
        max_reads = std::max(max_reads, bstart[b + 1] - bstart[b]);
        if (need_bound) max_cont = std::max(max_cont, pack_bound(rs, bstart[b], bstart[b + 1] - bstart[b], opt.k));
    }

    const int workers = (int)std::max<size_t>(1, std::min<size_t>((size_t)std::max(opt.threads, 1), nbatches));
    const double t_alloc0 = now();
    std::vector<std::unique_ptr<DevSlots>> ds;
    bool reuse = false;
    if (early) {
        early_alloc.join();
        reuse = max_reads <= spec_reads && max_cont <= spec_cont;
        for (int rc : early_rc) reuse = reuse && rc == 0;
        if (reuse) { max_reads = spec_reads; max_cont = spec_cont; }
    }
    for (ccb_handle* h : devs) {
        if (!reuse)
            if (int rc = ccb_batch_alloc(h, (uint32_t)max_reads, max_cont, opt.extended)) { err = ccb_last_error(h); return rc; }
        ds.emplace_back(new DevSlots());
        ds.back()->h = h;
        for (int s = 0; s < ccb_batch_slots(h); s++) ds.back()->free_slots.push_back(s);
    }

    stats.seconds_alloc = now() - t_alloc0;
    std::mutex stat_mu;

    OrderedSink sink;
    sink.f = fopen(result_csv, "w");
    if (!sink.f) { err = std::string("Failed to create/open file result: ") + result_csv; return 1; }
    if (opt.dump_final5) sink.dump = fopen(opt.dump_final5, "wb");
    {
        std::string hdr; csv_header(hdr, opt.extended, names);
        fwrite(hdr.data(), 1, hdr.size(), sink.f);
    }

    std::atomic<size_t> next_batch(0);
    std::atomic<int> failed(0);
    std::mutex err_mu;
    auto fail = [&](int rc, const std::string& m) { std::lock_guard<std::mutex> g(err_mu); if (!failed) { failed = rc ? rc : 1; err = m; } };

    auto work = [&](int wid) {
        double c_pack = 0, c_wait = 0, c_fmt = 0, c_sink = 0;
        struct Fold { std::mutex& m; ClassifyStats& st; double &a, &b, &c, &d;
                      ~Fold() { std::lock_guard<std::mutex> g(m); st.cpu_pack += a; st.cpu_wait += b; st.cpu_format += c; st.cpu_sink += d; } }
            fold{stat_mu, stats, c_pack, c_wait, c_fmt, c_sink};
        for (;;) {
            // take a slot first, then a batch number: the oldest unwritten batch always owns a slot
            DevSlots& d = *ds[(size_t)wid % ds.size()];
            int slot;
            {
                std::unique_lock<std::mutex> g(d.mu);
                d.cv.wait(g, [&] { return !d.free_slots.empty(); });
                slot = d.free_slots.back(); d.free_slots.pop_back();
            }
            const size_t b = next_batch.fetch_add(1);
            auto release = [&] { { std::lock_guard<std::mutex> g(d.mu); d.free_slots.push_back(slot); } d.cv.notify_one(); };
            if (b >= nbatches || failed) { release(); return; }
            const size_t first = bstart[b], cnt = bstart[b + 1] - first;
            uint32_t* off; uint16_t *cont, *fin, *rows;
            ccb_batch_buffers(d.h, slot, &off, &cont, &fin, &rows);
            double t1 = now();
            const int64_t nc = pack_reads(rs, first, cnt, opt.k, off, cont, max_cont);
            c_pack += now() - t1; t1 = now();
            if (nc == -2) { fail(255, "ERROR: Batch overflow. Please increase the number of batches (-b <numberofbatches>)."); release(); sink.put(b, "", {}); continue; }
            if (nc < 0) { fail(255, "Bad container estimation. Abort."); release(); sink.put(b, "", {}); continue; }
            int rc;
            {
                std::lock_guard<std::mutex> g(d.submit_mu);
                rc = ccb_batch_submit(d.h, slot, (uint32_t)cnt, (uint64_t)nc, opt.extended ? CCB_EXTENDED : 0);
            }
            if (!rc) rc = ccb_batch_wait(d.h, slot);
            if (rc) { fail(rc, ccb_last_error(d.h)); release(); sink.put(b, "", {}); continue; }
            c_wait += now() - t1; t1 = now();
            std::string text;
            text.reserve(cnt * 48);
            csv_rows(text, rs, first, cnt, opt.k, opt.paired, opt.extended, names, fin, rows, opt.row_stride);
            std::vector<uint16_t> raw;
            if (sink.dump) raw.assign(fin, fin + cnt * 5);
            release();
            c_fmt += now() - t1; t1 = now();
            sink.put(b, std::move(text), std::move(raw));
            c_sink += now() - t1;
        }
    };
    std::vector<std::thread> th;
    for (int w = 0; w < workers; w++) th.emplace_back(work, w);
    for (auto& t : th) t.join();
    const bool io_err = sink.io_error || fclose(sink.f) != 0;
    if (sink.dump) fclose(sink.dump);
    // the slots stay with the handles: the next file of a list reuses them (ccb_batch_alloc), ccb_destroy releases them
    stats.seconds_total = now() - t0;
    if (failed) return failed;
    if (io_err) { err = std::string("Failed to write ") + result_csv; return 1; }
    return 0;
}

}  // namespace ccbh
