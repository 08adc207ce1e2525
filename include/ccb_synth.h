/*
 * ccb_synth.h -- definition of the synthetic benchmark workload.
 *
 * Counter-based, so any base of any genome and any read can be produced
 * independently on the device (bench.py) and on the host (oracle checks):
 *
 *   genome g, position p : code = (mix(seed_g + (p>>5)) >> 2*(p&31)) & 3
 *                          in the reference's x-encoding A=3 C=2 G=1 T=0
 *   read i               : 90.0 % a window of a uniform genome / position / strand
 *                                 with 1 % iid substitutions,
 *                           9.9 % uniform random sequence (misses),
 *                           0.1 % genome window with one N at a uniform position
 *
 * (SURVEY D2 names xoshiro256**; a sequential generator cannot be evaluated per
 * element on 148 SMs, so the same distribution is drawn from splitmix64 used as
 * a counter-based hash.)
 */
#ifndef CCB_SYNTH_H
#define CCB_SYNTH_H
#include <stdint.h>

#if defined(__CUDACC__)
#define CCB_SYNTH_FN static __host__ __device__ __forceinline__
#else
#define CCB_SYNTH_FN static inline
#endif

CCB_SYNTH_FN uint64_t ccb_sm64(uint64_t z)
{
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

CCB_SYNTH_FN uint64_t ccb_genome_seed(uint64_t seed, uint32_t g)
{
    return ccb_sm64(seed ^ ((uint64_t)(g + 1) * 0xD1342543DE82EF95ull));
}

/* x-encoding code (0..3) of base p of the genome with seed gs */
CCB_SYNTH_FN uint32_t ccb_genome_code(uint64_t gs, uint64_t p)
{
    uint64_t w = ccb_sm64(gs + (p >> 5) * 0x9E3779B97F4A7C15ull);
    return (uint32_t)(w >> (2 * (p & 31))) & 3u;
}

typedef struct ccb_read_plan {
    uint32_t kind;      /* 0 genome window, 1 random, 2 genome window with one N */
    uint32_t genome;
    uint64_t pos;
    uint32_t reverse;
    uint32_t n_pos;     /* position of the N (kind 2) */
    uint64_t rs;        /* per-read stream seed */
} ccb_read_plan;

CCB_SYNTH_FN ccb_read_plan ccb_plan_read(uint64_t seed, uint64_t i, uint32_t n_genomes,
                                         uint64_t genome_len, uint32_t read_len)
{
    ccb_read_plan p;
    p.rs = ccb_sm64(seed ^ (i * 0xA24BAED4963EE407ull));
    uint64_t a = ccb_sm64(p.rs + 1), b = ccb_sm64(p.rs + 2), c = ccb_sm64(p.rs + 3);
    uint32_t u = (uint32_t)(a % 1000u);
    p.kind = u < 900u ? 0u : (u < 999u ? 1u : 2u);
    p.genome = (uint32_t)((a >> 20) % n_genomes);
    p.pos = b % (genome_len - read_len + 1);
    p.reverse = (uint32_t)(c & 1u);
    p.n_pos = (uint32_t)((c >> 8) % read_len);
    return p;
}

/* code of base j of read described by plan: 0..3, or 4 for 'N' */
CCB_SYNTH_FN uint32_t ccb_read_code(const ccb_read_plan* p, uint64_t seed, uint32_t j, uint32_t read_len)
{
    uint64_t r = ccb_sm64(p->rs + 16 + j);
    if (p->kind == 1u) return (uint32_t)(r & 3u);
    if (p->kind == 2u && j == p->n_pos) return 4u;
    uint64_t gs = ccb_genome_seed(seed, p->genome);
    uint32_t code = p->reverse ? 3u - ccb_genome_code(gs, p->pos + (read_len - 1 - j))
                               : ccb_genome_code(gs, p->pos + j);
    if ((r >> 8) % 100u == 0u) code = (uint32_t)((r >> 40) & 3u);     /* 1 % substitutions */
    return code;
}

/* "TGCA"[code], 'N' for 4 */
CCB_SYNTH_FN char ccb_code_char(uint32_t code) { return code > 3u ? 'N' : "TGCA"[code]; }

/* ---- the other configurations of SURVEY D2 (C3 pairs, C4 long reads) --------------------------------
 *   kind 1  pairs: a fragment of `insert` bases of a uniform genome / position / strand; mate 1 = its first
 *           read_len bases, mate 2 = the reverse complement of its last read_len bases (opposite strands);
 *           1 % iid substitutions per mate; 9.9 % of the pairs are random sequence, 0.1 % carry one N in
 *           mate 1.  The read is "<mate1>N<mate2>", what mergePairedFiles writes (src/file.cc:205-268).
 *   kind 2  long reads: read_len bases walked off a genome window with per-base errors -- ins_pm per mille
 *           insertions (a random base, the genome does not advance), del_pm deletions (one genome base
 *           skipped), sub_pm substitutions; 10 % of the reads are random sequence; no N.
 * Light databases (C3) sample the genomes: the non-overlapping k-mers at positions 0, k, 2k, ... of which
 * every gap-th is kept (src/CuCLARK_hh.hh:711-767) = one k-mer every gap*k bases. */
typedef struct ccb_workload {
    uint32_t kind;       /* 1 pairs, 2 long reads */
    uint32_t read_len;   /* bases per mate / per read */
    uint32_t insert;     /* pairs: fragment length (>= read_len) */
    uint32_t sub_pm, ins_pm, del_pm;    /* per mille per base (pairs: sub_pm only) */
} ccb_workload;

/* bytes of the longest read text (no terminator) */
CCB_SYNTH_FN uint32_t ccb_synth2_max_len(const ccb_workload* w) { return w->kind == 1u ? 2u * w->read_len + 1u : w->read_len; }
/* genome bases a long read may consume */
CCB_SYNTH_FN uint64_t ccb_synth2_span(const ccb_workload* w)
{
    return w->kind == 1u ? (uint64_t)w->insert : (uint64_t)w->read_len + w->read_len / 8u + 8u;
}

/* text of read i (bases as characters, 'N' between / inside mates); returns its length */
CCB_SYNTH_FN uint32_t ccb_synth2_text(const ccb_workload* w, uint64_t reads_seed, uint64_t db_seed, uint64_t i,
                                      uint32_t n_genomes, uint64_t genome_len, char* out)
{
    const uint64_t rs = ccb_sm64(reads_seed ^ (i * 0xA24BAED4963EE407ull));
    const uint64_t a = ccb_sm64(rs + 1), b = ccb_sm64(rs + 2), c = ccb_sm64(rs + 3);
    const uint32_t u = (uint32_t)(a % 1000u);
    const uint32_t L = w->read_len;
    const uint64_t span = ccb_synth2_span(w);
    const uint64_t gs = ccb_genome_seed(db_seed, (uint32_t)((a >> 20) % n_genomes));
    const uint64_t pos = b % (genome_len - span + 1);
    const uint32_t rev = (uint32_t)(c & 1u);
    if (w->kind == 1u) {
        const uint32_t kind = u < 900u ? 0u : (u < 999u ? 1u : 2u);
        const uint32_t n_pos = (uint32_t)((c >> 8) % L);
        for (uint32_t m = 0; m < 2u; m++) {
            for (uint32_t j = 0; j < L; j++) {
                const uint64_t r = ccb_sm64(rs + 16 + m * (L + 1u) + j);
                uint32_t code;
                if (kind == 1u) code = (uint32_t)(r & 3u);
                else {
                    /* mate (m ^ rev) == 0 reads the fragment forward from its start, the other one backward from its end */
                    code = (m ^ rev) == 0u ? ccb_genome_code(gs, pos + j) : 3u - ccb_genome_code(gs, pos + span - 1 - j);
                    if ((r >> 8) % 1000u < w->sub_pm) code = (uint32_t)((r >> 40) & 3u);
                    if (kind == 2u && m == 0u && j == n_pos) code = 4u;
                }
                out[m * (L + 1u) + j] = ccb_code_char(code);
            }
        }
        out[L] = 'N';
        return 2u * L + 1u;
    }
    {
        const int random_read = u >= 900u;
        uint64_t gp = 0;                                    /* genome bases consumed */
        for (uint32_t j = 0; j < L; j++) {
            const uint64_t r = ccb_sm64(rs + 16 + j);
            const uint32_t e = (uint32_t)(r % 1000u);
            uint32_t code;
            if (random_read || e < w->ins_pm) code = (uint32_t)((r >> 12) & 3u);
            else {
                if (e < w->ins_pm + w->del_pm && gp + 2 < span) gp++;
                if (gp >= span) gp = span - 1;
                code = rev ? 3u - ccb_genome_code(gs, pos + span - 1 - gp) : ccb_genome_code(gs, pos + gp);
                gp++;
                if ((r >> 20) % 1000u < w->sub_pm) code = (uint32_t)((r >> 40) & 3u);
            }
            out[j] = ccb_code_char(code);
        }
        return L;
    }
}

#endif
