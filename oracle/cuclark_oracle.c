/*
 * cuclark_oracle.c -- CPU ORACLE (test infrastructure, NOT product code).
 * See cuclark_oracle.h.  Plain C11; every function cites the reference
 * file:line it restates.  Nothing under jn_cuclark_b200/ links this file.
 */
#define _GNU_SOURCE
#include "cuclark_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ===================================================================== */
/* A3/A4: encoding                                                        */
/* ===================================================================== */

/* src/CuCLARK_hh.hh:286-300: m_rTable (x-encoding, complement code) with the
 * non-nucleotide classes of m_table.  A=3 C=2 G=1 T/U=0. */
int ccbo_rcode(unsigned char c)
{
    switch (c) {
    case 'A': case 'a': return 3;
    case 'C': case 'c': return 2;
    case 'G': case 'g': return 1;
    case 'T': case 't': case 'U': case 'u': return 0;
    case '>':  return -2;
    case '\n': return -10;
    default:   return -1;
    }
}

/* src/CuClarkDB.cu:1195-1203: reverse the 2-bit groups of the word, then
 * complement and shift down so the k bases sit in the low 2k bits. */
uint64_t ccbo_revcomp(uint64_t x, int k)
{
    uint64_t r = x;
    r = ((r >> 2)  & 0x3333333333333333ULL) | ((r & 0x3333333333333333ULL) << 2);
    r = ((r >> 4)  & 0x0F0F0F0F0F0F0F0FULL) | ((r & 0x0F0F0F0F0F0F0F0FULL) << 4);
    r = ((r >> 8)  & 0x00FF00FF00FF00FFULL) | ((r & 0x00FF00FF00FF00FFULL) << 8);
    r = ((r >> 16) & 0x0000FFFF0000FFFFULL) | ((r & 0x0000FFFF0000FFFFULL) << 16);
    r = (r >> 32) | (r << 32);
    return (~r) >> (64 - 2 * k);
}

/* src/CuClarkDB.cu:1206 */
uint64_t ccbo_canonical(uint64_t x, int k)
{
    uint64_t r = ccbo_revcomp(x, k);
    return x < r ? x : r;
}

/* src/main.cc:251-286: t_b = (size_t)(log(HTSIZE)/log(4.0)); k<=t_b+8 -> u16,
 * k<=t_b+16 -> u32, else u64. */
int ccbo_key_bytes(int k, uint64_t htsize)
{
    size_t t_b = (size_t)(log((double)htsize) / log(4.0));
    if ((size_t)k <= t_b + 8)  return 2;
    if ((size_t)k <= t_b + 16) return 4;
    return 8;
}

/* ===================================================================== */
/* targets file                                                           */
/* ===================================================================== */

/* src/file.cc:63-87: separators ' ' ',' '\n' '\t' '\r', at most max tokens */
static int split_line(const char *line, int max, char out[][4096])
{
    int n = 0; size_t t = 0, len = strlen(line);
    while (t < len && n < max) {
        while (t < len && (line[t] == ' ' || line[t] == ',' || line[t] == '\n' ||
                           line[t] == '\t' || line[t] == '\r')) t++;
        size_t s = t;
        while (t < len && !(line[t] == ' ' || line[t] == ',' || line[t] == '\n' ||
                            line[t] == '\t' || line[t] == '\r')) t++;
        if (t > s) {
            size_t l = t - s; if (l > 4095) l = 4095;
            memcpy(out[n], line + s, l); out[n][l] = 0; n++;
        }
    }
    return n;
}

/* src/CuCLARK_hh.hh:1811-1863: one target per line, labels collected in
 * first-appearance order (:1825-1829).  A third column (centromere labels,
 * :1844-1861) is outside the hot path and rejected. */
int ccbo_targets_read(const char *path, ccbo_targets *t)
{
    memset(t, 0, sizeof *t);
    FILE *f = fopen(path, "r");
    if (!f) return -1;
    char *line = NULL; size_t cap = 0;
    static char tok[3][4096];
    while (getline(&line, &cap, f) != -1) {
        int n = split_line(line, 3, tok);
        if (n == 0) continue;
        if (n < 2) { fclose(f); free(line); return -2; }   /* :1842 Missing label */
        if (n > 2) { fclose(f); free(line); return -3; }
        int li = -1;
        for (int i = 0; i < t->n_labels; i++)
            if (strcmp(t->labels[i], tok[1]) == 0) { li = i; break; }
        if (li < 0) {
            t->labels = realloc(t->labels, sizeof(char *) * (t->n_labels + 1));
            t->labels[t->n_labels] = strdup(tok[1]);
            li = t->n_labels++;
        }
        t->files = realloc(t->files, sizeof(char *) * (t->n_files + 1));
        t->file_label = realloc(t->file_label, sizeof(uint16_t) * (t->n_files + 1));
        t->files[t->n_files] = strdup(tok[0]);
        t->file_label[t->n_files] = (uint16_t)li;
        t->n_files++;
    }
    free(line); fclose(f);
    return 0;
}

void ccbo_targets_free(ccbo_targets *t)
{
    for (int i = 0; i < t->n_files; i++) free(t->files[i]);
    for (int i = 0; i < t->n_labels; i++) free(t->labels[i]);
    free(t->files); free(t->file_label); free(t->labels);
    memset(t, 0, sizeof *t);
}

/* src/CuCLARK_hh.hh:581-592 */
void ccbo_db_name(char *out, size_t cap, const char *folder, int k, int n_labels,
                  uint64_t htsize, unsigned min_t, int light, unsigned gap)
{
    if (light)
        snprintf(out, cap, "%s/db_central_k%lu_t%lu_s%lu_m%lu_light_%lu.tsk", folder,
                 (unsigned long)k, (unsigned long)n_labels, (unsigned long)htsize,
                 (unsigned long)min_t, (unsigned long)gap);
    else
        snprintf(out, cap, "%s/db_central_k%lu_t%lu_s%lu_m%lu.tsk", folder,
                 (unsigned long)k, (unsigned long)n_labels, (unsigned long)htsize,
                 (unsigned long)min_t);
}

/* ===================================================================== */
/* A8: k-mer extraction from target FASTA                                 */
/* ===================================================================== */

static void occ_push(ccbo_occ_vec *v, uint64_t c, uint16_t label)
{
    if (v->n == v->cap) {
        v->cap = v->cap ? v->cap * 2 : (1u << 16);
        v->v = realloc(v->v, v->cap * sizeof(ccbo_occ));
    }
    v->v[v->n].c = c; v->v[v->n].label = label; v->n++;
}

/* full : src/CuCLARK_hh.hh:1129-1198.  The reference rolls the forward word
 *        (_km_r, :1149-1150) for the first k bases and afterwards the
 *        reverse-complement word (_km_f, :1140-1141); addElement
 *        canonicalises either (HashTableStorage_hh.hh:425-435), so the set of
 *        canonical values equals "every window of every ACGT run".
 * light: src/CuCLARK_hh.hh:711-767.  Non-overlapping windows: the window is
 *        cleared after each complete k-mer (:727); only windows whose running
 *        per-file index iter satisfies iter % gap == 0 are added (:725).
 * Shared rules: '\n' skipped (:737/:1170); any other non-ACGTU byte resets the
 * window (:740/:1173); '>' resets and skips to end of line (:743-757/:1176-1188).
 * Only FASTA targets (first byte '>', :711/:1129) are restated. */
int ccbo_extract_fasta(const unsigned char *txt, size_t len, int k, int light,
                       unsigned gap, uint16_t label, ccbo_occ_vec *out)
{
    if (len == 0 || txt[0] != '>') return -1;
    const uint64_t mask = (k == 32) ? ~0ULL : ((1ULL << (2 * k)) - 1);
    uint64_t x = 0;      /* forward word, x-encoding, first base most significant */
    int cpt = 0;         /* bases in the window */
    uint64_t iter = 0;   /* :709 */
    size_t i = 0;
    while (i < len) {
        int code = ccbo_rcode(txt[i]);
        if (code >= 0) {
            x = ((x << 2) | (uint64_t)code) & mask;
            if (light) {
                if (cpt == k - 1) {
                    if (iter % gap == 0) occ_push(out, ccbo_canonical(x, k), label);
                    x = 0; cpt = 0; iter++;
                } else cpt++;
            } else {
                if (cpt < k) cpt++;
                if (cpt == k) occ_push(out, ccbo_canonical(x, k), label);
            }
            i++; continue;
        }
        if (code == -10) { i++; continue; }
        if (code == -1)  { x = 0; cpt = 0; i++; continue; }
        /* '>' */
        x = 0; cpt = 0;
        while (i < len && txt[i] != '\n') i++;
        i++;
    }
    return 0;
}

/* ===================================================================== */
/* A8/A9: discriminative filter + on-disk layout                          */
/* ===================================================================== */

static int occ_cmp(const void *a, const void *b)
{
    const ccbo_occ *x = a, *y = b;
    if (x->c != y->c) return x->c < y->c ? -1 : 1;
    return (int)x->label - (int)y->label;
}

typedef struct { uint64_t r, q; uint16_t label; } ent_t;
static int ent_cmp(const void *a, const void *b)
{
    const ent_t *x = a, *y = b;
    if (x->r != y->r) return x->r < y->r ? -1 : 1;
    if (x->q != y->q) return x->q < y->q ? -1 : 1;
    return 0;
}

/* addElement (HashTableStorage_hh.hh:445-460): first sight inserts with
 * Multiplicity 1, Count 1; a later sight under the same label string leaves
 * Multiplicity alone, any other label string raises it by 1 or 2
 * (hashTable_hh.hh:401-409) -- so Multiplicity==1 <=> one distinct label.
 * lElement::AddToCount (dataType.hh:341-342) saturates Count at 254.
 * RemoveCommon (HashTableStorage_hh.hh:244): keep iff Multiplicity==1 and
 * Count > min_t.  Layout (hashTable_hh.hh:488-531): bucket r = c % HTSIZE,
 * key = c / HTSIZE, ascending key inside a bucket (sortall(2), :204-216),
 * buckets larger than 255 are a fatal error (:499-507). */
int ccbo_db_from_occ(ccbo_occ_vec *occ, int k, uint64_t htsize, unsigned min_t,
                     ccbo_db *db)
{
    memset(db, 0, sizeof *db);
    db->htsize = htsize; db->k = k; db->key_bytes = ccbo_key_bytes(k, htsize);
    qsort(occ->v, occ->n, sizeof(ccbo_occ), occ_cmp);
    ent_t *e = malloc(sizeof(ent_t) * (occ->n ? occ->n : 1));
    size_t ne = 0;
    for (size_t i = 0; i < occ->n;) {
        size_t j = i; int one_label = 1; unsigned count = 0;
        while (j < occ->n && occ->v[j].c == occ->v[i].c) {
            if (occ->v[j].label != occ->v[i].label) one_label = 0;
            if (count == 0) count = 1; else if (count + 1 < 255) count++;
            j++;
        }
        if (one_label && count > min_t) {
            e[ne].r = occ->v[i].c % htsize; e[ne].q = occ->v[i].c / htsize;
            e[ne].label = occ->v[i].label; ne++;
        }
        i = j;
    }
    free(occ->v); occ->v = NULL; occ->n = occ->cap = 0;
    qsort(e, ne, sizeof(ent_t), ent_cmp);
    db->n = ne;
    db->sz = calloc(htsize, 1);
    db->keys = malloc(sizeof(uint64_t) * (ne ? ne : 1));
    db->labels = malloc(sizeof(uint16_t) * (ne ? ne : 1));
    for (size_t i = 0; i < ne; i++) {
        if (db->sz[e[i].r] == 255) { free(e); ccbo_db_free(db); return -2; }
        db->sz[e[i].r]++;
        db->keys[i] = e[i].q; db->labels[i] = e[i].label;
    }
    free(e);
    return 0;
}

/* src/CuCLARK_hh.hh:691-1329 (makeSpecificTargetSets), FASTA targets only */
int ccbo_db_build(const ccbo_targets *t, int k, uint64_t htsize, int light,
                  unsigned gap, unsigned min_t, ccbo_db *db)
{
    ccbo_occ_vec occ = {0};
    for (int f = 0; f < t->n_files; f++) {
        FILE *fp = fopen(t->files[f], "rb");
        if (!fp) continue;                      /* :700-704 "Failed to open", skip */
        fseek(fp, 0, SEEK_END); long sz = ftell(fp); fseek(fp, 0, SEEK_SET);
        unsigned char *buf = malloc(sz > 0 ? (size_t)sz : 1);
        size_t got = fread(buf, 1, (size_t)sz, fp); fclose(fp);
        int rc = ccbo_extract_fasta(buf, got, k, light, gap, t->file_label[f], &occ);
        free(buf);
        if (rc) { free(occ.v); return rc; }
    }
    return ccbo_db_from_occ(&occ, k, htsize, min_t, db);
}

/* src/hashTable_hh.hh:474-546 */
int ccbo_db_write(const ccbo_db *db, const char *prefix)
{
    char p[4200]; FILE *f;
    snprintf(p, sizeof p, "%s.sz", prefix);
    if (!(f = fopen(p, "wb"))) return -1;
    fwrite(db->sz, 1, db->htsize, f); fclose(f);
    snprintf(p, sizeof p, "%s.ky", prefix);
    if (!(f = fopen(p, "wb"))) return -1;
    for (uint64_t i = 0; i < db->n; i++) {
        if (db->key_bytes == 2) { uint16_t v = (uint16_t)db->keys[i]; fwrite(&v, 2, 1, f); }
        else if (db->key_bytes == 4) { uint32_t v = (uint32_t)db->keys[i]; fwrite(&v, 4, 1, f); }
        else fwrite(&db->keys[i], 8, 1, f);
    }
    fclose(f);
    snprintf(p, sizeof p, "%s.lb", prefix);
    if (!(f = fopen(p, "wb"))) return -1;
    fwrite(db->labels, 2, db->n, f); fclose(f);
    return 0;
}

/* src/CuClarkDB.cu:464-741: read all HTSIZE sizes (:493-496), pick buckets
 * (-s sampling keeps every sampling-th non-empty bucket, :503-513; skipped
 * buckets get size 0 in the pointer array, :596-617), stream keys/labels
 * (:671-741). */
int ccbo_db_read(const char *prefix, uint64_t htsize, int key_bytes, int k,
                 unsigned sampling, ccbo_db *db)
{
    memset(db, 0, sizeof *db);
    db->htsize = htsize; db->key_bytes = key_bytes; db->k = k;
    char p[4200]; FILE *fs, *fk, *fl;
    snprintf(p, sizeof p, "%s.sz", prefix); fs = fopen(p, "rb");
    snprintf(p, sizeof p, "%s.ky", prefix); fk = fopen(p, "rb");
    snprintf(p, sizeof p, "%s.lb", prefix); fl = fopen(p, "rb");
    if (!fs || !fk || !fl) { if (fs) fclose(fs); if (fk) fclose(fk); if (fl) fclose(fl); return -1; }
    uint8_t *fsz = malloc(htsize);
    if (fread(fsz, 1, htsize, fs) != htsize) { fclose(fs); fclose(fk); fclose(fl); free(fsz); return -2; }
    fclose(fs);
    db->sz = calloc(htsize, 1);
    const int all = sampling <= 1;
    uint64_t n_keep = 0, n_all = 0; uint32_t nonzero = 0;
    for (uint64_t b = 0; b < htsize; b++) if (fsz[b]) {
        nonzero++; n_all += fsz[b];
        if (all || (nonzero % sampling) == 0) { db->sz[b] = fsz[b]; n_keep += fsz[b]; }
    }
    db->n = n_keep;
    db->keys = malloc(sizeof(uint64_t) * (n_keep ? n_keep : 1));
    db->labels = malloc(sizeof(uint16_t) * (n_keep ? n_keep : 1));
    uint64_t w = 0;
    for (uint64_t b = 0; b < htsize; b++) {
        for (unsigned j = 0; j < fsz[b]; j++) {
            uint64_t key = 0; uint16_t lb = 0;
            if (fread(&key, (size_t)key_bytes, 1, fk) != 1 || fread(&lb, 2, 1, fl) != 1) {
                fclose(fk); fclose(fl); free(fsz); ccbo_db_free(db); return -3;
            }
            if (db->sz[b]) { db->keys[w] = key; db->labels[w] = lb; w++; }
        }
    }
    fclose(fk); fclose(fl); free(fsz);
    return 0;
}

void ccbo_db_free(ccbo_db *db)
{
    free(db->sz); free(db->keys); free(db->labels); free(db->ptr);
    memset(db, 0, sizeof *db);
}

/* src/hashTable_hh.hh:359-396 == src/CuClarkDB.cu:1190-1254: canonicalise,
 * q = c / HTSIZE, r = c % HTSIZE, early-out on the bucket's first/last key,
 * ascending linear scan. */
/* bucket pointers = exclusive prefix sum of the sizes (src/CuClarkDB.cu:589-617);
 * built on first use because it is 4 B x HTSIZE (6.4 GB in full mode). */
void ccbo_db_index(ccbo_db *db)
{
    if (db->ptr) return;
    db->ptr = malloc(sizeof(uint32_t) * (db->htsize + 1));
    uint32_t acc = 0;
    for (uint64_t b = 0; b < db->htsize; b++) { db->ptr[b] = acc; acc += db->sz[b]; }
    db->ptr[db->htsize] = acc;
}

int ccbo_db_find(const ccbo_db *db, uint64_t x, uint16_t *label)
{
    if (!db->ptr) ccbo_db_index((ccbo_db *)db);
    uint64_t c = ccbo_canonical(x, db->k);
    uint64_t q = c / db->htsize, r = c - q * db->htsize;
    uint64_t b = db->ptr[r], e = db->ptr[r + 1];
    if (e == b) return 0;
    if (db->keys[b] > q || db->keys[e - 1] < q) return 0;
    for (uint64_t i = b; db->keys[i] <= q; i++)
        if (db->keys[i] == q) { *label = db->labels[i]; return 1; }
    return 0;
}

/* ===================================================================== */
/* A1: read indexing (one batch covering the whole file)                  */
/* ===================================================================== */

static void reads_push(ccbo_reads *r, size_t *cap, uint64_t ns, uint64_t ne,
                       uint64_t ss, uint64_t se, uint64_t len)
{
    if (r->n == *cap) {
        *cap = *cap ? *cap * 2 : 1024;
        r->name_s = realloc(r->name_s, *cap * 8); r->name_e = realloc(r->name_e, *cap * 8);
        r->seq_s = realloc(r->seq_s, *cap * 8);   r->seq_e = realloc(r->seq_e, *cap * 8);
        r->len = realloc(r->len, *cap * 8);
    }
    r->name_s[r->n] = ns; r->name_e[r->n] = ne; r->seq_s[r->n] = ss;
    r->seq_e[r->n] = se; r->len[r->n] = len; r->n++;
}

static unsigned char at(const unsigned char *m, size_t nb, size_t p) { return p < nb ? m[p] : 0; }
static int is_sep(unsigned char c) { return c == ' ' || c == '\t' || c == '\n'; } /* :302 */

/* FASTA: src/CuCLARK_hh.hh:1364-1394.  FASTQ: :1496-1524 (four lines per
 * record, no validation).  Reads past the end of the map are guarded here
 * (the reference relies on mmap zero padding; a NUL is not a separator). */
int ccbo_index_reads(const unsigned char *map, size_t nb, ccbo_reads *r)
{
    memset(r, 0, sizeof *r);
    size_t cap = 0;
    if (nb == 0) return -1;
    if (map[0] == '>') {
        r->is_fasta = 1;
        size_t i = 1;                               /* :1364 i past the first '>' */
        for (;;) {
            uint64_t ns = i;                        /* :1369 */
            while (i < nb && !is_sep(at(map, nb, ++i))) {}   /* :1370 pre-increment */
            uint64_t ne = i;
            while (i < nb && map[i++] != '\n') {}   /* :1374 */
            uint64_t ss = i, se = i, lines = 0;
            while (i < nb && map[i] != '>') {       /* :1379-1387 */
                while (i < nb && map[i] != '\n') i++;
                lines++; se = i++;
            }
            uint64_t len = (uint64_t)0 - lines + (se - ss + 1);   /* :1388-1389 */
            reads_push(r, &cap, ns, ne, ss, se, len);
            if (i >= nb) break;                     /* :1390 */
            i++;                                    /* :1392 skip '>' */
        }
        return 0;
    }
    if (map[0] == '@') {
        r->is_fasta = 0;
        size_t i = 1;                               /* :1407 */
        for (;;) {
            uint64_t ns = i;
            while (i < nb && !is_sep(at(map, nb, ++i))) {}   /* :1499 */
            uint64_t ne = i;
            while (i < nb && map[i++] != '\n') {}   /* :1503 */
            uint64_t ss = i;
            while (i < nb && map[i] != '\n') i++;   /* :1508 */
            uint64_t se = i++;                      /* :1512 */
            reads_push(r, &cap, ns, ne, ss, se, se - ss);
            while (i < nb && map[i++] != '\n') {}   /* :1515 third line */
            while (i < nb && map[i++] != '\n') {}   /* :1518 fourth line */
            if ((++i) >= nb) break;                 /* :1521 */
        }
        return 0;
    }
    return -1;                                      /* :1537 */
}

void ccbo_reads_free(ccbo_reads *r)
{
    free(r->name_s); free(r->name_e); free(r->seq_s); free(r->seq_e); free(r->len);
    memset(r, 0, sizeof *r);
}

/* ===================================================================== */
/* A2: packing                                                            */
/* ===================================================================== */

/* src/CuCLARK_hh.hh:1629-1715, statement for statement.  A "part" is a maximal
 * run of ACGTU/acgtu bytes ('\n' skipped, :1673-1677); it is stored as
 * [len:u16][ceil(len/8) containers], 8 bases per u16, first base in the two
 * most significant bits, last container left-aligned (:1682/:1694).  A part
 * shorter than k is overwritten by the next one (:1646-1650) or dropped at
 * the end of the read (:1700-1703).  Reads shorter than k emit nothing
 * (:1632). */
int64_t ccbo_pack(const unsigned char *map, const ccbo_reads *r, size_t first,
                  size_t count, int k, uint32_t *offsets, uint16_t *cont, size_t cap)
{
    uint16_t kc = 0; size_t cur = 0; int new_seq = 1;
    size_t cc = 0, part_begin;
    const size_t per = 8;
    for (size_t lr = 0; lr < count; lr++) {
        size_t rd = first + lr;
        size_t ic = r->len[rd] < (uint64_t)k ? r->seq_e[rd] : r->seq_s[rd];
        part_begin = cc;
        offsets[lr] = (uint32_t)part_begin;
        if (cc + (r->seq_e[rd] - ic) / 8 + (r->seq_e[rd] - ic) / (size_t)(k ? k : 1) + 8 > cap) return -1;
        while (ic < r->seq_e[rd]) {
            int code = ccbo_rcode(map[ic]);
            if (code >= 0) {
                if (new_seq) {
                    cont[cc++] = 0;
                    if (cont[part_begin] < k) { cc = part_begin + 1; cont[part_begin] = 0; }
                    else part_begin = cc - 1;
                    new_seq = 0;
                }
                kc = (uint16_t)(kc << 2); kc ^= (uint16_t)code;
                cur++; ic++;
                if (cur == per) { cont[cc++] = kc; cont[part_begin] = (uint16_t)(cont[part_begin] + cur); cur = 0; }
                continue;
            }
            if (map[ic] == '\n') { ic++; continue; }
            if (cur > 0) {
                kc = (uint16_t)(kc << (2 * (per - cur)));
                cont[cc++] = kc; cont[part_begin] = (uint16_t)(cont[part_begin] + cur);
            }
            kc = 0; cur = 0; new_seq = 1; ic++;
        }
        if (cur > 0) {
            kc = (uint16_t)(kc << (2 * (per - cur)));
            cont[cc++] = kc; cont[part_begin] = (uint16_t)(cont[part_begin] + cur);
        }
        /* :1700: the reference reads cont[part_begin] even when the read stored
         * nothing (then it is stale memory and either outcome leaves cc as is) */
        if (cc > part_begin && cont[part_begin] < k) cc = part_begin;
        kc = 0; cur = 0; new_seq = 1;
    }
    if (cc > 0xFFFFFFFFull) return -2;              /* :1709 */
    offsets[count] = (uint32_t)cc;
    return (int64_t)cc;
}

/* ===================================================================== */
/* A3-A6: query + vote                                                    */
/* ===================================================================== */

typedef struct { uint16_t t; uint32_t c; } tc_t;
static int tc_cmp(const void *a, const void *b)
{ return (int)((const tc_t *)a)->t - (int)((const tc_t *)b)->t; }

/* One read.  Parts are walked like src/CuClarkDB.cu:1042-1047; the k-mer at
 * offset j of a part is the 2k-bit window of the container bit stream
 * (:1063-1083); each is probed with queryElement (:1104); hits[target]++ in
 * u16 (:1106-1108).  The sparse row lists non-zero targets ascending
 * (:1122-1182); resultKernel (:1370-1405) scans it in that order. */
static void query_one(ccbo_find_fn find, void *ctx, int k, const uint16_t *c, uint32_t beg,
                      uint32_t end, uint16_t *out5, uint16_t *row, int row_stride,
                      uint64_t *lookups)
{
    const uint64_t mask = (k == 32) ? ~0ULL : ((1ULL << (2 * k)) - 1);
    tc_t *acc = NULL; size_t na = 0, ca = 0;
    uint32_t p = beg;
    while (p < end) {
        uint32_t len = c[p];
        uint32_t first = ++p;
        p += (len - 1) / 8 + 1;
        uint64_t x = 0;
        for (uint32_t j = 0; j < len; j++) {
            uint32_t code = (c[first + j / 8] >> (2 * (7 - j % 8))) & 3;
            x = ((x << 2) | code) & mask;
            if (j + 1 >= (uint32_t)k) {
                uint16_t lb;
                (*lookups)++;
                if (find(ctx, x, &lb)) {
                    size_t a = 0;
                    for (; a < na; a++) if (acc[a].t == lb) break;
                    if (a == na) {
                        if (na == ca) { ca = ca ? 2 * ca : 16; acc = realloc(acc, ca * sizeof(tc_t)); }
                        acc[na].t = lb; acc[na].c = 0; na++;
                    }
                    acc[a].c++;
                }
            }
        }
    }
    qsort(acc, na, sizeof(tc_t), tc_cmp);
    uint16_t best = 0, s_best = 0, ib = 0, isb = 0, sum = 0;
    for (size_t a = 0; a < na; a++) {
        uint16_t sc = (uint16_t)acc[a].c;              /* u16 counters */
        if (sc > best) { s_best = best; isb = ib; best = sc; ib = (uint16_t)(acc[a].t + 1); }
        else if (sc > s_best) { s_best = sc; isb = (uint16_t)(acc[a].t + 1); }
        sum = (uint16_t)(sum + sc);
    }
    out5[0] = sum; out5[1] = ib; out5[2] = best; out5[3] = isb; out5[4] = s_best;
    if (row) {
        size_t fit = (size_t)(row_stride - 1) / 2;
        size_t n = na < fit ? na : fit;
        memset(row, 0, sizeof(uint16_t) * (size_t)row_stride);
        row[0] = (uint16_t)n;
        for (size_t a = 0; a < n; a++) { row[1 + 2 * a] = acc[a].t; row[2 + 2 * a] = (uint16_t)acc[a].c; }
    }
    free(acc);
}

static int db_find_cb(void *ctx, uint64_t x, uint16_t *label)
{ return ccbo_db_find((const ccbo_db *)ctx, x, label); }

int ccbo_query_packed_cb(ccbo_find_fn find, void *ctx, int k, const uint32_t *offsets,
                         const uint16_t *containers, size_t n_reads,
                         uint16_t *out5, uint16_t *rows, int row_stride,
                         uint64_t *n_lookups, int threads)
{
    uint64_t total = 0;
#ifdef _OPENMP
    if (threads > 1) {
#pragma omp parallel for schedule(dynamic, 256) reduction(+:total) num_threads(threads)
        for (long i = 0; i < (long)n_reads; i++) {
            uint64_t l = 0;
            query_one(find, ctx, k, containers, offsets[i], offsets[i + 1], out5 + 5 * i,
                      rows ? rows + (size_t)row_stride * i : NULL, row_stride, &l);
            total += l;
        }
    } else
#endif
    {
        (void)threads;
        for (size_t i = 0; i < n_reads; i++)
            query_one(find, ctx, k, containers, offsets[i], offsets[i + 1], out5 + 5 * i,
                      rows ? rows + (size_t)row_stride * i : NULL, row_stride, &total);
    }
    if (n_lookups) *n_lookups = total;
    return 0;
}

int ccbo_query_packed(const ccbo_db *db, int k, const uint32_t *offsets,
                      const uint16_t *containers, size_t n_reads, int n_targets,
                      uint16_t *out5, uint16_t *rows, int row_stride,
                      uint64_t *n_lookups, int threads)
{
    (void)n_targets;
    ccbo_db_index((ccbo_db *)db);
    return ccbo_query_packed_cb(db_find_cb, (void *)db, k, offsets, containers, n_reads,
                                out5, rows, row_stride, n_lookups, threads);
}

/* ===================================================================== */
/* A7: CSV                                                                */
/* ===================================================================== */

/* src/CuCLARK_hh.hh:1945-2122.  Header :1952-1967; normal row :2096-2118;
 * extended row :2006-2051 (one column per target, zeros filled in). */
int ccbo_write_csv(const char *path, const unsigned char *map, const ccbo_reads *r,
                   int k, int paired, int extended, int n_targets, char **names,
                   const uint16_t *out5, const uint16_t *rows, int row_stride)
{
    FILE *f = fopen(path, "w");
    if (!f) return -1;
    fputs("Object_ID", f);
    if (extended) for (int t = 1; t <= n_targets; t++) fprintf(f, ",%s", names[t]);
    fputs(",Gamma,Assignment,Score,Confidence\n", f);
    for (size_t i = 0; i < r->n; i++) {
        unsigned total = out5[5 * i], ib = out5[5 * i + 1], best = out5[5 * i + 2],
                 s_best = out5[5 * i + 4];
        char name[CCBO_OBJECTNAMEMAX];
        size_t nl = (size_t)(r->name_e[i] - r->name_s[i]);
        if (nl >= CCBO_OBJECTNAMEMAX) nl = CCBO_OBJECTNAMEMAX - 1;      /* :2103 */
        memcpy(name, map + r->name_s[i], nl); name[nl] = 0;
        /* strncpy semantics (:2104): an embedded NUL ends the name */
        uint32_t norm = paired ? (uint32_t)r->len[i] - CCBO_NBN : (uint32_t)r->len[i]; /* :2107, ITYPE=u32 */
        double gamma = (double)total / (((double)norm - (double)(size_t)k) + 1.0);     /* :2110 */
        double delta = (double)(best + s_best);
        delta = (delta < 0.001) ? 0 : ((double)best) / delta;                           /* :2111-2112 */
        fputs(name, f);
        if (extended) {
            const uint16_t *row = rows + (size_t)row_stride * i;
            unsigned w = 0;
            for (unsigned h = 0; h < row[0]; h++) {
                unsigned ti = row[1 + 2 * h], sc = row[2 + 2 * h];
                for (; w < ti; w++) fputs(",0", f);
                fprintf(f, ",%d", (int)sc); w++;
            }
            for (; w < (unsigned)n_targets; w++) fputs(",0", f);
        }
        fprintf(f, ",%g,%s,%u,%g\n", gamma, names[ib], best, delta);
    }
    fclose(f);
    return 0;
}

/* ===================================================================== */
/* A10: paired-end join                                                   */
/* ===================================================================== */

static int get_line(FILE *f, char **buf, size_t *cap)
{   /* src/file.cc:123-142: strip one trailing '\n' */
    ssize_t n = getline(buf, cap, f);
    if (n < 0) return 0;
    if (n > 0 && (*buf)[n - 1] == '\n') (*buf)[n - 1] = 0;
    return 1;
}

/* first token under separators ' ' '/' '\t' and the delimiter '@'
 * (src/file.cc:211-213,231; tokeniser :89-121) */
static void first_tok(const char *s, char *out, size_t cap)
{
    size_t t = 0, len = strlen(s), o = 0;
    while (t < len && (s[t] == ' ' || s[t] == '/' || s[t] == '\t' || s[t] == '@')) t++;
    while (t < len && !(s[t] == ' ' || s[t] == '/' || s[t] == '\t' || s[t] == '@') && o + 1 < cap)
        out[o++] = s[t++];
    out[o] = 0;
}

/* src/file.cc:205-268: ">id\n<R1>N<R2>\n" per pair; FASTQ only */
int ccbo_merge_paired(const char *f1, const char *f2, const char *outp)
{
    FILE *a = fopen(f1, "r"), *b = fopen(f2, "r");
    if (!a || !b) { if (a) fclose(a); if (b) fclose(b); return -1; }
    char *l1 = NULL, *l2 = NULL; size_t c1 = 0, c2 = 0;
    if (!get_line(a, &l1, &c1) || !get_line(b, &l2, &c2) || l1[0] != l2[0] || l1[0] != '@') {
        fclose(a); fclose(b); free(l1); free(l2); return -2;
    }
    rewind(a); rewind(b);
    FILE *o = fopen(outp, "wb");
    if (!o) { fclose(a); fclose(b); free(l1); free(l2); return -3; }
    int rc = 0;
    static char t1[4096], t2[4096];
    while (get_line(a, &l1, &c1) && get_line(b, &l2, &c2)) {
        if (l1[0] == '@' && l2[0] == '@') {
            first_tok(l1, t1, sizeof t1); first_tok(l2, t2, sizeof t2);
            if (strcmp(t1, t2) != 0) { rc = -4; break; }
            fprintf(o, ">%s\n", t1);
            if (get_line(a, &l1, &c1) && get_line(b, &l2, &c2)) {
                fprintf(o, "%sN%s\n", l1, l2);
                if (get_line(a, &l1, &c1) && get_line(b, &l2, &c2))
                    if (get_line(a, &l1, &c1) && get_line(b, &l2, &c2)) continue;
            } else { rc = -5; break; }
        }
    }
    fclose(a); fclose(b); fclose(o); free(l1); free(l2);
    return rc;
}

/* ===================================================================== */
/* synthetic benchmark workload (include/ccb_synth.h) on the CPU          */
/* ===================================================================== */
#include "../include/ccb_synth.h"

/* Database of the first n_genomes synthetic genomes, built with the same rules
 * as any FASTA target (ccbo_extract_fasta + ccbo_db_from_occ). */
int ccbo_synth_db(uint64_t db_seed, uint32_t n_genomes, uint64_t genome_len, uint32_t n_labels,
                  int k, uint64_t htsize, ccbo_db *db)
{
    return ccbo_synth_db_ex(db_seed, n_genomes, genome_len, n_labels, k, htsize, 0, 0, db);
}

/* light != 0: the light-mode scan (every gap-th non-overlapping k-mer, ccbo_extract_fasta) over the same genomes */
int ccbo_synth_db_ex(uint64_t db_seed, uint32_t n_genomes, uint64_t genome_len, uint32_t n_labels,
                     int k, uint64_t htsize, int light, unsigned gap, ccbo_db *db)
{
    ccbo_occ_vec occ = {0};
    unsigned char *txt = malloc(genome_len + 8);
    if (!txt) return -1;
    txt[0] = '>'; txt[1] = 'g'; txt[2] = '\n';
    for (uint32_t g = 0; g < n_genomes; g++) {
        uint64_t gs = ccb_genome_seed(db_seed, g);
        for (uint64_t p = 0; p < genome_len; p++) txt[3 + p] = (unsigned char)ccb_code_char(ccb_genome_code(gs, p));
        int rc = ccbo_extract_fasta(txt, 3 + genome_len, k, light, gap, (uint16_t)(g % n_labels), &occ);
        if (rc) { free(txt); free(occ.v); return rc; }
    }
    free(txt);
    return ccbo_db_from_occ(&occ, k, htsize, 0, db);
}

/* FASTA text of synthetic reads [first, first+n): ">r<i>\n<seq>\n" */
int ccbo_synth_reads_fasta(const char *path, uint64_t reads_seed, uint64_t db_seed, uint64_t first,
                           uint32_t n, uint32_t read_len, uint32_t n_genomes, uint64_t genome_len)
{
    FILE *f = fopen(path, "wb");
    if (!f) return -1;
    char *line = malloc(read_len + 2);
    for (uint32_t i = 0; i < n; i++) {
        ccb_read_plan p = ccb_plan_read(reads_seed, first + i, n_genomes, genome_len, read_len);
        for (uint32_t j = 0; j < read_len; j++) line[j] = ccb_code_char(ccb_read_code(&p, db_seed, j, read_len));
        line[read_len] = '\n';
        fprintf(f, ">r%lu\n", (unsigned long)(first + i));
        fwrite(line, 1, read_len + 1, f);
    }
    free(line); fclose(f);
    return 0;
}

/* FASTA text of the reads of a ccb_workload (include/ccb_synth.h: pairs joined by N, long reads with indels) */
int ccbo_synth2_reads_fasta(const char *path, const ccb_workload *w, uint64_t reads_seed, uint64_t db_seed, uint64_t first,
                            uint32_t n, uint32_t n_genomes, uint64_t genome_len)
{
    FILE *f = fopen(path, "wb");
    if (!f) return -1;
    char *line = malloc(ccb_synth2_max_len(w) + 2);
    for (uint32_t i = 0; i < n; i++) {
        uint32_t len = ccb_synth2_text(w, reads_seed, db_seed, first + i, n_genomes, genome_len, line);
        line[len] = '\n';
        fprintf(f, ">r%lu\n", (unsigned long)(first + i));
        fwrite(line, 1, len + 1, f);
    }
    free(line); fclose(f);
    return 0;
}
