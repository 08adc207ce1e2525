/*
 * ref_harness.cc -- the REFERENCE's own CPU lookup, driven from the command
 * line (test infrastructure, NOT product code).
 *
 * Compiled by oracle/Makefile against the reference headers where they lie
 * (-I/root/reference/src): the database is loaded with the reference's
 * EHashtable::Read (src/HashTableStorage_hh.hh:134-140 ->
 * src/hashTable_hh.hh:549-829) and every k-mer is probed with the reference's
 * EHashtable::queryElement(uint64,ILBL&) (src/HashTableStorage_hh.hh:123-126
 * -> src/hashTable_hh.hh:359-396).  The reference has no CPU classify loop
 * (SURVEY.md section 8, C2), so read parsing, part splitting and the vote come
 * from the oracle restatement (cuclark_oracle.c) through its probe callback.
 * HTSIZE is a compile-time constant of the reference, so there are two
 * binaries: ref_harness (full) and ref_harness-l (light parameters).
 *
 * Used (a) to validate oracle A4/A5/A9 against reference code and (b) as the
 * `--impl reference` / cpu_baseline("reference") timing arm of bench.py.
 */
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#include "HashTableStorage_hh.hh"   /* reference, via -I */
#include "cuclark_oracle.h"

static double now_s()
{
    struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

template <typename T> static int find_cb(void *ctx, uint64_t x, uint16_t *label)
{
    EHashtable<T, rElement> *ht = (EHashtable<T, rElement> *)ctx;
    ILBL l = 0;
    if (ht->queryElement(x, l)) { *label = l; return 1; }
    return 0;
}

struct Args {
    int k = 31, threads = 1, extended = 0, light = 0, reps = 1;
    unsigned gap = 0, min_t = 0, sampling = 1;
    const char *targets = 0, *folder = 0, *obj = 0, *obj2 = 0, *res = 0, *dump = 0;
};

template <typename T>
static int run(const Args &a, const ccbo_targets &tg, const char *prefix)
{
    std::vector<std::string> labels, labels_c;
    for (int i = 0; i < tg.n_labels; i++) labels.push_back(tg.labels[i]);
    double t0 = now_s();
    EHashtable<T, rElement> *ht = new EHashtable<T, rElement>(a.k, labels, labels_c);
    size_t fsize = 0;
    if (!ht->Read(prefix, fsize, a.threads, a.sampling, false)) { fprintf(stderr, "Failed to find the database.\n"); return 1; }
    double t1 = now_s();
    fprintf(stderr, "ref_harness: database loaded in %.1f s (HTSIZE=%lu)\n", t1 - t0, (unsigned long)HTSIZE);

    char merged[4400]; const char *in = a.obj; int paired = 0;
    if (a.obj2) {
        snprintf(merged, sizeof merged, "%s_ConcatenatedByCLARK.fa", a.obj);
        mergePairedFiles(a.obj, a.obj2, merged);        /* reference src/file.cc:205-268 */
        in = merged; paired = 1;
    }
    int fd = open(in, O_RDONLY);
    struct stat st;
    if (fd < 0 || fstat(fd, &st) || st.st_size == 0) { fprintf(stderr, "Failed to open %s\n", in); return 1; }
    const unsigned char *map = (const unsigned char *)mmap(0, (size_t)st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
    ccbo_reads rd;
    if (ccbo_index_reads(map, (size_t)st.st_size, &rd)) { fprintf(stderr, "Failed to recognize the format of the file.\n"); return 255; }
    size_t cap = (size_t)st.st_size / 4 + 16 * rd.n + 64;
    uint32_t *off = (uint32_t *)malloc(sizeof(uint32_t) * (rd.n + 1));
    uint16_t *cont = (uint16_t *)malloc(sizeof(uint16_t) * cap);
    if (ccbo_pack(map, &rd, 0, rd.n, a.k, off, cont, cap) < 0) { fprintf(stderr, "pack failed\n"); return 1; }
    int row_stride = 2 * MAXHITS + 2;
    if (a.extended && row_stride < 2 * tg.n_labels + 2) row_stride = 2 * tg.n_labels + 2;
    uint16_t *out5 = (uint16_t *)calloc(rd.n * 5 + 1, 2);
    uint16_t *rows = a.extended ? (uint16_t *)calloc(rd.n * (size_t)row_stride + 1, 2) : 0;
    uint64_t lookups = 0; double best = 1e30;
    for (int rep = 0; rep < a.reps; rep++) {
        double q0 = now_s();
        ccbo_query_packed_cb(find_cb<T>, ht, a.k, off, cont, rd.n, out5, rows, row_stride, &lookups, a.threads);
        double q1 = now_s();
        if (q1 - q0 < best) best = q1 - q0;
        /* machine-readable line consumed by bench.py */
        printf("REFTIME reads=%lu lookups=%lu seconds=%.6f threads=%d\n",
               (unsigned long)rd.n, (unsigned long)lookups, q1 - q0, a.threads);
    }
    fflush(stdout);
    if (a.res) {
        std::vector<char *> names(tg.n_labels + 1);
        names[0] = (char *)"NA";
        for (int i = 0; i < tg.n_labels; i++) names[i + 1] = tg.labels[i];
        std::string p = std::string(a.res) + ".csv";
        ccbo_write_csv(p.c_str(), map, &rd, a.k, paired, a.extended, tg.n_labels, names.data(), out5, rows, row_stride);
    }
    if (a.dump) { FILE *f = fopen(a.dump, "wb"); fwrite(out5, 2, rd.n * 5, f); fclose(f); }
    if (a.obj2) remove(merged);
    return 0;
}

int main(int argc, char **argv)
{
    Args a;
    for (int i = 1; i < argc; i++) {
        const char *s = argv[i];
        if (!strcmp(s, "-k") && i + 1 < argc) a.k = atoi(argv[++i]);
        else if (!strcmp(s, "-t") && i + 1 < argc) a.min_t = (unsigned)atoi(argv[++i]);
        else if (!strcmp(s, "-T") && i + 1 < argc) a.targets = argv[++i];
        else if (!strcmp(s, "-D") && i + 1 < argc) a.folder = argv[++i];
        else if (!strcmp(s, "-O") && i + 1 < argc) a.obj = argv[++i];
        else if (!strcmp(s, "-P") && i + 2 < argc) { a.obj = argv[++i]; a.obj2 = argv[++i]; }
        else if (!strcmp(s, "-R") && i + 1 < argc) a.res = argv[++i];
        else if (!strcmp(s, "-n") && i + 1 < argc) a.threads = atoi(argv[++i]);
        else if (!strcmp(s, "-g") && i + 1 < argc) a.gap = (unsigned)atoi(argv[++i]);
        else if (!strcmp(s, "-s") && i + 1 < argc) a.sampling = (unsigned)atoi(argv[++i]);
        else if (!strcmp(s, "--extended")) a.extended = 1;
        else if (!strcmp(s, "--dump") && i + 1 < argc) a.dump = argv[++i];
        else if (!strcmp(s, "--reps") && i + 1 < argc) a.reps = atoi(argv[++i]);
        else { fprintf(stderr, "Failed to recognize option: %s\n", s); return 1; }
    }
    a.light = (HTSIZE == LHTSIZE);                       /* src/main.cc:214-226 */
    if (a.light) { if (!a.gap) a.gap = 4; a.k = 27; a.sampling = 1; } else a.gap = 0;
    if (!a.targets || !a.folder || !a.obj) { fprintf(stderr, "usage: ref_harness -k K -T targets -D dir -O reads [-R out] [--dump f] [-n threads]\n"); return 1; }
    ccbo_targets tg;
    if (ccbo_targets_read(a.targets, &tg)) { fprintf(stderr, "Failed to open targets data in file: %s\n", a.targets); return 1; }
    char fol[4096]; snprintf(fol, sizeof fol, "%s", a.folder);
    size_t fl = strlen(fol);
    if (fl && fol[fl - 1] != '/') { fol[fl] = '/'; fol[fl + 1] = 0; }
    char prefix[4400];
    ccbo_db_name(prefix, sizeof prefix, fol, a.k, tg.n_labels, HTSIZE, a.min_t, a.light, a.gap);
    int kb = ccbo_key_bytes(a.k, HTSIZE);
    if (kb == 2) return run<T16>(a, tg, prefix);
    if (kb == 4) return run<T32>(a, tg, prefix);
    return run<T64>(a, tg, prefix);
}
