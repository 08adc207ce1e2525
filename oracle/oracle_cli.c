/*
 * oracle_cli.c -- command-line driver of the CPU ORACLE (test infrastructure).
 *
 * Mirrors the flag set of the reference classifier (src/main.cc:103-212) for
 * the subset the oracle restates, so a test can run
 *     oracle_cli [--light] -k 31 -T targets.txt -D db/ -O reads.fa -R out
 * next to the reference binary and diff the CSVs.  Extra flags:
 *     --htsize N     override HTSIZE (tests use small tables)
 *     --dump FILE    also write the raw u16[5]-per-read results
 *     --build-only   stop after the database files exist
 *     --time         print lookups/s of the query loop to stderr
 */
#define _GNU_SOURCE
#include "cuclark_oracle.h"
#include <fcntl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

static double now_s(void)
{
    struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

static int file_exists(const char *p) { FILE *f = fopen(p, "r"); if (!f) return 0; fclose(f); return 1; }

int main(int argc, char **argv)
{
    int k = 31, light = 0, extended = 0, build_only = 0, threads = 1, do_time = 0;
    unsigned gap = 0, min_t = 0, sampling = 1;
    uint64_t htsize = 0;
    const char *targets = NULL, *folder = NULL, *obj = NULL, *obj2 = NULL, *res = NULL, *dump = NULL;
    for (int i = 1; i < argc; i++) {
        const char *a = argv[i];
        if (!strcmp(a, "-k") && i + 1 < argc) k = atoi(argv[++i]);
        else if (!strcmp(a, "-t") && i + 1 < argc) min_t = (unsigned)atoi(argv[++i]);
        else if (!strcmp(a, "-T") && i + 1 < argc) targets = argv[++i];
        else if (!strcmp(a, "-D") && i + 1 < argc) folder = argv[++i];
        else if (!strcmp(a, "-O") && i + 1 < argc) obj = argv[++i];
        else if (!strcmp(a, "-P") && i + 2 < argc) { obj = argv[++i]; obj2 = argv[++i]; }
        else if (!strcmp(a, "-R") && i + 1 < argc) res = argv[++i];
        else if (!strcmp(a, "-n") && i + 1 < argc) threads = atoi(argv[++i]);
        else if (!strcmp(a, "-b") && i + 1 < argc) ++i;
        else if (!strcmp(a, "-d") && i + 1 < argc) ++i;
        else if (!strcmp(a, "-g") && i + 1 < argc) gap = (unsigned)atoi(argv[++i]);
        else if (!strcmp(a, "-s") && i + 1 < argc) sampling = (unsigned)atoi(argv[++i]);
        else if (!strcmp(a, "--light")) light = 1;
        else if (!strcmp(a, "--extended")) extended = 1;
        else if (!strcmp(a, "--htsize") && i + 1 < argc) htsize = strtoull(argv[++i], NULL, 10);
        else if (!strcmp(a, "--dump") && i + 1 < argc) dump = argv[++i];
        else if (!strcmp(a, "--build-only")) build_only = 1;
        else if (!strcmp(a, "--time")) do_time = 1;
        else { fprintf(stderr, "Failed to recognize option: %s\n", a); return 1; }
    }
    /* src/main.cc:214-226 */
    if (light) { if (gap == 0) gap = 4; k = 27; sampling = 1; } else gap = 0;
    if (!htsize) htsize = light ? CCBO_HTSIZE_LIGHT : CCBO_HTSIZE_FULL;
    if (!targets || !folder || (!build_only && (!obj || !res))) {
        fprintf(stderr, "usage: oracle_cli [--light] -k K -T targets -D dir -O reads -R out\n");
        return 1;
    }
    ccbo_targets tg;
    if (ccbo_targets_read(targets, &tg)) { fprintf(stderr, "Failed to open targets data in file: %s\n", targets); return 1; }

    char fol[4096]; snprintf(fol, sizeof fol, "%s", folder);
    size_t fl = strlen(fol);
    if (fl && fol[fl - 1] != '/') { fol[fl] = '/'; fol[fl + 1] = 0; }   /* main.cc:245-247 */
    char prefix[4400], p[4500];
    ccbo_db_name(prefix, sizeof prefix, fol, k, tg.n_labels, htsize, min_t, light, gap);
    int key_bytes = ccbo_key_bytes(k, htsize);

    snprintf(p, sizeof p, "%s.sz", prefix);
    if (!file_exists(p)) {
        ccbo_db db;
        fprintf(stderr, "oracle: building database %s\n", prefix);
        int rc = ccbo_db_build(&tg, k, htsize, light, gap, min_t, &db);
        if (rc) { fprintf(stderr, "oracle: db build failed (%d)\n", rc); return 1; }
        if (ccbo_db_write(&db, prefix)) { fprintf(stderr, "oracle: db write failed\n"); return 1; }
        fprintf(stderr, "%lu %d-mers successfully stored in database.\n", (unsigned long)db.n, k);
        ccbo_db_free(&db);
    }
    if (build_only) return 0;

    ccbo_db db;
    if (ccbo_db_read(prefix, htsize, key_bytes, k, sampling, &db)) { fprintf(stderr, "Failed to find the database.\n"); return 1; }

    char merged[4400]; const char *in = obj; int paired = 0;
    if (obj2) {
        snprintf(merged, sizeof merged, "%s_ConcatenatedByCLARK.fa", obj);   /* CuCLARK_hh.hh:445 */
        if (ccbo_merge_paired(obj, obj2, merged)) { fprintf(stderr, "oracle: paired merge failed\n"); return 1; }
        in = merged; paired = 1;
    }
    int fd = open(in, O_RDONLY);
    struct stat st;
    if (fd < 0 || fstat(fd, &st) || st.st_size == 0) { fprintf(stderr, "Failed to open %s\n", in); return 1; }
    const unsigned char *map = mmap(0, (size_t)st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
    if (map == MAP_FAILED) { fprintf(stderr, "Failed to mmapping the file.\n"); return 1; }

    ccbo_reads rd;
    if (ccbo_index_reads(map, (size_t)st.st_size, &rd)) { fprintf(stderr, "Failed to recognize the format of the file.\n"); return 255; }
    size_t cap = (size_t)st.st_size / 4 + 16 * rd.n + 64;
    uint32_t *off = malloc(sizeof(uint32_t) * (rd.n + 1));
    uint16_t *cont = malloc(sizeof(uint16_t) * cap);
    int64_t nc = ccbo_pack(map, &rd, 0, rd.n, k, off, cont, cap);
    if (nc < 0) { fprintf(stderr, "oracle: pack failed\n"); return 1; }
    int max_hits = light ? CCBO_MAXHITS_LIGHT : CCBO_MAXHITS_FULL;
    int row_stride = 2 * max_hits + 2;                                       /* CuCLARK_hh.hh:1586-1589 */
    if (extended && row_stride < 2 * tg.n_labels + 2) row_stride = 2 * tg.n_labels + 2; /* oracle never truncates */
    uint16_t *out5 = calloc(rd.n * 5 + 1, 2);
    uint16_t *rows = extended ? calloc(rd.n * (size_t)row_stride + 1, 2) : NULL;
    uint64_t lookups = 0;
    double t0 = now_s();
    ccbo_query_packed(&db, k, off, cont, rd.n, tg.n_labels, out5, rows, row_stride, &lookups, threads);
    double t1 = now_s();
    if (do_time)
        fprintf(stderr, "oracle: %lu reads, %lu lookups, %.3f s, %.3e lookups/s (%d thread(s))\n",
                (unsigned long)rd.n, (unsigned long)lookups, t1 - t0, lookups / (t1 - t0), threads);

    char **names = malloc(sizeof(char *) * (size_t)(tg.n_labels + 1));
    names[0] = "NA";                                                          /* CuCLARK_hh.hh:1874 */
    for (int i = 0; i < tg.n_labels; i++) names[i + 1] = tg.labels[i];
    snprintf(p, sizeof p, "%s.csv", res);                                     /* :541-543 */
    if (ccbo_write_csv(p, map, &rd, k, paired, extended, tg.n_labels, names, out5, rows, row_stride)) {
        fprintf(stderr, "Failed to create/open file result: %s\n", p); return 1;
    }
    if (dump) { FILE *f = fopen(dump, "wb"); fwrite(out5, 2, rd.n * 5, f); fclose(f); }
    if (obj2) remove(merged);
    return 0;
}
