// cli_main.cpp -- bin/cuCLARK and bin/cuCLARK-l (compiled with -DCCB_LIGHT).
//
// Drop-in for the reference classifier binaries at the process boundary
// (SURVEY.md section 8, B1): the flag set, defaults, messages and exit codes of
// reference src/main.cc:71-290, the file handling of CuCLARK::run / runSimple
// (src/CuCLARK_hh.hh:384-574) and the result file <R>.csv.  scripts/
// classify_metagenome.sh and `kent -c` call it unchanged.  `-m <n>` is accepted
// and ignored (north-star text; the reference has no such flag).
#include "../../include/cuclark_b200.h"
#include "hostpipe.hpp"

#include <fcntl.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/time.h>
#include <unistd.h>
#include <zlib.h>

#include <algorithm>
#include <atomic>
#include <mutex>
#include <iostream>
#include <string>
#include <thread>
#include <vector>

using std::cerr;
using std::cout;
using std::endl;
using std::string;
using std::vector;

// src/parameters.hh:37-48 / src/parameters_light_hh:37-49
#ifdef CCB_LIGHT
static const uint64_t HTSIZE = 57777779ull;
static const int MAXHITS = 23;
static const bool kLight = true;
#else
static const uint64_t HTSIZE = 1610612741ull;
static const int MAXHITS = 15;
static const bool kLight = false;
#endif
static const int MAXK = 32;
static const unsigned SFACTORMAX = 30;
#define CCB_CLI_VERSION "1.1.3-b200"

static void printUsage()
{   // option list of src/main.cc:38-68
    cout << "\n";
    cout << (kLight ? "cuCLARK-l" : "cuCLARK") << " -- B200 build: classification of metagenomic reads on CUDA sm_100a\n\n";
    cout << "./" << (kLight ? "cuCLARK-l" : "cuCLARK") << " -k <kmerSize> -T <fileTargets> -D <directoryDB/> -O <fileObjects> -R <fileResults> -n <numberofthreads> ...\n\n";
    cout << "Definitions of parameters (cf. README file, for details):\n\n";
    cout << "-k <kmerSize>,       \t k-mer length:\tinteger, >= 2 and <= 32 (in version 1.*).\n";
    cout << "-t <minFreqTarget>,  \t minimum of k-mer frequency/occurence for the discriminative k-mers:  integer, >=0.\n";
    cout << "-T <fileTargets>,    \t filename of the targets definition: text.\n";
    cout << "-D <directoryDB/>,   \t directory of the database : text.\n";
    cout << "-O <fileObjects>,    \t file containing objects: text.\n";
    cout << "-P <file1> <file2>,  \t Paired-end fastq files: texts.\n";
    cout << "-R <fileResults>,    \t file to store results:  text. Results are stored in CSV format in the file <fileResults>.csv (the extension \".csv\" is automatically added to the filename).\n";
    cout << "-n <numberofthreads>,\t number of threads:\tinteger >= 1.\n";
    cout << "-b <numberofbatches>,\t number of batches:\tinteger >= number of threads.\n";
    cout << "-d <numberofdevices>,\t number of CUDA devices to use:\tinteger >= 1.\n";
    cout << "--tsk,               \t to request a detailed creation of the database (target specific k-mers files).\n";
    cout << "-g <iteration>,      \t gap or number of non-overlapping k-mers to pass for the database creation (for CuCLARK-l only). The default value is 4.\n";
    cout << "-s <factor>,         \t sampling factor value (for CuCLARK only).\n";
    cout << "--extended,          \t extended output (one column of hit counts per target).\n";
    cout << "--verbose,           \t progress messages.\n";
    cout << "--gzipped,           \t the objects are gzip files (inflated in memory; classify_metagenome.sh strips the flag and gunzips itself).\n";
    cout << "--sharded,           \t with -d N: ONE copy of the database, cut into N shards over the N devices (k-mers travel over NVLink),\n";
    cout << "                     \t instead of N copies; chosen automatically when a copy does not fit a device's memory.\n";
    cout << "\n";
    cout << "--help,              \t to print help/options.\n";
    cout << "--version,           \t to print the version info.\n";
    cout << endl;
}

static bool validFile(const char* f)
{   // src/file.cc:275-283
    FILE* fd = fopen(f, "r");
    if (!fd) return false;
    fclose(fd);
    return true;
}

static bool read_line(FILE* f, string& out)
{   // getLineFromFile (src/file.cc:121-139): one trailing '\n' removed
    char* line = NULL; size_t cap = 0;
    ssize_t n = getline(&line, &cap, f);
    if (n < 0) { free(line); out.clear(); return false; }
    out.assign(line, (size_t)n);
    free(line);
    if (!out.empty() && out.back() == '\n') out.pop_back();
    return true;
}

static void split(const string& s, const char* seps, size_t max_el, vector<string>& out)
{
    out.clear();
    size_t i = 0;
    while (i < s.size() && out.size() < max_el) {
        while (i < s.size() && strchr(seps, s[i]) && s[i]) i++;
        size_t j = i;
        while (j < s.size() && !(strchr(seps, s[j]) && s[j])) j++;
        if (j > i) out.push_back(s.substr(i, j - i));
        i = j;
    }
}

struct Targets {
    vector<string> names;          // [0] = "NA", then labels in order of first appearance (src/CuCLARK_hh.hh:1874-1878)
    size_t n_labels = 0;
    vector<std::pair<string, uint16_t>> files;   // target file -> 0-based label index (m_targetsID, :1830-1840)
};

// getTargetsData (src/CuCLARK_hh.hh:1790-1901): "<file> <label> [<label_c>]" per line
static void read_targets(const char* path, Targets& t)
{
    FILE* f = fopen(path, "r");
    if (!f) { cerr << "Failed to open targets data in file: " << path << endl; exit(-1); }
    string line; vector<string> el;
    while (read_line(f, line)) {            // every listed file must exist (:1800-1811)
        split(line, " \t\n\r", 1, el);
        if (el.empty()) continue;
        if (!validFile(el[0].c_str())) { cerr << "Failed to open file: " << el[0] << " defined in " << path << endl; exit(-1); }
    }
    rewind(f);
    vector<string> labels, labels_c;
    while (read_line(f, line)) {
        split(line, " ,\n\t\r", 3, el);
        if (el.empty()) continue;
        if (el.size() < 2) { cerr << " Missing label for " << el[0] << endl; exit(-1); }
        if (std::find(labels.begin(), labels.end(), el[1]) == labels.end()) labels.push_back(el[1]);
        t.files.emplace_back(el[0], (uint16_t)(std::find(labels.begin(), labels.end(), el[1]) - labels.begin()));
        if (el.size() > 2 && std::find(labels_c.begin(), labels_c.end(), el[2]) == labels_c.end()) labels_c.push_back(el[2]);
    }
    fclose(f);
    if (!labels_c.empty()) t.files.clear();       // a third column (label_c) is outside the rebuilt path: never build from it
    t.names.clear();
    t.names.push_back("NA");
    for (auto& l : labels) t.names.push_back(l);
    for (auto& l : labels_c) t.names.push_back(l);
    t.n_labels = labels.size() + labels_c.size();
}

struct Cli {
    size_t k = 31, cpu = 1, iterKmers = 0, batches = 1, devices = 0;
    unsigned minT = 0, sfactor = 1;
    bool tsk = false, ext = false, verbose = false, sharded = false, gzipped = false;
    const char *targets = NULL, *folder = NULL, *objects = NULL, *objects2 = NULL, *results = NULL;
};

struct Engine {
    Cli cli;
    Targets tg;
    vector<ccb_handle*> devs;
    int key_bytes = 4;
    string dbname;

    bool sharded = false;          // one table cut over the devices (ccb_group_create) instead of a copy per device

    // k-mers of the database on disk (0 when it is still to be built)
    uint64_t db_kmers() const
    {
        struct stat sb;
        return stat((db_prefix() + ".ky").c_str(), &sb) == 0 ? (uint64_t)sb.st_size / (uint64_t)key_bytes : 0;
    }

    void open_devices()
    {
        int ndev = ccb_device_count();
        if (ndev <= 0) { cerr << "No CUDA devices found. Abort.\n"; exit(1); }                       // src/CuClarkDB.cu:126-129
        // CCB_SHARD_SAME_DEVICE=1 (tests on a one-GPU box): -d N --sharded may list a device more than once
        const bool same_dev = cli.sharded && getenv("CCB_SHARD_SAME_DEVICE") && (size_t)ndev < cli.devices;
        if ((size_t)ndev < cli.devices && !same_dev) { cerr << cli.devices << " CUDA devices requested. Insufficient devices found. Abort.\n"; exit(1); }
        const int have = ndev;
        if (cli.devices > 0) { cerr << "Using " << cli.devices << " CUDA devices as requested.\n"; ndev = (int)cli.devices; }
        const int workers = (int)std::max<size_t>(cli.cpu, (size_t)ndev);
        // The reference cuts a database that does not fit one device into parts and swaps them through the GPUs
        // (src/CuClarkDB.cu:531-559, 776-815).  Here: one shard per device, all resident (--sharded, or by itself when a
        // copy of the table -- 16 bytes per k-mer at the default fill -- would not fit beside the batch buffers).
        sharded = cli.sharded && ndev > 1;
        if (!sharded && ndev > 1 && cli.sfactor <= 1) {
            uint64_t fr = 0, tot = 0;
            const uint64_t need = db_kmers() * 16ull + (4ull << 30);
            if (ccb_device_memory(0, &fr, &tot) == 0 && need > fr) {
                sharded = true;
                cerr << "The database does not fit one device (" << need / 1000000000.0 << " GB needed, " << fr / 1000000000.0
                     << " GB free): sharding it over the " << ndev << " devices.\n";
            }
        }
        if (sharded) {
            vector<int> dv;
            for (int d = 0; d < ndev; d++) dv.push_back(d % have);
            ccb_config c;
            memset(&c, 0, sizeof c);
            c.k = (int)cli.k; c.n_targets = (int)std::max<size_t>(tg.n_labels, 1); c.max_hits = MAXHITS;
            c.n_slots = workers + ndev; c.load_factor = 0; c.verbose = cli.verbose;
            ccb_handle* h = NULL;
            if (ccb_group_create(dv.data(), ndev, &c, &h)) { cerr << ccb_last_error(NULL) << endl; exit(1); }
            if (cli.verbose) cerr << "Database sharded over " << ndev << " devices (k-mer exchange over peer memory).\n";
            devs.push_back(h);
            return;
        }
        for (int d = 0; d < ndev; d++) {
            ccb_config c;
            memset(&c, 0, sizeof c);
            c.device = d; c.k = (int)cli.k; c.n_targets = (int)std::max<size_t>(tg.n_labels, 1); c.max_hits = MAXHITS;
            c.n_slots = (workers + ndev - 1) / ndev + 1; c.load_factor = 0; c.verbose = cli.verbose;
            ccb_handle* h = NULL;
            if (ccb_create(&c, &h)) { cerr << ccb_last_error(NULL) << endl; exit(1); }
            devs.push_back(h);
        }
    }

    string db_prefix() const
    {   // getdbName (src/CuCLARK_hh.hh:581-592)
        char name[4096];
        if (kLight) snprintf(name, sizeof name, "%s/db_central_k%lu_t%lu_s%lu_m%lu_light_%lu.tsk", cli.folder, cli.k, tg.n_labels, (size_t)HTSIZE, (size_t)cli.minT, cli.iterKmers);
        else        snprintf(name, sizeof name, "%s/db_central_k%lu_t%lu_s%lu_m%lu.tsk", cli.folder, cli.k, tg.n_labels, (size_t)HTSIZE, (size_t)cli.minT);
        return name;
    }

    void load_db()
    {   // loadSpecificTargetSets (src/CuCLARK_hh.hh:599-632)
        dbname = db_prefix();
        bool present = true;
        for (const char* e : {".sz", ".ky", ".lb"}) present = present && validFile((dbname + e).c_str());
        if (!present && build_db()) return;          // built on the device and already resident
        if (cli.verbose) cerr << "Loading database [" << dbname << ".*] (s=" << cli.sfactor << ")..." << endl;
        struct timeval t0, t1;
        gettimeofday(&t0, NULL);
        vector<std::thread> th; vector<int> rc(devs.size(), 0);
        for (size_t d = 0; d < devs.size(); d++)
            th.emplace_back([&, d] { uint64_t n = 0; rc[d] = ccb_db_load_files(devs[d], dbname.c_str(), HTSIZE, key_bytes, cli.sfactor, &n); });
        for (auto& t : th) t.join();
        for (size_t d = 0; d < devs.size(); d++)
            if (rc[d]) { cerr << ccb_last_error(devs[d]) << endl; exit(1); }
        if (cli.verbose) { gettimeofday(&t1, NULL); cerr << "Loading time: " << (t1.tv_sec - t0.tv_sec) + (t1.tv_usec - t0.tv_usec) / 1e6 << " s\n"; }
    }

    // makeSpecificTargetSets (src/CuCLARK_hh.hh:691-1329): the database is absent, build it from the
    // target FASTA files on device 0, write <dbname>.sz/.ky/.lb, then load it everywhere.
    // returns true when the table is already resident (one device, no -s sampling).
    bool build_db()
    {
        if (tg.files.empty()) { cerr << "Failed to find the database." << endl; exit(-1); }
        cerr << "Starting the creation of the database of targets specific " << cli.k << "-mers from input files..." << endl;
        uint64_t total = 0;
        for (auto& f : tg.files) { struct stat sb; if (stat(f.first.c_str(), &sb) == 0) total += (uint64_t)sb.st_size; }
        ccb_handle* h = devs[0];
        if (ccb_build_begin(h, HTSIZE, key_bytes, kLight ? 1 : 0, (unsigned)cli.iterKmers, cli.minT, total)) { cerr << ccb_last_error(h) << endl; exit(1); }
        std::atomic<size_t> next{0}, done{0};
        std::mutex io;
        bool failed = false;
        auto work = [&] {
            for (size_t t = next++; t < tg.files.size(); t = next++) {
                const string& fn = tg.files[t].first;
                int fd = open(fn.c_str(), O_RDONLY);
                struct stat sb;
                if (fd == -1 || fstat(fd, &sb) == -1) { std::lock_guard<std::mutex> g(io); cerr << "Failed to open " << fn << endl; if (fd != -1) close(fd); continue; }
                if (sb.st_size == 0) { close(fd); continue; }
                void* map = mmap(0, (size_t)sb.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
                if (map == MAP_FAILED) { close(fd); std::lock_guard<std::mutex> g(io); cerr << "Failed to open " << fn << endl; continue; }
                int rc = ccb_build_add_fasta(h, (const unsigned char*)map, (uint64_t)sb.st_size, tg.files[t].second);
                munmap(map, (size_t)sb.st_size); close(fd);
                std::lock_guard<std::mutex> g(io);
                if (rc) { cerr << fn << ": " << ccb_last_error(h) << endl; failed = true; }
                cerr << "\r Progress report: (" << ++done << "/" << tg.files.size() << ")              ";
            }
        };
        vector<std::thread> th;
        for (size_t i = 0; i < std::max<size_t>(1, std::min<size_t>(cli.cpu, tg.files.size())); i++) th.emplace_back(work);
        for (auto& t : th) t.join();
        if (failed) exit(1);
        const bool resident = devs.size() == 1 && cli.sfactor <= 1 && !sharded;
        ccb_build_info bi;
        memset(&bi, 0, sizeof bi);
        cerr << endl << "Creating database in disk..." << endl;
        if (ccb_build_end(h, dbname.c_str(), resident ? 1 : 0, &bi)) { cerr << ccb_last_error(h) << endl; exit(1); }
        cerr << bi.bases << " nt read in total." << endl;
        cerr << "Mother Hashtable successfully built. " << bi.distinct << " " << cli.k << "-mers stored." << endl;
        cerr << bi.stored << " " << cli.k << "-mers successfully stored in database." << endl;
        return resident;
    }

    // The anchor index (no counterpart in the reference, results are unchanged): the target sequences are kept on
    // every device so that read windows which match a target base by base are answered without a table probe
    // (include/cuclark_b200.h "Anchor index").  On by default whenever the target FASTA files are there to build it from
    // and fit (at most 4 Gbp of text per device, memory permitting; not with a sharded table); CCB_ANCHOR=0 turns it
    // off, CCB_ANCHOR=1 reports why it could not be built.  Any failure only costs the speed-up.
    void build_anchor()
    {
        const char* env = getenv("CCB_ANCHOR");
        if ((env && *env == '0') || tg.files.empty() || sharded) return;
        const bool loud = (env && *env && *env != '0') || cli.verbose;
        {   // room for it beside the table and the batch slots?  (text 1.25 B + sampled k-mers 2 B + build scratch 1 B per base)
            uint64_t total_b = 0, fr = 0, tot = 0;
            for (auto& f : tg.files) { struct stat sb; if (stat(f.first.c_str(), &sb) == 0) total_b += (uint64_t)sb.st_size; }
            for (size_t d = 0; d < devs.size(); d++)
                if (ccb_device_memory((int)d, &fr, &tot) == 0 && total_b * 9 / 2 + (6ull << 30) > fr) {
                    if (loud) cerr << "Anchor index not built: not enough device memory left beside the database." << endl;
                    return;
                }
        }
        struct timeval t0, t1;
        gettimeofday(&t0, NULL);
        uint64_t total = 0;
        for (auto& f : tg.files) { struct stat sb; if (stat(f.first.c_str(), &sb) == 0) total += (uint64_t)sb.st_size; }
        for (ccb_handle* h : devs)
            if (ccb_anchor_begin(h, total)) {
                if (loud) cerr << "Anchor index not built: " << ccb_last_error(h) << endl;
                for (ccb_handle* x : devs) ccb_anchor_drop(x);
                return;
            }
        std::atomic<size_t> next{0};
        std::atomic<bool> failed{false};
        auto work = [&] {
            for (size_t t = next++; t < tg.files.size() && !failed; t = next++) {
                const string& fn = tg.files[t].first;
                int fd = open(fn.c_str(), O_RDONLY);
                struct stat sb;
                if (fd == -1 || fstat(fd, &sb) == -1 || sb.st_size == 0) { if (fd != -1) close(fd); continue; }
                void* map = mmap(0, (size_t)sb.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
                if (map == MAP_FAILED) { close(fd); continue; }
                for (ccb_handle* h : devs)
                    if (ccb_anchor_add_fasta(h, (const unsigned char*)map, (uint64_t)sb.st_size, tg.files[t].second)) failed = true;
                munmap(map, (size_t)sb.st_size); close(fd);
            }
        };
        vector<std::thread> th;
        for (size_t i = 0; i < std::max<size_t>(1, std::min<size_t>(cli.cpu, tg.files.size())); i++) th.emplace_back(work);
        for (auto& t : th) t.join();
        for (ccb_handle* h : devs) {
            ccb_anchor_info ai;
            memset(&ai, 0, sizeof ai);
            if (failed) { ccb_anchor_end(h, NULL); ccb_anchor_drop(h); if (loud) cerr << "Anchor index not built: " << ccb_last_error(h) << endl; continue; }
            if (ccb_anchor_end(h, &ai)) { if (loud) cerr << "Anchor index not built: " << ccb_last_error(h) << endl; continue; }
            if (cli.verbose) {
                gettimeofday(&t1, NULL);
                cerr << "Anchor index: " << ai.bases << " bases, " << ai.entries << " sampled k-mers, " << ai.bytes / 1e9 << " GB, "
                     << (t1.tv_sec - t0.tv_sec) + (t1.tv_usec - t0.tv_usec) / 1e6 << " s" << endl;
            }
        }
    }

    // runSimple (src/CuCLARK_hh.hh:513-574) on a text already in memory or mapped
    void classify(const char* shown_name, const uint8_t* map, size_t nb, const char* result_prefix, bool paired)
    {
        string res = string(result_prefix) + ".csv";
        ccbh::ClassifyOptions o;
        o.k = (int)cli.k; o.threads = (int)cli.cpu; o.paired = paired; o.extended = cli.ext; o.verbose = cli.verbose;
        o.row_stride = 2 * MAXHITS + 2; o.dump_final5 = getenv("CCB_DUMP_FINAL5");
        ccbh::ClassifyStats st; string err;
        struct timeval a, b;
        gettimeofday(&a, NULL);
        cerr << (cli.ext ? "Writing extended results... " : "Writing results... ") << endl;
        int rc = ccbh::classify_text(devs, map, nb, res.c_str(), tg.names, o, st, err);
        if (rc) { cerr << err << endl; exit(rc == 255 ? -1 : 1); }
        cerr << "Done." << endl;
        gettimeofday(&b, NULL);
        double diff = (b.tv_sec - a.tv_sec) + (b.tv_usec - a.tv_usec) / 1000000.0;
        char buf[256];                                                                   // printSpeedStats (:1931-1939)
        snprintf(buf, sizeof buf, "Done in %.1fs (%zu reads/min, %zu reads)\n", diff, (size_t)(((double)st.reads) / diff * 60.0), (size_t)st.reads);
        cerr << buf << "Results: " << res << "\n";
        if (cli.verbose) {
            snprintf(buf, sizeof buf, "  host pipeline: index %.3f s, slots %.3f s, total %.3f s, %zu batches; worker seconds: index %.3f, pack %.3f, device wait %.3f, rows %.3f, writer %.3f\n",
                     st.seconds_index, st.seconds_alloc, st.seconds_total, (size_t)st.batches, st.cpu_index, st.cpu_pack, st.cpu_wait, st.cpu_format, st.cpu_sink);
            cerr << buf;
        }
        (void)shown_name;
    }

    // --gzipped: the whole file inflated in memory (zlib; concatenated members are fine, a plain text passes through).
    // The reference leaves this to classify_metagenome.sh, which copies the file, gunzips the copy and hands the binary
    // the text (scripts/classify_metagenome.sh:118-137); through the script nothing changes, directly it saves the copy.
    static bool inflate_file(const char* file, vector<char>& out)
    {
        gzFile g = gzopen(file, "rb");
        if (!g) return false;
        gzbuffer(g, 1u << 20);
        struct stat sb;
        out.clear();
        if (stat(file, &sb) == 0) out.reserve((size_t)sb.st_size * 4);
        size_t n = 0;
        for (;;) {
            if (out.size() < n + (8u << 20)) out.resize(std::max(out.size() * 2, n + (size_t)(16u << 20)));
            const int r = gzread(g, out.data() + n, (unsigned)std::min<size_t>(out.size() - n, 1u << 30));
            if (r < 0) { gzclose(g); return false; }
            if (r == 0) break;
            n += (size_t)r;
        }
        gzclose(g);
        out.resize(n);
        return true;
    }

    void run_file(const char* file, const char* result_prefix)
    {
        cerr << "Classifying: " << file << "\n";
        if (cli.gzipped) {
            vector<char> text;
            if (!inflate_file(file, text) || text.empty()) { cerr << "Failed to uncompress input objects." << endl; return; }
            classify(file, (const uint8_t*)text.data(), text.size(), result_prefix, false);
            return;
        }
        int fd = open(file, O_RDONLY);
        struct stat sb;
        if (fd == -1 || fstat(fd, &sb) != 0 || sb.st_size == 0) { cerr << "Failed to open " << file << endl; if (fd != -1) close(fd); return; }
        void* map = mmap(0, (size_t)sb.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
        if (map == MAP_FAILED) { close(fd); cerr << "Failed to mmapping the file." << endl; return; }
        madvise(map, (size_t)sb.st_size, MADV_SEQUENTIAL);
        classify(file, (const uint8_t*)map, (size_t)sb.st_size, result_prefix, false);
        munmap(map, (size_t)sb.st_size);
        close(fd);
    }

    void run_pair(const char* f1, const char* f2, const char* result_prefix)
    {   // the reference writes <f1>_ConcatenatedByCLARK.fa, classifies it as FASTA and deletes it
        // (src/CuCLARK_hh.hh:442-452); the same text is built in memory here.
        ccbh::TextBuf text; string err;
        if (cli.gzipped) {
            vector<char> a, b;
            if (!inflate_file(f1, a) || !inflate_file(f2, b) || a.empty() || b.empty()) { cerr << "Failed to uncompress input objects." << endl; exit(1); }
            if (ccbh::join_pairs_texts((const uint8_t*)a.data(), a.size(), (const uint8_t*)b.data(), b.size(), text, err, (int)cli.cpu)) { cerr << err << endl; exit(1); }
        } else if (ccbh::join_pairs(f1, f2, text, err, (int)cli.cpu)) { cerr << err << endl; exit(1); }
        cerr << "Classifying: " << f1 << "_ConcatenatedByCLARK.fa" << "\n";
        if (text.empty()) { cerr << "Failed to open " << f1 << "_ConcatenatedByCLARK.fa" << endl; return; }
        classify(f1, (const uint8_t*)text.data(), text.size(), result_prefix, true);
    }

    static bool is_reads_file(const char* path)
    {   // src/CuCLARK_hh.hh:398-416: first line starts with '>' or '@', or has exactly two tokens
        FILE* f = fopen(path, "r");
        string line;
        if (!f) return true;
        read_line(f, line);
        fclose(f);
        vector<string> el;
        split(line, " \t,", (size_t)-1, el);
        return (!line.empty() && (line[0] == '>' || line[0] == '@')) || el.size() == 2;
    }

    void run()
    {
        const bool results_exists = validFile(cli.results);
        if (!cli.objects2) {
            if (!results_exists || is_reads_file(cli.objects)) { run_file(cli.objects, cli.results); return; }
            FILE* rf = fopen(cli.results, "r"); FILE* of = fopen(cli.objects, "r");       // lists of files (:419-428)
            string o, r;
            while (read_line(of, o) && read_line(rf, r)) run_file(o.c_str(), r.c_str());
            fclose(rf); fclose(of);
            return;
        }
        if (!results_exists || is_reads_file(cli.objects)) { run_pair(cli.objects, cli.objects2, cli.results); return; }
        FILE* rf = fopen(cli.results, "r"); FILE* f1 = fopen(cli.objects, "r"); FILE* f2 = fopen(cli.objects2, "r");
        string a, b, r;
        while (read_line(f1, a) && read_line(f2, b) && read_line(rf, r)) run_pair(a.c_str(), b.c_str(), r.c_str());
        fclose(rf); fclose(f1); fclose(f2);
    }
};

int main(int argc, char** argv)
{
    if (argc == 2) {
        string val(argv[1]);
        if (val == "--help" || val == "--HELP") { printUsage(); return 0; }
        if (val == "--version" || val == "--VERSION") {
            cout << "Version: " << CCB_CLI_VERSION << " (" << ccb_version() << "; command line of cuCLARK 1.1, R. Kobus / CLARK 1.1.3, R. Ounit)" << endl;
            return 0;
        }
    }
    if (argc < 6) {
        cerr << "To run " << argv[0] << ", at least four  parameters are necessary:\n";
        cerr << "filename of the targets definition, directory of database, filename for objects, filename for results." << endl;
        printUsage();
        return -1;
    }
    Cli c;
    auto need = [&](int& i, const char* msg) { if (++i >= argc) { cerr << msg << endl; exit(1); } return argv[i]; };
    for (int i = 1; i < argc; i++) {
        string val(argv[i]);
        if (val == "-k") {
            c.k = (size_t)atoi(need(i, "Please specify the k-mer length!"));
            if (c.k <= 1 || c.k > (size_t)MAXK) { cerr << "The k-mer length should be in [2," << MAXK << "]." << endl; exit(1); }
            continue;
        }
        if (val == "-t") {
            c.minT = (unsigned)atoi(need(i, "Please specify the minimum frequency (targets)!"));
            if (c.minT >= 65536) { cerr << "The min k-mer frequency should be in [0,65535]." << endl; exit(1); }
            continue;
        }
        if (val == "-n") {
            c.cpu = (size_t)atoi(need(i, "Please specify the number of threads!"));
            if (c.batches < c.cpu) c.batches = c.cpu;
            if (c.cpu < 1) { cerr << "The number of threads should be higher than 0." << endl; exit(1); }
            continue;
        }
        if (val == "--tsk") { c.tsk = true; continue; }
        if (val == "--extended") { c.ext = true; continue; }
        if (val == "-T") {
            c.targets = need(i, "Please specify the targets!");
            if (!validFile(c.targets)) { cerr << "Failed to find/read the file of the targets definition: " << c.targets << endl; exit(1); }
            continue;
        }
        if (val == "-O") {
            c.objects = need(i, "Please specify the objects!");
            if (!validFile(c.objects)) { cerr << "Failed to find/read the filename of objects: " << c.objects << endl; exit(1); }
            continue;
        }
        if (val == "-P") {
            if (i + 2 >= argc) { cerr << "Please specify the paired-end reads!" << endl; exit(1); }
            c.objects = argv[++i]; c.objects2 = argv[++i];
            if (!validFile(c.objects)) { cerr << "Failed to find/read " << c.objects << endl; exit(1); }
            if (!validFile(c.objects2)) { cerr << "Failed to find/read " << c.objects2 << endl; exit(1); }
            continue;
        }
        if (val == "-D") {
            c.folder = need(i, "Please specify the database directory!");
            if (!validFile(c.folder)) { cerr << "Failed to find/read the directory:  " << c.folder << endl; exit(1); }
            continue;
        }
        if (val == "-R") { c.results = need(i, "Please specify where to store results!"); continue; }
        if (val == "-g") {
            c.iterKmers = (size_t)atoi(need(i, "Please specify a gap value!"));
            if (c.iterKmers < 4) { cerr << "The gap value should be >= 4." << endl; exit(1); }
            continue;
        }
        if (val == "-s") {
            c.sfactor = (unsigned)atoi(need(i, "Please specify a sampling factor value!"));
            if (c.sfactor < 2 || c.sfactor > SFACTORMAX) { cerr << "The sampling factor value should be in the interval [2," << SFACTORMAX << "]." << endl; exit(1); }
            continue;
        }
        if (val == "-b") {
            c.batches = (size_t)atoi(need(i, "Please specify the number of batches!"));
            if (c.batches < c.cpu) { cerr << "The number of batches should be higher than the number of threads." << endl; exit(1); }
            continue;
        }
        if (val == "-d") {
            c.devices = (size_t)atoi(need(i, "Please specify the number of devices to use!"));
            if (c.devices < 1) { cerr << "The number of devices should be higher than 0." << endl; exit(1); }
            continue;
        }
        if (val == "--verbose") { c.verbose = true; continue; }
        if (val == "--sharded") { c.sharded = true; continue; }
        // `kent -c -P f1 f2 ...` ends its command line with --light, and classify_metagenome.sh:104-137 forwards everything
        // after the second mate untouched (FILE2 is never reset): the reference's own cuCLARK-l then stops at
        // "Failed to recognize option: --light".  The light binary takes its own name as a no-op instead.
        if (val == "--light" && kLight) continue;
        if (val == "--gzipped") { c.gzipped = true; continue; }
        if (val == "-m") { need(i, "Please specify the mode!"); continue; }      // tolerated, ignored
        cerr << "Failed to recognize option: " << val << endl;
        exit(1);
    }
    if (kLight) { if (c.iterKmers == 0) c.iterKmers = 4; c.k = 27; c.sfactor = 1; }     // src/main.cc:214-222
    else c.iterKmers = 0;
    if (!c.targets || !c.folder || !c.objects || !c.results) {
        cerr << "Failed to run " << argv[0] << ": at least four  parameters are necessary";
        cerr << ": file of targets, directory of database, file of objects, file for results." << endl;
        printUsage();
        exit(1);
    }
    string folder(c.folder);
    if (folder[folder.size() - 1] != '/') folder.push_back('/');
    c.folder = folder.c_str();

    // key width by k (src/main.cc:251-286)
    const size_t t_b = (size_t)(log((double)HTSIZE) / log(4.0));
    Engine e;
    e.cli = c;
    e.key_bytes = c.k <= t_b + 8 ? 2 : (c.k <= t_b + 16 ? 4 : 8);
    read_targets(c.targets, e.tg);
    if (c.verbose) {                                                                       // CuCLARK::print (:1906-1926)
        if (c.minT > 0) cerr << "Minimum k-mers occurences in Targets is set to " << c.minT << endl;
        if (c.tsk) cerr << "Creation of targets specific k-mers files requested " << endl;
        if (kLight) cerr << "Using light database in RAM (" << c.iterKmers << ")" << endl;
        if (c.sfactor > 2) cerr << "Sampling factor is " << c.sfactor << endl;
    }
    e.open_devices();
    // a large reads file: pin the batch slots while the database loads (hostpipe.hpp reserve_slots)
    std::thread slots;
    {
        struct stat sb;
        if (stat(c.objects, &sb) == 0 && (uint64_t)sb.st_size >= (256ull << 20))
            slots = std::thread([&] { ccbh::reserve_slots(e.devs, (int)e.cli.k, e.cli.ext); });
    }
    e.load_db();
    if (slots.joinable()) slots.join();
    e.build_anchor();
    e.run();
    for (ccb_handle* h : e.devs) ccb_destroy(h);
    return 0;
}
