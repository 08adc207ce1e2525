// group.cu -- one process, several GPUs, ONE table sharded over them.
//
// Replaces the reference's handling of a database larger than one device: bucket-range parts swapped through the
// GPUs with every batch re-queried per part and the sparse rows merged (CuClarkDB::read part sizing
// src/CuClarkDB.cu:531-559, swapDbParts :776-815, queryBatch's cross-device merge :909-928).  Here shard g stays
// resident on GPU g and a batch makes one trip:
//
//   home GPU (slot s lives on GPU s mod G)      every owner GPU o                         home GPU
//   H2D -> route (shard_block.cu): k-mers   ->  probe this slot's region of its inbox  -> vote -> D2H
//          stored into the owners' inboxes      (k_probe_inbox), labels stored into
//          through peer mappings (NVLink)       the home GPU's label box (NVLink)
//
// The three stages of a slot are chained with events across the devices' streams; the host enqueues all of it in
// ccb_batch_submit and never synchronises in between.  Slots are independent, so one slot's route overlaps another
// slot's probe.  The group is reached through an ordinary ccb_handle (the "leader", ccb_group_create): the batch calls
// of include/cuclark_b200.h work on it unchanged, which is how bin/cuCLARK --sharded drives it.
#include "capi_internal.cuh"
#include <thread>

using namespace ccb;

#define CU(h, x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) \
    return ccb_fail(h, (int)e_, "%s failed: %s (%s:%d)", #x, cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)

namespace ccb {

struct GroupSlot {
    int home = 0;
    cudaStream_t stream = nullptr;                    // home device
    cudaEvent_t routed = nullptr, done = nullptr;
    uint32_t* h_offsets = nullptr; uint16_t* h_cont = nullptr; uint16_t* h_final5 = nullptr; uint16_t* h_rows = nullptr;   // pinned
    uint32_t* d_offsets = nullptr; uint16_t* d_cont = nullptr; uint16_t* d_final5 = nullptr; uint16_t* d_rows = nullptr;   // home
    uint16_t* d_labbox = nullptr;                     // home: G regions of cap labels
    ccb_shardws* ws = nullptr;                        // home
    // per owner
    std::vector<uint64_t*> inbox;                     // owner o: this slot's region, cap k-mers
    std::vector<unsigned long long*> count;           // owner o: k-mers in the region
    std::vector<cudaStream_t> ostream;
    std::vector<cudaEvent_t> probed;
    bool in_flight = false;
    uint32_t n_reads = 0; int flags = 0;
};

struct Group {
    std::vector<ccb_handle*> member;                  // member[g] holds shard g on devices[g]
    std::vector<GroupSlot> slots;
    uint64_t cap = 0;
    uint32_t max_reads = 0; uint64_t max_cont = 0; bool extended = false;
};

}  // namespace ccb

static int G_of(const ccb_handle* h) { return (int)h->group->member.size(); }

extern "C" int ccb_group_create(const int* devices, int n_devices, const ccb_config* cfg, ccb_handle** out)
{
    if (!devices || !cfg || !out) return ccb_fail(nullptr, -1, "ccb_group_create: null argument");
    if (n_devices < 1 || n_devices > kMaxShards) return ccb_fail(nullptr, -60, "ccb_group_create: 1 to %d devices", kMaxShards);
    ccb_config c0 = *cfg; c0.device = devices[0];
    ccb_handle* leader = nullptr;
    int rc = ccb_create(&c0, &leader);
    if (rc) return rc;
    Group* g = new Group();
    leader->group = g;
    for (int i = 0; i < n_devices; i++) {
        ccb_config c = *cfg; c.device = devices[i]; c.n_slots = 1;
        ccb_handle* m = nullptr;
        rc = ccb_create(&c, &m);
        if (!rc) rc = ccb_shard_config(m, n_devices, i);
        if (rc) { ccb_fail(nullptr, rc, "%s", m ? ccb_last_error(m) : ccb_last_error(nullptr)); if (m) ccb_destroy(m); ccb_destroy(leader); return rc; }
        g->member.push_back(m);
    }
    // every device maps every other device's memory: the route kernel stores into the owners' inboxes, the probe
    // kernel into the home device's label box
    for (int i = 0; i < n_devices; i++) {
        cudaSetDevice(devices[i]);
        for (int j = 0; j < n_devices; j++) {
            if (devices[i] == devices[j]) continue;          // (tests put several shards on one device)
            int can = 0;
            cudaDeviceCanAccessPeer(&can, devices[i], devices[j]);
            if (!can) { ccb_fail(nullptr, -66, "devices %d and %d cannot map each other's memory (no peer access)", devices[i], devices[j]); ccb_destroy(leader); return -66; }
            cudaError_t e = cudaDeviceEnablePeerAccess(devices[j], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { ccb_fail(nullptr, (int)e, "cudaDeviceEnablePeerAccess: %s", cudaGetErrorString(e)); ccb_destroy(leader); return (int)e; }
            cudaGetLastError();
        }
    }
    *out = leader;
    return 0;
}

extern "C" int ccb_group_size(const ccb_handle* h) { return h && h->group ? G_of(h) : 0; }
extern "C" ccb_handle* ccb_group_member(ccb_handle* h, int i)
{
    return h && h->group && i >= 0 && i < G_of(h) ? h->group->member[(size_t)i] : nullptr;
}

namespace ccb {

static void group_free_slots(ccb_handle* h)
{
    Group* g = h->group;
    for (GroupSlot& s : g->slots) {
        ccb_handle* hm = g->member[(size_t)s.home];
        cudaSetDevice(hm->cfg.device);
        if (s.stream) cudaStreamSynchronize(s.stream);
        cudaFreeHost(s.h_offsets); cudaFreeHost(s.h_cont); cudaFreeHost(s.h_final5); cudaFreeHost(s.h_rows);
        cudaFree(s.d_offsets); cudaFree(s.d_cont); cudaFree(s.d_final5); cudaFree(s.d_rows); cudaFree(s.d_labbox);
        ccb_shardws_destroy(hm, s.ws);
        if (s.routed) cudaEventDestroy(s.routed);
        if (s.done) cudaEventDestroy(s.done);
        if (s.stream) cudaStreamDestroy(s.stream);
        for (size_t o = 0; o < s.inbox.size(); o++) {
            cudaSetDevice(g->member[o]->cfg.device);
            if (o < s.ostream.size() && s.ostream[o]) { cudaStreamSynchronize(s.ostream[o]); cudaStreamDestroy(s.ostream[o]); }
            cudaFree(s.inbox[o]);
            if (o < s.count.size()) cudaFree(s.count[o]);
            if (o < s.probed.size() && s.probed[o]) cudaEventDestroy(s.probed[o]);
        }
    }
    g->slots.clear();
    g->max_reads = 0; g->max_cont = 0;
    cudaGetLastError();
}

void group_destroy(ccb_handle* h)
{
    if (!h->group) return;
    group_free_slots(h);
    for (ccb_handle* m : h->group->member) ccb_destroy(m);
    delete h->group; h->group = nullptr;
}

int group_batch_free(ccb_handle* h) { group_free_slots(h); return 0; }

int group_db_load_files(ccb_handle* h, const char* prefix, uint64_t htsize, int key_bytes, unsigned sampling, uint64_t* n_loaded)
{
    Group* g = h->group;
    const size_t G = g->member.size();
    std::vector<int> rc(G, 0); std::vector<uint64_t> n(G, 0);
    std::vector<std::thread> th;
    for (size_t i = 0; i < G; i++)
        th.emplace_back([&, i] { rc[i] = ccb_db_load_files(g->member[i], prefix, htsize, key_bytes, sampling, &n[i]); });
    for (auto& t : th) t.join();
    for (size_t i = 0; i < G; i++) if (rc[i]) return ccb_fail(h, rc[i], "%s", ccb_last_error(g->member[i]));
    if (n_loaded) *n_loaded = n[0];                     // k-mers in the files (each shard keeps its share)
    h->db_ready = true;
    return 0;
}

int group_db_get_info(const ccb_handle* h, ccb_db_info* info)
{
    memset(info, 0, sizeof *info);
    for (ccb_handle* m : h->group->member) {
        ccb_db_info x;
        if (ccb_db_get_info(m, &x)) return -1;
        info->entries += x.entries; info->bytes += x.bytes; info->tombstones += x.tombstones; info->buckets = x.buckets;
        if (x.max_displacement > info->max_displacement) info->max_displacement = x.max_displacement;
    }
    return 0;
}

int group_batch_alloc(ccb_handle* h, uint32_t max_reads, uint64_t max_containers, int extended)
{
    Group* g = h->group;
    const int G = G_of(h);
    if (max_containers > 0xFFFFFFFFull)
        return ccb_fail(h, -30, "ERROR: Batch overflow. Please increase the number of batches (-b <numberofbatches>).");
    if (!g->slots.empty() && (int)g->slots.size() == h->cfg.n_slots && g->max_reads >= max_reads && g->max_cont >= max_containers &&
        g->extended == (extended != 0)) {
        bool idle = true;
        for (const GroupSlot& s : g->slots) idle = idle && !s.in_flight;
        if (idle) return 0;
    }
    group_free_slots(h);
    g->max_reads = max_reads; g->max_cont = max_containers; g->extended = extended != 0;
    // region size: a container holds 8 bases, so a batch has at most 8 k-mer windows per container; mix2k spreads them
    // evenly over the owners.  Two owners: the worst case fits.  More: twice the even share (a batch of one repeated
    // k-mer would overflow: flagged, the batch fails, nothing is read or written out of bounds).
    const uint64_t worst = 8ull * (max_containers ? max_containers : 1) + 64;
    uint64_t cap = G <= 2 ? worst : worst * 2 / (uint64_t)G;
    if (const char* e = getenv("CCB_SHARD_CAP")) cap = strtoull(e, nullptr, 10);
    cap = (cap + 4096 + 15) & ~15ull;
    if (cap > 0xFFFFFFF0ull) cap = 0xFFFFFFF0ull;
    g->cap = cap;
    g->slots.resize((size_t)h->cfg.n_slots);
    const size_t so = ((size_t)max_reads + 1) * 4, sc = (size_t)(max_containers ? max_containers : 1) * 2,
                 sf = (size_t)(max_reads ? max_reads : 1) * 10, sr = (size_t)(max_reads ? max_reads : 1) * h->row_stride * 2;
    int si = 0;
    for (GroupSlot& s : g->slots) {
        s.home = si++ % G;
        ccb_handle* hm = g->member[(size_t)s.home];
        cudaError_t e = cudaSetDevice(hm->cfg.device);
        DeviceNodeBind near_gpu(hm->cfg.device);                  // the slot's pinned buffers next to its home GPU
        auto ok = [&](cudaError_t x) { if (e == cudaSuccess) e = x; };
        ok(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
        ok(cudaEventCreateWithFlags(&s.routed, cudaEventDisableTiming));
        ok(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming | cudaEventBlockingSync));
        ok(cudaMallocHost(&s.h_offsets, so)); ok(cudaMallocHost(&s.h_cont, sc)); ok(cudaMallocHost(&s.h_final5, sf));
        ok(cudaMalloc(&s.d_offsets, so)); ok(cudaMalloc(&s.d_cont, sc)); ok(cudaMalloc(&s.d_final5, sf));
        ok(cudaMalloc(&s.d_labbox, (size_t)G * cap * 2));
        if (extended) { ok(cudaMallocHost(&s.h_rows, sr)); ok(cudaMalloc(&s.d_rows, sr)); }
        int rc = e == cudaSuccess ? ccb_shardws_create(hm, max_reads ? max_reads : 1, G, cap, &s.ws) : 0;
        s.inbox.assign((size_t)G, nullptr); s.count.assign((size_t)G, nullptr); s.ostream.assign((size_t)G, nullptr); s.probed.assign((size_t)G, nullptr);
        for (int o = 0; o < G && e == cudaSuccess && !rc; o++) {
            ok(cudaSetDevice(g->member[(size_t)o]->cfg.device));
            ok(cudaMalloc(&s.inbox[(size_t)o], cap * 8)); ok(cudaMalloc(&s.count[(size_t)o], 8));
            ok(cudaStreamCreateWithFlags(&s.ostream[(size_t)o], cudaStreamNonBlocking));
            ok(cudaEventCreateWithFlags(&s.probed[(size_t)o], cudaEventDisableTiming));
        }
        if (e != cudaSuccess || rc) {
            std::string why = rc ? ccb_last_error(hm) : cudaGetErrorString(e);
            group_free_slots(h);
            if (e == cudaErrorMemoryAllocation)
                return ccb_fail(h, (int)e, "ERROR: Out of GPU memory.\nPlease increase the number of batches (-b <numberofbatches>).");
            return ccb_fail(h, rc ? rc : (int)e, "batch allocation failed: %s", why.c_str());
        }
    }
    return 0;
}

int group_batch_buffers(ccb_handle* h, int slot, uint32_t** offsets, uint16_t** containers, uint16_t** final5, uint16_t** rows)
{
    Group* g = h->group;
    if (slot < 0 || slot >= (int)g->slots.size()) return ccb_fail(h, -31, "bad slot %d", slot);
    GroupSlot& s = g->slots[(size_t)slot];
    if (offsets) *offsets = s.h_offsets;
    if (containers) *containers = s.h_cont;
    if (final5) *final5 = s.h_final5;
    if (rows) *rows = s.h_rows;
    return 0;
}

int group_batch_submit(ccb_handle* h, int slot, uint32_t n_reads, uint64_t n_containers, int flags)
{
    Group* g = h->group;
    const int G = G_of(h);
    if (slot < 0 || slot >= (int)g->slots.size()) return ccb_fail(h, -31, "bad slot %d", slot);
    for (ccb_handle* m : g->member) if (!m->db_ready) return ccb_fail(h, -32, "Failed to find the database.");   // shards may have been loaded one by one
    if (n_reads > g->max_reads || n_containers > g->max_cont) return ccb_fail(h, -33, "batch larger than allocated (%u reads, %llu containers)", n_reads, (unsigned long long)n_containers);
    if ((flags & CCB_EXTENDED) && !g->extended) return ccb_fail(h, -34, "extended results requested but ccb_batch_alloc(extended=0)");
    GroupSlot& s = g->slots[(size_t)slot];
    ccb_handle* hm = g->member[(size_t)s.home];
    CU(h, cudaSetDevice(hm->cfg.device));
    if (s.in_flight) { CU(h, cudaEventSynchronize(s.done)); s.in_flight = false; }
    s.n_reads = n_reads; s.flags = flags;
    if (n_reads == 0) { CU(h, cudaEventRecord(s.done, s.stream)); s.in_flight = true; return 0; }
    CU(h, cudaMemcpyAsync(s.d_offsets, s.h_offsets, ((size_t)n_reads + 1) * 4, cudaMemcpyHostToDevice, s.stream));
    if (n_containers) CU(h, cudaMemcpyAsync(s.d_cont, s.h_cont, (size_t)n_containers * 2, cudaMemcpyHostToDevice, s.stream));
    void* dst[kMaxShards]; void* cnt[kMaxShards]; const void* src[kMaxShards];
    for (int o = 0; o < G; o++) { dst[o] = s.inbox[(size_t)o]; cnt[o] = s.count[(size_t)o]; src[o] = s.d_labbox + (size_t)o * g->cap; }
    int rc = ccb_shard_route2(hm, s.ws, s.d_offsets, s.d_cont, n_reads, (int64_t)n_containers, dst, cnt, s.stream);
    if (rc) return ccb_fail(h, rc, "%s", ccb_last_error(hm));
    CU(h, cudaEventRecord(s.routed, s.stream));
    for (int o = 0; o < G; o++) {
        ccb_handle* ho = g->member[(size_t)o];
        CU(h, cudaSetDevice(ho->cfg.device));
        CU(h, cudaStreamWaitEvent(s.ostream[(size_t)o], s.routed, 0));
        void* lab[1] = {(void*)(s.d_labbox + (size_t)o * g->cap)};
        rc = ccb_shard_probe_inbox(ho, 1, s.inbox[(size_t)o], g->cap, s.count[(size_t)o], lab, nullptr, s.ostream[(size_t)o]);
        if (rc) return ccb_fail(h, rc, "%s", ccb_last_error(ho));
        CU(h, cudaEventRecord(s.probed[(size_t)o], s.ostream[(size_t)o]));
    }
    CU(h, cudaSetDevice(hm->cfg.device));
    for (int o = 0; o < G; o++) CU(h, cudaStreamWaitEvent(s.stream, s.probed[(size_t)o], 0));
    rc = ccb_shard_vote2(hm, s.ws, n_reads, src, s.d_final5, (flags & CCB_EXTENDED) ? s.d_rows : nullptr, s.stream);
    if (rc) return ccb_fail(h, rc, "%s", ccb_last_error(hm));
    CU(h, cudaMemcpyAsync(s.h_final5, s.d_final5, (size_t)n_reads * 10, cudaMemcpyDeviceToHost, s.stream));
    if (flags & CCB_EXTENDED) CU(h, cudaMemcpyAsync(s.h_rows, s.d_rows, (size_t)n_reads * h->row_stride * 2, cudaMemcpyDeviceToHost, s.stream));
    CU(h, cudaEventRecord(s.done, s.stream));
    s.in_flight = true;
    return 0;
}

static int group_harvest(ccb_handle* h, GroupSlot& s)
{
    Group* g = h->group;
    ccb_handle* hm = g->member[(size_t)s.home];
    s.in_flight = false;
    { std::lock_guard<std::mutex> lk(h->mu); h->stats.reads += s.n_reads; }
    if (s.n_reads == 0) return 0;
    int err = 0;
    int rc = ccb_shardws_error(hm, s.ws, s.stream, &err);
    if (rc) return ccb_fail(h, rc, "%s", ccb_last_error(hm));
    if (err) return ccb_fail(h, -65, "ERROR: the k-mers of a batch for one GPU exceed its exchange region (%llu).\nPlease increase the number of batches (-b <numberofbatches>).",
                             (unsigned long long)g->cap);
    return 0;
}

int group_batch_wait(ccb_handle* h, int slot)
{
    Group* g = h->group;
    if (slot < 0 || slot >= (int)g->slots.size()) return ccb_fail(h, -31, "bad slot %d", slot);
    GroupSlot& s = g->slots[(size_t)slot];
    if (!s.in_flight) return 0;
    CU(h, cudaSetDevice(g->member[(size_t)s.home]->cfg.device));
    CU(h, cudaEventSynchronize(s.done));
    return group_harvest(h, s);
}

int group_batch_poll(ccb_handle* h, int slot)
{
    Group* g = h->group;
    if (slot < 0 || slot >= (int)g->slots.size()) return ccb_fail(h, -31, "bad slot %d", slot);
    GroupSlot& s = g->slots[(size_t)slot];
    if (!s.in_flight) return 0;
    CU(h, cudaSetDevice(g->member[(size_t)s.home]->cfg.device));
    cudaError_t e = cudaEventQuery(s.done);
    if (e == cudaErrorNotReady) return 1;
    if (e != cudaSuccess) return ccb_fail(h, (int)e, "cudaEventQuery: %s", cudaGetErrorString(e));
    return group_harvest(h, s);
}

int group_sync(ccb_handle* h)
{
    for (ccb_handle* m : h->group->member) { CU(h, cudaSetDevice(m->cfg.device)); CU(h, cudaDeviceSynchronize()); }
    return 0;
}

}  // namespace ccb
