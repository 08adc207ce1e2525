// probe_fill.cu -- what does a random 32-byte read of HBM cost on this part, by request flavour?
// (measurement tool, not on the product path)
//
//   tools/probe_fill [GiB=8] [log2 probes=28]
// prints G requests/s per flavour; run under
//   ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum --clock-control none
// for the DRAM bytes each flavour moves per request.
//
// flavours: 0 ld.global.nc.L1::no_allocate (the product's load_bucket)   1-3 with .L2::64B/128B/256B
//           4 ld.relaxed.sys   5 ld.volatile (2 x 16 B)   6 cp.async.bulk 32 B -> smem
//           7 TMA 1-D tile of 32 B, tensor map with L2 promotion NONE    8 the same with promotion 64B
//           9 TMA tile of 32 B, promotion NONE, cache hint evict_first
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

static __device__ __forceinline__ uint64_t sm64(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ull; x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull; return x ^ (x >> 31);
}
static __device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int MODE>
__global__ void __launch_bounds__(256) k_fill_ld(const ulonglong4* tbl, uint64_t n_sectors, uint64_t n_probes, unsigned long long* sink)
{
    const uint64_t stride = gridDim.x * (uint64_t)blockDim.x;
    unsigned long long acc = 0;
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n_probes; i += 8 * stride) {
        uint64_t e[8][4];
#pragma unroll
        for (int u = 0; u < 8; u++) {
            const ulonglong4* p = tbl + __umul64hi(sm64(i + u * stride), n_sectors);
            e[u][2] = e[u][3] = 0;
            if (MODE == 0) asm volatile("ld.global.nc.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(e[u][0]), "=l"(e[u][1]), "=l"(e[u][2]), "=l"(e[u][3]) : "l"(p));
            if (MODE == 1) asm volatile("ld.global.nc.L1::no_allocate.L2::64B.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(e[u][0]), "=l"(e[u][1]), "=l"(e[u][2]), "=l"(e[u][3]) : "l"(p));
            if (MODE == 2) asm volatile("ld.global.nc.L1::no_allocate.L2::128B.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(e[u][0]), "=l"(e[u][1]), "=l"(e[u][2]), "=l"(e[u][3]) : "l"(p));
            if (MODE == 3) asm volatile("ld.global.nc.L1::no_allocate.L2::256B.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(e[u][0]), "=l"(e[u][1]), "=l"(e[u][2]), "=l"(e[u][3]) : "l"(p));
            if (MODE == 4) asm volatile("ld.relaxed.sys.global.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(e[u][0]), "=l"(e[u][1]), "=l"(e[u][2]), "=l"(e[u][3]) : "l"(p));
            if (MODE == 5) {
                const ulonglong2* p2 = reinterpret_cast<const ulonglong2*>(p);
                asm volatile("ld.volatile.global.v2.u64 {%0,%1}, [%2];" : "=l"(e[u][0]), "=l"(e[u][1]) : "l"(p2));
                asm volatile("ld.volatile.global.v2.u64 {%0,%1}, [%2];" : "=l"(e[u][2]), "=l"(e[u][3]) : "l"(p2 + 1));
            }
        }
#pragma unroll
        for (int u = 0; u < 8; u++) acc += e[u][0] ^ e[u][1] ^ e[u][2] ^ e[u][3];
    }
    if (acc == 0x123456789abcdefull) *sink = acc;
}

// every lane asks for LPT random 32-byte pieces per step with async copies into shared memory, one mbarrier per warp
template <int MODE>
__global__ void __launch_bounds__(128) k_fill_async(const ulonglong4* tbl, const __grid_constant__ CUtensorMap tmap,
                                                     uint64_t n_sectors, uint64_t n_probes, unsigned long long* sink)
{
    constexpr int LPT = 2, WARPS = 4;
    __shared__ alignas(128) unsigned char buf[WARPS][32 * LPT][128];
    __shared__ alignas(8) unsigned long long bar[WARPS];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint64_t stride = gridDim.x * (uint64_t)blockDim.x;
    const uint64_t tid = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    const uint64_t steps = n_probes / (stride * LPT);
    const uint32_t bar_a = smem_u32(&bar[w]);
    if (lane == 0) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(bar_a));
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    uint64_t pol = 0;
    if (MODE == 9) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    unsigned long long acc = 0;
    uint32_t parity = 0;
    for (uint64_t st = 0; st < steps; st++) {
        if (lane == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar_a), "r"(32 * LPT * 32) : "memory");
        __syncwarp();
#pragma unroll
        for (int u = 0; u < LPT; u++) {
            const uint64_t sec = __umul64hi(sm64((st * LPT + u) * stride + tid), n_sectors);
            const uint32_t dst = smem_u32(&buf[w][lane * LPT + u][0]);
            if (MODE == 6)
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 32, [%2];"
                             :: "r"(dst), "l"(tbl + sec), "r"(bar_a) : "memory");
            else if (MODE == 9)
                asm volatile("cp.async.bulk.tensor.1d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%2}], [%3], %4;"
                             :: "r"(dst), "l"(&tmap), "r"((int)(sec * 8)), "r"(bar_a), "l"(pol) : "memory");
            else
                asm volatile("cp.async.bulk.tensor.1d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2}], [%3];"
                             :: "r"(dst), "l"(&tmap), "r"((int)(sec * 8)), "r"(bar_a) : "memory");
        }
        uint32_t done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(bar_a), "r"(parity) : "memory");
        parity ^= 1;
#pragma unroll
        for (int u = 0; u < LPT; u++) {
            const ulonglong2* q = (const ulonglong2*)&buf[w][lane * LPT + u][0];
            acc += q[0].x ^ q[1].y;
        }
        __syncwarp();
    }
    if (acc == 0x123456789abcdefull) *sink = acc;
}

typedef CUresult (*EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char** argv)
{
    const double gib = argc > 1 ? atof(argv[1]) : 8.0;
    const int lg = argc > 2 ? atoi(argv[2]) : 28;
    uint64_t bytes = (uint64_t)(gib * (1ull << 30));
    if (bytes > (1ull << 33)) bytes = 1ull << 33;                 // 2^31 u32 elements: one tensor-map dimension
    const uint64_t n_sectors = bytes / 32, n_probes = 1ull << lg;
    void* tbl = nullptr; unsigned long long* sink = nullptr;
    if (cudaMalloc(&tbl, bytes) != cudaSuccess || cudaMalloc(&sink, 8) != cudaSuccess) { fprintf(stderr, "cudaMalloc failed\n"); return 1; }
    cudaMemset(tbl, 0x5a, bytes);
    EncodeTiled enc = nullptr;
    cudaDriverEntryPointQueryResult qr;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&enc, cudaEnableDefault, &qr);
    CUtensorMap tm[2]; bool have_tm = false;
    if (enc) {
        have_tm = true;
        for (int v = 0; v < 2; v++) {
            cuuint64_t gdim[1] = {bytes / 4}; cuuint64_t gstr[1] = {0}; cuuint32_t box[1] = {8}; cuuint32_t estr[1] = {1};
            CUresult r = enc(&tm[v], CU_TENSOR_MAP_DATA_TYPE_UINT32, 1, tbl, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_NONE, v ? CU_TENSOR_MAP_L2_PROMOTION_L2_64B : CU_TENSOR_MAP_L2_PROMOTION_NONE,
                             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) { fprintf(stderr, "cuTensorMapEncodeTiled(%d) failed: %d\n", v, (int)r); have_tm = false; }
        }
    }
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    const ulonglong4* t = (const ulonglong4*)tbl;
    for (int mode = 0; mode <= 9; mode++) {
        if (mode >= 7 && !have_tm) continue;
        float best = 1e30f;
        for (int rep = 0; rep < 3; rep++) {
            cudaEventRecord(a);
            switch (mode) {
            case 0: k_fill_ld<0><<<148 * 8, 256>>>(t, n_sectors, n_probes, sink); break;
            case 1: k_fill_ld<1><<<148 * 8, 256>>>(t, n_sectors, n_probes, sink); break;
            case 2: k_fill_ld<2><<<148 * 8, 256>>>(t, n_sectors, n_probes, sink); break;
            case 3: k_fill_ld<3><<<148 * 8, 256>>>(t, n_sectors, n_probes, sink); break;
            case 4: k_fill_ld<4><<<148 * 8, 256>>>(t, n_sectors, n_probes, sink); break;
            case 5: k_fill_ld<5><<<148 * 8, 256>>>(t, n_sectors, n_probes, sink); break;
            case 6: k_fill_async<6><<<148 * 4, 128>>>(t, tm[0], n_sectors, n_probes, sink); break;
            case 7: k_fill_async<7><<<148 * 4, 128>>>(t, tm[0], n_sectors, n_probes, sink); break;
            case 8: k_fill_async<8><<<148 * 4, 128>>>(t, tm[1], n_sectors, n_probes, sink); break;
            case 9: k_fill_async<9><<<148 * 4, 128>>>(t, tm[0], n_sectors, n_probes, sink); break;
            }
            cudaEventRecord(b);
            cudaError_t e = cudaEventSynchronize(b);
            if (e != cudaSuccess) { printf("{\"mode\": %d, \"error\": \"%s\"}\n", mode, cudaGetErrorString(e)); return 2; }
            float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
        }
        printf("{\"mode\": %d, \"gib\": %.1f, \"probes\": %llu, \"ms\": %.3f, \"g_requests_per_s\": %.2f}\n", mode, bytes / 1073741824.0,
               (unsigned long long)n_probes, best, n_probes / (best * 1e6));
        fflush(stdout);
    }
    return 0;
}
