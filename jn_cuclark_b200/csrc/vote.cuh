// vote.cuh -- result epilogue shared by the dense exact paths (query.cu, shard_block.cu).
#pragma once
#include "query.cuh"

namespace ccb {

// top-2 under (count desc, target asc) + sum over a CTA's dense counters, then the result row
static __device__ __forceinline__ void dense_result(const QueryArgs& a, uint32_t r, const uint32_t* cnt, uint32_t n_targets,
                                             uint32_t (&s_key)[2][256], uint32_t (&s_sum)[256])
{
    uint32_t k1 = 0, k2 = 0, sum = 0;
    for (uint32_t i = threadIdx.x; i < n_targets; i += blockDim.x) {
        uint32_t c16 = cnt[i] & 0xFFFFu;
        if (!c16) continue;
        sum += c16;
        uint32_t key = (c16 << 16) | (0xFFFFu - i);
        if (key > k1) { k2 = k1; k1 = key; } else if (key > k2) k2 = key;
    }
    s_key[0][threadIdx.x] = k1; s_key[1][threadIdx.x] = k2; s_sum[threadIdx.x] = sum;
    __syncthreads();
    for (int o = 128; o; o >>= 1) {
        if ((int)threadIdx.x < o) {
            uint32_t a1 = s_key[0][threadIdx.x], a2 = s_key[1][threadIdx.x];
            uint32_t b1 = s_key[0][threadIdx.x + o], b2 = s_key[1][threadIdx.x + o];
            uint32_t m1 = max(a1, b1);
            uint32_t m2 = max(min(a1, b1), max(a2, b2));
            s_key[0][threadIdx.x] = m1; s_key[1][threadIdx.x] = m2;
            s_sum[threadIdx.x] += s_sum[threadIdx.x + o];
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        k1 = s_key[0][0]; k2 = s_key[1][0];
        uint16_t* o5 = a.out5 + (size_t)r * 5;
        o5[0] = (uint16_t)s_sum[0];
        o5[1] = (uint16_t)(k1 ? (0xFFFFu - (k1 & 0xFFFFu)) + 1u : 0u);
        o5[2] = (uint16_t)(k1 >> 16);
        o5[3] = (uint16_t)(k2 ? (0xFFFFu - (k2 & 0xFFFFu)) + 1u : 0u);
        o5[4] = (uint16_t)(k2 >> 16);
        if (a.rows) {
            uint16_t* row = a.rows + (size_t)r * a.row_stride;
            int n = 0, fit = (a.row_stride - 1) / 2;
            for (uint32_t i = 0; i < n_targets && n < fit; i++)
                if (cnt[i] & 0xFFFFu) { row[1 + 2 * n] = (uint16_t)i; row[2 + 2 * n] = (uint16_t)cnt[i]; n++; }
            row[0] = (uint16_t)n;
        }
    }
}

}  // namespace ccb
