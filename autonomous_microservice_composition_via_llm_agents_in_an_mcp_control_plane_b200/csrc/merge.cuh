// merge.cuh -- K5: merge of per-shard top-k lists after the all-gather.
// Each row shard returns its exact local top-k sorted by (score_f64, id); the global top-k is the
// k smallest of their union under the same key, so sharding cannot change the answer.
// One CTA per query; S*k <= MERGE_MAX entries are sorted in shared memory.
#pragma once
#include "common.cuh"

#define MERGE_THREADS 256
#define MERGE_MAX 4096

__global__ void __launch_bounds__(MERGE_THREADS)
merge_shards_kernel(const double *__restrict__ scores64, const int32_t *__restrict__ ids, int n_shards,
                    int64_t shard_stride_scores /* doubles between two shards' score blocks */,
                    int64_t shard_stride_ids /* int32s between two shards' id blocks */,
                    int n_queries, int k, int npow2, int32_t *__restrict__ out_ids,
                    float *__restrict__ out_scores, double *__restrict__ out_scores64) {
    extern __shared__ unsigned char smem_raw[];
    u64 *sk = reinterpret_cast<u64 *>(smem_raw);
    int *sv = reinterpret_cast<int *>(sk + npow2);
    int q = blockIdx.x;
    int total = n_shards * k;
    for (int i = threadIdx.x; i < npow2; i += blockDim.x) {
        u64 key = ~0ull; int id = MCP_PAD_ROW;
        if (i < total) {
            int s = i / k, j = i - s * k;
            int64_t o = (int64_t)q * k + j;
            int32_t v = ids[(int64_t)s * shard_stride_ids + o];
            if (v >= 0) { key = f64_ord(scores64[(int64_t)s * shard_stride_scores + o]); id = v; }
        }
        sk[i] = key; sv[i] = id;
    }
    cta_bitonic_sort_pairs(sk, sv, npow2);
    for (int i = threadIdx.x; i < k; i += blockDim.x) {
        int64_t o = (int64_t)q * k + i;
        bool real = sv[i] != MCP_PAD_ROW;
        double s = real ? ord_f64(sk[i]) : CUDART_INF;
        out_ids[o] = real ? sv[i] : -1;
        out_scores[o] = (float)s;
        if (out_scores64) out_scores64[o] = s;
    }
}
