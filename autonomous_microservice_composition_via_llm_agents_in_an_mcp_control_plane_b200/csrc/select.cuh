// select.cuh -- K3: top-k' selection of coarse candidates.
//
// Every coarse scan leaves, per query, L lists of 64-bit keys (fp32 score image | row).  A
// list-reduction kernel keeps the k' smallest keys of a group of lists; it is applied until one
// list per query is left.  Selection is an MSB-first radix select in shared memory (keys are
// unique, so the k'-th key is well defined and ties resolve to the lowest row for free).
//
// Invariant the exactness certificate relies on (rescore.cuh): every key that any stage drops
// is larger than every key that survives to the final list.
#pragma once
#include "common.cuh"

#define SEL_THREADS 512
#define SEL_MAX 4096   // keys one CTA selects from (32 KB of shared memory)

struct SelectSmem {
    u64 keys[SEL_MAX];
    uint32_t hist[256];
    int digit, need, cnt, out_count;
};

// Returns the largest key that belongs to the kk smallest of keys[0..n).  Block-uniform.
__device__ inline u64 cta_radix_threshold(SelectSmem &s, int n, int kk) {
    if (kk >= n) return ~0ull;
    u64 prefix = 0, mask = 0;
    int need = kk;
    for (int shift = 56; shift >= 0; shift -= 8) {
        for (int i = threadIdx.x; i < 256; i += blockDim.x) s.hist[i] = 0;
        __syncthreads();
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            u64 key = s.keys[i];
            if ((key & mask) == prefix) atomicAdd(&s.hist[(uint32_t)(key >> shift) & 255u], 1u);
        }
        __syncthreads();
        if (threadIdx.x < 32) {
            int lane = threadIdx.x;
            uint32_t c[8], sum = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) { c[j] = s.hist[lane * 8 + j]; sum += c[j]; }
            uint32_t incl = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t t = __shfl_up_sync(MCP_FULL, incl, o);
                if (lane >= o) incl += t;
            }
            uint32_t excl = incl - sum;
            if (excl < (uint32_t)need && (uint32_t)need <= incl) {
                uint32_t run = excl;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    if ((uint32_t)need <= run + c[j]) {
                        s.digit = lane * 8 + j; s.need = need - (int)run; s.cnt = (int)c[j];
                        break;
                    }
                    run += c[j];
                }
            }
        }
        __syncthreads();
        int d = s.digit, cnt = s.cnt;
        need = s.need;
        prefix |= (u64)d << shift;
        mask |= 0xFFull << shift;
        if (cnt == need)                       // the whole bin is taken: no need to refine
            return prefix | (shift ? ((1ull << shift) - 1ull) : 0ull);
        __syncthreads();
    }
    return prefix;
}

// keep the kk smallest of s.keys[0..n) -> out[0..kk) (unordered); returns nothing, *out_cnt = kk'
__device__ inline void cta_select_write(SelectSmem &s, int n, int kk, u64 *out, int *out_cnt) {
    if (threadIdx.x == 0) s.out_count = 0;
    __syncthreads();
    u64 thr = cta_radix_threshold(s, n, kk);
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        u64 key = s.keys[i];
        if (key <= thr) out[atomicAdd(&s.out_count, 1)] = key;
    }
    __syncthreads();
    if (threadIdx.x == 0) *out_cnt = s.out_count;
}

// Stage A (after the fp32 streaming scan): scores[q][n] -> lists[q][P][cap]
// grid (P, Q); slice p covers rows [p*SEL_MAX, ...).
__global__ void __launch_bounds__(SEL_THREADS)
select_from_scores_kernel(const float *__restrict__ scores, int64_t score_stride, int64_t n_rows,
                          int kprime, u64 *__restrict__ lists, int *__restrict__ list_cnt, int cap) {
    __shared__ SelectSmem s;
    int p = blockIdx.x, q = blockIdx.y, P = gridDim.x;
    int64_t r0 = (int64_t)p * SEL_MAX;
    int n = (int)min((int64_t)SEL_MAX, n_rows - r0);
    const float *sc = scores + (int64_t)q * score_stride + r0;
    for (int i = threadIdx.x; i < n; i += blockDim.x)
        s.keys[i] = coarse_key(sc[i], (uint32_t)(r0 + i));
    __syncthreads();
    int64_t l = (int64_t)q * P + p;
    cta_select_write(s, n, min(kprime, n), lists + l * cap, list_cnt + l);
}

// Stage B: lists[q][L_in][cap_in] -> lists[q][L_out][cap_out]; group g reduces G consecutive
// input lists (G * cap_in <= SEL_MAX, G <= SEL_THREADS).  grid (L_out, Q).
// Lists are usually far from full: when all of a query's lists together fit one CTA, group 0 takes them all and
// the other groups hand back empty lists (the launch shape stays static, the work does not).
__global__ void __launch_bounds__(SEL_THREADS)
select_from_lists_kernel(const u64 *__restrict__ in, const int *__restrict__ in_cnt, int L_in,
                         int cap_in, int G, int kprime, u64 *__restrict__ out,
                         int *__restrict__ out_cnt, int cap_out, int raw_in) {
    __shared__ SelectSmem s;
    __shared__ int base[SEL_THREADS + 1];      // exclusive prefix of the list lengths
    __shared__ int wsum[SEL_THREADS / 32];
    const int g = blockIdx.x, q = blockIdx.y, L_out = gridDim.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // list lengths, one per thread, and their block-wide exclusive scan: all lists of the query when there are
    // at most SEL_THREADS of them (lets group 0 take everything), else just this group's
    int l0 = g * G, l1 = min(l0 + G, L_in);
    const bool whole = L_in <= SEL_THREADS;
    const int lb = whole ? 0 : l0, nl = whole ? L_in : l1 - l0;
    const int mycnt = (int)threadIdx.x < nl ? min(in_cnt[(int64_t)q * L_in + lb + threadIdx.x], cap_in) : 0;
    int incl = mycnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(MCP_FULL, incl, o); if (lane >= o) incl += t; }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        int w = lane < SEL_THREADS / 32 ? wsum[lane] : 0, wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(MCP_FULL, wi, o); if (lane >= o) wi += t; }
        if (lane < SEL_THREADS / 32) wsum[lane] = wi - w;
    }
    __syncthreads();
    base[threadIdx.x] = wsum[warp] + incl - mycnt;             // base[i] = keys before list lb + i
    if (threadIdx.x == SEL_THREADS - 1) base[SEL_THREADS] = wsum[warp] + incl;
    __syncthreads();
    if (whole && base[SEL_THREADS] <= SEL_MAX) {
        if (g != 0) { if (threadIdx.x == 0) out_cnt[(int64_t)q * L_out + g] = 0; return; }
        l0 = 0; l1 = L_in;
    }
    const int off = base[l0 - lb], n = base[l1 - lb] - off;
    for (int l = l0 + warp; l < l1; l += SEL_THREADS / 32) {      // one warp per list
        const int b = base[l - lb] - off, c = base[l - lb + 1] - base[l - lb];
        const u64 *src = in + ((int64_t)q * L_in + l) * cap_in;
        if (raw_in)   // K2 appends raw (fp32 bits, row) pairs: build the ordered key here
            for (int i = lane; i < c; i += 32) {
                const u64 raw = src[i];
                s.keys[b + i] = ((u64)f32_ord(__uint_as_float((uint32_t)(raw >> 32))) << 32) | (raw & 0xFFFFFFFFull);
            }
        else
            for (int i = lane; i < c; i += 32) s.keys[b + i] = src[i];
    }
    __syncthreads();
    const int64_t lo = (int64_t)q * L_out + g;
    cta_select_write(s, n, min(kprime, n), out + lo * cap_out, out_cnt + lo);
}
