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

#define SEL_HISTS 8    // private histograms (warp w uses copy w % 8): early digits are nearly constant, one copy would serialise
struct SelectSmem {
    u64 keys[SEL_MAX];
    uint32_t hist[SEL_HISTS][256];
    u64 kmin, kmax;
    int digit, need, cnt, out_count;
};

// Returns the largest key that belongs to the kk smallest of keys[0..n).  Block-uniform.
__device__ inline u64 cta_radix_threshold(SelectSmem &s, int n, int kk) {
    if (kk >= n) return ~0ull;
    // bytes on which every key agrees need no pass: find them from the block-wide min and max
    u64 mn = ~0ull, mx = 0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) { const u64 key = s.keys[i]; mn = key < mn ? key : mn; mx = key > mx ? key : mx; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const u64 a = __shfl_xor_sync(MCP_FULL, mn, o), b = __shfl_xor_sync(MCP_FULL, mx, o);
        mn = a < mn ? a : mn; mx = b > mx ? b : mx;
    }
    if (threadIdx.x == 0) { s.kmin = ~0ull; s.kmax = 0; }
    __syncthreads();
    if ((threadIdx.x & 31) == 0) { atomicMin(&s.kmin, mn); atomicMax(&s.kmax, mx); }
    __syncthreads();
    const u64 diff = s.kmin ^ s.kmax;
    int shift0 = 56;
    while (shift0 > 0 && (diff >> shift0) == 0) shift0 -= 8;
    u64 prefix = shift0 == 56 ? 0ull : (s.kmin >> (shift0 + 8)) << (shift0 + 8);
    u64 mask = shift0 == 56 ? 0ull : ~0ull << (shift0 + 8);
    int need = kk;
    uint32_t *myhist = s.hist[(threadIdx.x >> 5) % SEL_HISTS];
    for (int shift = shift0; shift >= 0; shift -= 8) {
        for (int i = threadIdx.x; i < SEL_HISTS * 256; i += blockDim.x) (&s.hist[0][0])[i] = 0;
        __syncthreads();
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            u64 key = s.keys[i];
            if ((key & mask) == prefix) atomicAdd(&myhist[(uint32_t)(key >> shift) & 255u], 1u);
        }
        __syncthreads();
        if (threadIdx.x < 32) {
            int lane = threadIdx.x;
            uint32_t c[8], sum = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                uint32_t t = 0;
#pragma unroll
                for (int h = 0; h < SEL_HISTS; ++h) t += s.hist[h][lane * 8 + j];
                c[j] = t; sum += t;
            }
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

// Stage B: lists[q][L_in][cap_in] -> lists[q][L_out][cap_out]; CTA g reduces the G consecutive input lists
// [g*G, g*G+G) (G <= SEL_THREADS) to their kprime smallest keys.  grid (L_out, Q).
// Lists are usually far from full, so one CTA normally takes all of a query's lists at once; when their keys do
// not fit the SEL_MAX-key buffer they are consumed in chunks, the best kprime so far riding along.
// raw_in: the lists hold raw (fp32 score bits, row) pairs as the GEMM epilogue appends them.
// tau_pub (raw lists only): the scan's final shared bound per query, (launch epoch << 32 | float bits).  At least kprime keys
// of the query's lists lie at or below it, so keys above it are dropped as they are gathered -- the lists were filled under
// looser, earlier bounds and hold several times what the final bound admits (one query over 1 M rows: ~15 K keys in ~300
// lists, a few thousand under the bound).  Survivors are compacted into the buffer, chunk after chunk, and selected once.
__global__ void __launch_bounds__(SEL_THREADS)
select_from_lists_kernel(const u64 *__restrict__ in, const int *__restrict__ in_cnt, int L_in,
                         int cap_in, int G, int kprime, u64 *__restrict__ out,
                         int *__restrict__ out_cnt, int cap_out, int raw_in,
                         const u64 *__restrict__ tau_pub, uint32_t epoch) {
    __shared__ SelectSmem s;
    __shared__ int base[SEL_THREADS + 1];      // exclusive prefix of the list lengths
    __shared__ int wsum[SEL_THREADS / 32];
    __shared__ int s_lend, s_fill;
    pdl_trigger();
    pdl_wait();                                                // the scan that wrote the lists has completed
    const int g = blockIdx.x, q = blockIdx.y, L_out = gridDim.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int l0 = g * G, nl = min(G, L_in - l0);
    float tau = CUDART_INF_F;
    if (tau_pub && raw_in) {
        const u64 wv = tau_pub[q];
        if ((uint32_t)(wv >> 32) == epoch) tau = __uint_as_float((uint32_t)wv);
    }
    const int mycnt = (int)threadIdx.x < nl ? min(in_cnt[(int64_t)q * L_in + l0 + threadIdx.x], cap_in) : 0;
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
    base[threadIdx.x] = wsum[warp] + incl - mycnt;             // base[i] = keys before list l0 + i
    if (threadIdx.x == SEL_THREADS - 1) base[SEL_THREADS] = wsum[warp] + incl;
    __syncthreads();
    const u64 *lists = in + ((int64_t)q * L_in + l0) * cap_in;
    const int64_t lo = (int64_t)q * L_out + g;
    u64 *dst = out + lo * cap_out;
    int carry = 0, ls = 0;
    for (;;) {
        // lists [ls, le) are the longest run whose keys fit behind the carried ones (normally: all of them)
        if (threadIdx.x == 0) {
            int le = nl;
            if (carry + base[nl] - base[ls] > SEL_MAX) {       // base[] is monotone: binary search for the last list that fits
                int a = ls + 1, b = nl;                        // invariant: list run [ls, a) fits, [ls, b) does not
                while (b - a > 1) { const int mid = (a + b) >> 1; if (carry + base[mid] - base[ls] <= SEL_MAX) a = mid; else b = mid; }
                le = a;
            }
            s_lend = le;
            s_fill = carry;
        }
        __syncthreads();
        const int le = s_lend, b0 = base[ls], m = base[le] - b0;
        // flat gather, at most SEL_MAX / SEL_THREADS = 8 keys per thread: first every thread finds where its keys
        // live (binary search over the prefix sums, shared memory only), then all its global loads go out together
        {
            const u64 *src[SEL_MAX / SEL_THREADS];
#pragma unroll
            for (int u = 0; u < SEL_MAX / SEL_THREADS; ++u) {
                const int i = threadIdx.x + u * SEL_THREADS;
                src[u] = nullptr;
                if (i < m) {
                    const int e = b0 + i;
                    int a = ls, b = le;                        // list l with base[l] <= e < base[l+1]
                    while (b - a > 1) { const int mid = (a + b) >> 1; if (base[mid] <= e) a = mid; else b = mid; }
                    src[u] = lists + (int64_t)a * cap_in + (e - base[a]);
                }
            }
            u64 x[SEL_MAX / SEL_THREADS];
#pragma unroll
            for (int u = 0; u < SEL_MAX / SEL_THREADS; ++u) x[u] = src[u] ? __ldg(src[u]) : 0ull;
#pragma unroll
            for (int u = 0; u < SEL_MAX / SEL_THREADS; ++u) {
                const u64 v = x[u];
                if (raw_in) {
                    // at or below the final bound, or NaN: compacted behind what is held (one shared-memory atomic per warp)
                    const float sc = __uint_as_float((uint32_t)(v >> 32));
                    const bool keep = src[u] != nullptr && !(sc > tau);
                    const uint32_t bal = __ballot_sync(MCP_FULL, keep);
                    int wbase = 0;
                    if (lane == 0 && bal) wbase = atomicAdd(&s_fill, __popc(bal));
                    wbase = __shfl_sync(MCP_FULL, wbase, 0);
                    if (keep) s.keys[wbase + __popc(bal & ((1u << lane) - 1u))] = ((u64)f32_ord(sc) << 32) | (v & 0xFFFFFFFFull);
                } else if (src[u])
                    s.keys[carry + threadIdx.x + u * SEL_THREADS] = v;
            }
        }
        __syncthreads();
        const int n = raw_in ? s_fill : carry + m;
        if (le < nl && n <= SEL_MAX / 2) {                     // room left: keep gathering, select once at the end
            carry = n; ls = le;
            __syncthreads();                                   // s_fill / s_lend are rewritten at the top
            continue;
        }
        cta_select_write(s, n, min(kprime, n), dst, out_cnt + lo);
        if (le >= nl) break;
        __syncthreads();
        carry = min(kprime, n);
        for (int i = threadIdx.x; i < carry; i += SEL_THREADS) s.keys[i] = dst[i];   // written by this CTA just above
        ls = le;
        __syncthreads();
    }
}
