// rescore.cuh -- K4: exact rescoring of the oversampled candidates, final (score,id) sort,
// exactness certificate, and the fp64 full scan the certificate falls back to.
//
// Why: ids must be bit-exact against the fp64 oracle, but the coarse scans (fp32 tree sums,
// TF32 tensor-core products) order near neighbours differently.  So the coarse pass keeps
// k' > k candidates, this stage recomputes their scores with the oracle's own arithmetic
// (common.cuh exact_*), sorts by (score, id) and takes k.  The certificate then proves no row
// outside the candidate set could have entered the top-k:
//     every dropped row j has coarse score c_j >= T  (T = worst surviving coarse score,
//     select.cuh invariant) and exact score s_j >= c_j - eps  (eps = coarse error bound)
//     =>  if the k-th exact score s_(k) < T - eps the answer is the true top-k.
// Queries that fail are re-answered by the fp64 scan, so "bit-exact" is a guarantee.
#pragma once
#include "common.cuh"
#include "prep.cuh"

#define RESCORE_THREADS 128
#define FINAL_THREADS 256
#define FINAL_MAX 512        // max k' (power of two)
#define EXSEL_THREADS 1024

struct CertParams {
    int metric;
    int path;              // MCP_PATH_*
    int complete;          // candidate set == all rows: nothing was dropped
    double rel;            // relative coarse error bound of the dot product / squared distance
    double abs_pen;        // multiplier of pen_absmax for fp32 rounding of the final add
};

// ---- K4a: exact rescoring ------------------------------------------------------------------------
// One CTA = one query x 32 candidates.  The fp64 chain of a candidate (dim dependent FMAs, index order) is
// the latency floor.  On B200 the fp64 pipe is narrow (tools/fp64_latency.cu, tools/chain_bench.cu: a dependent DFMA
// every ~8-16 cycles, F2F.F64.F32 ~12 cycles per warp instruction, both on the same pipe, ~0.4 warp instructions per
// clock per SM), so the chain carries as little as possible: the query is converted to fp64 once per CTA
// by all threads and kept in shared memory; all 128 threads stream the 32 candidate rows through a 3-stage
// cp.async ring (64 floats per row per stage, 256 contiguous bytes per row); the 32 lanes of warp 0 each walk
// their own row out of shared memory, converting only the row element.  grid (ceil(stride / 32), Q).
#define RS_ROWS 32
#define RS_COLS 64
#define RS_PITCH (RS_COLS + 4)        // fp32 ring row pitch (floats): 16-byte aligned rows, conflict-free LDS.128
#define RS_STAGES 5                   // 4 chunks (~2.4 us of chain time) of row gathers in flight per CTA; 55.5 KB smem at dim 1536 -> 4 CTAs / SM
#define RS_SMEM_BYTES(dim) (sizeof(double) * (((dim) + 3) & ~3) + sizeof(float) * RS_STAGES * RS_ROWS * RS_PITCH + 16)
__device__ __forceinline__ void cp_async16(void *dst, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__global__ void __launch_bounds__(RESCORE_THREADS)
rescore_kernel(const float *__restrict__ rows, int dim, int64_t ld,
               const float *__restrict__ queries, const double *__restrict__ nq2,
               const double *__restrict__ ne2, const float *__restrict__ penalty, int metric,
               const u64 *__restrict__ cand, const int *__restrict__ cand_cnt, int cand_stride,
               u64 *__restrict__ exact_key, int *__restrict__ exact_row) {
    extern __shared__ __align__(16) unsigned char rs_raw[];
    const int dimp = (dim + 3) & ~3;                       // rows are padded with zeros up to ld >= dimp
    double *sq = reinterpret_cast<double *>(rs_raw);       // [dimp] the query, already fp64
    float *tile = reinterpret_cast<float *>(sq + dimp);    // [RS_STAGES][RS_ROWS][RS_PITCH]
    __shared__ int s_row[RS_ROWS];
    const int q = blockIdx.y, j0 = blockIdx.x * RS_ROWS;
    const int cnt = min(cand_cnt[q], cand_stride);
    if (j0 >= cand_stride) return;
    const float *qv = queries + (int64_t)q * dim;
    {   // query -> fp64 in shared memory; the loads of a thread are issued together, not one per loop trip
        float t[8];
        for (int i0 = 0; i0 < dimp; i0 += 8 * RESCORE_THREADS) {
#pragma unroll
            for (int u = 0; u < 8; ++u) { const int i = i0 + u * RESCORE_THREADS + threadIdx.x; t[u] = i < dim ? __ldg(qv + i) : 0.f; }
#pragma unroll
            for (int u = 0; u < 8; ++u) { const int i = i0 + u * RESCORE_THREADS + threadIdx.x; if (i < dimp) sq[i] = (double)(t[u]); }
        }
    }
    if (threadIdx.x < RS_ROWS) {
        int j = j0 + threadIdx.x;
        s_row[threadIdx.x] = (j < cnt) ? (int)(cand[(int64_t)q * cand_stride + j] & 0xFFFFFFFFull) : -1;
    }
    __syncthreads();
    const int nchunks = (dimp + RS_COLS - 1) / RS_COLS;
    auto issue = [&](int c) {
        if (c < nchunks) {
            const int c0 = c * RS_COLS, w4 = min(RS_COLS, dimp - c0) >> 2;
            float *dst = tile + (size_t)(c % RS_STAGES) * RS_ROWS * RS_PITCH;
#pragma unroll
            for (int u = 0; u < (RS_ROWS * RS_COLS / 4) / RESCORE_THREADS; ++u) {
                const int f = threadIdx.x + u * RESCORE_THREADS, rr = f / (RS_COLS / 4), c4 = f % (RS_COLS / 4);
                const int row = s_row[rr];
                if (row >= 0 && c4 < w4) cp_async16(dst + rr * RS_PITCH + c4 * 4, rows + (int64_t)row * ld + c0 + c4 * 4);
            }
        }
        cp_async_commit();
    };
    for (int c = 0; c < RS_STAGES - 1; ++c) issue(c);
    double acc = 0.0;
    // The chain warp differs from CTA to CTA: warp w sits on scheduler w % 4, and the F2F pipe of one scheduler
    // cannot feed the chains of every CTA resident on the SM.
    const int cw = (blockIdx.x + blockIdx.y) & 3, cl = threadIdx.x & 31;
    const bool chain = (threadIdx.x >> 5) == cw;
    for (int c = 0; c < nchunks; ++c) {
        cp_async_wait<RS_STAGES - 2>();
        __syncthreads();                                   // chunk c has landed; everyone is done with chunk c-1's stage
        issue(c + RS_STAGES - 1);
        if (chain) {
            const int c0 = c * RS_COLS, w = min(RS_COLS, dimp - c0);
            const float *e = tile + (size_t)(c % RS_STAGES) * RS_ROWS * RS_PITCH + cl * RS_PITCH;
            const double *qq = sq + c0;
            if (metric == 0) {
#pragma unroll 4
                for (int i = 0; i < w; i += 4) {
                    const float4 v = *reinterpret_cast<const float4 *>(e + i);
                    const double2 u0 = *reinterpret_cast<const double2 *>(qq + i), u1 = *reinterpret_cast<const double2 *>(qq + i + 2);
                    acc = fma(u0.x, (double)(v.x), acc); acc = fma(u0.y, (double)(v.y), acc);
                    acc = fma(u1.x, (double)(v.z), acc); acc = fma(u1.y, (double)(v.w), acc);
                }
            } else {
#pragma unroll 4
                for (int i = 0; i < w; i += 4) {
                    const float4 v = *reinterpret_cast<const float4 *>(e + i);
                    const double2 u0 = *reinterpret_cast<const double2 *>(qq + i), u1 = *reinterpret_cast<const double2 *>(qq + i + 2);
                    double d;
                    d = __dsub_rn(u0.x, (double)(v.x)); acc = __dadd_rn(acc, __dmul_rn(d, d));
                    d = __dsub_rn(u0.y, (double)(v.y)); acc = __dadd_rn(acc, __dmul_rn(d, d));
                    d = __dsub_rn(u1.x, (double)(v.z)); acc = __dadd_rn(acc, __dmul_rn(d, d));
                    d = __dsub_rn(u1.y, (double)(v.w)); acc = __dadd_rn(acc, __dmul_rn(d, d));
                }
            }
        }
    }
    if (chain && j0 + cl < cand_stride) {
        const int64_t o = (int64_t)q * cand_stride + j0 + cl;
        const int row = s_row[cl];
        if (row < 0) { exact_key[o] = ~0ull; exact_row[o] = MCP_PAD_ROW; }
        else {
            double s = exact_finish(acc, metric, nq2[q], ne2[row]);
            if (penalty) s = s + (double)penalty[row];
            exact_key[o] = f64_ord(s);
            exact_row[o] = row;
        }
    }
}

// ---- K4b: final sort + outputs + certificate; one CTA per query --------------------------------
__global__ void __launch_bounds__(FINAL_THREADS)
finalize_kernel(const u64 *__restrict__ exact_key, const int *__restrict__ exact_row,
                const u64 *__restrict__ cand, const int *__restrict__ cand_cnt, int cand_stride,
                int npow2, int k, int64_t id_base, const double *__restrict__ nq2,
                const unsigned char *__restrict__ qflag, const TableConsts *__restrict__ consts,
                CertParams cp, int32_t *__restrict__ out_ids, float *__restrict__ out_scores,
                double *__restrict__ out_scores64, int *__restrict__ flag_list, int *__restrict__ n_flag) {
    __shared__ u64 sk[FINAL_MAX];
    __shared__ int sv[FINAL_MAX];
    __shared__ unsigned int s_tmax;
    int q = blockIdx.x;
    int cnt = min(cand_cnt[q], cand_stride);
    if (threadIdx.x == 0) s_tmax = 0;
    __syncthreads();
    unsigned int tmax = 0;
    for (int i = threadIdx.x; i < npow2; i += blockDim.x) {
        if (i < cnt) {
            int64_t o = (int64_t)q * cand_stride + i;
            sk[i] = exact_key[o]; sv[i] = exact_row[o];
            tmax = max(tmax, (unsigned int)(cand[o] >> 32));
        } else { sk[i] = ~0ull; sv[i] = MCP_PAD_ROW; }
    }
    atomicMax(&s_tmax, tmax);
    cta_bitonic_sort_pairs(sk, sv, npow2);
    for (int i = threadIdx.x; i < k; i += blockDim.x) {
        int64_t o = (int64_t)q * k + i;
        bool real = (i < cnt) && sv[i] != MCP_PAD_ROW;
        double s = real ? ord_f64(sk[i]) : CUDART_INF;
        out_ids[o] = real ? (int32_t)(id_base + sv[i]) : -1;
        out_scores[o] = (float)s;
        if (out_scores64) out_scores64[o] = s;
    }
    if (threadIdx.x == 0) {
        bool pass;
        if (qflag[q]) pass = false;                          // degenerate query: always the fp64 scan
        else if (cp.complete) pass = true;
        else if (s_tmax == 0xFFFFFFFFu) pass = true;         // worst survivor is NaN: every dropped row is NaN with a larger id
        else if (cnt < k) pass = false;                      // cannot happen when !complete; be safe
        else {
            double T = (double)ord_f32(s_tmax);
            double sk_k = ord_f64(sk[k - 1]);                // NaN if fewer than k numeric scores
            double qn = sqrt(nq2[q]);
            double en = sqrt(__longlong_as_double((long long)consts->max_norm2_bits));
            double pen_abs = (double)consts->pen_absmax, pen_max = (double)consts->pen_max;
            double eps;
            if (cp.metric == 0) {
                eps = cp.rel + cp.abs_pen * (2.0 + pen_abs);
            } else if (cp.path == 2) {
                // squared distance assembled from norms and a TF32 dot: |d2c - d2| <= delta
                double delta = 2.0 * cp.rel * qn * en + 4.0 * 5.96e-8 * (qn + en) * (qn + en);
                double x0 = T - pen_max; x0 = x0 > 0.0 ? x0 * x0 : 0.0;   // every dropped row has coarse d2 >= x0
                eps = sqrt(x0) - sqrt(fmax(x0 - delta, 0.0)) + cp.abs_pen * (sqrt(x0) + pen_abs);
            } else {
                eps = cp.rel * (qn + en) + cp.abs_pen * ((qn + en) + pen_abs);
            }
            pass = (sk_k == sk_k) && (sk_k < T - eps);
        }
        if (!pass) flag_list[atomicAdd(n_flag, 1)] = q;
    }
}

// ---- K4 fused: exact rescoring + query norm + final sort + certificate; one CTA per query ------------------------
// Replaces prep_queries_kernel + rescore_kernel + finalize_kernel on the search path (three launches, two global round
// trips).  The CTA has GP candidate warps (32 candidates each; more than RF_MAXG * 32 candidates take several passes)
// and one extra warp whose lane 0 walks the query against itself: that chain *is* the exact fp64 squared norm of the
// query, in index order, and it is exactly as long as a candidate's chain, so it costs nothing on the critical path.
// All threads stream the candidate rows through the cp.async ring; then the CTA sorts, writes the top k and checks
// the certificate.  nq2 / qflag are also written out for the fp64 fallback scan that may follow.
#define RF_MAXG 4
#define RF_STAGES 4
#define RF_SMEM_BYTES(dim, gp) (sizeof(double) * (((dim) + 3) & ~3) + sizeof(float) * RF_STAGES * (gp) * RS_ROWS * RS_PITCH + 16)

__global__ void __launch_bounds__(32 * (RF_MAXG + 1))
rescore_finalize_kernel(const float *__restrict__ rows, int dim, int64_t ld, const float *__restrict__ queries,
                        const double *__restrict__ ne2, const float *__restrict__ penalty, int metric,
                        const u64 *__restrict__ cand, const int *__restrict__ cand_cnt, int cand_stride,
                        int npow2, int k, int64_t id_base, const TableConsts *__restrict__ consts, CertParams cp,
                        double *__restrict__ nq2_out, unsigned char *__restrict__ qflag_out,
                        int32_t *__restrict__ out_ids, float *__restrict__ out_scores, double *__restrict__ out_scores64,
                        int *__restrict__ flag_list, int *__restrict__ n_flag) {
    extern __shared__ __align__(16) unsigned char rf_raw[];
    __shared__ u64 sk[FINAL_MAX];
    __shared__ int sv[FINAL_MAX];
    __shared__ int s_row[RF_MAXG * RS_ROWS];
    __shared__ unsigned int s_tmax;
    __shared__ int s_bad;
    __shared__ double s_nq2;
    const int nthreads = blockDim.x, GP = nthreads / 32 - 1;       // candidate warps per pass
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int dimp = (dim + 3) & ~3;
    double *sq = reinterpret_cast<double *>(rf_raw);                // [dimp] the query in fp64
    float *tile = reinterpret_cast<float *>(sq + dimp);             // [RF_STAGES][GP * RS_ROWS][RS_PITCH]
    const int q = blockIdx.x;
    const int cnt = min(cand_cnt[q], cand_stride);
    const float *qv = queries + (int64_t)q * dim;
    if (threadIdx.x == 0) { s_tmax = 0; s_bad = 0; }
    __syncthreads();
    {   // query -> fp64 in shared memory (loads batched), non-finite detection
        bool bad = false;
        float t[8];
        for (int i0 = 0; i0 < dimp; i0 += 8 * nthreads) {
#pragma unroll
            for (int u = 0; u < 8; ++u) { const int i = i0 + u * nthreads + threadIdx.x; t[u] = i < dim ? __ldg(qv + i) : 0.f; }
#pragma unroll
            for (int u = 0; u < 8; ++u) { const int i = i0 + u * nthreads + threadIdx.x; if (i < dimp) { sq[i] = (double)t[u]; bad |= !isfinite(t[u]); } }
        }
        if (bad) s_bad = 1;
    }
    unsigned int tmax = 0;
    for (int i = threadIdx.x; i < npow2; i += nthreads) {
        sk[i] = ~0ull; sv[i] = MCP_PAD_ROW;
        if (i < cnt) tmax = max(tmax, (unsigned int)(cand[(int64_t)q * cand_stride + i] >> 32));
    }
    atomicMax(&s_tmax, tmax);
    const int nchunks = (dimp + RS_COLS - 1) / RS_COLS;
    const int n_groups = (cnt + RS_ROWS - 1) / RS_ROWS;
    double nq2_acc = 0.0;                                           // lane 0 of the extra warp
    for (int g0 = 0; g0 < max(n_groups, 1); g0 += GP) {
        __syncthreads();                                            // previous pass has left the ring and s_row
        if (warp < GP) {
            const int j = (g0 + warp) * RS_ROWS + lane;
            s_row[warp * RS_ROWS + lane] = (j < cnt) ? (int)(cand[(int64_t)q * cand_stride + j] & 0xFFFFFFFFull) : -1;
        }
        __syncthreads();
        auto issue = [&](int c) {
            if (c < nchunks) {
                const int c0 = c * RS_COLS, w4 = min(RS_COLS, dimp - c0) >> 2;
                float *dst = tile + (size_t)(c % RF_STAGES) * GP * RS_ROWS * RS_PITCH;
                for (int f = threadIdx.x; f < GP * RS_ROWS * (RS_COLS / 4); f += nthreads) {
                    const int rr = f / (RS_COLS / 4), c4 = f % (RS_COLS / 4);
                    const int row = s_row[rr];
                    if (row >= 0 && c4 < w4) cp_async16(dst + rr * RS_PITCH + c4 * 4, rows + (int64_t)row * ld + c0 + c4 * 4);
                }
            }
            cp_async_commit();
        };
        for (int c = 0; c < RF_STAGES - 1; ++c) issue(c);
        double acc = 0.0;
        for (int c = 0; c < nchunks; ++c) {
            cp_async_wait<RF_STAGES - 2>();
            __syncthreads();                                        // chunk c has landed; everyone is done with chunk c-1's stage
            issue(c + RF_STAGES - 1);
            const int c0 = c * RS_COLS, w = min(RS_COLS, dimp - c0);
            const double *qq = sq + c0;
            if (warp < GP) {
                const float *e = tile + (size_t)(c % RF_STAGES) * GP * RS_ROWS * RS_PITCH + (warp * RS_ROWS + lane) * RS_PITCH;
                if (metric == 0) {
#pragma unroll 4
                    for (int i = 0; i < w; i += 4) {
                        const float4 v = *reinterpret_cast<const float4 *>(e + i);
                        const double2 u0 = *reinterpret_cast<const double2 *>(qq + i), u1 = *reinterpret_cast<const double2 *>(qq + i + 2);
                        acc = fma(u0.x, (double)v.x, acc); acc = fma(u0.y, (double)v.y, acc);
                        acc = fma(u1.x, (double)v.z, acc); acc = fma(u1.y, (double)v.w, acc);
                    }
                } else {
#pragma unroll 4
                    for (int i = 0; i < w; i += 4) {
                        const float4 v = *reinterpret_cast<const float4 *>(e + i);
                        const double2 u0 = *reinterpret_cast<const double2 *>(qq + i), u1 = *reinterpret_cast<const double2 *>(qq + i + 2);
                        double d;
                        d = __dsub_rn(u0.x, (double)v.x); acc = __dadd_rn(acc, __dmul_rn(d, d));
                        d = __dsub_rn(u0.y, (double)v.y); acc = __dadd_rn(acc, __dmul_rn(d, d));
                        d = __dsub_rn(u1.x, (double)v.z); acc = __dadd_rn(acc, __dmul_rn(d, d));
                        d = __dsub_rn(u1.y, (double)v.w); acc = __dadd_rn(acc, __dmul_rn(d, d));
                    }
                }
            } else if (g0 == 0 && lane == 0) {
                // the query against itself: sum q_i^2 in index order == the oracle's nq2 (exact products)
#pragma unroll 4
                for (int i = 0; i < w; i += 2) {
                    const double2 u = *reinterpret_cast<const double2 *>(qq + i);
                    nq2_acc = fma(u.x, u.x, nq2_acc); nq2_acc = fma(u.y, u.y, nq2_acc);
                }
            }
        }
        if (g0 == 0 && warp == GP && lane == 0) s_nq2 = nq2_acc;
        __syncthreads();
        if (warp < GP) {
            const int j = (g0 + warp) * RS_ROWS + lane;
            const int row = s_row[warp * RS_ROWS + lane];
            if (j < npow2 && row >= 0) {
                double sc = exact_finish(acc, metric, s_nq2, ne2[row]);
                if (penalty) sc = sc + (double)penalty[row];
                sk[j] = f64_ord(sc); sv[j] = row;
            }
        }
    }
    cta_bitonic_sort_pairs(sk, sv, npow2);
    for (int i = threadIdx.x; i < k; i += nthreads) {
        const int64_t o = (int64_t)q * k + i;
        const bool real = (i < cnt) && sv[i] != MCP_PAD_ROW;
        const double s = real ? ord_f64(sk[i]) : CUDART_INF;
        out_ids[o] = real ? (int32_t)(id_base + sv[i]) : -1;
        out_scores[o] = (float)s;
        if (out_scores64) out_scores64[o] = s;
    }
    if (threadIdx.x == 0) {
        const double nq2 = s_nq2;
        const bool odd = (nq2 != 0.0) && (nq2 < 1e-30 || nq2 > 1e30);
        const unsigned char flag = (unsigned char)((s_bad ? 2 : 0) | ((odd || nq2 == 0.0 || s_bad) ? 1 : 0));
        nq2_out[q] = nq2;
        qflag_out[q] = flag;
        bool pass;
        if (flag) pass = false;                              // degenerate query: always the fp64 scan
        else if (cp.complete) pass = true;
        else if (s_tmax == 0xFFFFFFFFu) pass = true;         // worst survivor is NaN: every dropped row is NaN with a larger id
        else if (cnt < k) pass = false;                      // cannot happen when !complete; be safe
        else {
            const double T = (double)ord_f32(s_tmax);
            const double sk_k = ord_f64(sk[k - 1]);          // NaN if fewer than k numeric scores
            const double qn = sqrt(nq2);
            const double en = sqrt(__longlong_as_double((long long)consts->max_norm2_bits));
            const double pen_abs = (double)consts->pen_absmax, pen_max = (double)consts->pen_max;
            double eps;
            if (cp.metric == 0) {
                eps = cp.rel + cp.abs_pen * (2.0 + pen_abs);
            } else if (cp.path == 2) {
                const double delta = 2.0 * cp.rel * qn * en + 4.0 * 5.96e-8 * (qn + en) * (qn + en);
                double x0 = T - pen_max; x0 = x0 > 0.0 ? x0 * x0 : 0.0;
                eps = sqrt(x0) - sqrt(fmax(x0 - delta, 0.0)) + cp.abs_pen * (sqrt(x0) + pen_abs);
            } else {
                eps = cp.rel * (qn + en) + cp.abs_pen * ((qn + en) + pen_abs);
            }
            pass = (sk_k == sk_k) && (sk_k < T - eps);
        }
        if (!pass) flag_list[atomicAdd(n_flag, 1)] = q;
    }
}

// ---- K4c: fp64 scan of every row for the flagged queries ---------------------------------------
// grid (blocks, max_fallback); slot f serves query flag_list[slot0 + f]; slots past *n_flag exit.
__global__ void __launch_bounds__(256)
exact_scan_kernel(const float *__restrict__ rows, int64_t n_rows, int dim, int64_t ld,
                  const float *__restrict__ queries, const double *__restrict__ nq2,
                  const double *__restrict__ ne2, const float *__restrict__ penalty, int metric,
                  const int *__restrict__ flag_list, const int *__restrict__ n_flag, int slot0,
                  u64 *__restrict__ exact_all /* [slots][n_rows] */) {
    extern __shared__ float sqv[];
    int f = blockIdx.y;
    if (slot0 + f >= *n_flag) return;
    int q = flag_list[slot0 + f];
    const float *qv = queries + (int64_t)q * dim;
    for (int i = threadIdx.x; i < dim; i += blockDim.x) sqv[i] = qv[i];
    __syncthreads();
    double qn2 = nq2[q];
    u64 *out = exact_all + (int64_t)f * n_rows;
    int d4 = dim & ~3;
    for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < n_rows; r += (int64_t)gridDim.x * blockDim.x) {
        const float *e = rows + r * ld;
        ExactAcc a; a.acc = 0.0;
        if (metric == 0) {
            for (int i = 0; i < d4; i += 4) {
                float4 v = *reinterpret_cast<const float4 *>(e + i);
                a.cos_step(sqv[i], v.x); a.cos_step(sqv[i + 1], v.y);
                a.cos_step(sqv[i + 2], v.z); a.cos_step(sqv[i + 3], v.w);
            }
            for (int i = d4; i < dim; ++i) a.cos_step(sqv[i], e[i]);
        } else {
            for (int i = 0; i < d4; i += 4) {
                float4 v = *reinterpret_cast<const float4 *>(e + i);
                a.l2_step(sqv[i], v.x); a.l2_step(sqv[i + 1], v.y);
                a.l2_step(sqv[i + 2], v.z); a.l2_step(sqv[i + 3], v.w);
            }
            for (int i = d4; i < dim; ++i) a.l2_step(sqv[i], e[i]);
        }
        double s = exact_finish(a.acc, metric, qn2, ne2[r]);
        if (penalty) s = s + (double)penalty[r];
        out[r] = f64_ord(s);
    }
}

// ---- K4d: exact top-k of one fp64 score vector; one CTA per slot ---------------------------------
// Radix select on the 64-bit score image (values may repeat), then rows with the boundary value
// are taken in row order, which is the "ties -> lowest id" rule.
__global__ void __launch_bounds__(EXSEL_THREADS)
exact_select_kernel(const u64 *__restrict__ exact_all, int64_t n_rows, int k, int64_t id_base,
                    const int *__restrict__ flag_list, int *__restrict__ n_flag, int slot0,
                    int32_t *__restrict__ out_ids, float *__restrict__ out_scores,
                    double *__restrict__ out_scores64, int *__restrict__ status /* NULL, or [failures, unresolved, ticket] */) {
    __shared__ uint32_t hist[256];
    __shared__ int s_digit, s_cnt, s_warp_base[EXSEL_THREADS / 32 + 1], s_taken, s_nless;
    __shared__ long long s_need;
    __shared__ u64 sk[MCP_MAX_K_POW2];
    __shared__ int sv[MCP_MAX_K_POW2];
    int f = blockIdx.x;
    const int nf = *n_flag;
    // the last CTA to get here folds this batch's flag count into the running status and clears it for the next batch
    // (every CTA has read n_flag by then: the ticket is taken after the read)
    if (status && threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(&status[2], 1) == (int)gridDim.x - 1) {
            status[0] += nf;
            status[1] += nf > (int)gridDim.x ? nf - (int)gridDim.x : 0;
            status[2] = 0;
            *n_flag = 0;
        }
    }
    if (slot0 + f >= nf) return;
    int q = flag_list[slot0 + f];
    const u64 *E = exact_all + (int64_t)f * n_rows;
    int kk = (int)min((int64_t)k, n_rows);
    u64 prefix = 0, mask = 0, thr = ~0ull;
    long long need = kk;
    bool whole_bin = (kk >= n_rows);
    if (!whole_bin) {
        for (int shift = 56; shift >= 0; shift -= 8) {
            for (int i = threadIdx.x; i < 256; i += blockDim.x) hist[i] = 0;
            __syncthreads();
            for (int64_t i = threadIdx.x; i < n_rows; i += blockDim.x) {
                u64 key = E[i];
                if ((key & mask) == prefix) atomicAdd(&hist[(uint32_t)(key >> shift) & 255u], 1u);
            }
            __syncthreads();
            if (threadIdx.x == 0) {
                long long run = 0;
                for (int d = 0; d < 256; ++d) {
                    if (need <= run + (long long)hist[d]) { s_digit = d; s_need = need - run; s_cnt = (int)hist[d]; break; }
                    run += hist[d];
                }
            }
            __syncthreads();
            prefix |= (u64)s_digit << shift;
            mask |= 0xFFull << shift;
            need = s_need;
            if ((long long)s_cnt == need) { thr = prefix | (shift ? ((1ull << shift) - 1ull) : 0ull); whole_bin = true; break; }
            __syncthreads();
        }
        if (!whole_bin) thr = prefix;    // boundary value repeats: `need` of its rows are wanted, lowest rows first
    }
    for (int i = threadIdx.x; i < MCP_MAX_K_POW2; i += blockDim.x) { sk[i] = ~0ull; sv[i] = MCP_PAD_ROW; }
    if (threadIdx.x == 0) { s_nless = 0; s_taken = 0; }
    __syncthreads();
    // pass 1: everything strictly better than the boundary (or <= when the whole bin is taken)
    for (int64_t i = threadIdx.x; i < n_rows; i += blockDim.x) {
        u64 key = E[i];
        if (whole_bin ? (key <= thr) : (key < thr)) {
            int p = atomicAdd(&s_nless, 1);
            sk[p] = key; sv[p] = (int)i;
        }
    }
    __syncthreads();
    if (!whole_bin) {
        // pass 2: boundary-valued rows in row order until `need` are taken
        int base_out = s_nless;
        int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        for (int64_t c0 = 0; c0 < n_rows; c0 += blockDim.x) {
            int64_t i = c0 + threadIdx.x;
            bool hit = (i < n_rows) && (E[i] == thr);
            unsigned int b = __ballot_sync(MCP_FULL, hit);
            if (lane == 0) s_warp_base[warp] = __popc(b);
            __syncthreads();
            if (threadIdx.x == 0) {
                int run = s_taken;
                for (int w = 0; w < EXSEL_THREADS / 32; ++w) { int c = s_warp_base[w]; s_warp_base[w] = run; run += c; }
                s_taken = run;
            }
            __syncthreads();
            if (hit) {
                int p = s_warp_base[warp] + __popc(b & ((1u << lane) - 1u));
                if (p < need) { sk[base_out + p] = thr; sv[base_out + p] = (int)i; }
            }
            bool done = s_taken >= need;
            __syncthreads();
            if (done) break;
        }
    }
    cta_bitonic_sort_pairs(sk, sv, MCP_MAX_K_POW2);
    for (int i = threadIdx.x; i < k; i += blockDim.x) {
        int64_t o = (int64_t)q * k + i;
        bool real = sv[i] != MCP_PAD_ROW;
        double s = real ? ord_f64(sk[i]) : CUDART_INF;
        out_ids[o] = real ? (int32_t)(id_base + sv[i]) : -1;
        out_scores[o] = (float)s;
        if (out_scores64) out_scores64[o] = s;
    }
}
