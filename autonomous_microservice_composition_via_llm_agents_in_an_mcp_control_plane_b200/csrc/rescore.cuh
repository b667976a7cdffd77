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
    double rel;            // relative coarse error bound of the dot product / squared distance that does not depend on the query
    int kind;              // tensor-core scans: 1 tf32, 2 fp16 operands (0: the fp32 streaming scan, `rel` is the whole bound)
    double abs_pen;        // multiplier of pen_absmax for fp32 rounding of the final add
    int debug;             // option gemm_debug >> 8 (timing experiments, wrong results): 1 no chain, 2 no gathers
};

#define RS_ROWS 32                    // candidates per warp (lane = candidate)
__device__ __forceinline__ void cp_async16(void *dst, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// ---- K4 fused: exact rescoring + query norm + final sort + certificate; one CTA per query ------------------------
// Replaces prep_queries_kernel + rescore_kernel + finalize_kernel on the search path (three launches, two global round
// trips).  Warp roles (GP + RF_NPROD + 1 warps):
//   * GP candidate warps, lane = candidate (more than GP * 32 candidates take several passes): wait for a chunk of
//     64 columns on the stage's `full` mbarrier, walk it in index order, release the stage on `empty`.  They issue
//     nothing but their chain -- an ncu capture of the version where all threads shared the copy loop showed that
//     loop costing as much as the chains, and the chain (about 15 cycles per element) is the kernel's floor.
//   * RF_NPROD producer warps: 16 lanes per row issue cp.async (16 bytes each, 256 contiguous bytes per row and chunk) into
//     a ring of RF_STAGES chunk tiles and hand completion to `full` (cp.async.mbarrier.arrive).  256-byte
//     cp.async.bulk copies were tried here and ran at about one copy per 44 cycles per SM -- 1.5x slower overall.
//   * one warp whose lane 0 walks the query against itself: that chain *is* the exact fp64 squared norm of the
//     query, in index order, and it is as long as a candidate's chain, so it costs nothing on the critical path.
// Then the CTA sorts, writes the top k and checks the certificate.  nq2 / qflag are also written out for the fp64
// fallback scan that may follow.
#define RF_MAXG 4
#define RF_COLS 64                    // columns per chunk: 256 contiguous bytes of a row per gather.  tools/gather_bench.cu: alone, 256-byte
                                      // pieces gather at 4.5 TB/s over the GPU and 512-byte pieces at 6.1 TB/s, but inside this kernel
                                      // 512-byte pieces (3 stages to keep two CTAs per SM) measured slower: 45.6 vs 35.8 us
#define RF_PITCH (RF_COLS + 4)
#define RF_NPROD 2                     // producer warps (each keeps its own copies in flight)
// ring shapes (template arguments): <stages, chunks of gathers in flight behind the one being handed over>
//   <5, 3>  87 KB at GP = 2 -> two CTAs per SM.  A deeper ring for the one-CTA-per-SM case (<10, 8>) measured no faster.
// dynamic shared memory: query in fp64 | ring | final sort keys (u64) and rows (int), npow2 each
#define RF_SMEM_BYTES(dim, gp, stages, npow2) (sizeof(double) * (((dim) + 3) & ~3) + sizeof(float) * (stages) * (gp) * RS_ROWS * RF_PITCH + 64 + 12 * (size_t)(npow2))
#define RF_SMEM_LIMIT (216 * 1024)    // dynamic shared memory the kernel may ask for (227 KB per CTA less its static arrays)

// One candidate's walk over one chunk (w columns, a multiple of 4), in index order.  The compiler left to itself puts
// each LDS right before the conversion that needs it, and each conversion right before its FMA, so every group of four
// columns pays the shared-memory and conversion latencies on top of the four dependent FMAs (29 cycles per column
// measured in the kernel).  Here the loads run two groups ahead of the chain and the conversions one group ahead, in
// registers, so the chain sees only the FMA -> FMA dependency.  The look-ahead reads up to 11 floats / doubles past
// the chunk: still inside the CTA's shared memory (next row of the tile, the tile behind the query), never used.
template <int METRIC>
__device__ __forceinline__ double rf_chain(const float *__restrict__ e, const double *__restrict__ qq, int w, double acc) {
    float4 v2 = *reinterpret_cast<const float4 *>(e + 4);
    double2 a2 = *reinterpret_cast<const double2 *>(qq + 4), b2 = *reinterpret_cast<const double2 *>(qq + 6);
    double2 a1 = *reinterpret_cast<const double2 *>(qq), b1 = *reinterpret_cast<const double2 *>(qq + 2);
    double d0, d1, d2, d3;
    { const float4 v1 = *reinterpret_cast<const float4 *>(e); d0 = (double)v1.x; d1 = (double)v1.y; d2 = (double)v1.z; d3 = (double)v1.w; }
#pragma unroll 2
    for (int i = 0; i < w; i += 4) {
        const float4 v3 = *reinterpret_cast<const float4 *>(e + i + 8);
        const double2 a3 = *reinterpret_cast<const double2 *>(qq + i + 8), b3 = *reinterpret_cast<const double2 *>(qq + i + 10);
        double n0, n1, n2, n3;     // volatile: keeps the conversions here, a whole group ahead of their FMAs (nvcc sinks plain casts to their use)
        asm volatile("cvt.f64.f32 %0, %1;" : "=d"(n0) : "f"(v2.x));
        asm volatile("cvt.f64.f32 %0, %1;" : "=d"(n1) : "f"(v2.y));
        asm volatile("cvt.f64.f32 %0, %1;" : "=d"(n2) : "f"(v2.z));
        asm volatile("cvt.f64.f32 %0, %1;" : "=d"(n3) : "f"(v2.w));
        if (METRIC == 0) {
            acc = fma(a1.x, d0, acc); acc = fma(a1.y, d1, acc); acc = fma(b1.x, d2, acc); acc = fma(b1.y, d3, acc);
        } else {
            double t;
            t = __dsub_rn(a1.x, d0); acc = __dadd_rn(acc, __dmul_rn(t, t));
            t = __dsub_rn(a1.y, d1); acc = __dadd_rn(acc, __dmul_rn(t, t));
            t = __dsub_rn(b1.x, d2); acc = __dadd_rn(acc, __dmul_rn(t, t));
            t = __dsub_rn(b1.y, d3); acc = __dadd_rn(acc, __dmul_rn(t, t));
        }
        d0 = n0; d1 = n1; d2 = n2; d3 = n3;
        a1 = a2; b1 = b2; a2 = a3; b2 = b3; v2 = v3;
    }
    return acc;
}

template <int RF_STAGES, int RF_DEPTH>
__global__ void __launch_bounds__(32 * (RF_MAXG + RF_NPROD + 1))
rescore_finalize_kernel(const float *__restrict__ rows, int dim, int64_t ld, const float *__restrict__ queries,
                        const double *__restrict__ ne2, const float *__restrict__ penalty, int metric,
                        const u64 *__restrict__ cand, const int *__restrict__ cand_cnt, int cand_stride,
                        int npow2, int k, int64_t id_base, const TableConsts *__restrict__ consts, CertParams cp,
                        double *__restrict__ nq2_out, unsigned char *__restrict__ qflag_out,
                        int32_t *__restrict__ out_ids, float *__restrict__ out_scores, double *__restrict__ out_scores64,
                        int *__restrict__ flag_list, int *__restrict__ n_flag,
                        const unsigned int *__restrict__ drop_min /* NULL, or [q]: best coarse score the scan dropped outside its lists */,
                        const float *__restrict__ rho_q /* NULL, or [q][2]: the query's rounding residual ratios (tf32, fp16) */,
                        u64 *__restrict__ bound_key, int *__restrict__ bound_row /* [q]: for a flagged query, the k-th best exact (score, row)
                                                                                    pair among its candidates -- what the fp64 scan filters with */) {
    extern __shared__ __align__(16) unsigned char rf_raw[];
    __shared__ int s_row[RF_MAXG * RS_ROWS];
    __shared__ const float *s_rowptr[RF_MAXG * RS_ROWS];
    __shared__ __align__(8) u64 s_full[RF_STAGES], s_empty[RF_STAGES];
    __shared__ unsigned int s_tmax;
    __shared__ int s_bad;
    __shared__ double s_nq2;
    const int nthreads = blockDim.x, GP = nthreads / 32 - RF_NPROD - 1;   // candidate warps per pass
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int dimp = (dim + 3) & ~3;
    double *sq = reinterpret_cast<double *>(rf_raw);                // [dimp] the query in fp64
    float *tile = reinterpret_cast<float *>(sq + dimp);             // [RF_STAGES][GP * RS_ROWS][RF_PITCH]
    const size_t stage_floats = (size_t)GP * RS_ROWS * RF_PITCH;
    u64 *sk = reinterpret_cast<u64 *>(tile + RF_STAGES * stage_floats + 16);   // [npow2] (8-byte aligned: every term above is)
    int *sv = reinterpret_cast<int *>(sk + npow2);                             // [npow2]
    const int q = blockIdx.x;
    pdl_trigger();
    pdl_wait();                                                     // the selection that wrote the candidate lists has completed
    const int cnt = min(cand_cnt[q], cand_stride);
    const float *qv = queries + (int64_t)q * dim;
    if (threadIdx.x == 0) {
        s_tmax = 0; s_bad = 0;
        for (int s = 0; s < RF_STAGES; ++s) { mbar_init(&s_full[s], 32 * RF_NPROD); mbar_init(&s_empty[s], GP); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    const int nchunks = (dimp + RF_COLS - 1) / RF_COLS;
    const int n_groups = (cnt + RS_ROWS - 1) / RS_ROWS;
    auto load_rows = [&](int g0) {                                  // candidate warps: this pass's rows
        if (warp < GP) {
            const int j = (g0 + warp) * RS_ROWS + lane;
            const int row = (j < cnt) ? (int)(cand[(int64_t)q * cand_stride + j] & 0xFFFFFFFFull) : -1;
            s_row[warp * RS_ROWS + lane] = row;
            s_rowptr[warp * RS_ROWS + lane] = rows + (int64_t)max(row, 0) * ld;
        }
    };
    load_rows(0);
    __syncthreads();                                                // barriers initialised, first pass's rows known
    const bool producer = warp >= GP && warp < GP + RF_NPROD;
    if (!producer) {
        // Everyone but the producer (which is already gathering rows) brings the query in: fp64 in shared memory
        // (loads batched), non-finite detection, and the worst surviving coarse score for the certificate.
        const int nt = nthreads - 32 * RF_NPROD, t = threadIdx.x - (warp > GP ? 32 * RF_NPROD : 0);
        bool bad = false;
        float f[8];
        for (int i0 = 0; i0 < dimp; i0 += 8 * nt) {
#pragma unroll
            for (int u = 0; u < 8; ++u) { const int i = i0 + u * nt + t; f[u] = i < dim ? __ldg(qv + i) : 0.f; }
#pragma unroll
            for (int u = 0; u < 8; ++u) { const int i = i0 + u * nt + t; if (i < dimp) { sq[i] = (double)f[u]; bad |= !isfinite(f[u]); } }
        }
        if (bad) s_bad = 1;
        unsigned int tmax = 0;
        for (int i = t; i < npow2; i += nt) {
            sk[i] = ~0ull; sv[i] = MCP_PAD_ROW;
            if (i < cnt) tmax = max(tmax, (unsigned int)(cand[(int64_t)q * cand_stride + i] >> 32));
        }
        atomicMax(&s_tmax, tmax);
        asm volatile("bar.sync 1, %0;" ::"r"(nt) : "memory");       // the query is complete for its readers
    }
    double nq2_acc = 0.0;                                           // lane 0 of the last warp
    int it0 = 0;                                                    // chunks issued by earlier passes: the ring's phases run on
    for (int g0 = 0; g0 < max(n_groups, 1); g0 += GP, it0 += nchunks) {
        if (g0 > 0) {
            __syncthreads();                                        // previous pass has left s_row
            load_rows(g0);
            __syncthreads();
        }
        double acc = 0.0, ne2_row = 0.0, pen_row = 0.0;
        if (producer) {
            // ---- producer: lanes 0-15 take even rows, 16-31 odd rows; lane & 15 = which 16 bytes of the 256
            const int n_valid = max(0, min(cnt - g0 * RS_ROWS, GP * RS_ROWS));     // valid rows are a prefix of s_row
            constexpr int LPR = RF_COLS / 4, RPI = 32 / LPR;               // lanes per row, rows per iteration
            const int half = lane / LPR, c4 = (lane % LPR) * 4;
            auto issue = [&](int c) {
                if (c >= nchunks) return;
                const int it = it0 + c, s = it % RF_STAGES, n = it / RF_STAGES;
                const int c0 = c * RF_COLS, w = min(RF_COLS, dimp - c0);
                if (n > 0) mbar_wait(&s_empty[s], (n - 1) & 1);
                float *dst = tile + (size_t)s * stage_floats + c4;
                if (c4 < w && !(cp.debug & 2)) {
#pragma unroll 4
                    for (int rr = half + RPI * (warp - GP); rr < n_valid; rr += RPI * RF_NPROD) cp_async16(dst + rr * RF_PITCH, s_rowptr[rr] + c0 + c4);
                }
                cp_async_commit();
            };
            // Completion is relayed by hand: the warp keeps RF_DEPTH chunks in flight (wait_group's watermark), and once
            // chunk c - RF_DEPTH is known complete every producer lane arrives on its `full` barrier (release; the
            // candidate warps' try_wait acquires).  cp.async.mbarrier.arrive.noinc gave the same timing.
            for (int c = 0; c < nchunks + RF_DEPTH; ++c) {
                if (c < nchunks) issue(c); else cp_async_commit();
                if (c >= RF_DEPTH) {
                    cp_async_wait<RF_DEPTH>();
                    mbar_arrive(&s_full[(it0 + c - RF_DEPTH) % RF_STAGES]);
                }
            }
        } else if (warp == GP + RF_NPROD) {
            if (g0 == 0 && lane == 0) {
                // the query against itself: sum q_i^2 in index order == the oracle's nq2 (exact products)
#pragma unroll 8
                for (int i = 0; i < dimp; i += 2) {
                    const double2 u = *reinterpret_cast<const double2 *>(sq + i);
                    nq2_acc = fma(u.x, u.x, nq2_acc); nq2_acc = fma(u.y, u.y, nq2_acc);
                }
                s_nq2 = nq2_acc;
            }
        } else {
            // ---- candidate warps: lane = candidate, columns in index order
            const int my_row = s_row[warp * RS_ROWS + lane];
            const bool live = my_row >= 0;
            if (live) { ne2_row = ne2[my_row]; pen_row = penalty ? (double)penalty[my_row] : 0.0; }   // in flight under the first chunk
            for (int c = 0; c < nchunks; ++c) {
                const int it = it0 + c, s = it % RF_STAGES, n = it / RF_STAGES;
                mbar_wait(&s_full[s], n & 1);
                const int c0 = c * RF_COLS, w = min(RF_COLS, dimp - c0);
                const double *qq = sq + c0;
                const float *e = tile + (size_t)s * stage_floats + (warp * RS_ROWS + lane) * RF_PITCH;
                if (live && !(cp.debug & 1)) acc = (metric == 0) ? rf_chain<0>(e, qq, w, acc) : rf_chain<1>(e, qq, w, acc);
                __syncwarp();
                if (lane == 0) mbar_arrive(&s_empty[s]);
            }
        }
        __syncthreads();
        if (warp < GP) {
            const int j = (g0 + warp) * RS_ROWS + lane;
            const int row = s_row[warp * RS_ROWS + lane];
            if (j < npow2 && row >= 0) {
                double sc = exact_finish(acc, metric, s_nq2, ne2_row);
                if (penalty) sc = sc + pen_row;
                sk[j] = f64_ord(sc); sv[j] = row;
            }
        }
    }
    if (npow2 <= 2 * 32 * 3) {
        // small lists: rank sort (every thread counts the entries ahead of its own; no barrier per stage, where the
        // bitonic network below takes 21 barriers for 64 entries).  Pads are identical, so the index breaks ties.
        __syncthreads();
        u64 mk[2]; int mv[2], rank[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const int i = threadIdx.x + u * nthreads;
            rank[u] = 0;
            if (i < npow2) {
                mk[u] = sk[i]; mv[u] = sv[i];
                for (int j = 0; j < npow2; ++j) {
                    const u64 kj = sk[j]; const int vj = sv[j];
                    rank[u] += (kj < mk[u] || (kj == mk[u] && (vj < mv[u] || (vj == mv[u] && j < i)))) ? 1 : 0;
                }
            }
        }
        __syncthreads();
#pragma unroll
        for (int u = 0; u < 2; ++u)
            if (threadIdx.x + u * nthreads < npow2) { sk[rank[u]] = mk[u]; sv[rank[u]] = mv[u]; }
        __syncthreads();
    } else {
        cta_bitonic_sort_pairs(sk, sv, npow2);
    }
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
        // every row outside the candidate list scored (coarsely) at least T: the worst survivor, or -- when the scan's
        // bootstrap tile dropped rows without listing them -- the best score such a row can have had
        const unsigned int dropped = drop_min ? drop_min[q] : 0xFFFFFFFFu;
        const unsigned int tmax_eff = min(s_tmax, dropped);
        bool pass;
        if (flag) pass = false;                              // degenerate query: always the fp64 scan
        else if (cp.complete && dropped == 0xFFFFFFFFu) pass = true;
        else if (tmax_eff == 0xFFFFFFFFu) pass = true;       // worst survivor is NaN: every dropped row is NaN with a larger id
        else if (cnt < k) pass = false;                      // cannot happen when !complete; be safe
        else {
            const double T = (double)ord_f32(tmax_eff);
            const double sk_k = ord_f64(sk[k - 1]);          // NaN if fewer than k numeric scores
            const double qn = sqrt(nq2);
            const double en = sqrt(__longlong_as_double((long long)consts->max_norm2_bits));
            const double pen_abs = (double)consts->pen_absmax, pen_max = (double)consts->pen_max;
            // Tensor-core scans multiply rounded operands q^ = q - g, e^ = e - d:  q.e - q^.e^ = g.e + q^.d, so
            // |q.e - q^.e^| <= |g||e| + |q^||d| <= (rho_q + (1 + rho_q) rho_e) |q||e|  (Cauchy-Schwarz on the residuals; rho_e is
            // the table's maximum, rho_q this query's own).  cp.rel adds the fp32 accumulation inside the tensor core and the
            // rounding of the two inverse norms.
            double rel = cp.rel;
            if (cp.kind && rho_q) {
                // read from the device copy of the table constants: a row upsert enqueued ahead of this search may have raised it
                const double rq = (double)rho_q[2 * q + (cp.kind - 1)];
                const double re = (double)__uint_as_float(cp.kind == 2 ? consts->rho_f16 : consts->rho_tf32);
                rel += 1.001 * (rq + (1.0 + rq) * re);
            }
            double eps;
            if (cp.metric == 0) {
                eps = rel + cp.abs_pen * (2.0 + pen_abs);
            } else if (cp.path == 2 || cp.path == 4) {
                const double delta = 2.0 * rel * qn * en + 4.0 * 5.96e-8 * (qn + en) * (qn + en);
                double x0 = T - pen_max; x0 = x0 > 0.0 ? x0 * x0 : 0.0;
                eps = sqrt(x0) - sqrt(fmax(x0 - delta, 0.0)) + cp.abs_pen * (sqrt(x0) + pen_abs);
            } else {
                eps = rel * (qn + en) + cp.abs_pen * ((qn + en) + pen_abs);
            }
            pass = (sk_k == sk_k) && (sk_k < T - eps);
        }
        if (!pass) {
            flag_list[atomicAdd(n_flag, 1)] = q;
            // the true k-th pair cannot exceed the k-th pair of any k rows; pads (fewer than k candidates) mean "no bound"
            bound_key[q] = sk[k - 1]; bound_row[q] = sv[k - 1];
        }
    }
}

// ---- K4c lives in exact_scan.cuh (it uses the TMA wrappers of gemm_tf32.cuh); its output descriptors: ------------------
#define XS_SPARSE_CAP 2048
struct XsSparse {
    const u64 *bound_key;      // [query] ordered fp64 image of the bound score (all-ones: no bound)
    const int *bound_row;      // [query] row of the bound pair
    u64 *keys;                 // [slot][cap]
    int *rows;                 // [slot][cap]
    int *cnt;                  // [slot] appended so far (may exceed cap: the list overflowed); zero on entry
    int cap;
};

// ---- sparse lists -> top k, one CTA per flagged query; folds the status words -------------------------------------
// status: [0] certificate failures, [1] unresolved (a sparse list overflowed: the bound let more than `cap` rows through --
// only data with thousands of rows inside the coarse error band of the k-th score does that), [2] ticket.  The last CTA
// mirrors the totals into host-mapped memory, so the host needs no copy to learn them.  unres: [0] count, [1..] global query
// numbers of the unresolved queries (the host entry point re-answers exactly those with the dense scan).
__global__ void __launch_bounds__(256)
exact_pick_kernel(const int *__restrict__ flag_list, const int *__restrict__ n_flag, XsSparse sp, int k, int64_t id_base,
                  int q_base, int32_t *__restrict__ out_ids, float *__restrict__ out_scores, double *__restrict__ out_scores64,
                  int *__restrict__ status, volatile int *host_status, int *__restrict__ unres, int unres_cap) {
    __shared__ u64 sk[XS_SPARSE_CAP];
    __shared__ int sv[XS_SPARSE_CAP];
    pdl_trigger();
    pdl_wait();
    const int nf = *n_flag;
    for (int f = blockIdx.x; f < nf; f += gridDim.x) {
        const int q = flag_list[f];
        const int c = sp.cnt[f];
        __syncthreads();                             // everyone has read the count (and left sk / sv of the previous turn)
        if (threadIdx.x == 0) sp.cnt[f] = 0;         // ready for the next launch
        if (c > sp.cap) {
            if (threadIdx.x == 0) {
                atomicAdd(&status[1], 1);
                const int p = atomicAdd(&unres[0], 1);
                if (p < unres_cap) unres[1 + p] = q_base + q;
            }
            continue;
        }
        int np2 = 2;
        while (np2 < c) np2 <<= 1;
        np2 = max(np2, 2);
        for (int i = threadIdx.x; i < np2; i += blockDim.x) {
            const bool real = i < c;
            sk[i] = real ? sp.keys[(int64_t)f * sp.cap + i] : ~0ull;
            sv[i] = real ? sp.rows[(int64_t)f * sp.cap + i] : MCP_PAD_ROW;
        }
        cta_bitonic_sort_pairs(sk, sv, np2);
        for (int i = threadIdx.x; i < k; i += blockDim.x) {
            const int64_t o = (int64_t)q * k + i;
            const bool real = i < c && sv[i] != MCP_PAD_ROW;
            const double s = real ? ord_f64(sk[i]) : CUDART_INF;
            out_ids[o] = real ? (int32_t)(id_base + sv[i]) : -1;
            out_scores[o] = (float)s;
            if (out_scores64) out_scores64[o] = s;
        }
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const int t = atomicAdd(&status[2], 1);
        if (t == (int)gridDim.x - 1) {
            status[2] = 0;
            __threadfence();
            status[0] += nf;
            // nothing flagged: the totals the host last saw still stand (the host zeroes its mirror and the device counters together)
            if (host_status && nf > 0) { host_status[0] = status[0]; host_status[1] = *(volatile int *)&status[1]; __threadfence_system(); }
        }
    }
}

// ---- K4d: exact top-k of one fp64 score vector; one CTA per slot ---------------------------------
// Radix select on the 64-bit score image (values may repeat), then rows with the boundary value
// are taken in row order, which is the "ties -> lowest id" rule.  Every thread of an EXSEL_THREADS-wide CTA calls it.
__device__ __forceinline__ void exact_select_body(const u64 *E /* may have been written by other CTAs of this launch: read through L2 */, int64_t n_rows, int k, int64_t id_base, int q,
                                                  int32_t *__restrict__ out_ids, float *__restrict__ out_scores,
                                                  double *__restrict__ out_scores64) {
    __shared__ uint32_t hist[256];
    __shared__ int s_digit, s_cnt, s_warp_base[EXSEL_THREADS / 32 + 1], s_taken, s_nless;
    __shared__ long long s_need;
    __shared__ u64 sk[MCP_MAX_K_POW2];
    __shared__ int sv[MCP_MAX_K_POW2];
    int kk = (int)min((int64_t)k, n_rows);
    u64 prefix = 0, mask = 0, thr = ~0ull;
    long long need = kk;
    bool whole_bin = (kk >= n_rows);
    if (!whole_bin) {
        for (int shift = 56; shift >= 0; shift -= 8) {
            for (int i = threadIdx.x; i < 256; i += blockDim.x) hist[i] = 0;
            __syncthreads();
            for (int64_t i = threadIdx.x; i < n_rows; i += blockDim.x) {
                u64 key = __ldcg(E + i);
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
        u64 key = __ldcg(E + i);
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
            bool hit = (i < n_rows) && (__ldcg(E + i) == thr);
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

__global__ void __launch_bounds__(EXSEL_THREADS)
exact_select_kernel(const u64 *__restrict__ exact_all, int64_t n_rows, int k, int64_t id_base,
                    const int *__restrict__ flag_list, int *__restrict__ n_flag, int slot0,
                    int32_t *__restrict__ out_ids, float *__restrict__ out_scores,
                    double *__restrict__ out_scores64) {
    const int f = blockIdx.x;
    if (slot0 + f >= *n_flag) return;
    exact_select_body(exact_all + (int64_t)f * n_rows, n_rows, k, id_base, flag_list[slot0 + f], out_ids, out_scores, out_scores64);
}

