// exact_scan.cuh -- K4c: fp64 scan of every row for the flagged queries.
//
// The certificate's fallback, the explicit path = exact_fp64 and the allocation-free entry point all run this kernel.
// It streams the table once per pass for up to S flagged queries ("slots") at a time:
//   * one producer warp moves [XS_ROWS rows x 32 columns] boxes of the fp32 table into a ring of shared-memory stages with
//     TMA (one instruction and 32 KB per stage, 128-byte swizzle), mbarrier full / empty per stage -- the consumers never
//     meet at a block barrier.  The ring takes whatever shared memory the pass's queries leave: 6 stages for one or two
//     flagged queries, 4 for six.  (First version: cp.async by all threads, three stages, two __syncthreads per chunk:
//     1.9 TB/s, 38 % of the samples waiting at the barrier.)
//   * consumer thread = one row of the tile: it walks ITS row in index order, one fp64 chain per slot (the oracle's
//     arithmetic and order, so scores are bit-equal), reading the queries as fp64 from shared memory (warp-wide
//     broadcasts).  A chain is ~15 cycles per element and cannot be split without changing the rounding, so the rate for
//     one flagged query is set by how many rows an SM walks at once: 256.  The swizzle (16-byte chunk index ^ row % 8)
//     makes the eight rows of a quarter-warp hit eight different bank groups.
// Output, dense mode: the ordered fp64 image of every score, [slot][n_rows] (exact_select_kernel then picks the top k).
// Output, sparse mode: only rows whose (score, row) pair is at or below the query's bound -- the k-th best exact pair among
// the candidates the coarse pass found, which the true k-th pair cannot exceed -- appended to a short per-slot list.  A
// handful of rows per query instead of n_rows scores, so the fallback needs no n_rows-sized scratch and serves any number
// of flagged queries in one launch (the kernel loops over them S at a time).
#pragma once
#include "gemm_tf32.cuh"
#include "rescore.cuh"

#define XS_ROWS 256
#define XS_COLS 32                       // one 128-byte line of a row per box
#define XS_MAX_STAGES 6
#define XS_MIN_STAGES 2
#define XS_CONSUMERS XS_ROWS
#define XS_THREADS (XS_CONSUMERS + 32)
#define XS_MAXS 6
#define XS_STAGE_BYTES (XS_ROWS * XS_COLS * 4)
#define XS_TILE_BYTES ((size_t)XS_MIN_STAGES * XS_STAGE_BYTES + 1024)      // the least the ring needs, + slack for the 1024-byte alignment of the swizzle
// dynamic shared memory of a launch: everything the kernel may have (the ring grows into what S queries do not use)
#define XS_SMEM_BYTES(dim, S) ((size_t)XS_SMEM_LIMIT)

// one row's walk over one box (w columns, a multiple of 4) for NSL slots; e = the row's 128 bytes in the (swizzled) stage
template <int NSL>
__device__ __forceinline__ void xs_chunk(const unsigned char *__restrict__ e, uint32_t swz, const double *__restrict__ sq, int dimp, int w,
                                         int metric, double (&acc)[XS_MAXS]) {
    for (int j = 0; j < w; j += 4) {
        const float4 x = *reinterpret_cast<const float4 *>(e + ((((uint32_t)j >> 2) ^ swz) << 4));
        const double d0 = (double)x.x, d1 = (double)x.y, d2 = (double)x.z, d3 = (double)x.w;
#pragma unroll
        for (int v = 0; v < NSL; ++v) {
            const double *qq = sq + (size_t)v * dimp + j;
            const double2 qa = *reinterpret_cast<const double2 *>(qq), qb = *reinterpret_cast<const double2 *>(qq + 2);
            if (metric == 0) {
                acc[v] = fma(qa.x, d0, acc[v]); acc[v] = fma(qa.y, d1, acc[v]);
                acc[v] = fma(qb.x, d2, acc[v]); acc[v] = fma(qb.y, d3, acc[v]);
            } else {
                double t;
                t = __dsub_rn(qa.x, d0); acc[v] = __dadd_rn(acc[v], __dmul_rn(t, t));
                t = __dsub_rn(qa.y, d1); acc[v] = __dadd_rn(acc[v], __dmul_rn(t, t));
                t = __dsub_rn(qb.x, d2); acc[v] = __dadd_rn(acc[v], __dmul_rn(t, t));
                t = __dsub_rn(qb.y, d3); acc[v] = __dadd_rn(acc[v], __dmul_rn(t, t));
            }
        }
    }
}

__global__ void __launch_bounds__(XS_THREADS, 1)
exact_scan2_kernel(const __grid_constant__ CUtensorMap tm, int64_t n_rows, int dim,
                   const float *__restrict__ queries, const double *__restrict__ nq2,
                   const double *__restrict__ ne2, const float *__restrict__ penalty, int metric,
                   const int *__restrict__ flag_list, const int *__restrict__ n_flag, int slot0, int S,
                   u64 *__restrict__ dense /* [S][n_rows] for slots slot0 .. slot0+S-1, or NULL: sparse */, XsSparse sp) {
    extern __shared__ __align__(16) unsigned char xs_raw[];
    __shared__ __align__(8) u64 s_full[XS_MAX_STAGES], s_empty[XS_MAX_STAGES];
    pdl_trigger();
    pdl_wait();                                      // the kernel that flagged queries has completed
    const int nf_all = *n_flag;
    const int g_end = dense ? min(nf_all, slot0 + S) : nf_all;
    if (slot0 >= g_end) return;                      // nothing flagged: the common case
    const int dimp = (dim + 3) & ~3;
    // the most queries any pass of this launch holds decides how much of the shared memory is ring
    const int s_eff = min(S, g_end - slot0);
    const uint32_t pad = (1024u - (smem_u32(xs_raw) & 1023u)) & 1023u;
    const int NS = (int)min((size_t)XS_MAX_STAGES, ((size_t)XS_SMEM_LIMIT - pad - sizeof(double) * (size_t)s_eff * dimp) / XS_STAGE_BYTES);
    unsigned char *ring = xs_raw + pad;                                                   // [NS][XS_ROWS][128 B], swizzled
    double *sq = reinterpret_cast<double *>(ring + (size_t)NS * XS_STAGE_BYTES);          // [s_eff][dimp] the pass's queries in fp64
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < NS; ++s) { mbar_init(&s_full[s], 1); mbar_init(&s_empty[s], XS_CONSUMERS / 32); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int nchunks = (dim + XS_COLS - 1) / XS_COLS;
    const int64_t n_tiles = (n_rows + XS_ROWS - 1) / XS_ROWS;
    const int my_tiles = (int64_t)blockIdx.x < n_tiles ? (int)((n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0;
    const int n_it = my_tiles * nchunks;             // boxes per pass
    const int n_pass = (g_end - slot0 + S - 1) / S;
    if (warp == XS_CONSUMERS / 32) {
        // ===== producer: the same boxes once per pass, the ring's phases running on across passes =====
        if (lane == 0) {
            asm volatile("prefetch.tensormap [%0];" ::"l"(&tm) : "memory");
            int64_t git = 0;
            for (int p = 0; p < n_pass; ++p)
                for (int it = 0; it < n_it; ++it, ++git) {
                    const int st = (int)(git % NS);
                    const uint32_t n = (uint32_t)(git / NS);
                    if (n > 0) mbar_wait(&s_empty[st], (n - 1) & 1u);
                    const int ti = it / nchunks, c = it - ti * nchunks;
                    const int64_t r0 = ((int64_t)blockIdx.x + (int64_t)ti * gridDim.x) * XS_ROWS;
                    mbar_expect_tx(&s_full[st], XS_STAGE_BYTES);                          // rows / columns past the table are zero-filled: always a whole box
                    tma_load_2d(ring + (size_t)st * XS_STAGE_BYTES, &tm, c * XS_COLS, (int)r0, &s_full[st]);
                }
        }
        return;
    }
    // ===== consumers =====
    const int row_l = threadIdx.x;
    const uint32_t swz = (uint32_t)(row_l & 7);
    int64_t git = 0;
    for (int s0 = slot0; s0 < g_end; s0 += S) {
        const int sc = min(S, g_end - s0);           // slots of this pass
        asm volatile("bar.sync 1, %0;" ::"n"(XS_CONSUMERS) : "memory");      // the previous pass has left sq
        for (int u = 0; u < sc; ++u) {
            const float *qv = queries + (int64_t)flag_list[s0 + u] * dim;
            for (int i = threadIdx.x; i < dimp; i += XS_CONSUMERS) sq[(size_t)u * dimp + i] = i < dim ? (double)qv[i] : 0.0;
        }
        asm volatile("bar.sync 1, %0;" ::"n"(XS_CONSUMERS) : "memory");
        const int myn = sc;                          // every thread serves all slots of the pass
        int myq[XS_MAXS]; double mynq2[XS_MAXS]; u64 mybk[XS_MAXS]; int mybr[XS_MAXS];
#pragma unroll
        for (int v = 0; v < XS_MAXS; ++v) {
            myq[v] = v < myn ? flag_list[s0 + v] : 0;
            mynq2[v] = v < myn ? nq2[myq[v]] : 1.0;
            mybk[v] = (v < myn && !dense) ? sp.bound_key[myq[v]] : ~0ull;
            mybr[v] = (v < myn && !dense) ? sp.bound_row[myq[v]] : 0x7FFFFFFF;
        }
        double acc[XS_MAXS] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
        for (int it = 0; it < n_it; ++it, ++git) {
            const int st = (int)(git % NS);
            mbar_wait(&s_full[st], (uint32_t)(git / NS) & 1u);
            const int ti = it / nchunks, c = it - ti * nchunks;
            const int c0 = c * XS_COLS, w = min(XS_COLS, dimp - c0);
            const unsigned char *e = ring + (size_t)st * XS_STAGE_BYTES + row_l * 128;
            if (c == 0) {
#pragma unroll
                for (int v = 0; v < XS_MAXS; ++v) acc[v] = 0.0;
            }
            // one instantiation per slot count: with a run-time count the compiler predicates all six bodies and five of them are
            // issued for nothing when one query is flagged (measured: 70 instructions per four columns instead of 11)
            switch (myn) {
                case 1: xs_chunk<1>(e, swz, sq + c0, dimp, w, metric, acc); break;
                case 2: xs_chunk<2>(e, swz, sq + c0, dimp, w, metric, acc); break;
                case 3: xs_chunk<3>(e, swz, sq + c0, dimp, w, metric, acc); break;
                case 4: xs_chunk<4>(e, swz, sq + c0, dimp, w, metric, acc); break;
                case 5: xs_chunk<5>(e, swz, sq + c0, dimp, w, metric, acc); break;
                case 6: xs_chunk<6>(e, swz, sq + c0, dimp, w, metric, acc); break;
                default: break;
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&s_empty[st]);                         // this warp has read the stage
            if (c == nchunks - 1) {
                const int64_t row = ((int64_t)blockIdx.x + (int64_t)ti * gridDim.x) * XS_ROWS + row_l;
                if (row < n_rows && myn > 0) {
                    const double ne2r = ne2[row], pr = penalty ? (double)penalty[row] : 0.0;
#pragma unroll
                    for (int v = 0; v < XS_MAXS; ++v) {
                        if (v < myn) {
                            double sco = exact_finish(acc[v], metric, mynq2[v], ne2r);
                            if (penalty) sco = sco + pr;
                            const u64 key = f64_ord(sco);
                            const int slot = s0 + v;
                            if (dense) __stcg(dense + (int64_t)(slot - slot0) * n_rows + row, key);
                            else if (key < mybk[v] || (key == mybk[v] && row <= (int64_t)mybr[v])) {
                                const int p = atomicAdd(&sp.cnt[slot], 1);
                                if (p < sp.cap) { sp.keys[(int64_t)slot * sp.cap + p] = key; sp.rows[(int64_t)slot * sp.cap + p] = (int)row; }
                            }
                        }
                    }
                }
            }
        }
    }
}
