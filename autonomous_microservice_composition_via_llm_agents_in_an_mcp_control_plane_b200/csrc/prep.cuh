// prep.cuh -- K0: table / query preparation.
//   * per-row fp64 squared norm accumulated in index order (bit-identical to the oracle's ne2),
//     fp32 inverse norm for the coarse cosine score, non-finite detection, max norm;
//   * the same for a query batch;
//   * penalty statistics (max, max |.|) the certificate needs.
#pragma once
#include "common.cuh"

struct TableConsts {
    u64 max_norm2_bits;      // bit image of the largest row squared norm (non-negative doubles order as integers)
    unsigned int bad_rows;   // rows holding NaN / Inf
    unsigned int odd_rows;   // non-zero rows whose norm is outside [1e-15, 1e15]: coarse fp32 scoring is not trusted
    float pen_max;           // max penalty (signed), 0 when no penalty
    float pen_absmax;        // max |penalty|
    unsigned int bad_pen;    // non-finite penalties
};

#define PREP_ROWS 128   // rows per CTA (= threads)
#define PREP_COLS 32    // floats staged per row per step (one 128-byte line)

// Rows are staged through shared memory so that HBM is read in full 128-byte lines while each
// thread still walks *its own* row in index order (the order is what makes the norm exact-match
// the oracle).
template <bool IS_QUERY>
__global__ void __launch_bounds__(PREP_ROWS)
prep_rows_kernel(const float *__restrict__ rows, int64_t n_rows, int dim, int64_t ld,
                 double *__restrict__ norm2, float *__restrict__ inv_norm,
                 unsigned char *__restrict__ qflag, TableConsts *consts) {
    __shared__ float tile[PREP_ROWS][PREP_COLS + 1];
    int64_t r0 = (int64_t)blockIdx.x * PREP_ROWS;
    int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    double acc = 0.0;
    bool bad = false;
    for (int c0 = 0; c0 < dim; c0 += PREP_COLS) {
        // warp w stages rows w, w+4, ... ; lane = column
        for (int rr = warp; rr < PREP_ROWS; rr += PREP_ROWS / 32) {
            int64_t r = r0 + rr;
            float v = 0.f;
            if (r < n_rows && c0 + lane < dim) v = rows[r * ld + c0 + lane];
            tile[rr][lane] = v;
        }
        __syncthreads();
        int lim = min(PREP_COLS, dim - c0);
        for (int j = 0; j < lim; ++j) {
            float v = tile[threadIdx.x][j];
            bad |= !isfinite(v);
            acc = fma((double)v, (double)v, acc);
        }
        __syncthreads();
    }
    int64_t r = r0 + threadIdx.x;
    if (r < n_rows) {
        norm2[r] = acc;
        inv_norm[r] = (float)(1.0 / sqrt(acc));       // zero row -> +inf -> coarse score NaN, like the exact one
        bool odd = (acc != 0.0) && (acc < 1e-30 || acc > 1e30);
        if (IS_QUERY) {
            // a query the coarse pass cannot be trusted on goes straight to the fp64 scan
            qflag[r] = (unsigned char)((bad ? 2 : 0) | ((odd || acc == 0.0) ? 1 : 0));
        } else {
            if (bad) atomicAdd(&consts->bad_rows, 1u);
            else {
                if (odd) atomicAdd(&consts->odd_rows, 1u);
                atomicMax(&consts->max_norm2_bits, (u64)__double_as_longlong(acc));
            }
        }
    }
}

// Query batches are small (<= a few thousand vectors), so the table kernel above (128 rows per CTA) would leave
// the GPU empty: here one CTA owns one query, all threads pull it into shared memory at once, and thread 0 walks
// it in index order (the fp64 chain is the latency floor: dim dependent FMAs).
#define PREPQ_THREADS 128
__global__ void __launch_bounds__(PREPQ_THREADS)
prep_queries_kernel(const float *__restrict__ queries, int dim, double *__restrict__ norm2,
                    float *__restrict__ inv_norm, unsigned char *__restrict__ qflag, float *__restrict__ norm2f) {
    extern __shared__ __align__(16) float sq_prep[];
    __shared__ int s_bad;
    const int q = blockIdx.x;
    const float *src = queries + (int64_t)q * dim;
    if (threadIdx.x == 0) s_bad = 0;
    __syncthreads();
    bool bad = false;
    for (int i = threadIdx.x; i < dim; i += PREPQ_THREADS) {
        float v = src[i];
        bad |= !isfinite(v);
        sq_prep[i] = v;
    }
    if (bad) s_bad = 1;
    __syncthreads();
    if (threadIdx.x == 0) {
        double acc = 0.0;
        int d4 = dim & ~3;
        for (int i = 0; i < d4; i += 4) {
            float4 v = *reinterpret_cast<const float4 *>(sq_prep + i);
            acc = fma((double)v.x, (double)v.x, acc); acc = fma((double)v.y, (double)v.y, acc);
            acc = fma((double)v.z, (double)v.z, acc); acc = fma((double)v.w, (double)v.w, acc);
        }
        for (int i = d4; i < dim; ++i) acc = fma((double)sq_prep[i], (double)sq_prep[i], acc);
        norm2[q] = acc;
        inv_norm[q] = (float)(1.0 / sqrt(acc));
        if (norm2f) norm2f[q] = (float)acc;
        bool odd = (acc != 0.0) && (acc < 1e-30 || acc > 1e30);
        qflag[q] = (unsigned char)((s_bad ? 2 : 0) | ((odd || acc == 0.0 || s_bad) ? 1 : 0));
    }
}

// fp32 inverse norms of a query batch, one warp per query: all the coarse kernels need.  The exact fp64 norm
// (prep_queries_kernel: a dim-long dependent chain) is only needed by the rescoring stage, so it runs on a side
// stream while the coarse scan is under way.
__global__ void __launch_bounds__(128)
prep_queries_fast_kernel(const float *__restrict__ queries, int n_queries, int dim, float *__restrict__ inv_norm,
                         float *__restrict__ norm2f, int *__restrict__ n_flag /* NULL, or the certificate's flag count to clear */,
                         unsigned int *__restrict__ drop_min /* NULL, or [n_queries]: reset to "nothing dropped" */) {
    if (n_flag && blockIdx.x == 0 && threadIdx.x == 0) *n_flag = 0;     // first kernel of a search: nothing is flagged yet
    const int q = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (q >= n_queries) return;
    if (drop_min && lane == 0) drop_min[q] = 0xFFFFFFFFu;
    const float *src = queries + (int64_t)q * dim;
    float acc = 0.f;
    for (int i = lane; i < dim; i += 32) { const float v = src[i]; acc = fmaf(v, v, acc); }
    acc = warp_sum(acc);
    if (lane == 0) { inv_norm[q] = 1.0f / sqrtf(acc); norm2f[q] = acc; }   // zero query -> inf -> NaN scores, like the exact path
}

__global__ void penalty_stats_kernel(const float *__restrict__ pen, int64_t n, TableConsts *consts) {
    float mx = -CUDART_INF_F, amx = 0.f;
    bool bad = false;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        float p = pen[i];
        bad |= !isfinite(p);
        mx = fmaxf(mx, p); amx = fmaxf(amx, fabsf(p));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mx = fmaxf(mx, __shfl_xor_sync(MCP_FULL, mx, o));
        amx = fmaxf(amx, __shfl_xor_sync(MCP_FULL, amx, o));
    }
    if ((threadIdx.x & 31) == 0) {
        // float max through the ordered integer image (handles negatives)
        atomicMax((unsigned int *)&consts->pen_max, f32_ord(mx));
        atomicMax((unsigned int *)&consts->pen_absmax, f32_ord(amx));
    }
    if (bad) atomicAdd(&consts->bad_pen, 1u);
}
// pen_max / pen_absmax are accumulated as ordered images; this turns them back into floats.
__global__ void penalty_stats_finish_kernel(TableConsts *consts) {
    consts->pen_max = ord_f32(*(unsigned int *)&consts->pen_max);
    consts->pen_absmax = ord_f32(*(unsigned int *)&consts->pen_absmax);
}
