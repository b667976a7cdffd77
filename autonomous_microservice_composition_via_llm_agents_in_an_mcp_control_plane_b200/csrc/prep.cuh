// prep.cuh -- K0: table / query preparation.
//   * per-row fp64 squared norm accumulated in index order (bit-identical to the oracle's ne2),
//     fp32 inverse norm for the coarse cosine score, non-finite detection, max norm;
//   * the same for a query batch;
//   * penalty statistics (max, max |.|) the certificate needs.
#pragma once
#include <cuda_fp16.h>

#include "common.cuh"

struct TableConsts {
    u64 max_norm2_bits;      // bit image of the largest row squared norm (non-negative doubles order as integers)
    unsigned int bad_rows;   // rows holding NaN / Inf
    unsigned int odd_rows;   // non-zero rows whose norm is outside [1e-15, 1e15]: coarse fp32 scoring is not trusted
    float pen_max;           // max penalty (signed), 0 when no penalty
    float pen_absmax;        // max |penalty|
    unsigned int bad_pen;    // non-finite penalties
    // Largest relative rounding residual of a row under the two tensor-core operand formats (float bit images; non-negative
    // floats order as integers): rho = |e - round(e)|_2 / |e|_2, rounded up.  With the same quantity of the query this bounds
    // the coarse dot product rigorously (Cauchy-Schwarz on the residuals, see CertParams), about 2.4x tighter on real
    // embeddings than the worst-case per-element bound and valid for any data.
    unsigned int rho_tf32;   // fp32 -> tf32 as kind::tf32 reads it: the low 13 mantissa bits are dropped
    unsigned int rho_f16;    // fp32 -> fp16 (row scaled by a power of two, round to nearest even): the shadow table
};

// rounding residual helpers (TableConsts::rho_*): the tf32 image of an fp32 value as kind::tf32 reads it, and an upper bound
// of sqrt(sum delta^2 / |v|^2)
__device__ __forceinline__ float tf32_trunc(float v) { return __uint_as_float(__float_as_uint(v) & 0xFFFFE000u); }
__device__ __forceinline__ float rho_up(float sum_d2, double norm2) {
    if (!(norm2 > 0.0)) return 0.f;                              // zero row: every score is NaN, no bound needed
    const double r = sqrt((double)sum_d2 * (1.0 + 1.0 / 1024.0) / norm2) * (1.0 + 1e-6);
    return __double2float_ru(r);
}
__device__ __forceinline__ float pow2_scale_for(float vmax) {
    // s = 2^(14 - floor(log2 vmax)), exponent clamped so that s and 1/s stay normal fp32 numbers
    if (!(vmax > 0.f) || !isfinite(vmax)) return 1.f;
    int e = (int)((__float_as_uint(vmax) >> 23) & 0xFF) - 127;
    e = max(-100, min(100, 14 - e));
    return __uint_as_float((uint32_t)(e + 127) << 23);
}

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
// It also writes what the tensor-core scans need besides: the query's rounding residual ratios (rho_q[2q] for tf32, [2q+1] for
// fp16; same definition as TableConsts::rho_*), and -- q_out != NULL -- the operand image of the query with pitch ldq (zero
// padded): fp16 of the query scaled by a power of two like a shadow row (as_f16; 1 / scale goes to q_scale_inv), or an fp32
// copy (the pitch a TMA tensor map needs when dim % 4 != 0).
__global__ void __launch_bounds__(128)
prep_queries_fast_kernel(const float *__restrict__ queries, int n_queries, int dim, float *__restrict__ inv_norm,
                         float *__restrict__ norm2f, int *__restrict__ n_flag /* NULL, or the certificate's flag count to clear */,
                         unsigned int *__restrict__ drop_min /* NULL, or [n_queries]: reset to "nothing dropped" */,
                         float *__restrict__ rho_q /* NULL, or [n_queries][2] */, void *__restrict__ q_out, int ldq, int as_f16,
                         float *__restrict__ q_scale_inv /* [n_queries], with as_f16 */,
                         int n_pad /* rows of q_out, >= n_queries: the rows past the batch are written as zeros */) {
    if (n_flag && blockIdx.x == 0 && threadIdx.x == 0) *n_flag = 0;     // first kernel of a search: nothing is flagged yet
    const int q = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (q >= n_queries) {
        // The operand image is padded to a multiple of 8 rows: a TMA box that ends inside an 8-row swizzle group of the
        // tensor (7 queries, 9 queries ...) loaded measurably slower (K2 over 1 M rows: 0.65 ms against 0.51 ms for 8 or 16).
        if (q_out && q < n_pad)
            for (int i = lane; i < ldq; i += 32) {
                if (as_f16) reinterpret_cast<__half *>(q_out)[(int64_t)q * ldq + i] = __float2half_rn(0.f);
                else reinterpret_cast<float *>(q_out)[(int64_t)q * ldq + i] = 0.f;
            }
        return;
    }
    if (drop_min && lane == 0) drop_min[q] = 0xFFFFFFFFu;
    const float *src = queries + (int64_t)q * dim;
    float acc = 0.f, dt = 0.f, vmax = 0.f;
    for (int i = lane; i < dim; i += 32) {
        const float v = src[i];
        acc = fmaf(v, v, acc);
        vmax = fmaxf(vmax, fabsf(v));
        const float et = v - tf32_trunc(v);
        dt = fmaf(et, et, dt);
    }
    acc = warp_sum(acc); dt = warp_sum(dt);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) vmax = fmaxf(vmax, __shfl_xor_sync(MCP_FULL, vmax, o));
    const float sc = pow2_scale_for(vmax), sci = 1.0f / sc;
    float dh = 0.f;
    const int lim = q_out ? ldq : dim;
    for (int i = lane; i < lim; i += 32) {                             // second walk: the query is in L1 by now
        const float v = i < dim ? src[i] : 0.f;
        const float x = v * sc;
        const __half h = __float2half_rn(x);
        const float eh = (x - __half2float(h)) * sci;
        dh = fmaf(eh, eh, dh);
        if (q_out) {
            if (as_f16) reinterpret_cast<__half *>(q_out)[(int64_t)q * ldq + i] = h;
            else reinterpret_cast<float *>(q_out)[(int64_t)q * ldq + i] = v;
        }
    }
    dh = warp_sum(dh);
    if (lane == 0) {
        inv_norm[q] = 1.0f / sqrtf(acc); norm2f[q] = acc;              // zero query -> inf -> NaN scores, like the exact path
        if (q_scale_inv) q_scale_inv[q] = sci;
        if (rho_q) {
            // acc is an fp32 tree sum of the norm: shrink it by its own error bound so that the ratio stays an upper bound
            const double n2 = (double)acc * (1.0 - 1e-5);
            rho_q[2 * q] = rho_up(dt, n2); rho_q[2 * q + 1] = rho_up(dh, n2);
        }
    }
}

// ---- fp16 shadow of the table + rounding residuals ------------------------------------------------------------------
// One warp per row.  The shadow row is fp16(e * s) with s a power of two chosen per row so that the largest element
// lands in [2^14, 2^15): every row uses fp16's full precision (11 bits, round to nearest even: relative error <= 2^-11,
// half of what truncation to tf32 leaves) whatever its norm, and 1 / s is handed to the scan, which folds it into the
// row's inverse norm -- exact, both are powers of two away from the unscaled values.  Pitch ldh elements (a multiple of
// 8; pad columns zero).  The squared residuals of both operand formats are summed on the way: rho^2 = sum(delta^2) / |e|^2;
// the 1 + 2^-10 factor in rho_up covers the fp32 summation of the (exactly representable) residual squares, so the
// stored value is an upper bound.
__global__ void __launch_bounds__(256)
shadow_rows_kernel(const float *__restrict__ rows, int64_t n_rows, int dim, int64_t ld, const double *__restrict__ norm2,
                   __half *__restrict__ shadow /* NULL: residual statistics only */, int64_t ldh,
                   float *__restrict__ scale_inv /* [n_rows] 1 / s, or NULL with shadow */, TableConsts *consts) {
    const int lane = threadIdx.x & 31;
    const int64_t r = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (r >= n_rows) return;
    const float *e = rows + r * ld;
    const bool vec = (ld & 3) == 0 && ((uintptr_t)rows & 15) == 0;
    auto load4 = [&](int c0, float (&v)[4]) {
        if (vec && c0 + 3 < dim) { const float4 t = *reinterpret_cast<const float4 *>(e + c0); v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w; }
        else {
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = c0 + j < dim ? e[c0 + j] : 0.f;
        }
    };
    float vmax = 0.f, dt = 0.f;
    for (int c0 = lane * 4; c0 < (int)ldh; c0 += 128) {
        float v[4];
        load4(c0, v);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            vmax = fmaxf(vmax, fabsf(v[j]));
            const float et = v[j] - tf32_trunc(v[j]);            // exact in fp32
            dt = fmaf(et, et, dt);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) vmax = fmaxf(vmax, __shfl_xor_sync(MCP_FULL, vmax, o));
    const float sc = pow2_scale_for(vmax), sci = 1.0f / sc;      // powers of two: both exact
    float dh = 0.f;
    for (int c0 = lane * 4; c0 < (int)ldh; c0 += 128) {          // second walk over the row: L1 / L2 hits
        float v[4];
        load4(c0, v);
        __half h[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float x = v[j] * sc;                           // exact
            h[j] = __float2half_rn(x);
            const float eh = (x - __half2float(h[j])) * sci;     // exact: the residual of a rounding is representable
            dh = fmaf(eh, eh, dh);
        }
        if (shadow) *reinterpret_cast<uint2 *>(shadow + r * ldh + c0) = *reinterpret_cast<const uint2 *>(h);
    }
    dh = warp_sum(dh); dt = warp_sum(dt);
    if (lane == 0) {
        if (scale_inv) scale_inv[r] = sci;
        const double n2 = norm2[r];
        if (isfinite(dh) && isfinite(dt)) {                      // rows with NaN / Inf are rejected by prep_rows_kernel
            atomicMax(&consts->rho_f16, __float_as_uint(rho_up(dh, n2)));
            atomicMax(&consts->rho_tf32, __float_as_uint(rho_up(dt, n2)));
        }
    }
}

// ---- row upsert: staged rows -> table, norms, shadow, residual maxima, in one kernel --------------------------------
// One warp per updated row i (target row ids[i], or first_row + i when ids is NULL).  The warp copies the staged row into
// the table, lane 0 walks it in index order for the fp64 squared norm (the oracle's order; the row is in L1 by then), and
// the shadow row, its scale and the residual maxima follow as in shadow_rows_kernel.  Runs on the stream the searches use:
// a search enqueued before it sees the old row and its old norms, one enqueued after it the new ones, nobody waits on the host.
__global__ void __launch_bounds__(256)
upsert_rows_kernel(const float *__restrict__ stage, int64_t lds, const int64_t *__restrict__ ids, int64_t first_row, int64_t n_upd,
                   float *__restrict__ rows, int dim, int64_t ld, double *__restrict__ norm2, float *__restrict__ inv_norm,
                   __half *__restrict__ shadow /* or NULL */, int64_t ldh, float *__restrict__ scale_inv, TableConsts *consts) {
    const int lane = threadIdx.x & 31;
    const int64_t i = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (i >= n_upd) return;
    const int64_t r = ids ? ids[i] : first_row + i;
    const float *src = stage + i * lds;
    float *dst = rows + r * ld;
    float vmax = 0.f, dt = 0.f;
    for (int c = lane; c < (int)ld; c += 32) {
        const float v = c < dim ? src[c] : 0.f;                  // pad columns of the table stay zero
        dst[c] = v;
        vmax = fmaxf(vmax, fabsf(v));
        const float et = v - tf32_trunc(v);
        dt = fmaf(et, et, dt);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) vmax = fmaxf(vmax, __shfl_xor_sync(MCP_FULL, vmax, o));
    const float sc = pow2_scale_for(vmax), sci = 1.0f / sc;
    float dh = 0.f;
    for (int c = lane; c < (int)ldh; c += 32) {
        const float x = (c < dim ? src[c] : 0.f) * sc;
        const __half h = __float2half_rn(x);
        const float eh = (x - __half2float(h)) * sci;
        dh = fmaf(eh, eh, dh);
        if (shadow) shadow[r * ldh + c] = h;
    }
    dh = warp_sum(dh); dt = warp_sum(dt);
    double n2 = 0.0;
    if (lane == 0) {
        for (int c = 0; c < dim; ++c) { const double v = (double)src[c]; n2 = fma(v, v, n2); }
        norm2[r] = n2;
        inv_norm[r] = (float)(1.0 / sqrt(n2));
        if (scale_inv) scale_inv[r] = sci;
        atomicMax(&consts->max_norm2_bits, (u64)__double_as_longlong(n2));
        if (isfinite(dh) && isfinite(dt)) {
            atomicMax(&consts->rho_f16, __float_as_uint(rho_up(dh, n2)));
            atomicMax(&consts->rho_tf32, __float_as_uint(rho_up(dt, n2)));
        }
    }
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
