// scan_fp32.cuh -- K1: memory-bound fp32 streaming distance + telemetry score (small batches).
//
// One warp owns one table row at a time and reads it with coalesced 128-bit loads that bypass
// L1 (each row is touched once per scan); the QT queries of the pass live in shared memory and
// are re-read from there for every row.  For dim = 1536 the 12 loads of a lane are all issued
// before the first FMA (NJ = 12), which with >= 16 resident warps per SM keeps ~100 KB in
// flight per SM -- what the ~6.5 TB/s HBM3e needs by Little's law.
//
// Output: the final coarse score (distance + penalty) of every (query,row) as fp32:
// 4 bytes written per 6144 bytes read, so materialising it costs 0.07 % of the traffic per
// query and keeps selection (select.cuh) out of the streaming loop.
//
// Error bound used by the certificate (rescore.cuh): every lane sums dim/32 products
// sequentially, then 5 shuffle adds, so |fl(dot) - dot| <= (dim/32 + 6) u * sum|q_i e_i|
// <= (dim/32 + 6) u |q||e|, u = 2^-24.
#pragma once
#include "common.cuh"

#define SCAN_THREADS 256

template <int METRIC>
__device__ __forceinline__ void scan_accum(float &acc, const float4 e, const float4 q) {
    if (METRIC == 0) {
        acc = fmaf(e.x, q.x, acc); acc = fmaf(e.y, q.y, acc);
        acc = fmaf(e.z, q.z, acc); acc = fmaf(e.w, q.w, acc);
    } else {
        float d;
        d = q.x - e.x; acc = fmaf(d, d, acc); d = q.y - e.y; acc = fmaf(d, d, acc);
        d = q.z - e.z; acc = fmaf(d, d, acc); d = q.w - e.w; acc = fmaf(d, d, acc);
    }
}

// QT queries per pass; NJ = float4 per lane per row when known at compile time (dim == NJ*128), 0 = any dim % 4 == 0
template <int QT, int NJ, int METRIC>
__global__ void __launch_bounds__(SCAN_THREADS)
scan_fp32_kernel(const float *__restrict__ rows, int64_t n_rows, int dim, int64_t ld,
                 const float *__restrict__ queries,   // [QT][dim] (this pass)
                 const float *__restrict__ inv_nq,    // [QT]
                 const float *__restrict__ inv_ne,    // [n_rows]
                 const float *__restrict__ penalty,   // [n_rows] or NULL
                 float *__restrict__ scores, int64_t score_stride, int n_q_valid) {
    extern __shared__ float4 sq[];                   // [QT][dim/4]
    const int dim4 = dim >> 2;
    for (int i = threadIdx.x; i < QT * dim4; i += blockDim.x) {
        int q = i / dim4;
        sq[i] = (q < n_q_valid) ? reinterpret_cast<const float4 *>(queries)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int64_t warps = (int64_t)gridDim.x * (SCAN_THREADS / 32);
    int64_t r = (int64_t)blockIdx.x * (SCAN_THREADS / 32) + (threadIdx.x >> 5);
    float my_inv_nq = (lane < n_q_valid) ? inv_nq[lane] : 0.f;

    for (; r < n_rows; r += warps) {
        const float4 *row = reinterpret_cast<const float4 *>(rows + r * ld);
        float acc[QT];
#pragma unroll
        for (int q = 0; q < QT; ++q) acc[q] = 0.f;
        if (NJ > 0) {
            float4 e[NJ > 0 ? NJ : 1];
#pragma unroll
            for (int j = 0; j < NJ; ++j) e[j] = ldg_stream(row + j * 32 + lane);
#pragma unroll
            for (int j = 0; j < NJ; ++j)
#pragma unroll
                for (int q = 0; q < QT; ++q) scan_accum<METRIC>(acc[q], e[j], sq[q * dim4 + j * 32 + lane]);
        } else {
            for (int j0 = 0; j0 < dim4; j0 += 128) {     // 4 loads in flight per lane
                float4 e[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    int j = j0 + u * 32 + lane;
                    e[u] = (j < dim4) ? ldg_stream(row + j) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    int j = j0 + u * 32 + lane;
                    if (j < dim4) {
#pragma unroll
                        for (int q = 0; q < QT; ++q) scan_accum<METRIC>(acc[q], e[u], sq[q * dim4 + j]);
                    }
                }
            }
        }
        float mine = 0.f;
#pragma unroll
        for (int q = 0; q < QT; ++q) {
            float v = warp_sum(acc[q]);
            if (lane == q) mine = v;
        }
        if (lane < n_q_valid) {
            float pen = penalty ? penalty[r] : 0.f;
            float s;
            if (METRIC == 0) s = (1.0f - mine * my_inv_nq * inv_ne[r]) + pen;
            else             s = sqrtf(mine) + pen;
            scores[(int64_t)lane * score_stride + r] = s;
        }
    }
}
