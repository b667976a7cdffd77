// gemm_tf32.cuh -- K2 placeholder until the tcgen05 kernel lands (next commit).
#pragma once
#include "common.cuh"
struct GemmPlan { bool valid = false; };
struct Buf;
static inline bool gemm_supported(int, int64_t) { return false; }
template <class B>
static int gemm_tf32_run(GemmPlan &, B &, cudaStream_t, int, const float *, int64_t, int, int64_t, const float *, int,
                         const float *, const double *, const float *, const double *, const float *, int, int,
                         u64 **, int **, int *, int *) {
    return -3;
}
