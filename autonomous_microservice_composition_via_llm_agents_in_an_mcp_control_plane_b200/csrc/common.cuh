// common.cuh -- shared device helpers: ordered keys, warp/CTA primitives, exact fp64 distance.
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

typedef unsigned long long u64;

#define MCP_WARP 32
#define MCP_FULL 0xFFFFFFFFu
#define MCP_PAD_ROW 0x7FFFFFFF

// ---------------------------------------------------------------------------------------------
// Order-preserving integer images of floats.  Smaller integer <=> better (smaller) score.
// NaN maps to the all-ones pattern so it ranks after +inf; -0 is folded into +0.
// ---------------------------------------------------------------------------------------------
__host__ __device__ __forceinline__ uint32_t f32_ord(float f) {
    if (f != f) return 0xFFFFFFFFu;
    f += 0.0f;
#ifdef __CUDA_ARCH__
    uint32_t b = __float_as_uint(f);
#else
    uint32_t b; memcpy(&b, &f, 4);
#endif
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__host__ __device__ __forceinline__ float ord_f32(uint32_t o) {
    uint32_t b = (o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o;   // all-ones -> 0x7FFFFFFF = NaN
#ifdef __CUDA_ARCH__
    return __uint_as_float(b);
#else
    float f; memcpy(&f, &b, 4); return f;
#endif
}
__host__ __device__ __forceinline__ u64 f64_ord(double d) {
    if (d != d) return ~0ull;
    d += 0.0;
#ifdef __CUDA_ARCH__
    u64 b = (u64)__double_as_longlong(d);
#else
    u64 b; memcpy(&b, &d, 8);
#endif
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__host__ __device__ __forceinline__ double ord_f64(u64 o) {
    u64 b = (o >> 63) ? (o & 0x7FFFFFFFFFFFFFFFull) : ~o;       // all-ones -> NaN
#ifdef __CUDA_ARCH__
    return __longlong_as_double((long long)b);
#else
    double d; memcpy(&d, &b, 8); return d;
#endif
}
// coarse candidate key: fp32 score image in the high word, local row number in the low word.
// Rows are unique, so keys are unique and "ties -> lowest id" is part of the integer order.
__device__ __forceinline__ u64 coarse_key(float score, uint32_t row) {
    return ((u64)f32_ord(score) << 32) | (u64)row;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(MCP_FULL, v, o);
    return v;
}

// streaming 128-bit load that does not allocate in L1 (table rows are read once per scan)
__device__ __forceinline__ float4 ldg_stream(const float4 *p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}

// ---------------------------------------------------------------------------------------------
// Exact distance, the oracle's arithmetic (oracle/oracle.c distance_f64): fp64, index order.
// a_i*b_i of two fp32 values is exact in fp64, so fma(a,b,acc) == acc + a*b rounded once.
// The L2 form rounds (a-b), (a-b)^2 and the sum separately: explicit _rn intrinsics keep
// nvcc from contracting them into an FMA.
// ---------------------------------------------------------------------------------------------
struct ExactAcc {
    double acc;
    __device__ __forceinline__ void cos_step(float a, float b) { acc = fma((double)a, (double)b, acc); }
    __device__ __forceinline__ void l2_step(float a, float b) {
        double d = __dsub_rn((double)a, (double)b);
        acc = __dadd_rn(acc, __dmul_rn(d, d));
    }
};
__device__ __forceinline__ double exact_finish(double acc, int metric, double nq2, double ne2) {
    if (metric == 0) {
        double sim = acc / sqrt(__dmul_rn(nq2, ne2));
        if (sim > 1.0) sim = 1.0;
        else if (sim < -1.0) sim = -1.0;
        return 1.0 - sim;
    }
    return sqrt(acc);
}

// ---------------------------------------------------------------------------------------------
// CTA-wide bitonic sort of (key,val) pairs in shared memory, ascending by (key, val).
// n must be a power of two; every thread of the block calls it.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ bool pair_greater(u64 ka, int va, u64 kb, int vb) {
    return ka > kb || (ka == kb && va > vb);
}
__device__ inline void cta_bitonic_sort_pairs(u64 *k, int *v, int n) {
    for (int size = 2; size <= n; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            __syncthreads();
            for (int t = threadIdx.x; t < (n >> 1); t += blockDim.x) {
                int lo = 2 * t - (t & (stride - 1));
                int hi = lo + stride;
                bool asc = ((lo & size) == 0);
                u64 ka = k[lo], kb = k[hi];
                int va = v[lo], vb = v[hi];
                if (pair_greater(ka, va, kb, vb) == asc) { k[lo] = kb; k[hi] = ka; v[lo] = vb; v[hi] = va; }
            }
        }
    }
    __syncthreads();
}

// ---------------------------------------------------------------------------------------------
// mbarrier wrappers (shared by the tensor-core scan and the rescoring ring) and the 1-D bulk copy
// global -> shared that reports its bytes to an mbarrier.  src, dst and bytes are multiples of 16.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(u64 *b, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(u64 *b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(u64 *b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_wait(u64 *b, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_load_1d(void *dst, const void *src, uint32_t bytes, u64 *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// ---------------------------------------------------------------------------------------------
// Programmatic dependent launch.  A kernel launched with the programmatic-stream-serialization attribute may be
// scheduled while its predecessor in the stream is still draining; it must call pdl_wait() before it touches anything
// the predecessor wrote (the call returns once that grid has completed and its writes are visible; it is a no-op in a
// kernel launched the ordinary way).  pdl_trigger() in the predecessor lets the dependent's CTAs take free SM
// resources early -- they then sit in pdl_wait().
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
