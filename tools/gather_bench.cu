// micro-benchmark: how fast can one SM pull scattered row pieces (the rescoring stage's access pattern) into shared
// memory?  Table: N rows x 1536 fp32 (6 KB rows).  Each CTA owns 64 random rows and walks them left to right in
// pieces of PIECE bytes per row, NSTAGE pieces-of-64-rows in flight.  Variants:
//   0  cp.async.cg 16 B per lane     1  cp.async.ca 16 B per lane     2  ld.global.v4 + st.shared
//   3  cp.async.bulk (one per row piece, mbarrier)
// Prints GB/s over the whole GPU for each (variant, piece, CTAs per SM).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
template <int VAR, int PIECE, int NSTAGE>
__global__ void __launch_bounds__(128) k(const float *tab, const int *ids, int dim, float *sink) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) u64 bar[NSTAGE];
    const int rows = 64, pf = PIECE / 4, pitch = pf + 4;
    float *ring = reinterpret_cast<float *>(smem);
    const int *my = ids + blockIdx.x * rows;
    const int npieces = dim / pf;
    if (VAR == 3 && threadIdx.x == 0) {
        for (int s = 0; s < NSTAGE; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar[s])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    float acc = 0.f;
    const int lpr = PIECE / 16;                          // lanes per row piece
    auto issue = [&](int c) {
        float *dst = ring + (size_t)(c % NSTAGE) * rows * pitch;
        if (VAR == 3) {
            if (threadIdx.x == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&bar[c % NSTAGE])), "r"(rows * PIECE) : "memory");
            __syncthreads();
            if (threadIdx.x < rows) {
                const float *src = tab + (size_t)my[threadIdx.x] * dim + c * pf;
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(s32(dst + threadIdx.x * pitch)), "l"(src), "r"(PIECE), "r"(s32(&bar[c % NSTAGE])) : "memory");
            }
            return;
        }
        for (int f = threadIdx.x; f < rows * lpr; f += 128) {
            const int r = f / lpr, l = f % lpr;
            const float *src = tab + (size_t)my[r] * dim + c * pf + l * 4;
            float *d = dst + r * pitch + l * 4;
            if (VAR == 0) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s32(d)), "l"(src) : "memory");
            else if (VAR == 1) asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(s32(d)), "l"(src) : "memory");
            else { float4 v = *reinterpret_cast<const float4 *>(src); *reinterpret_cast<float4 *>(d) = v; }
        }
        if (VAR < 2) asm volatile("cp.async.commit_group;" ::: "memory");
    };
    for (int c = 0; c < NSTAGE - 1 && c < npieces; ++c) issue(c);
    for (int c = 0; c < npieces; ++c) {
        if (c + NSTAGE - 1 < npieces) issue(c + NSTAGE - 1);
        else if (VAR < 2) asm volatile("cp.async.commit_group;" ::: "memory");
        if (VAR < 2) asm volatile("cp.async.wait_group %0;" ::"n"(NSTAGE - 1) : "memory");
        if (VAR == 3) {
            const uint32_t par = (c / NSTAGE) & 1;
            asm volatile("{\n\t.reg .pred p;\n\tW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@!p bra W;\n\t}" ::"r"(s32(&bar[c % NSTAGE])), "r"(par) : "memory");
        }
        __syncthreads();
        acc += ring[(size_t)(c % NSTAGE) * rows * pitch + (threadIdx.x % rows) * pitch + (threadIdx.x / rows)];
        __syncthreads();
    }
    if (acc == 1.2345f) sink[0] = acc;
}
template <int VAR, int PIECE, int NSTAGE>
void run(const float *tab, const int *ids, float *sink, int n_cta, const char *name) {
    const int dim = 1536;
    size_t smem = (size_t)NSTAGE * 64 * (PIECE / 4 + 4) * 4;
    cudaFuncSetAttribute(k<VAR, PIECE, NSTAGE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    k<VAR, PIECE, NSTAGE><<<n_cta, 128, smem>>>(tab, ids, dim, sink);
    cudaEventRecord(a);
    k<VAR, PIECE, NSTAGE><<<n_cta, 128, smem>>>(tab, ids, dim, sink);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    double bytes = (double)n_cta * 64 * dim * 4;
    printf("%-22s piece %4d B stages %2d ctas %4d smem %3zu KB : %7.1f us  %7.1f GB/s  (%s)\n", name, PIECE, NSTAGE, n_cta, smem >> 10, ms * 1e3, bytes / ms / 1e6,
           cudaGetErrorString(cudaGetLastError()));
}
int main() {
    const int N = 100000, dim = 1536;
    float *tab, *sink; int *ids;
    cudaMalloc(&tab, (size_t)N * dim * 4); cudaMemset(tab, 0, (size_t)N * dim * 4);
    cudaMalloc(&sink, 4);
    const int max_cta = 1024;
    int *h = (int *)malloc(max_cta * 64 * 4);
    unsigned x = 12345;
    for (int i = 0; i < max_cta * 64; ++i) { x = x * 1664525u + 1013904223u; h[i] = (x >> 8) % N; }
    cudaMalloc(&ids, max_cta * 64 * 4); cudaMemcpy(ids, h, max_cta * 64 * 4, cudaMemcpyHostToDevice);
    for (int n_cta : {64, 256, 592}) {
        run<0, 256, 5>(tab, ids, sink, n_cta, "cp.async.cg");
        run<1, 256, 5>(tab, ids, sink, n_cta, "cp.async.ca");
        run<2, 256, 5>(tab, ids, sink, n_cta, "ldg.v4 + sts");
        run<3, 256, 5>(tab, ids, sink, n_cta, "bulk copy");
        run<0, 512, 3>(tab, ids, sink, n_cta, "cp.async.cg");
        run<3, 512, 3>(tab, ids, sink, n_cta, "bulk copy");
        run<0, 1024, 3>(tab, ids, sink, n_cta, "cp.async.cg");
        run<3, 1024, 3>(tab, ids, sink, n_cta, "bulk copy");
        run<3, 2048, 1>(tab, ids, sink, n_cta, "bulk copy");
        run<3, 3072, 1>(tab, ids, sink, n_cta, "bulk copy");
    }
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
}
