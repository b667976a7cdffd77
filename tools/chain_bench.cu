// micro-benchmark of the rescoring chain: one warp, lane l walks row l of a smem tile (pitch 68 floats) against an
// fp64 query in smem.  Variants: (a) F2F on the chain, (b) row pre-converted to fp64 (no F2F), (c) a, 4 warps per SM sub-partition
#include <cstdio>
#include <cuda_runtime.h>
#define COLS 64
#define PITCH 68
__global__ void k_chain(double *o, int iters, long long *cyc, int variant) {
    __shared__ __align__(16) float tile[32 * PITCH];
    __shared__ __align__(16) double tiled[32 * (COLS + 2)];
    __shared__ __align__(16) double q[COLS];
    for (int i = threadIdx.x; i < 32 * PITCH; i += blockDim.x) tile[i] = 1.0f + i * 1e-6f;
    for (int i = threadIdx.x; i < 32 * (COLS + 2); i += blockDim.x) tiled[i] = 1.0 + i * 1e-6;
    for (int i = threadIdx.x; i < COLS; i += blockDim.x) q[i] = 0.5 + i * 1e-3;
    __syncthreads();
    int lane = threadIdx.x & 31;
    double acc = 0;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        if (variant == 0) {
            const float *e = tile + lane * PITCH;
#pragma unroll 4
            for (int i = 0; i < COLS; i += 4) {
                float4 v = *reinterpret_cast<const float4 *>(e + i);
                double2 u0 = *reinterpret_cast<const double2 *>(q + i), u1 = *reinterpret_cast<const double2 *>(q + i + 2);
                acc = fma(u0.x, (double)v.x, acc); acc = fma(u0.y, (double)v.y, acc);
                acc = fma(u1.x, (double)v.z, acc); acc = fma(u1.y, (double)v.w, acc);
            }
        } else {
            const double *e = tiled + lane * (COLS + 2);
#pragma unroll 4
            for (int i = 0; i < COLS; i += 2) {
                double2 v = *reinterpret_cast<const double2 *>(e + i);
                double2 u0 = *reinterpret_cast<const double2 *>(q + i);
                acc = fma(u0.x, v.x, acc); acc = fma(u0.y, v.y, acc);
            }
        }
    }
    long long t1 = clock64();
    o[threadIdx.x + blockIdx.x * blockDim.x] = acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
    double *o; long long *c, h;
    cudaMalloc(&o, 1 << 20); cudaMalloc(&c, 8);
    int iters = 200;
    for (int variant = 0; variant < 2; ++variant)
        for (int blocks : {1, 148 * 4})
            for (int threads : {32, 128}) {
                k_chain<<<blocks, threads>>>(o, iters, c, variant);
                cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
                printf("variant %d blocks %4d threads %3d : %.1f cycles / element\n", variant, blocks, threads, (double)h / (iters * COLS));
            }
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
}
