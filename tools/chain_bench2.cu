// micro-benchmark of the exact rescoring chain (lane = row, fp32 row tile and fp64 query in shared memory):
//   mode 0  F2F.F64.F32 on the chain (the conversion shares the fp64 pipe with the DFMA)
//   mode 1  integer widening of normal fp32 values (add, mad.wide, lop3), software-pipelined one group ahead
//   mode 2  mode 1 + detection of exponent 0 / 255 (the values the integer widening gets wrong)
//   mode 3  mode 0, one active lane per warp (latency)
// Block = 4 chain warps (one per SM sub-partition) + a 5th warp whose lane 0 walks the query against itself.
#include <cstdio>
#include <cuda_runtime.h>
#define COLS 64
#define PITCH 68
__device__ __forceinline__ double widen(unsigned b) {   // normal finite values only
    unsigned s = b + b;
    unsigned long long w;
    asm("mad.wide.u32 %0, %1, 0x10000000, 0x3800000000000000;" : "=l"(w) : "r"(s));
    unsigned hi = (unsigned)(w >> 32), lo = (unsigned)w;
    hi |= b & 0x80000000u;
    return __hiloint2double((int)hi, (int)lo);
}
__global__ void k(double *o, int iters, long long *cyc, int mode) {
    __shared__ __align__(16) float tile[4][32 * PITCH];
    __shared__ __align__(16) double q[COLS];
    for (int i = threadIdx.x; i < 4 * 32 * PITCH; i += blockDim.x) (&tile[0][0])[i] = 1.0f + i * 1e-6f;
    for (int i = threadIdx.x; i < COLS; i += blockDim.x) q[i] = 0.5 + i * 1e-3;
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    double acc = 0;
    unsigned mx = 0;
    long long t0 = clock64();
    if (warp < 4) {
        const float *e = tile[warp] + lane * PITCH;
        if (mode == 0 || (mode == 3 && lane == 0)) {
            for (int it = 0; it < iters; ++it) {
#pragma unroll 4
                for (int i = 0; i < COLS; i += 4) {
                    float4 v = *reinterpret_cast<const float4 *>(e + i);
                    double2 u0 = *reinterpret_cast<const double2 *>(q + i), u1 = *reinterpret_cast<const double2 *>(q + i + 2);
                    acc = fma(u0.x, (double)v.x, acc); acc = fma(u0.y, (double)v.y, acc);
                    acc = fma(u1.x, (double)v.z, acc); acc = fma(u1.y, (double)v.w, acc);
                }
            }
        } else if (mode == 1 || mode == 2) {
            double nx[8];
            auto conv = [&](int i) {
                uint4 a = *reinterpret_cast<const uint4 *>(e + i), b = *reinterpret_cast<const uint4 *>(e + i + 4);
                unsigned r[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    nx[u] = widen(r[u]);
                    if (mode == 2) mx = max(mx, (r[u] + r[u]) - 0x01000000u);
                }
            };
            for (int it = 0; it < iters; ++it) {
                conv(0);
#pragma unroll 1
                for (int i = 0; i < COLS; i += 8) {
                    double cu[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) cu[u] = nx[u];
                    conv((i + 8) & (COLS - 1));
                    const double2 *qq = reinterpret_cast<const double2 *>(q + i);
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        double2 w = qq[u];
                        acc = fma(w.x, cu[2 * u], acc); acc = fma(w.y, cu[2 * u + 1], acc);
                    }
                }
            }
        }
    } else if (lane == 0) {
        for (int it = 0; it < iters; ++it) {
#pragma unroll 4
            for (int i = 0; i < COLS; i += 2) {
                double2 u = *reinterpret_cast<const double2 *>(q + i);
                acc = fma(u.x, u.x, acc); acc = fma(u.y, u.y, acc);
            }
        }
    }
    long long t1 = clock64();
    o[threadIdx.x + blockIdx.x * blockDim.x] = acc + (double)mx;
    if (lane == 0 && blockIdx.x == 0) cyc[warp] = t1 - t0;
}
int main() {
    double *o; long long *c, h[8];
    cudaMalloc(&o, 1 << 22); cudaMalloc(&c, 64);
    int iters = 200;
    const char *names[] = {"F2F on the chain", "integer widening, pipelined", "integer widening + special detection", "F2F, one lane per warp (latency)"};
    for (int mode = 0; mode < 4; ++mode)
        for (int threads : {128, 160, 288}) {
            cudaMemset(c, 0, 64);
            k<<<148, threads>>>(o, iters, c, mode);
            cudaMemcpy(h, c, 64, cudaMemcpyDeviceToHost);
            printf("%-40s threads %3d : cycles/element per warp:", names[mode], threads);
            for (int w = 0; w < 5 && w < threads / 32; ++w) printf(" %.1f", (double)h[w] / (iters * COLS));
            printf("\n");
        }
    double ho[4];
    cudaMemcpy(ho, o, 32, cudaMemcpyDeviceToHost);
    printf("%s %.17g\n", cudaGetErrorString(cudaDeviceSynchronize()), ho[0]);
}
