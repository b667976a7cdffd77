// micro-benchmark: dependent-chain latency of the fp64 ops the exact rescoring chain is made of (B200, sm_100a)
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k_dfma(double *o, double a, double b, int n, long long *cyc) {
    double acc = a;
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < n; ++i) acc = fma(acc, b, a);
    long long t1 = clock64();
    o[threadIdx.x] = acc; if (threadIdx.x == 0) *cyc = t1 - t0;
}
__global__ void k_dadd(double *o, double a, double b, int n, long long *cyc) {
    double acc = a;
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < n; ++i) acc = __dadd_rn(acc, b);
    long long t1 = clock64();
    o[threadIdx.x] = acc; if (threadIdx.x == 0) *cyc = t1 - t0;
}
__global__ void k_f2f_chain(double *o, float a, int n, long long *cyc) {
    float x = a; double acc = 0;
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < n; ++i) { double d = (double)x; acc = __dadd_rn(acc, d); x = (float)acc; }
    long long t1 = clock64();
    o[threadIdx.x] = acc; if (threadIdx.x == 0) *cyc = t1 - t0;
}
__global__ void k_f2f_tput(double *o, const float *in, int n, long long *cyc) {
    double acc0 = 0, acc1 = 0, acc2 = 0, acc3 = 0;
    float x0 = in[0], x1 = in[1], x2 = in[2], x3 = in[3];
    long long t0 = clock64();
    for (int i = 0; i < n; ++i) {   // 4 independent conversions + xor to keep them alive
        acc0 += (double)x0; acc1 += (double)x1; acc2 += (double)x2; acc3 += (double)x3;
        x0 += 1.0f; x1 += 1.0f; x2 += 1.0f; x3 += 1.0f;
    }
    long long t1 = clock64();
    o[threadIdx.x] = acc0 + acc1 + acc2 + acc3; if (threadIdx.x == 0) *cyc = t1 - t0;
}
__global__ void k_ffma(float *o, float a, float b, int n, long long *cyc) {
    float acc = a;
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < n; ++i) acc = fmaf(acc, b, a);
    long long t1 = clock64();
    o[threadIdx.x] = acc; if (threadIdx.x == 0) *cyc = t1 - t0;
}
int main() {
    double *o; float *of; long long *c, h; float *in;
    cudaMalloc(&o, 8192); cudaMalloc(&of, 8192); cudaMalloc(&c, 8); cudaMalloc(&in, 64); cudaMemset(in, 0, 64);
    int n = 4096;
    for (int warps = 1; warps <= 8; warps *= 2) {
        k_dfma<<<1, 32 * warps>>>(o, 1.0, 0.999, n, c); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost); printf("warps %d  dfma chain %.1f cyc/op", warps, (double)h / n);
        k_dadd<<<1, 32 * warps>>>(o, 1.0, 0.5, n, c);   cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost); printf("  dadd chain %.1f", (double)h / n);
        k_f2f_chain<<<1, 32 * warps>>>(o, 1.0f, n, c);  cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost); printf("  f2f+dadd+d2f chain %.1f", (double)h / n);
        k_f2f_tput<<<1, 32 * warps>>>(o, in, n, c);     cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost); printf("  4x(f2f+dadd) indep %.1f cyc/iter", (double)h / n);
        k_ffma<<<1, 32 * warps>>>(of, 1.0f, 0.999f, n, c); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost); printf("  ffma chain %.1f\n", (double)h / n);
    }
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
