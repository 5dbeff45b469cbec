#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__global__ void k_dmma(double* out, int iters, double a, double b)
{
    double c[16];
    for (int i = 0; i < 16; i++) c[i] = threadIdx.x + i;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 8; u++) dmma(c[2 * u], c[2 * u + 1], a, b);
    }
    double s = 0;
    for (int i = 0; i < 16; i++) s += c[i];
    if (s == 1.2345) out[0] = s;
}
__global__ void k_dfma(double* out, int iters, double a, double b)
{
    double c[8];
    for (int i = 0; i < 8; i++) c[i] = threadIdx.x + i;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 8; u++) c[u] = fma(c[u], a, b);
    }
    double s = 0;
    for (int i = 0; i < 8; i++) s += c[i];
    if (s == 1.2345) out[0] = s;
}
int main()
{
    double* d;
    cudaMalloc(&d, 8);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0), cudaEventCreate(&e1);
    for (int warps : {1, 2, 4, 8, 16}) {
        int blocks = 148 * 4, threads = warps * 32, iters = 20000;
        float ms;
        k_dmma<<<blocks, threads>>>(d, 100, 1.0, 1e-9);
        cudaDeviceSynchronize();
        cudaEventRecord(e0);
        k_dmma<<<blocks, threads>>>(d, iters, 1.0, 1e-9);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        double fl = 2.0 * 256 * 8.0 * iters * blocks * warps;
        printf("DMMA  warps/block %2d: %.2f TFLOP/s (%.3f ms)\n", warps, fl / ms * 1e-9, ms);
        cudaEventRecord(e0);
        k_dfma<<<blocks, threads>>>(d, iters, 0.999, 1e-9);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        fl = 2.0 * 32 * 8.0 * iters * blocks * warps;
        printf("DFMA  warps/block %2d: %.2f TFLOP/s (%.3f ms)\n", warps, fl / ms * 1e-9, ms);
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
