// Measures the FP64 FMA issue peak of the device (DFMA/clk/SM) — the co-bound of the VBD colour pass.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/fp64_peak tools/fp64_peak.cu
#include <cstdio>
#include <cuda_runtime.h>
__global__ void dfma(double *out, int iters)
{
    double a[8];
    for (int k = 0; k < 8; ++k) a[k] = threadIdx.x * 1e-3 + k;
    const double b = 1.0000001, c = 1e-9;
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = fma(a[k], b, c);
    double s = 0;
    for (int k = 0; k < 8; ++k) s += a[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main()
{
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const int blocks = p.multiProcessorCount * 4, threads = 512, iters = 20000;
    double *out;
    cudaMalloc(&out, sizeof(double) * blocks * threads);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    dfma<<<blocks, threads>>>(out, 100);
    float best = 1e30f;
    for (int r = 0; r < 5; ++r) {
        cudaEventRecord(e0);
        dfma<<<blocks, threads>>>(out, iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        best = ms < best ? ms : best;
    }
    const double fmas = double(blocks) * threads * iters * 8;
    printf("{\"sms\": %d, \"dfma_per_s\": %.4e, \"fp64_tflops\": %.2f, \"ms\": %.3f}\n", p.multiProcessorCount, fmas / (best * 1e-3),
           2 * fmas / (best * 1e-3) / 1e12, best);
    return cudaGetLastError() != cudaSuccess;
}
