// Measures the FP64 FMA issue peak of the device (DFMA chains, no memory traffic): the ceiling
// that bounds the double-precision WENO5 kernel.  Prints one JSON line.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/fp64_peak tools/fp64_peak.cu
#include <cstdio>
#include <cuda_runtime.h>

template <typename T>
__global__ void fma_chains(T* out, int iters, T a, T b)
{
    T x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int r = 0; r < 8; r++) {
            x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
            x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}

template <typename T>
double run(int sms, int iters)
{
    const int blocks = sms * 8, threads = 256;
    T* out;
    cudaMalloc(&out, sizeof(T) * blocks * threads);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    fma_chains<T><<<blocks, threads>>>(out, iters / 10, T(0.999), T(0.001));
    cudaDeviceSynchronize();
    double best = 0;
    for (int rep = 0; rep < 5; rep++) {
        cudaEventRecord(e0);
        fma_chains<T><<<blocks, threads>>>(out, iters, T(0.999), T(0.001));
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        const double fmas = (double)blocks * threads * iters * 64.0;
        const double rate = fmas / (ms * 1e-3);
        if (rate > best) best = rate;
    }
    cudaFree(out);
    return best;
}

int main()
{
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, 0) != cudaSuccess) { printf("{\"error\": \"no device\"}\n"); return 1; }
    const double d = run<double>(p.multiProcessorCount, 4000);
    const double f = run<float>(p.multiProcessorCount, 8000);
    printf("{\"device\": \"%s\", \"sms\": %d, \"fp64_fma_per_s\": %.4e, \"fp64_tflops\": %.2f, \"fp32_fma_per_s\": %.4e, \"fp32_tflops\": %.2f, "
           "\"fp64_fma_per_clk_per_sm_at_1965MHz\": %.1f}\n",
           p.name, p.multiProcessorCount, d, 2 * d / 1e12, f, 2 * f / 1e12, d / p.multiProcessorCount / 1.965e9);
    return 0;
}
