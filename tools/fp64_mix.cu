// How much of the FP64 peak survives when the issue port is shared with other instructions?
// Per DFMA, R integer IMADs (independent chains) are interleaved; reports DFMA/s for R = 0..3 and for
// 16 and 32 resident warps per SM.  (The WENO5 kernel issues ~1.15 non-FP64 instructions per FP64 one.)
#include <cstdio>
#include <cuda_runtime.h>

template <int R>
__global__ void mix(double* out, int iters, double a, double b, int ia, int ib)
{
    double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    int j0 = threadIdx.x, j1 = j0 + 1, j2 = j0 + 2, j3 = j0 + 3, j4 = j0 + 4, j5 = j0 + 5, j6 = j0 + 6, j7 = j0 + 7;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
            x0 = fma(x0, a, b); if (R > 0) j0 = j0 * ia + ib; if (R > 1) j1 = j1 * ia + ib; if (R > 2) j2 = j2 * ia + ib;
            x1 = fma(x1, a, b); if (R > 0) j3 = j3 * ia + ib; if (R > 1) j4 = j4 * ia + ib; if (R > 2) j5 = j5 * ia + ib;
            x2 = fma(x2, a, b); if (R > 0) j6 = j6 * ia + ib; if (R > 1) j7 = j7 * ia + ib; if (R > 2) j0 = j0 * ia + ib;
            x3 = fma(x3, a, b); if (R > 0) j1 = j1 * ia + ib; if (R > 1) j2 = j2 * ia + ib; if (R > 2) j3 = j3 * ia + ib;
            x4 = fma(x4, a, b); if (R > 0) j4 = j4 * ia + ib; if (R > 1) j5 = j5 * ia + ib; if (R > 2) j6 = j6 * ia + ib;
            x5 = fma(x5, a, b); if (R > 0) j7 = j7 * ia + ib; if (R > 1) j0 = j0 * ia + ib; if (R > 2) j1 = j1 * ia + ib;
            x6 = fma(x6, a, b); if (R > 0) j2 = j2 * ia + ib; if (R > 1) j3 = j3 * ia + ib; if (R > 2) j4 = j4 * ia + ib;
            x7 = fma(x7, a, b); if (R > 0) j5 = j5 * ia + ib; if (R > 1) j6 = j6 * ia + ib; if (R > 2) j7 = j7 * ia + ib;
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7 + (double)(j0 + j1 + j2 + j3 + j4 + j5 + j6 + j7);
}

template <int R>
double run(int sms, int blocks_per_sm, int iters)
{
    const int blocks = sms * blocks_per_sm, threads = 256;
    double* out;
    cudaMalloc(&out, sizeof(double) * blocks * threads);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    mix<R><<<blocks, threads>>>(out, iters / 10, 0.999, 0.001, 3, 7);
    cudaDeviceSynchronize();
    double best = 0;
    for (int rep = 0; rep < 3; rep++) {
        cudaEventRecord(e0);
        mix<R><<<blocks, threads>>>(out, iters, 0.999, 0.001, 3, 7);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        const double rate = (double)blocks * threads * iters * 32.0 / (ms * 1e-3);
        if (rate > best) best = rate;
    }
    cudaFree(out);
    return best;
}

int main()
{
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    for (int bps = 2; bps <= 4; bps *= 2) {
        printf("{\"warps_per_sm\": %d, \"dfma_per_s\": {\"R0\": %.3e, \"R1\": %.3e, \"R2\": %.3e, \"R3\": %.3e}}\n", bps * 8,
               run<0>(p.multiProcessorCount, bps, 4000), run<1>(p.multiProcessorCount, bps, 4000),
               run<2>(p.multiProcessorCount, bps, 4000), run<3>(p.multiProcessorCount, bps, 4000));
    }
    return 0;
}
