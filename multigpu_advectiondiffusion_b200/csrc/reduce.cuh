// max|u| over a pitched field, for the dynamic time step dt = CFL*dx/max|u| of the reference's Matlab prototype
// (Matlab_Prototipes/InviscidBurgersNd/LFWENO5FDM3d.m:71).  One pass over the field: grid-stride
// rows, warp-shuffle and shared-memory reduction per CTA, one atomicMax per CTA on the IEEE bit pattern (ordered
// like the values for non-negative numbers).  HBM-bound: 8 B per cell, 1/8 of a time step's algorithmic traffic.
#pragma once
#include "common.cuh"

namespace adv {

template <typename REAL>
__global__ void __launch_bounds__(256)
absmax_kernel(const REAL* __restrict__ u, int nx, long long pitch, long long rows, unsigned long long* __restrict__ out)
{
    double m = 0.0;
    for (long long r = blockIdx.x; r < rows; r += gridDim.x) {
        const REAL* __restrict__ row = u + pitch * r;
        for (int i = threadIdx.x; i < nx; i += blockDim.x) m = fmax(m, fabs((double)row[i]));
    }
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    __shared__ double wm[8];
    if ((threadIdx.x & 31) == 0) wm[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; w++) m = fmax(m, wm[w]);
        atomicMax(out, (unsigned long long)__double_as_longlong(m));
    }
}

}  // namespace adv
