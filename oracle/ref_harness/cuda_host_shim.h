/* TEST INFRASTRUCTURE ONLY (oracle/_ref build).
 *
 * Prelude that lets the reference's CUDA kernel bodies compile as host C++
 * with plain g++ (no nvcc, no GPU): the CUDA built-in index variables become
 * host globals that a loop nest (launch()) sweeps, so every reference
 * __global__ function is executed one "thread" at a time in program order.
 * Valid because none of the replayed kernels use __shared__/__syncthreads.
 * Nothing from /root/reference is copied: its sources are read in place by
 * oracle/Makefile and piped into the compiler after this prelude.
 */
#ifndef CUDA_HOST_SHIM_H
#define CUDA_HOST_SHIM_H
#include <stdlib.h>
#include <stdio.h>
#include <string.h>
#include <math.h>
#include <cuda_runtime.h>

static uint3 threadIdx, blockIdx;
static dim3 blockDim, gridDim;
#define register

template <class Body>
static inline void launch(dim3 grid, dim3 block, Body body)
{
    gridDim = grid;
    blockDim = block;
    for (unsigned bz = 0; bz < grid.z; bz++)
    for (unsigned by = 0; by < grid.y; by++)
    for (unsigned bx = 0; bx < grid.x; bx++) {
        blockIdx.x = bx; blockIdx.y = by; blockIdx.z = bz;
        for (unsigned tz = 0; tz < block.z; tz++)
        for (unsigned ty = 0; ty < block.y; ty++)
        for (unsigned tx = 0; tx < block.x; tx++) {
            threadIdx.x = tx; threadIdx.y = ty; threadIdx.z = tz;
            body();
        }
    }
}
#define LAUNCH(grid, block, call) launch((grid), (block), [&]() { call; })
#endif
