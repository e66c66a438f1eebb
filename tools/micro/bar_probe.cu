// Developer microbenchmark: does one working warp run slower while 31 warps of its CTA wait at bar.sync?
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ long long clk() { long long c; asm volatile("mov.u64 %0, %%clock64;" : "=l"(c)); return c; }
__global__ void k(int iters, int chain, long long* out, int* gbuf)
{
    __shared__ int ring[64];
    __shared__ double dsum;
    if (threadIdx.x < 64) ring[threadIdx.x] = (threadIdx.x + 1) & 63;
    __syncthreads();
    long long t0 = clk();
    int idx = 0; double acc = 1.0;
    for (int it = 0; it < iters; it++)
    {
        if (threadIdx.x < 32)
        {
            for (int c = 0; c < chain; c++) idx = ring[idx];          // dependent LDS chain
            // a 5-level shuffle reduction of doubles, like the top-2 merge
            double v = acc + idx + threadIdx.x;
            for (int m = 16; m >= 1; m >>= 1) { double o = __shfl_xor_sync(0xffffffffu, v, m); v = o > v ? o : v; }
            acc = v * 0.5;
            if (threadIdx.x == 0) gbuf[it & 1023] = idx;               // a global store
        }
        __syncthreads();
        if (threadIdx.x == 0) dsum = acc;
        __syncthreads();
    }
    long long t1 = clk();
    if (threadIdx.x == 0) { out[0] = (t1 - t0) / iters; out[1] = idx + (int)acc; }
}
int main()
{
    long long* out; int* g; cudaMallocManaged(&out, 64); cudaMalloc(&g, 4096);
    for (int threads : {32, 256, 1024})
        for (int chain : {0, 20})
        {
            k<<<1, threads>>>(2000, chain, out, g); cudaDeviceSynchronize();
            printf("threads %4d chain %2d: %lld cycles/iter\n", threads, chain, out[0]);
        }
    return 0;
}
