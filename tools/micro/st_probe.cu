// Developer microbenchmark: cost (issue-side) of a burst of scattered global stores / prefetches from one warp.
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ long long clk() { long long c; asm volatile("mov.u64 %0, %%clock64;" : "=l"(c) :: "memory"); return c; }
__global__ void k(int mode, int active, int* a, int* b, double* c, float* d, int* e, int iters, long long* out)
{
    const int lane = threadIdx.x;
    long long tot = 0;
    for (int it = 0; it < iters; it++)
    {
        const int idx = ((lane * 7919 + it * 104729) & 0xffff) * 33 % 1000003;
        __syncwarp();
        const long long t0 = clk();
        if (lane < active)
        {
            if (mode >= 1) { a[idx] = it; }
            if (mode >= 2) { b[idx] = it; c[idx] = it; d[idx] = it; e[idx] = it; }
            if (mode >= 3) { asm volatile("prefetch.global.L1 [%0];" :: "l"(a + ((idx * 13) % 1000003))); asm volatile("prefetch.global.L1 [%0];" :: "l"(b + ((idx * 17) % 1000003))); }
        }
        __syncwarp();
        const long long t1 = clk();
        tot += t1 - t0;
        // let the stores drain a little: dependent ALU work
        int x = it;
        for (int k2 = 0; k2 < 200; k2++) x = x * 1664525 + 1013904223;
        if (x == 12345) a[0] = x;
    }
    if (lane == 0) out[0] = tot / iters;
}
int main()
{
    int *a, *b, *e; double* c; float* d; long long* out;
    cudaMalloc(&a, 4 << 20); cudaMalloc(&b, 4 << 20); cudaMalloc(&e, 4 << 20); cudaMalloc(&c, 8 << 20); cudaMalloc(&d, 4 << 20);
    cudaMallocManaged(&out, 64);
    for (int mode = 0; mode <= 3; mode++)
        for (int active : {1, 8, 32})
        {
            k<<<1, 32>>>(mode, active, a, b, c, d, e, 2000, out); cudaDeviceSynchronize();
            printf("mode %d (0 none, 1: 1 store, 2: 5 stores, 3: +2 prefetch) active lanes %2d: %lld cycles\n", mode, active, out[0]);
        }
    return 0;
}
