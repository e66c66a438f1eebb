// Developer microbenchmark: dependent-chain latency of warp primitives on one warp.
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ long long clk() { long long c; asm volatile("mov.u64 %0, %%clock64;" : "=l"(c) :: "memory"); return c; }
__global__ void k(long long* out, double* dsink)
{
    const int lane = threadIdx.x;
    unsigned x = lane * 2654435761u + 12345u;
    const int N = 512;
    long long t0, t1;
    t0 = clk();
    for (int i = 0; i < N; i++) x = __reduce_max_sync(0xffffffffu, x ^ lane) + i;
    t1 = clk(); if (lane == 0) out[0] = (t1 - t0) / N;
    t0 = clk();
    for (int i = 0; i < N; i++) x = __match_any_sync(0xffffffffu, (x + lane) & 7) + i;
    t1 = clk(); if (lane == 0) out[1] = (t1 - t0) / N;
    t0 = clk();
    for (int i = 0; i < N; i++) x = __ballot_sync(0xffffffffu, (x + lane) & 1) + i;
    t1 = clk(); if (lane == 0) out[2] = (t1 - t0) / N;
    t0 = clk();
    for (int i = 0; i < N; i++) x = __shfl_xor_sync(0xffffffffu, x, 1) + i;
    t1 = clk(); if (lane == 0) out[3] = (t1 - t0) / N;
    double d = x * 1e-3;
    t0 = clk();
    for (int i = 0; i < N; i++) d = d + 1.5;
    t1 = clk(); if (lane == 0) out[4] = (t1 - t0) / N;
    t0 = clk();
    for (int i = 0; i < N; i++) d = (d > 3.0 * i) ? d - 1.0 : d + 2.0;   // DSETP + select chain
    t1 = clk(); if (lane == 0) out[5] = (t1 - t0) / N;
    t0 = clk();
    for (int i = 0; i < N; i++) x = __match_any_sync(0xffffffffu, x + lane + i);   // all distinct
    t1 = clk(); if (lane == 0) out[6] = (t1 - t0) / N;
    dsink[lane] = d + x;
}
int main()
{
    long long* out; double* ds; cudaMallocManaged(&out, 128); cudaMalloc(&ds, 1024);
    k<<<1, 32>>>(out, ds); cudaDeviceSynchronize();
    printf("redux.max %lld | match.any(8 groups) %lld | ballot %lld | shfl %lld | DADD %lld | DSETP+sel+DADD %lld | match.any(distinct) %lld cycles\n",
           out[0], out[1], out[2], out[3], out[4], out[5], out[6]);
    return 0;
}
