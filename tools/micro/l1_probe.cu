// Developer microbenchmark (not product): latency of the load patterns the narrow-round controller
// depends on, on one SM.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o l1_probe l1_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ long long clk() { long long c; asm volatile("mov.u64 %0, %%clock64;" : "=l"(c)); return c; }

// mode 0: dependent plain loads of one address (L1 hit latency)
// mode 1: store then plain load of the same address, dependent chain
// mode 2: dependent ld.cg loads (L2 hit latency)
// mode 3: store to the line, then plain load of ANOTHER word of the same line
// mode 4: plain load chain over a 64 KB pointer ring (L1-resident after the first lap)
// mode 5: same ring, but a store to each visited line before moving on
__global__ void probe(int mode, int* buf, int iters, long long* out)
{
    if (threadIdx.x != 0) return;
    volatile int sink = 0;
    int idx = 0;
    // warm
    for (int k = 0; k < 64; k++) idx = buf[idx];
    idx = 0;
    long long t0 = clk();
    if (mode == 0)
        for (int k = 0; k < iters; k++) { int v; asm volatile("ld.global.ca.s32 %0, [%1];" : "=r"(v) : "l"(buf + idx)); idx = v; }
    else if (mode == 1)
        for (int k = 0; k < iters; k++) { asm volatile("st.global.s32 [%0], %1;" :: "l"(buf + idx), "r"(0) : "memory"); int v; asm volatile("ld.global.ca.s32 %0, [%1];" : "=r"(v) : "l"(buf + idx) : "memory"); idx = v; }
    else if (mode == 2)
        for (int k = 0; k < iters; k++) { int v; asm volatile("ld.global.cg.s32 %0, [%1];" : "=r"(v) : "l"(buf + idx)); idx = v; }
    else if (mode == 3)
        for (int k = 0; k < iters; k++) { asm volatile("st.global.s32 [%0], %1;" :: "l"(buf + idx + 1), "r"(0) : "memory"); int v; asm volatile("ld.global.ca.s32 %0, [%1];" : "=r"(v) : "l"(buf + idx) : "memory"); idx = v; }
    else if (mode == 4)
        for (int k = 0; k < iters; k++) { int v; asm volatile("ld.global.ca.s32 %0, [%1];" : "=r"(v) : "l"(buf + 32 + idx) : "memory"); idx = v; }
    else if (mode == 5)
        for (int k = 0; k < iters; k++) { int v; asm volatile("ld.global.ca.s32 %0, [%1];" : "=r"(v) : "l"(buf + 32 + idx) : "memory"); asm volatile("st.global.s32 [%0], %1;" :: "l"(buf + 32 + idx + 1), "r"(k) : "memory"); idx = v; }
    long long t1 = clk();
    sink = idx;
    out[mode] = (t1 - t0) / iters;
    out[16 + mode] = idx;
}

// barrier cost with 1024 threads
__global__ void bar_probe(int iters, long long* out)
{
    long long t0 = clk();
    for (int k = 0; k < iters; k++) __syncthreads();
    long long t1 = clk();
    if (threadIdx.x == 0) out[8] = (t1 - t0) / iters;
}

int main()
{
    int* buf; long long* out;
    cudaMalloc(&buf, 1 << 20); cudaMallocManaged(&out, 128);
    // ring: element at 32 + i*32 points to the next line (stride 128 B), 512 lines = 64 KB
    int h[1 << 18];
    for (int i = 0; i < (1 << 18); i++) h[i] = 0;
    for (int i = 0; i < 512; i++) h[32 + i * 32] = ((i + 1) % 512) * 32;
    cudaMemcpy(buf, h, sizeof h, cudaMemcpyHostToDevice);
    for (int m = 0; m < 6; m++) { probe<<<1, 32>>>(m, buf, 4096, out); cudaDeviceSynchronize(); cudaMemcpy(buf, h, sizeof h, cudaMemcpyHostToDevice); }
    bar_probe<<<1, 1024>>>(1000, out); cudaDeviceSynchronize();
    printf("L1 hit chain %lld | st+ld same addr %lld | ld.cg (L2) %lld | st other word + ld %lld | ring L1 %lld | ring with stores %lld | bar.sync 1024thr %lld cycles\n",
           out[0], out[1], out[2], out[3], out[4], out[5], out[8]);
    return 0;
}
