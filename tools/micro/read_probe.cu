// read_probe.cu — what a read-only streaming kernel reaches on this GPU (the bidding scan's ceiling).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o read_probe read_probe.cu && ./read_probe
// 1 GiB of floats, every variant sums them (one FADD per element, like the filter's FADD + FMNMX).
//   A  grid-stride tiles: a CTA takes contiguous tiles of threads x U x 16 bytes
//   B  warp per 64 KB row, U 512-byte pieces in flight per warp (k_bid's access pattern)
//   C  cp.async.bulk ring: one thread streams 1-D bulk copies into shared memory, the CTA sums them
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#define CK(x)                                                                                                          \
    do                                                                                                                 \
    {                                                                                                                  \
        cudaError_t e = (x);                                                                                           \
        if (e != cudaSuccess)                                                                                          \
        {                                                                                                              \
            printf("%s: %s\n", #x, cudaGetErrorString(e));                                                             \
            return 1;                                                                                                  \
        }                                                                                                              \
    } while (0)

__device__ __forceinline__ float4 ldnc(const float4* p)
{
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ float sum4(float4 v) { return (v.x + v.y) + (v.z + v.w); }

template <int U> __global__ void k_tiles(const float4* __restrict__ x, long long n4, float* out)
{
    const long long tile = (long long)blockDim.x * U;
    float acc = 0.f;
    for (long long base = (long long)blockIdx.x * tile; base < n4; base += (long long)gridDim.x * tile)
    {
        float4 v[U];
#pragma unroll
        for (int u = 0; u < U; u++)
        {
            const long long i = base + (long long)u * blockDim.x + threadIdx.x;
            v[u] = i < n4 ? ldnc(x + i) : make_float4(0, 0, 0, 0);
        }
#pragma unroll
        for (int u = 0; u < U; u++)
            acc += sum4(v[u]);
    }
    if (acc == 123.456f)
        *out = acc;
}

// warp per row of row4 float4s; rows dealt round-robin to warps
template <int U> __global__ void k_rows(const float4* __restrict__ x, int rows, int row4, float* out)
{
    const int lane = threadIdx.x & 31;
    const int warp = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int nwarps = gridDim.x * (blockDim.x >> 5);
    float acc = 0.f;
    for (int r = warp; r < rows; r += nwarps)
    {
        const float4* row = x + (long long)r * row4;
        for (int base = 0; base < row4; base += 32 * U)
        {
            float4 v[U];
#pragma unroll
            for (int u = 0; u < U; u++)
                v[u] = ldnc(row + base + u * 32 + lane);
#pragma unroll
            for (int u = 0; u < U; u++)
                acc += sum4(v[u]);
        }
    }
    if (acc == 123.456f)
        *out = acc;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* b)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity)
{
    uint32_t ok;
    do
    {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                     : "=r"(ok)
                     : "r"(smem_u32(b)), "r"(parity)
                     : "memory");
    } while (!ok);
}
__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* b)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src),
                 "r"(bytes), "r"(smem_u32(b))
                 : "memory");
}

// kStage bytes per stage, kN stages; chunks dealt round-robin to CTAs; consumers: all threads
template <int kStage, int kN> __global__ void k_bulk(const char* __restrict__ x, long long bytes, float* out)
{
    extern __shared__ __align__(128) char smem[];
    __shared__ uint64_t full[kN], empty[kN];
    const int tid = threadIdx.x;
    if (tid == 0)
    {
        for (int s = 0; s < kN; s++)
        {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], blockDim.x >> 5);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const long long chunks = bytes / kStage;
    const long long mine = (chunks - blockIdx.x + gridDim.x - 1) / gridDim.x;
    float acc = 0.f;
    if (tid == 0)
        for (int k = 0; k < kN && k < mine; k++)
        {
            mbar_expect_tx(&full[k], kStage);
            bulk_load(smem + (size_t)k * kStage, x + ((long long)blockIdx.x + (long long)k * gridDim.x) * kStage, kStage, &full[k]);
        }
    for (long long k = 0; k < mine; k++)
    {
        const int s = (int)(k % kN);
        const uint32_t ph = (uint32_t)((k / kN) & 1);
        mbar_wait(&full[s], ph);
        const float4* p = reinterpret_cast<const float4*>(smem + (size_t)s * kStage);
#pragma unroll 4
        for (int i = tid; i < kStage / 16; i += blockDim.x)
            acc += sum4(p[i]);
        __syncwarp();
        if ((tid & 31) == 0)
            mbar_arrive(&empty[s]);
        if (tid == 0 && k + kN < mine)
        {
            mbar_wait(&empty[s], ph);
            mbar_expect_tx(&full[s], kStage);
            bulk_load(smem + (size_t)s * kStage, x + ((long long)blockIdx.x + (k + kN) * gridDim.x) * kStage, kStage, &full[s]);
        }
    }
    if (acc == 123.456f)
        *out = acc;
}

template <typename F> float time_it(F f, int reps = 10)
{
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    float best = 1e9f, sum = 0.f;
    for (int r = 0; r < reps + 3; r++)
    {
        cudaEventRecord(a);
        f();
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms;
        cudaEventElapsedTime(&ms, a, b);
        if (r >= 3)
        {
            best = ms < best ? ms : best;
            sum += ms;
        }
    }
    printf("min %.4f ms %.0f GB/s | mean %.4f ms %.0f GB/s\n", best, 1073.741824f / best, sum / reps, 1073.741824f / (sum / reps));
    return best;
}

int main()
{
    const long long bytes = 1ll << 30;
    float *x, *out;
    CK(cudaMalloc(&x, bytes));
    CK(cudaMalloc(&out, 4));
    CK(cudaMemset(x, 0, bytes));
    const long long n4 = bytes / 16;
    const float4* x4 = reinterpret_cast<const float4*>(x);
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    printf("SMs %d\n", sms);
    for (int threads : {256, 512, 1024})
        for (int per_sm : {1, 2, 4, 8})
        {
            if (threads * per_sm > 2048)
                continue;
            const int grid = sms * per_sm;
            printf("A tiles threads %4d ctas/SM %d U4: ", threads, per_sm);
            time_it([&] { k_tiles<4><<<grid, threads>>>(x4, n4, out); });
            printf("A tiles threads %4d ctas/SM %d U8: ", threads, per_sm);
            time_it([&] { k_tiles<8><<<grid, threads>>>(x4, n4, out); });
        }
    printf("A tiles threads 256 grid = n/tile (one tile per CTA) U4: ");
    time_it([&] { k_tiles<4><<<(int)(n4 / (256 * 4)), 256>>>(x4, n4, out); });
    printf("A tiles threads 256 grid = n/tile (one tile per CTA) U8: ");
    time_it([&] { k_tiles<8><<<(int)(n4 / (256 * 8)), 256>>>(x4, n4, out); });
    for (int threads : {512, 1024})
    {
        const int grid = sms * (1024 / threads);
        printf("B rows 16384 x 64 KB, warp per row, %d threads, U2: ", threads);
        time_it([&] { k_rows<2><<<grid, threads>>>(x4, 16384, 4096, out); });
        printf("B rows 16384 x 64 KB, warp per row, %d threads, U4: ", threads);
        time_it([&] { k_rows<4><<<grid, threads>>>(x4, 16384, 4096, out); });
        printf("B rows 16384 x 64 KB, warp per row, %d threads, U8: ", threads);
        time_it([&] { k_rows<8><<<grid, threads>>>(x4, 16384, 4096, out); });
    }
    printf("B rows 32768 x 32 KB (balanced waves), 1024 threads, U4: ");
    time_it([&] { k_rows<4><<<sms, 1024>>>(x4, 32768, 2048, out); });
    {
        CK(cudaFuncSetAttribute(k_bulk<16384, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384 * 8));
        CK(cudaFuncSetAttribute(k_bulk<32768, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768 * 6));
        CK(cudaFuncSetAttribute(k_bulk<8192, 12>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8192 * 12));
        CK(cudaFuncSetAttribute(k_bulk<16384, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384 * 4));
        for (int threads : {256, 512})
        {
            printf("C bulk 16 KB x 8 stages, 1 CTA/SM, %d threads: ", threads);
            time_it([&] { k_bulk<16384, 8><<<sms, threads, 16384 * 8>>>((const char*)x, bytes, out); });
            printf("C bulk 32 KB x 6 stages, 1 CTA/SM, %d threads: ", threads);
            time_it([&] { k_bulk<32768, 6><<<sms, threads, 32768 * 6>>>((const char*)x, bytes, out); });
            printf("C bulk 8 KB x 12 stages, 2 CTA/SM, %d threads: ", threads);
            time_it([&] { k_bulk<8192, 12><<<sms * 2, threads, 8192 * 12>>>((const char*)x, bytes, out); });
            printf("C bulk 16 KB x 4 stages, 3 CTA/SM, %d threads: ", threads);
            time_it([&] { k_bulk<16384, 4><<<sms * 3, threads, 16384 * 4>>>((const char*)x, bytes, out); });
        }
    }
    CK(cudaDeviceSynchronize());
    CK(cudaGetLastError());
    // reference: device-to-device copy of the same buffer (read + write bytes)
    float* y;
    CK(cudaMalloc(&y, bytes));
    printf("cudaMemcpy D2D 1 GiB (read+write = 2 GiB): ");
    float ms = time_it([&] { cudaMemcpyAsync(y, x, bytes, cudaMemcpyDeviceToDevice); });
    printf("   -> %.0f GB/s read+write\n", 2 * 1073.741824f / ms);
    return 0;
}
