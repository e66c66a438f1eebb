/*
 * b200lsap_gen.h — counter-based synthetic benefit generator shared by host, device and tests.
 *
 * value(seed, problem, i, j) = (k + 1) * 2^-24,  k = top 24 bits of
 *     splitmix64( splitmix64(seed ^ problem * GOLD2) + (i * cols + j) )
 * so every coefficient lies on the 2^-24 grid in (0, 1]: never the 0 "no edge" sentinel of
 * the reference (Auction.h:134) and exactly representable in fp32.  The same (seed, problem,
 * index) gives the same bits on the CPU, in numpy (tests/gen_np.py) and in the CUDA generator
 * kernel, so a 25 GB matrix never has to cross PCIe to be checked.
 *
 * The reference's own driver recipe (LSAP.cpp:26-33: Random()+1 divided by the max) is
 * std::rand based and not reproducible across machines; it is not used for parity.
 */
#ifndef B200LSAP_GEN_H_
#define B200LSAP_GEN_H_

#include <stdint.h>

#if defined(__CUDACC__)
#define B200LSAP_HD __host__ __device__ __forceinline__
#else
#define B200LSAP_HD static inline
#endif

B200LSAP_HD uint64_t b200lsap_splitmix64(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

B200LSAP_HD uint64_t b200lsap_gen_key(uint64_t seed, uint64_t problem)
{
    return b200lsap_splitmix64(seed ^ (problem * 0xD1B54A32D192ED03ull));
}

/* k in [0, 2^24) */
B200LSAP_HD uint32_t b200lsap_gen_k(uint64_t key, uint64_t index)
{
    return (uint32_t)(b200lsap_splitmix64(key + index) >> 40);
}

B200LSAP_HD float b200lsap_gen_value(uint64_t key, uint64_t index)
{
    return (float)(b200lsap_gen_k(key, index) + 1u) * (1.0f / 16777216.0f);
}

#endif /* B200LSAP_GEN_H_ */
