// philox.cuh -- Philox4x32-10 (Salmon, Moraes, Dror, Shaw, SC'11), counter-based: the draw for
// (seed, aircraft id, episode, block) is a pure function, so resets need no generator state in HBM
// and are independent of how the fleet is sharded across GPUs.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace acb {

struct Philox {
    uint32_t key0, key1, c0, c1, c2;
    __host__ __device__ Philox(uint64_t seed, uint64_t aircraft_id, uint32_t episode)
        : key0((uint32_t)seed), key1((uint32_t)(seed >> 32)), c0((uint32_t)aircraft_id), c1((uint32_t)(aircraft_id >> 32)),
          c2(episode) {}
    __host__ __device__ __forceinline__ void block(uint32_t b, uint32_t (&out)[4]) const {
        uint32_t x0 = c0, x1 = c1, x2 = c2, x3 = b, k0 = key0, k1 = key1;
#pragma unroll
        for (int r = 0; r < 10; ++r) {
#ifdef __CUDA_ARCH__
            const uint32_t hi0 = __umulhi(0xD2511F53u, x0), lo0 = 0xD2511F53u * x0;
            const uint32_t hi1 = __umulhi(0xCD9E8D57u, x2), lo1 = 0xCD9E8D57u * x2;
#else
            const uint64_t p0 = (uint64_t)0xD2511F53u * x0, p1 = (uint64_t)0xCD9E8D57u * x2;
            const uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0, hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
#endif
            const uint32_t n0 = hi1 ^ x1 ^ k0, n2 = hi0 ^ x3 ^ k1;
            x0 = n0; x1 = lo1; x2 = n2; x3 = lo0;
            k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
        }
        out[0] = x0; out[1] = x1; out[2] = x2; out[3] = x3;
    }
};

// 53-bit uniform in [0,1): the construction numpy's random_sample uses on two 32-bit words
__host__ __device__ __forceinline__ double u53(uint32_t a, uint32_t b) {
    return ((double)(a >> 5) * 67108864.0 + (double)(b >> 6)) / 9007199254740992.0;
}

// RandomState.uniform(lo, hi) = lo + (hi - lo) * u, each op individually rounded in every build
__device__ __forceinline__ double uniform_lh(double lo, double hi, double u) {
    return __dadd_rn(lo, __dmul_rn(__dsub_rn(hi, lo), u));
}

}  // namespace acb
