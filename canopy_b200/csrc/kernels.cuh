// Shared device helpers of the Monte Carlo hot path (sm_100a): Philox4x32-10 and cache-hinted 128-bit accesses.
// The kernels live in sampler.cuh (north-star subsystem 2), walk.cuh (subsystem 3) and small.cuh (both, fused, for small
// models); all three are ABSENT in the reference.
//
// Value buffer layout (HBM):  vals[slab][node][half][lane] as uint4 -- node row = kRowBytes = 1 KB = 128 words x 64 samples
// (plan.h).  Sample s of the run lives in slab s/8192, word (s%8192)/64, bit s%64.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "plan.h"

namespace canopy {

// ----------------------------------------------------------------------------------------------
// Philox4x32-10.  DESIGN.md SPEC v1 "RNG"; constants as cuRAND curand_philox4x32_x.h:88-91.
// One round = 2 IMAD.WIDE.U32 + 2 LOP3; the key schedule is uniform across the grid.
// ----------------------------------------------------------------------------------------------
constexpr uint32_t kPhiloxM0 = 0xD2511F53u, kPhiloxM1 = 0xCD9E8D57u;
constexpr uint32_t kPhiloxW0 = 0x9E3779B9u, kPhiloxW1 = 0xBB67AE85u;

struct PhiloxKeys { uint32_t k0[10], k1[10]; };  // precomputed round keys (host side, passed by value)

__device__ __forceinline__ void philox_round(uint32_t& c0, uint32_t& c1, uint32_t& c2, uint32_t& c3,
                                             uint32_t k0, uint32_t k1) {
    const uint64_t p0 = (uint64_t)kPhiloxM0 * c0;
    const uint64_t p1 = (uint64_t)kPhiloxM1 * c2;
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    c1 = (uint32_t)p1; c3 = (uint32_t)p0; c0 = n0; c2 = n2;
}

__device__ __forceinline__ uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                               const PhiloxKeys& K) {
#pragma unroll
    for (int r = 0; r < 10; ++r) philox_round(c0, c1, c2, c3, K.k0[r], K.k1[r]);
    return make_uint4(c0, c1, c2, c3);
}

__device__ __forceinline__ uint4 ld_cg(const uint4* p) { return __ldcg(p); }
__device__ __forceinline__ void st_cg(uint4* p, uint4 v) { __stcg(p, v); }

}  // namespace canopy
