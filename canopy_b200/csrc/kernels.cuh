// Device kernels of the Monte Carlo hot path, hand-written for sm_100a.
//
//   sampler_sliced_kernel / sampler_kernel   north-star subsystem 2 (ABSENT in the reference):
//                       counter-based Philox4x32-10 variates, compared against the per-event threshold (MSB-first on
//                       bit planes in spec 2, whole in spec 1; spec 3 draws the count and positions of the ones of a
//                       128-sample group for rare events), 64 samples per word, one 128-bit coalesced store per lane.
//   gate_ring_kernel    north-star subsystem 3 (ABSENT in the reference):
//                       one CTA walks every layer of the DAG for one slab; child rows stream through per-warp
//                       shared-memory rings filled by TMA bulk copies, bit-sliced adders for k-of-n, fused
//                       popcount tallies reduced with warp REDUX and one 64-bit RED per (gate, slab).
//   finalize_kernel     maps device gate tallies to the caller-visible outputs (roots, per-gate) in place
//                       in the buffer that the NCCL all-reduce consumes.
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

// Output word 0 only / words 0 and 1 only (the tail variate of SPEC v2): the last round's M0 product is dead.
__device__ __forceinline__ uint2 philox4x32_10_xy(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                  const PhiloxKeys& K) {
#pragma unroll
    for (int r = 0; r < 9; ++r) philox_round(c0, c1, c2, c3, K.k0[r], K.k1[r]);
    const uint64_t p1 = (uint64_t)kPhiloxM1 * c2;
    return make_uint2((uint32_t)(p1 >> 32) ^ c1 ^ K.k0[9], (uint32_t)p1);
}
__device__ __forceinline__ uint32_t philox4x32_10_x(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                    const PhiloxKeys& K) {
    return philox4x32_10_xy(c0, c1, c2, c3, K).x;
}

// PTX shl.b32 clamps shift amounts above 31 (result 0); `<<` in C++ would be undefined there.
__device__ __forceinline__ uint32_t shl_clamp(uint32_t v, uint32_t n) {
    uint32_t r;
    asm("shl.b32 %0, %1, %2;" : "=r"(r) : "r"(v), "r"(n));
    return r;
}

__device__ __forceinline__ uint4 ld_cg(const uint4* p) { return __ldcg(p); }
__device__ __forceinline__ void st_cg(uint4* p, uint4 v) { __stcg(p, v); }

}  // namespace canopy
