// Spec 4 ("classed sparse"): class of an event and its cumulative count table.
// One definition for host (lowering, parity hook) and device (the table kernel): exact integer arithmetic, so every
// implementation of the spec holds the same 64-bit words.  The oracle and the Python restatement under tests/ are
// written separately from the spec text on purpose.
#pragma once
#include <cstdint>

#include "canopy.h"

#if defined(__CUDACC__)
#define CANOPY_HD __host__ __device__ __forceinline__
#else
#define CANOPY_HD inline
#endif

namespace canopy {

// Leading zero bits of a 64-bit threshold (thr64 != 0).
CANOPY_HD uint32_t clz64(uint64_t x) {
#if defined(__CUDA_ARCH__)
    return (uint32_t)__clzll((long long)x);
#else
    return (uint32_t)__builtin_clzll(x);
#endif
}

// Group size exponent k of an event with 64-bit threshold t (p = t / 2^64), or 0 for a dense event (spec 2 planes).
// Sparse iff p < 2^-Z; then k = min(KMAX, z + 2) with z the leading zeros of t, so that a group of 2^k samples holds
// fewer than four ones on average.
CANOPY_HD uint32_t sparse_class_k(uint64_t t) {
    if (t == 0) return 0;
    const uint32_t z = clz64(t);
    if (z < CANOPY_SPARSE_MIN_ZEROS) return 0;
    const uint32_t k = z + 2u;
    return k < CANOPY_SPARSE_MAX_LOG2 ? k : (uint32_t)CANOPY_SPARSE_MAX_LOG2;
}

// Cumulative table of Binomial(2^k, t / 2^64) in 0.64 fixed point, CANOPY_SPARSE_CDF_LEN entries:
//   q = 2^64 - t,  r = floor(t 2^64 / q)                       (the odds p / (1 - p), 0.64 fixed point)
//   a_0 = q^(2^k) by k truncating squarings                     ((1 - p)^(2^k))
//   a_{j+1} = floor( floor(a_j r / 2^51) (2^k - j) / (j + 1) ) >> 13
//   cdf_j = min(2^64 - 1, a_0 + ... + a_j)
CANOPY_HD void sparse_cdf(uint64_t t, uint32_t k, uint64_t* cdf) {
    typedef unsigned __int128 u128;
    const uint64_t q = 0ull - t;                               // 2^64 - t, t >= 1
    const uint64_t r = (uint64_t)((((u128)t) << 64) / q);      // t < 2^59: the quotient fits 64 bits
    uint64_t a = q;
    for (uint32_t i = 0; i < k; ++i) a = (uint64_t)(((u128)a * a) >> 64);
    u128 sum = 0;
    const uint64_t n = 1ull << k;
    for (uint32_t j = 0; j < CANOPY_SPARSE_CDF_LEN; ++j) {
        sum += a;
        cdf[j] = (sum >> 64) ? ~0ull : (uint64_t)sum;
        const u128 y = ((u128)a * r) >> 51;
        a = (uint64_t)(((y * (n - j)) / (j + 1u)) >> 13);
    }
}

// Position fields: k bits each, floor(32 / k) per 32-bit Philox output word.
CANOPY_HD uint32_t sparse_fields_per_word(uint32_t k) { return 32u / k; }

}  // namespace canopy
