// Monte Carlo engine: kernels + per-device state + the batch loop.  sm_100a only, no CPU fallback.
// See kernels.cuh for the kernel inventory and DESIGN.md for layout, spec and roofline accounting.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include <cuda.h>  // CUtensorMap (types only; the encoder is fetched through cudaGetDriverEntryPoint)

#include "kernels.cuh"

namespace canopy {

#define CUDA_CHECK(expr)                                                                              \
    do {                                                                                              \
        cudaError_t _e = (expr);                                                                      \
        if (_e != cudaSuccess)                                                                        \
            throw Error(CANOPY_E_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e));           \
    } while (0)

// ================================================================================================
// Sampler kernel.  One warp per (basic event, slab); lane r owns words 2r and 2r+1 of the row, i.e.
// 128 consecutive samples = 32 Philox calls (SPEC v1: call index q = s >> 2, lane of the call = s & 3).
// All 32 calls are independent, so ILP comes for free; the only memory traffic is one LDG of the
// threshold per warp and one 16-byte store per lane (512 B per warp, fully coalesced).
// ================================================================================================
// bits = 2*bits + (u < thr) in two instructions: `sub.cc (thr-1) - u` leaves the hardware carry set exactly
// when u <= thr-1 (the subtraction is thr-1 + ~u + 1, carry = no borrow), and addc shifts it in.
// thr >= 1 always holds: events whose threshold quantises to 0 are folded to constants by the lowering.
__device__ __forceinline__ uint32_t shift_in_lt(uint32_t bits, uint32_t u, uint32_t thr_m1) {
    asm("{ .reg .u32 t; sub.cc.u32 t, %2, %1; addc.u32 %0, %0, %0; }" : "+r"(bits) : "r"(u), "r"(thr_m1));
    return bits;
}

template <int BITS, int PACK>
__global__ void __launch_bounds__(256)
sampler_kernel(uint4* __restrict__ vals, const uint32_t* __restrict__ ev_orig,
               const uint32_t* __restrict__ thr32, const uint64_t* __restrict__ thr64,
               uint32_t n_ev, uint32_t n_nodes, uint64_t first_slab, PhiloxKeys K) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t ev = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (ev >= n_ev) return;
    const uint32_t blk = blockIdx.y;  // 4096-sample block of the batch: block h of slab blk / kRowHalves
    const uint32_t e = __ldg(ev_orig + ev);
    const uint64_t w0 = (first_slab * kRowHalves + blk) * 64u + 2 * lane;  // first global word of this lane
    uint32_t out[4];
    if (BITS == 32) {
        const uint32_t thr = __ldg(thr32 + ev);
        const uint64_t q = w0 * 16;  // 16 calls per 64-sample word; q % 32 == 0 so q_hi is fixed below
        const uint32_t q_lo = (uint32_t)q, q_hi = (uint32_t)(q >> 32);
#pragma unroll
        for (int h = 0; h < 4; ++h) {
            // bits = 2*bits + (u < thr): the borrow of u - thr shifted in through the carry chain, two
            // ALU instructions per sample. Samples are visited last-to-first so sample i lands on bit i.
            uint32_t bits = 0;
#pragma unroll
            for (int c = 7; c >= 0; --c) {
                const uint4 r = philox4x32_10(q_lo + h * 8 + c, q_hi, e, 0u, K);
                if (PACK == 0) {
                    bits = (bits << 4) | (uint32_t)(r.x < thr) | ((uint32_t)(r.y < thr) << 1) |
                           ((uint32_t)(r.z < thr) << 2) | ((uint32_t)(r.w < thr) << 3);
                } else {
                    bits = shift_in_lt(bits, r.w, thr - 1); bits = shift_in_lt(bits, r.z, thr - 1);
                    bits = shift_in_lt(bits, r.y, thr - 1); bits = shift_in_lt(bits, r.x, thr - 1);
                }
            }
            out[h] = bits;
        }
    } else {
        const uint64_t thr = __ldg(thr64 + ev);
        const uint64_t q = w0 * 32;  // 32 calls per word, 2 samples per call
        const uint32_t q_lo = (uint32_t)q, q_hi = (uint32_t)(q >> 32);
#pragma unroll 1
        for (int h = 0; h < 4; ++h) {
            uint32_t bits = 0;
#pragma unroll
            for (int c = 0; c < 16; ++c) {
                const uint4 r = philox4x32_10(q_lo + h * 16 + c, q_hi, e, 1u, K);
                const uint64_t u0 = ((uint64_t)r.y << 32) | r.x, u1 = ((uint64_t)r.w << 32) | r.z;
                bits |= (uint32_t)(u0 < thr) << (2 * c);
                bits |= (uint32_t)(u1 < thr) << (2 * c + 1);
            }
            out[h] = bits;
        }
    }
    uint4* row = vals + ((size_t)(blk / kRowHalves) * n_nodes + ev) * kRowVec + (blk % kRowHalves) * 32u;
    st_cg(row + lane, make_uint4(out[0], out[1], out[2], out[3]));
}

// ================================================================================================
// Sliced sampler (SPEC v2).  Same geometry as above -- one warp per (event, slab), lane r owns the 128
// samples of group g = slab*32 + r -- but the comparison u < thr runs MSB-first on bit planes:
//   head: H = CANOPY_SLICED_HEAD_BITS Philox calls ctr = (g, e, 0x100+k); word j of call k is bit (H-1-k)
//         of the variates' top H bits for samples 32j..32j+31 of the group.  Per plane and word: a lane
//         still undecided (all higher bits equal to the threshold's) becomes "less" where thr has 1 and u
//         has 0, "greater" where thr has 0 and u has 1.  The threshold is warp-uniform, so the per-plane
//         branch is uniform.
//   tail: the samples of the group still undecided after H planes (probability 2^-H each) take their low
//         32-H (64-H) bits, in ascending sample order, from the output words of the group's tail calls
//         ctr = (g, e, 0x200+n) (resp. 0x300+n): four (two) undecided samples per call, so one tail call
//         per lane covers all but 3.5e-3 of the groups (H = 7) and is issued without divergence.
// H + 1 Philox calls per 128 samples instead of 32 (64 for 64-bit thresholds).
// ================================================================================================
constexpr int kHeadBits = CANOPY_SLICED_HEAD_BITS;

#ifndef CANOPY_SLICED_MIN_BLOCKS
#define CANOPY_SLICED_MIN_BLOCKS 6
#endif
__device__ __forceinline__ uint32_t bcast0(uint32_t v) { return __shfl_sync(0xFFFFFFFFu, v, 0); }  // provably warp-uniform

// Variate queue of the tail: the output words of one tail call, consumed in order (x, y, z, w).
template <int BITS> struct TailQueue;
template <> struct TailQueue<32> {
    static constexpr uint32_t kPerCall = 4, kCtr = 0x200u;
    uint32_t q0, q1, q2, q3;
    __device__ __forceinline__ void fill(uint4 t) { q0 = t.x; q1 = t.y; q2 = t.z; q3 = t.w; }
    __device__ __forceinline__ bool pop_lt(uint32_t tail32, uint64_t) { const bool lt = q0 < tail32; q0 = q1; q1 = q2; q2 = q3; return lt; }
};
template <> struct TailQueue<64> {
    static constexpr uint32_t kPerCall = 2, kCtr = 0x300u;
    uint64_t q0, q1;
    __device__ __forceinline__ void fill(uint4 t) { q0 = ((uint64_t)t.y << 32) | t.x; q1 = ((uint64_t)t.w << 32) | t.z; }
    __device__ __forceinline__ bool pop_lt(uint32_t, uint64_t tail64) { const bool lt = q0 < tail64; q0 = q1; return lt; }
};

#ifndef CANOPY_SLICED_SLABS_PER_WARP
#define CANOPY_SLICED_SLABS_PER_WARP 8
#endif
// Spec 3, sparse events: per 128-sample group ONE Philox call, ctr = (g, e, 0x400).  Words 0-1
// are a 64-bit variate U; the number of ones of the group is N = #{j : U >= cdf_j} (Binomial(128, p) by inversion);
// 7-bit fields of words 2-3 (then of calls 0x401, 0x402, ...: 16 fields each) are the positions, a field naming a
// position already taken is skipped.  One lane = one group, as in the plane path; the count loop is warp-uniform
// (table entries are broadcast loads), the placement loop runs max-N-of-the-warp times.
__device__ __forceinline__ void sparse_rows(uint4* __restrict__ vals, const uint64_t* __restrict__ tab, uint32_t ev, uint32_t e,
                                            uint32_t n_nodes, uint64_t first_slab, uint32_t n_slabs, const PhiloxKeys& K) {
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t c0 = __ldg(tab), c1 = __ldg(tab + 1);
    const uint32_t n_blk = n_slabs * kRowHalves;
    for (uint32_t blk = blockIdx.y; blk < n_blk; blk += gridDim.y) {
        uint4* row = vals + ((size_t)(blk / kRowHalves) * n_nodes + ev) * kRowVec + (blk % kRowHalves) * 32u + lane;
        const uint64_t g = (first_slab * kRowHalves + blk) * 32u + lane;
        const uint32_t g_lo = (uint32_t)g, g_hi = (uint32_t)(g >> 32);
        const uint4 x = philox4x32_10(g_lo, g_hi, e, 0x400u, K);
        const uint64_t U = ((uint64_t)x.y << 32) | x.x;
        uint32_t n = (U >= c0 ? 1u : 0u) + (U >= c1 ? 1u : 0u);
#ifdef CANOPY_SPARSE_BSEARCH  // tuning experiment, not measured yet: per-lane binary search over entries 1..31 for
                              // events where a second one is common (cdf_1 < 3/4), instead of the warp-uniform linear walk
        if (c1 < 0xC000000000000000ull) {
            uint32_t m = 0;   // largest index in 1..31 with cdf[m] <= U (0: none) = number of such entries, the table being sorted
#pragma unroll
            for (uint32_t step = 16; step; step >>= 1)
                if (U >= __ldg(tab + m + step)) m += step;
            n = (U >= c0 ? 1u : 0u) + m;
        } else
#endif
        if (__any_sync(0xFFFFFFFFu, U >= c1)) {
#pragma unroll 1
            for (uint32_t j = 2; j < CANOPY_SPARSE_CDF_LEN; ++j) {
                const bool ge = U >= __ldg(tab + j);
                if (!__any_sync(0xFFFFFFFFu, ge)) break;
                n += ge ? 1u : 0u;
            }
        }
        uint32_t r0 = 0, r1 = 0, r2 = 0, r3 = 0;
        // f: the unread fields of the current field word (next field in its low 7 bits), avail of them left; field word fw
        // is word 2 + fw of call 0x400 for fw < 2, else word (fw - 2) & 3 of call 0x401 + ((fw - 2) >> 2)
        uint32_t f = x.z, avail = 4, fw = 0;
        uint4 ext = x;
#pragma unroll 1
        while (n) {
            if (avail == 0) {
                avail = 4; ++fw;
                if (fw == 1u) f = x.w;
                else {
                    const uint32_t k = fw - 2u;
                    if ((k & 3u) == 0u) ext = philox4x32_10(g_lo, g_hi, e, 0x401u + (k >> 2), K);
                    f = (k & 3u) == 0u ? ext.x : (k & 3u) == 1u ? ext.y : (k & 3u) == 2u ? ext.z : ext.w;
                }
            }
            const uint32_t pos = f & 127u;
            f >>= 7; --avail;
            // branch-free test-and-set: shl clamps shift counts outside 0..31 to a zero result, so exactly one of the four
            // masks is non-zero; OR-ing a bit that is already set changes nothing, so only the count depends on `hit`
            const uint32_t b0 = shl_clamp(1u, pos), b1 = shl_clamp(1u, pos - 32u), b2 = shl_clamp(1u, pos - 64u), b3 = shl_clamp(1u, pos - 96u);
            const uint32_t hit = (r0 & b0) | (r1 & b1) | (r2 & b2) | (r3 & b3);
            r0 |= b0; r1 |= b1; r2 |= b2; r3 |= b3;
            n -= hit == 0u ? 1u : 0u;
        }
        st_cg(row, make_uint4(r0, r1, r2, r3));
    }
}

// SPARSE (spec 3): the warps take the events in the order ev_perm (sliced events first); an event with fewer than
// one expected 1 per 128-sample group (thr < 2^(32 - CANOPY_SPARSE_MIN_ZEROS)) takes sparse_rows() instead of the planes.
template <int BITS, bool SPARSE>
__global__ void __launch_bounds__(256, CANOPY_SLICED_MIN_BLOCKS)
sampler_sliced_kernel(uint4* __restrict__ vals, const uint32_t* __restrict__ ev_orig,
                      const uint32_t* __restrict__ thr32, const uint64_t* __restrict__ thr64,
                      uint32_t n_ev, uint32_t n_nodes, uint64_t first_slab, uint32_t n_slabs, PhiloxKeys Kparam,
                      const uint32_t* __restrict__ ev_perm, const uint64_t* __restrict__ cdf) {
#ifdef CANOPY_KEYS_IN_REGS  // tuning knob: round keys in vector registers instead of uniform registers
    PhiloxKeys K;
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        asm volatile("mov.u32 %0, %1;" : "=r"(K.k0[i]) : "r"(Kparam.k0[i]));
        asm volatile("mov.u32 %0, %1;" : "=r"(K.k1[i]) : "r"(Kparam.k1[i]));
    }
#else
    const PhiloxKeys& K = Kparam;
#endif
    const uint32_t lane = threadIdx.x & 31;
    uint32_t ev = blockIdx.x * (blockDim.x >> 5) + bcast0(threadIdx.x >> 5);
    if (ev >= n_ev) return;
    if (SPARSE) ev = bcast0(__ldg(ev_perm + ev));
    // Everything that depends on the event only is loaded once per warp and broadcast, so the compiler keeps
    // it in uniform registers and the per-plane branch on the threshold bit is a uniform branch.
    const uint32_t e = bcast0(__ldg(ev_orig + ev));
    if (SPARSE && BITS == 32) {
        const uint32_t t = bcast0(__ldg(thr32 + ev));
        if ((t >> (32 - CANOPY_SPARSE_MIN_ZEROS)) == 0) {
            sparse_rows(vals, cdf + (size_t)ev * CANOPY_SPARSE_CDF_LEN, ev, e, n_nodes, first_slab, n_slabs, K);
            return;
        }
    }
    uint32_t head;                 // top H bits of the threshold, left-aligned in a 32-bit word
    uint32_t tail32 = 0; uint64_t tail64 = 0;  // (thr mod 2^(BITS-H)) << H: compared with the raw tail variate
    if (BITS == 32) { const uint32_t t = bcast0(__ldg(thr32 + ev)); head = t; tail32 = t << kHeadBits; }
    else {
        const uint64_t t = __ldg(thr64 + ev);
        const uint32_t lo = bcast0((uint32_t)t), hi = bcast0((uint32_t)(t >> 32));
        head = hi; tail64 = (((uint64_t)hi << 32) | lo) << kHeadBits;
    }
    const uint32_t trans = head ^ (head >> 1);  // bit k (MSB first) set iff threshold bits k and k-1 differ
    // the warp walks 4096-sample blocks blockIdx.y, +gridDim.y, ... of the batch: block h of slab blk / kRowHalves
    const uint32_t n_blk = n_slabs * kRowHalves;
    for (uint32_t blk = blockIdx.y; blk < n_blk; blk += gridDim.y) {
        uint4* row = vals + ((size_t)(blk / kRowHalves) * n_nodes + ev) * kRowVec + (blk % kRowHalves) * 32u + lane;
        const uint64_t g = (first_slab * kRowHalves + blk) * 32u + lane;  // 128-sample group of this lane
        const uint32_t g_lo = (uint32_t)g, g_hi = (uint32_t)(g >> 32);
        // Branch-free head, two LOP3 per word and plane.  With m_k = all-ones iff bit k (MSB first) of the
        // threshold is set:   u_{k+1} = u_k & ~(x_k ^ m_k)   (stay undecided where the variate bit equals the
        // threshold bit), and a sample decided at plane k is "less" iff m_k.  The decided sets u_k ^ u_{k+1}
        // are disjoint, so   r = XOR_k m_k & (u_k ^ u_{k+1}) = XOR_k (m_k ^ m_{k-1}) & u_k  ^  m_{H-1} & u_H
        // (m_{-1} = 0): one LOP3 on (r, u_k, c_k) with the warp-uniform transition mask c_k.
        uint32_t u[4] = {~0u, ~0u, ~0u, ~0u};  // undecided samples
        uint32_t r[4] = {0, 0, 0, 0};          // decided "u < thr"
#pragma unroll
        for (int k = 0; k < kHeadBits; ++k) {
            const uint4 x = philox4x32_10(g_lo, g_hi, e, 0x100u + k, K);
            const uint32_t m = (uint32_t)((int32_t)(head << k) >> 31);
            const uint32_t c = (uint32_t)((int32_t)(trans << k) >> 31);
            r[0] ^= c & u[0]; r[1] ^= c & u[1]; r[2] ^= c & u[2]; r[3] ^= c & u[3];
            u[0] &= ~(x.x ^ m); u[1] &= ~(x.y ^ m); u[2] &= ~(x.z ^ m); u[3] &= ~(x.w ^ m);
        }
        {   // closing term m_{H-1} & u_H (r is then exactly the decided-less set, disjoint from u)
            const uint32_t m = (uint32_t)((int32_t)(head << (kHeadBits - 1)) >> 31);
            r[0] ^= m & u[0]; r[1] ^= m & u[1]; r[2] ^= m & u[2]; r[3] ^= m & u[3];
        }
        if (__any_sync(0xFFFFFFFFu, (u[0] | u[1] | u[2] | u[3]) != 0u)) {
            TailQueue<BITS> Q;
            Q.fill(philox4x32_10(g_lo, g_hi, e, TailQueue<BITS>::kCtr, K));  // tail call 0, whole warp
            uint32_t left = TailQueue<BITS>::kPerCall, call = 0;
#pragma unroll
            for (int w = 0; w < 4; ++w) {
#pragma unroll 1
                while (u[w]) {
                    const uint32_t b = u[w] & (0u - u[w]);  // lowest undecided sample of word w
                    u[w] ^= b;
                    if (left == 0) {  // more undecided samples than one call serves: rare (3.5e-3 per group)
                        Q.fill(philox4x32_10(g_lo, g_hi, e, TailQueue<BITS>::kCtr + (++call), K));
                        left = TailQueue<BITS>::kPerCall;
                    }
                    --left;
                    if (Q.pop_lt(tail32, tail64)) r[w] |= b;
                }
            }
        }
        st_cg(row, make_uint4(r[0], r[1], r[2], r[3]));
    }
}

// ================================================================================================
// Gate kernel (slab walker).
// ================================================================================================
__device__ __forceinline__ uint4 and4(uint4 a, uint4 b) { return make_uint4(a.x & b.x, a.y & b.y, a.z & b.z, a.w & b.w); }
__device__ __forceinline__ uint4 or4(uint4 a, uint4 b) { return make_uint4(a.x | b.x, a.y | b.y, a.z | b.z, a.w | b.w); }
__device__ __forceinline__ uint4 xor4(uint4 a, uint4 b) { return make_uint4(a.x ^ b.x, a.y ^ b.y, a.z ^ b.z, a.w ^ b.w); }
__device__ __forceinline__ uint4 xors4(uint4 a, uint32_t m) { return make_uint4(a.x ^ m, a.y ^ m, a.z ^ m, a.w ^ m); }
__device__ __forceinline__ uint4 andn4(uint4 a, uint4 b) { return make_uint4(a.x & ~b.x, a.y & ~b.y, a.z & ~b.z, a.w & ~b.w); }
__device__ __forceinline__ uint4 splat4(uint32_t v) { return make_uint4(v, v, v, v); }

// Load one child row segment (16 B for this lane), applying the edge complement (bit 31 of the code).
__device__ __forceinline__ uint4 load_child(const uint4* __restrict__ base, uint32_t code, uint32_t lane, bool ok,
                                            uint32_t identity) {
    if (!ok) return splat4(identity);
    const uint32_t row = code & 0x7FFFFFFFu;
    const uint32_t m = (uint32_t)((int32_t)code >> 31);
    return xors4(ld_cg(base + (size_t)row * kRowVec + lane), m);
}

#ifndef CANOPY_SIMPLE_UNROLL
#define CANOPY_SIMPLE_UNROLL 4
#endif
[[maybe_unused]] constexpr int kSimpleUnroll = CANOPY_SIMPLE_UNROLL;  // used by the 512-B-row A/B kernel only
template <int OP>  // kDevAnd / kDevOr / kDevXor
__device__ __forceinline__ uint4 eval_simple(const uint4* __restrict__ base, const uint32_t* __restrict__ child,
                                             uint32_t n, uint32_t lane, uint32_t first_ids) {
    constexpr uint32_t ident = OP == kDevAnd ? 0xFFFFFFFFu : 0u;
    uint4 acc = splat4(ident);
    for (uint32_t j0 = 0; j0 < n; j0 += 32) {
        const uint32_t cnt = min(32u, n - j0);
        const uint32_t my = j0 == 0 ? first_ids : (lane < cnt ? __ldg(child + j0 + lane) : 0u);
        for (uint32_t j = 0; j < cnt; j += kSimpleUnroll) {
            uint4 v[kSimpleUnroll];
#pragma unroll
            for (int u = 0; u < kSimpleUnroll; ++u) {
                const uint32_t code = __shfl_sync(0xFFFFFFFFu, my, (j + u) & 31);
                v[u] = load_child(base, code, lane, j + u < cnt, ident);
            }
#pragma unroll
            for (int u = 0; u < kSimpleUnroll; ++u)
                acc = OP == kDevAnd ? and4(acc, v[u]) : OP == kDevOr ? or4(acc, v[u]) : xor4(acc, v[u]);
        }
    }
    return acc;
}

// count-in-range with a bit-sliced ripple counter of P planes (P bits hold counts < 2^P).
// V is the per-lane vector: uint4 (one pass, 16 B per lane per child) for the common fan-ins, or
// uint32_t (four passes, one 32-bit component each) for fan-in >= 512 where 16 planes x 4 registers
// would not fit the register budget of the common path.
__device__ __forceinline__ uint32_t and4(uint32_t a, uint32_t b) { return a & b; }
__device__ __forceinline__ uint32_t or4(uint32_t a, uint32_t b) { return a | b; }
__device__ __forceinline__ uint32_t xor4(uint32_t a, uint32_t b) { return a ^ b; }
__device__ __forceinline__ uint32_t andn4(uint32_t a, uint32_t b) { return a & ~b; }
__device__ __forceinline__ uint2 and4(uint2 a, uint2 b) { return make_uint2(a.x & b.x, a.y & b.y); }
__device__ __forceinline__ uint2 or4(uint2 a, uint2 b) { return make_uint2(a.x | b.x, a.y | b.y); }
__device__ __forceinline__ uint2 xor4(uint2 a, uint2 b) { return make_uint2(a.x ^ b.x, a.y ^ b.y); }
__device__ __forceinline__ uint2 andn4(uint2 a, uint2 b) { return make_uint2(a.x & ~b.x, a.y & ~b.y); }
template <class V> __device__ __forceinline__ V splat(uint32_t v);
template <> __device__ __forceinline__ uint2 splat<uint2>(uint32_t v) { return make_uint2(v, v); }
template <> __device__ __forceinline__ uint4 splat<uint4>(uint32_t v) { return splat4(v); }
template <> __device__ __forceinline__ uint32_t splat<uint32_t>(uint32_t v) { return v; }

__device__ __forceinline__ void load_child_v(uint4& out, const uint4* __restrict__ base, uint32_t code, uint32_t lane, bool ok, int) {
    out = load_child(base, code, lane, ok, 0u);
}
__device__ __forceinline__ void load_child_v(uint32_t& out, const uint4* __restrict__ base, uint32_t code, uint32_t lane, bool ok, int comp) {
    if (!ok) { out = 0u; return; }
    const uint32_t row = code & 0x7FFFFFFFu, m = (uint32_t)((int32_t)code >> 31);
    out = __ldcg(reinterpret_cast<const uint32_t*>(base + (size_t)row * kRowVec + lane) + comp) ^ m;
}

__device__ __forceinline__ void load_child_v(uint2& out, const uint4* __restrict__ base, uint32_t code, uint32_t lane, bool ok, int comp) {
    if (!ok) { out = make_uint2(0u, 0u); return; }
    const uint32_t row = code & 0x7FFFFFFFu, m = (uint32_t)((int32_t)code >> 31);
    const uint2 v = __ldcg(reinterpret_cast<const uint2*>(base + (size_t)row * kRowVec + lane) + comp);
    out = make_uint2(v.x ^ m, v.y ^ m);
}

template <int P, class V>
__device__ __forceinline__ V eval_range(const uint4* __restrict__ base, const uint32_t* __restrict__ child,
                                        uint32_t n, uint32_t lo, uint32_t hi, uint32_t lane, int comp,
                                        uint32_t first_ids) {
    V pl[P];
#pragma unroll
    for (int p = 0; p < P; ++p) pl[p] = splat<V>(0u);
    for (uint32_t j0 = 0; j0 < n; j0 += 32) {
        const uint32_t cnt = min(32u, n - j0);
        const uint32_t my = j0 == 0 ? first_ids : (lane < cnt ? __ldg(child + j0 + lane) : 0u);
        constexpr int U = (P * (int)sizeof(V) > 64) ? 2 : 4;  // keep planes + loads inside the register budget
        for (uint32_t j = 0; j < cnt; j += U) {
            V v[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const uint32_t code = __shfl_sync(0xFFFFFFFFu, my, (j + u) & 31);
                load_child_v(v[u], base, code, lane, j + u < cnt, comp);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                V carry = v[u];
#pragma unroll
                for (int p = 0; p < P; ++p) {
                    const V t = and4(pl[p], carry);
                    pl[p] = xor4(pl[p], carry);
                    carry = t;
                }
            }
        }
    }
    // count >= lo  (MSB-first comparison against the uniform constant lo)
    V gt = splat<V>(0u), eq = splat<V>(0xFFFFFFFFu);
#pragma unroll
    for (int p = P - 1; p >= 0; --p) {
        if ((lo >> p) & 1u) eq = and4(eq, pl[p]);
        else { gt = or4(gt, and4(eq, pl[p])); eq = andn4(eq, pl[p]); }
    }
    V res = or4(gt, eq);
    if (hi < n) {  // count <= hi
        V lt = splat<V>(0u); eq = splat<V>(0xFFFFFFFFu);
#pragma unroll
        for (int p = P - 1; p >= 0; --p) {
            if ((hi >> p) & 1u) { lt = or4(lt, andn4(eq, pl[p])); eq = and4(eq, pl[p]); }
            else eq = andn4(eq, pl[p]);
        }
        res = and4(res, or4(lt, eq));
    }
    return res;
}

__device__ __forceinline__ uint32_t valid_mask32(uint64_t s, uint64_t begin, uint64_t end) {
    if (s + 32 <= begin || s >= end) return 0u;
    uint32_t m = 0xFFFFFFFFu;
    if (s < begin) m &= 0xFFFFFFFFu << (uint32_t)(begin - s);
    if (s + 32 > end) m &= 0xFFFFFFFFu >> (uint32_t)(s + 32 - end);
    return m;
}

constexpr int kGateThreads = 256;
#ifndef CANOPY_GATE_MIN_BLOCKS
#define CANOPY_GATE_MIN_BLOCKS 4
#endif
[[maybe_unused]] constexpr int kGateMinBlocks = CANOPY_GATE_MIN_BLOCKS;

#if CANOPY_ROW_HALVES == 1  // first-generation register-load kernel, kept for A/B on 512-byte rows only
__global__ void __launch_bounds__(kGateThreads, kGateMinBlocks)
gate_kernel(uint4* __restrict__ vals, const uint4* __restrict__ grec, const uint32_t* __restrict__ child,
            const uint32_t* __restrict__ layer_off, uint32_t n_layers, uint32_t n_nodes, uint32_t n_ev,
            unsigned long long* __restrict__ tallies, uint64_t first_slab, uint64_t begin, uint64_t end) {
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    const uint32_t slab = blockIdx.x;
    uint4* base = vals + (size_t)slab * n_nodes * kRowVec;
    const uint64_t s0 = (first_slab + slab) * (uint64_t)kSlabSamples + lane * 128u;
    const uint4 mask = make_uint4(valid_mask32(s0, begin, end), valid_mask32(s0 + 32, begin, end),
                                  valid_mask32(s0 + 64, begin, end), valid_mask32(s0 + 96, begin, end));
    for (uint32_t l = 0; l < n_layers; ++l) {
        const uint32_t g_end = __ldg(layer_off + l + 1);
        uint32_t g = __ldg(layer_off + l) + warp;
        // Metadata pipeline: the record of gate g+2W and the first 32 child ids of gate g+W are requested
        // before the child rows of gate g, so the record -> ids -> rows dependency never stalls the warp.
        uint4 rec = make_uint4(0, 0, 0, 0), rec1 = rec;
        uint32_t ids = 0;
        if (g < g_end) {
            rec = __ldg(grec + g);
            if (g + n_warps < g_end) rec1 = __ldg(grec + g + n_warps);
            if (lane < (rec.y & 0xFFFFu)) ids = __ldg(child + rec.x + lane);
        }
        for (; g < g_end; g += n_warps) {
            uint4 rec2 = make_uint4(0, 0, 0, 0);
            uint32_t ids1 = 0;
            if (g + 2 * n_warps < g_end) rec2 = __ldg(grec + g + 2 * n_warps);
            if (g + n_warps < g_end && lane < (rec1.y & 0xFFFFu)) ids1 = __ldg(child + rec1.x + lane);
            const uint32_t* ch = child + rec.x;
            const uint32_t n = rec.y & 0xFFFFu, op = rec.z & 0xFFu, flags = rec.z >> 8;
            uint4 r;
            if (op == kDevAnd) r = eval_simple<kDevAnd>(base, ch, n, lane, ids);
            else if (op == kDevOr) r = eval_simple<kDevOr>(base, ch, n, lane, ids);
            else if (op == kDevXor) r = eval_simple<kDevXor>(base, ch, n, lane, ids);
            else {
                const uint32_t lo = rec.w & 0xFFFFu, hi = rec.w >> 16;
                if (n < 8) r = eval_range<3, uint4>(base, ch, n, lo, hi, lane, 0, ids);
                else if (n < 64) r = eval_range<6, uint4>(base, ch, n, lo, hi, lane, 0, ids);
                else if (n < 512) {  // 9 planes: two passes of 8 bytes per lane
                    const uint2 a = eval_range<9, uint2>(base, ch, n, lo, hi, lane, 0, ids);
                    const uint2 b = eval_range<9, uint2>(base, ch, n, lo, hi, lane, 1, ids);
                    r = make_uint4(a.x, a.y, b.x, b.y);
                }
                else {
                    r.x = eval_range<16, uint32_t>(base, ch, n, lo, hi, lane, 0, ids);
                    r.y = eval_range<16, uint32_t>(base, ch, n, lo, hi, lane, 1, ids);
                    r.z = eval_range<16, uint32_t>(base, ch, n, lo, hi, lane, 2, ids);
                    r.w = eval_range<16, uint32_t>(base, ch, n, lo, hi, lane, 3, ids);
                }
            }
            if (flags & kFlagStore) st_cg(base + (size_t)(n_ev + g) * kRowVec + lane, r);
            if (flags & kFlagTally) {
                uint32_t c = __popc(r.x & mask.x) + __popc(r.y & mask.y) + __popc(r.z & mask.z) + __popc(r.w & mask.w);
                c = __reduce_add_sync(0xFFFFFFFFu, c);
                if (lane == 0 && c) atomicAdd(tallies + g, (unsigned long long)c);
            }
            rec = rec1; rec1 = rec2; ids = ids1;
        }
        __syncthreads();  // rows written in this layer are read by later layers of the same CTA only
    }
}

#endif

// ================================================================================================
// Ring gate kernel (slab walker, TMA-bulk staging).
//
// Same decomposition as gate_kernel -- one CTA walks every layer of one slab, __syncthreads() between
// layers -- but child rows are no longer loaded into registers.  Each warp owns a ring of kRingSlots
// 512-byte row slots in shared memory and streams the child rows of its share of the layer through it with
// `cp.async.bulk.shared::cluster.global` (SASS UBLKCP), one copy per row, each completing on the slot's
// own mbarrier.  The lowering cuts every layer into kGateWarps contiguous runs of equal cost
// (Lowered::split / split_child), so the rows a warp needs are simply child[c0..c1) in order, independent
// of gate boundaries: the producer side is a linear stream, the consumer side walks the gates and pops rows.
// A row is re-requested into a slot one pop after the slot was read, so kRingSlots-1 rows (7.5 KB per warp,
// 180 KB per SM at 3 CTAs) are in flight without holding a single register.
// Arguments of a gate come plain-first, complemented-last (GateRec::n_children >> 16 = #complemented), so
// the consumer needs no per-row flag.  Count gates use uint4 bit planes straight from shared memory
// (3/6/9 planes, one pass); fan-in >= kWideFanin count gates sit at the end of the layer and take the
// multi-pass register path of gate_kernel.
// ================================================================================================
#ifndef CANOPY_RING_GROUPS
#define CANOPY_RING_GROUPS 2
#endif
#ifndef CANOPY_GROUP_ROWS
#define CANOPY_GROUP_ROWS 4
#endif
constexpr uint32_t kGroupRows = CANOPY_GROUP_ROWS;         // rows per mbarrier phase (power of two, <= 8)
constexpr uint32_t kRingGroups = CANOPY_RING_GROUPS;      // groups per warp ring (power of two)
constexpr uint32_t kRingSlots = kGroupRows * kRingGroups; // rows per warp ring
constexpr uint32_t kRingSmemBytes = kGateWarps * kRingSlots * kRowBytes + kGateWarps * kRingGroups * 8;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }

__device__ __forceinline__ bool elect_one() {  // one lane of the (converged) warp, chosen by the hardware
    uint32_t pred;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xFFFFFFFF;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
    return v;
}

// One lane's share of a node row: kRowHalves x 16 bytes (block h of the row sits at h * 512 + lane * 16).
struct Row { uint4 h[kRowHalves]; };
#define CANOPY_ROW_OP(name, expr)                                           \
    __device__ __forceinline__ Row name(Row a, Row b) {                     \
        Row r;                                                              \
        _Pragma("unroll") for (uint32_t i = 0; i < kRowHalves; ++i) r.h[i] = expr; \
        return r;                                                           \
    }
CANOPY_ROW_OP(and4, and4(a.h[i], b.h[i]))
CANOPY_ROW_OP(or4, or4(a.h[i], b.h[i]))
CANOPY_ROW_OP(xor4, xor4(a.h[i], b.h[i]))
CANOPY_ROW_OP(andn4, andn4(a.h[i], b.h[i]))
#undef CANOPY_ROW_OP
__device__ __forceinline__ Row xors4(Row a, uint32_t m) {
    Row r;
#pragma unroll
    for (uint32_t i = 0; i < kRowHalves; ++i) r.h[i] = xors4(a.h[i], m);
    return r;
}
template <> __device__ __forceinline__ Row splat<Row>(uint32_t v) {
    Row r;
#pragma unroll
    for (uint32_t i = 0; i < kRowHalves; ++i) r.h[i] = splat4(v);
    return r;
}
__device__ __forceinline__ uint32_t popc_masked(Row r, Row m) {
    uint32_t c = 0;
#pragma unroll
    for (uint32_t i = 0; i < kRowHalves; ++i)
        c += __popc(r.h[i].x & m.h[i].x) + __popc(r.h[i].y & m.h[i].y) + __popc(r.h[i].z & m.h[i].z) + __popc(r.h[i].w & m.h[i].w);
    return c;
}

// Per-warp stream of child rows through the shared-memory ring.  Rows travel in groups of kGroupRows
// consecutive rows of the run: one mbarrier phase (one expect_tx, one try_wait) per group, so the
// ~90-cycle mbarrier round trip and the copy-issue sequence are paid once per 2 KB, not once per row.
// Every member except `lane_off`/`idreg`/`idnext` is warp-uniform and derived from values the compiler can
// prove uniform (kernel parameters, blockIdx, loop counters, shuffles from a uniform lane), so the copy
// operands live in uniform registers and UBLKCP issues without a per-lane serialisation loop.
template <bool G4>
struct RowRing {
    const uint4* base;          // this slab's rows
    const CUtensorMap* tmap;    // G4: tensor map of the value buffer, rows of kRowBytes
    uint32_t row0;              // G4: tensor row of node 0 of this slab
    const uint32_t* ids;        // child + c0: the warp's run of child codes for this layer
    uint32_t ring;              // shared address of slot 0 of this warp's ring
    uint32_t bars;              // shared address of mbarrier 0 of this warp's ring
    uint32_t lane_off;          // lane * 16
    uint32_t lane;
    uint32_t total;             // rows in the run
    uint32_t n_groups;          // groups in the run
    uint32_t k_cur;             // groups of the run handed to the consumer so far
    uint32_t gi_next;           // ring group the next consumed group sits in
    uint32_t phase;             // bit g = parity of the phase the consumer waits for next on ring group g
    uint32_t cur;               // shared address of this lane's 16 bytes of the next row to read
    uint32_t in_group;          // rows of the current group not read yet
    uint32_t idreg, idnext;     // child codes of rows [32k, 32k+32) of the run (and of the next 32), one per lane
#ifdef CANOPY_RING_DUMMY_U
    uint32_t dummy, dummy_zero;
#endif

    __device__ __forceinline__ void issue_group(uint32_t k, uint32_t gi) {  // request rows [4k, 4k+4) of the run into ring group gi
        const uint32_t i = k * kGroupRows;
        if ((i & 31u) == 8u) idnext = (i + 24u + lane < total) ? __ldg(ids + i + 24u + lane) : 0u;  // codes of the next 32 rows
        if ((i & 31u) == 0u && i) idreg = idnext;
        const uint32_t cnt = min(kGroupRows, total - i);
        uint32_t code[kGroupRows];
#pragma unroll
        for (uint32_t r = 0; r < kGroupRows; ++r) code[r] = __shfl_sync(0xFFFFFFFFu, idreg, (i + r) & 31u);
        if (G4) {
            // one tensor copy gathers the four rows of the group (tile::gather4: a column offset and four row
            // coordinates; rows land back to back).  A short last group repeats its first row: always 4 rows, 4 KB.
            static_assert(!G4 || kGroupRows == 4, "tile::gather4 moves four rows");
            uint32_t row[4];
#pragma unroll
            for (uint32_t r = 0; r < 4; ++r) row[r] = row0 + ((r < cnt ? code[r] : code[0]) & 0x7FFFFFFFu);
            if (elect_one()) {
                const uint32_t bar = bars + gi * 8u;
                const uint32_t dst = ring + gi * (kGroupRows * kRowBytes);
                mbar_expect_tx(bar, kGroupRows * kRowBytes);
                asm volatile("cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];"
                             ::"r"(dst), "l"(tmap), "r"(0), "r"(row[0]), "r"(row[1]), "r"(row[2]), "r"(row[3]), "r"(bar) : "memory");
            }
        } else if (elect_one()) {
            const uint32_t bar = bars + gi * 8u;
            const uint32_t dst = ring + gi * (kGroupRows * kRowBytes);
            mbar_expect_tx(bar, cnt * kRowBytes);
#pragma unroll
            for (uint32_t r = 0; r < kGroupRows; ++r)
                if (r < cnt) bulk_g2s(dst + r * kRowBytes, base + (size_t)(code[r] & 0x7FFFFFFFu) * kRowVec, kRowBytes, bar);
        }
    }
    __device__ __forceinline__ void begin(const uint4* base_, const uint32_t* ids_, uint32_t total_) {
        base = base_; ids = ids_; total = total_; k_cur = 0; in_group = 0;
        n_groups = (total + kGroupRows - 1) / kGroupRows;
        idreg = lane < total ? __ldg(ids + lane) : 0u; idnext = 0u;
        const uint32_t ng = min(n_groups, kRingGroups);
        for (uint32_t k = 0; k < ng; ++k) issue_group(k, (gi_next + k) & (kRingGroups - 1));
    }
    __device__ __forceinline__ void advance() {  // current group exhausted: recycle it, wait for the next one
        const uint32_t gi = gi_next;
        gi_next = (gi_next + 1) & (kRingGroups - 1);
        // the group read before this one is free now (its values have been consumed by the ALU): refill it
        if (k_cur && k_cur - 1 + kRingGroups < n_groups) issue_group(k_cur - 1 + kRingGroups, (gi - 1) & (kRingGroups - 1));
#ifdef CANOPY_RING_DUMMY_U
        mbar_wait(bars + gi * 8u + (dummy & dummy_zero), (phase >> gi) & 1u);
#else
        mbar_wait(bars + gi * 8u, (phase >> gi) & 1u);
#endif
        phase ^= 1u << gi;
        cur = ring + gi * (kGroupRows * kRowBytes) + lane_off;
        in_group = min(kGroupRows, total - k_cur * kGroupRows);
        ++k_cur;
    }
    __device__ __forceinline__ Row pop() {
        if (in_group == 0) advance();
#ifdef CANOPY_RING_DUMMY_U  // experiment: extra uniform-datapath instructions per row (3 per unit), see DESIGN.md 6
#pragma unroll
        for (int q = 0; q < CANOPY_RING_DUMMY_U; ++q) dummy = (dummy + 0x9E3779B9u) ^ (dummy >> 7);
#endif
        Row v;
#pragma unroll
        for (uint32_t h = 0; h < kRowHalves; ++h) v.h[h] = lds128(cur + h * 512u);
        cur += kRowBytes; --in_group;
        return v;
    }
};

template <int OP, class RR>
__device__ __forceinline__ Row ring_simple(RR& R, uint32_t n_plain, uint32_t n_comp) {
    Row acc = splat<Row>(OP == kDevAnd ? 0xFFFFFFFFu : 0u);
    for (uint32_t j = 0; j < n_plain; ++j) {
        const Row v = R.pop();
        acc = OP == kDevAnd ? and4(acc, v) : OP == kDevOr ? or4(acc, v) : xor4(acc, v);
    }
    for (uint32_t j = 0; j < n_comp; ++j) {
        const Row v = xors4(R.pop(), 0xFFFFFFFFu);
        acc = OP == kDevAnd ? and4(acc, v) : OP == kDevOr ? or4(acc, v) : xor4(acc, v);
    }
    return acc;
}

template <int P, class RR>
__device__ __forceinline__ Row ring_range(RR& R, uint32_t n_plain, uint32_t n_comp, uint32_t lo, uint32_t hi) {
    Row pl[P];
#pragma unroll
    for (int p = 0; p < P; ++p) pl[p] = splat<Row>(0u);
    const uint32_t n = n_plain + n_comp;
    for (uint32_t j = 0; j < n; ++j) {
        Row carry = xors4(R.pop(), j < n_plain ? 0u : 0xFFFFFFFFu);
#pragma unroll
        for (int p = 0; p < P; ++p) {
            const Row t = and4(pl[p], carry);
            pl[p] = xor4(pl[p], carry);
            carry = t;
        }
    }
    Row gt = splat<Row>(0u), eq = splat<Row>(0xFFFFFFFFu);
#pragma unroll
    for (int p = P - 1; p >= 0; --p) {
        if ((lo >> p) & 1u) eq = and4(eq, pl[p]);
        else { gt = or4(gt, and4(eq, pl[p])); eq = andn4(eq, pl[p]); }
    }
    Row res = or4(gt, eq);
    if (hi < n) {
        Row lt = splat<Row>(0u); eq = splat<Row>(0xFFFFFFFFu);
#pragma unroll
        for (int p = P - 1; p >= 0; --p) {
            if ((hi >> p) & 1u) { lt = or4(lt, andn4(eq, pl[p])); eq = and4(eq, pl[p]); }
            else eq = andn4(eq, pl[p]);
        }
        res = and4(res, or4(lt, eq));
    }
    return res;
}

#ifndef CANOPY_RING_MIN_BLOCKS
#define CANOPY_RING_MIN_BLOCKS (4 / CANOPY_ROW_HALVES)  // same bytes of rows in flight per SM for either row size
#endif
// G4: child rows are fetched four at a time with tile::gather4 tensor copies through `vmap` (a 2-D view of the value
// buffer, one tensor row per node row); otherwise one bulk copy per row (`vmap` unused).
template <bool G4>
#ifdef CANOPY_RING_MAXREG  // tuning knob: cap registers so one more sampler CTA fits beside the walk
__global__ void __maxnreg__(CANOPY_RING_MAXREG)
#else
__global__ void __launch_bounds__(kGateThreads, CANOPY_RING_MIN_BLOCKS)
#endif
gate_ring_kernel_t(const __grid_constant__ CUtensorMap vmap, uint4* __restrict__ vals, const uint4* __restrict__ grec,
                   const uint32_t* __restrict__ child,
                   const uint32_t* __restrict__ layer_off, const uint32_t* __restrict__ wide_off,
                   const uint32_t* __restrict__ split, const uint32_t* __restrict__ split_child,
                   uint32_t n_layers, uint32_t n_nodes, uint32_t n_ev, unsigned long long* __restrict__ tallies,
                   uint64_t first_slab, uint32_t n_slabs, uint64_t begin, uint64_t end,
                   uint32_t* __restrict__ walk_done, uint32_t parity, uint32_t wait_val) {
    extern __shared__ __align__(128) unsigned char smem[];
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = __shfl_sync(0xFFFFFFFFu, threadIdx.x >> 5, 0);  // broadcast: provably warp-uniform
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kGateWarps * kRingSlots * kRowBytes);
    if (threadIdx.x < kGateWarps * kRingGroups) mbar_init(smem_u32(bars + threadIdx.x), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    // Buffer hazard of the pipelined run, closed explicitly: the sampler this walk releases below (S of batch b+1 ...
    // of the NEXT walk: S(b+2)) rewrites the value buffer that the walk BEFORE this one, G(b-1) ... read.  Stream order
    // already implies that walk is over (this launch is an ordinary one); CUDA does not document that for a chain of
    // programmatic launches, so every walk CTA counts itself out in walk_done[parity] when it is done, and this walk
    // only releases its dependents after the previous walk's count is complete.  All CTAs of the previous walk are
    // resident or finished when this kernel starts, so the wait cannot deadlock; in practice it never spins.
    if (threadIdx.x == 0 && wait_val) {
        const volatile uint32_t* prev = walk_done + (parity ^ 1u);
        while (*prev < wait_val) __nanosleep(64);
        __threadfence();
    }
    __syncthreads();
    // Pipelined run: the sampler of the next batch is launched as a programmatic dependent of this walk and needs
    // nothing from it, so it may start as soon as every CTA of this (persistent, single-wave) grid is resident.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    RowRing<G4> R;
    R.tmap = &vmap; R.row0 = 0;
    R.lane = lane; R.lane_off = lane * 16u; R.gi_next = 0; R.phase = 0;
#ifdef CANOPY_RING_DUMMY_U
    R.dummy = blockIdx.x; R.dummy_zero = (n_layers >> 30) << 3;  // zero at run time, unknown at compile time
#endif
    R.ring = smem_u32(smem) + warp * (kRingSlots * kRowBytes);
    R.bars = smem_u32(bars) + warp * (kRingGroups * 8u);

    for (uint32_t slab = blockIdx.x; slab < n_slabs; slab += gridDim.x) {
        uint4* base = vals + (size_t)slab * n_nodes * kRowVec;
        R.row0 = slab * n_nodes;
        Row mask;
#pragma unroll
        for (uint32_t h = 0; h < kRowHalves; ++h) {
            const uint64_t s0 = (first_slab + slab) * (uint64_t)kSlabSamples + h * 4096u + lane * 128u;
            mask.h[h] = make_uint4(valid_mask32(s0, begin, end), valid_mask32(s0 + 32, begin, end),
                                   valid_mask32(s0 + 64, begin, end), valid_mask32(s0 + 96, begin, end));
        }
        for (uint32_t l = 0; l < n_layers; ++l) {
            const uint32_t* sp = split + l * (kGateWarps + 1) + warp;
            const uint32_t* sc = split_child + l * (kGateWarps + 1) + warp;
            const uint32_t g0 = __ldg(sp), g1 = __ldg(sp + 1);
            const uint32_t c0 = __ldg(sc), c1 = __ldg(sc + 1);
            bool stored = false;
            if (g0 < g1) {
                uint4 rec = __ldg(grec + g0);
                uint4 rec1 = g0 + 1 < g1 ? __ldg(grec + g0 + 1) : make_uint4(0, 0, 0, 0);
                R.begin(base, child + c0, c1 - c0);
                for (uint32_t g = g0; g < g1; ++g) {
                    const uint4 rec2 = g + 2 < g1 ? __ldg(grec + g + 2) : make_uint4(0, 0, 0, 0);
                    const uint32_t n = rec.y & 0xFFFFu, n_comp = rec.y >> 16, n_plain = n - n_comp;
                    const uint32_t op = rec.z & 0xFFu, flags = rec.z >> 8;
                    Row r;
                    if (op == kDevAnd) r = ring_simple<kDevAnd>(R, n_plain, n_comp);
                    else if (op == kDevOr) r = ring_simple<kDevOr>(R, n_plain, n_comp);
                    else if (op == kDevXor) r = ring_simple<kDevXor>(R, n_plain, n_comp);
                    else {
                        const uint32_t lo = rec.w & 0xFFFFu, hi = rec.w >> 16;
                        if (n < 8) r = ring_range<3>(R, n_plain, n_comp, lo, hi);
                        else if (n < 64) r = ring_range<6>(R, n_plain, n_comp, lo, hi);
                        else r = ring_range<9>(R, n_plain, n_comp, lo, hi);
                    }
                    if (flags & kFlagStore) {
#pragma unroll
                        for (uint32_t h = 0; h < kRowHalves; ++h) st_cg(base + (size_t)(n_ev + g) * kRowVec + h * 32u + lane, r.h[h]);
                        stored = true;
                    }
                    if (flags & kFlagTally) {
                        const uint32_t c = __reduce_add_sync(0xFFFFFFFFu, popc_masked(r, mask));
                        if (lane == 0 && c) atomicAdd(tallies + g, (unsigned long long)c);
                    }
                    rec = rec1; rec1 = rec2;
                }
            }
            // wide count gates (fan-in >= kWideFanin): multi-pass register path, dealt round-robin
            const uint32_t w_end = __ldg(layer_off + l + 1);
            for (uint32_t g = __ldg(wide_off + l) + warp; g < w_end; g += kGateWarps) {
                const uint4 rec = __ldg(grec + g);
                const uint32_t n = rec.y & 0xFFFFu, flags = rec.z >> 8, lo = rec.w & 0xFFFFu, hi = rec.w >> 16;
                const uint32_t* ch = child + rec.x;
                const uint32_t ids = lane < n ? __ldg(ch + lane) : 0u;
                Row r;
#pragma unroll
                for (uint32_t h = 0; h < kRowHalves; ++h) {
                    const uint4* bh = base + h * 32u;  // block h of every row
                    r.h[h].x = eval_range<16, uint32_t>(bh, ch, n, lo, hi, lane, 0, ids);
                    r.h[h].y = eval_range<16, uint32_t>(bh, ch, n, lo, hi, lane, 1, ids);
                    r.h[h].z = eval_range<16, uint32_t>(bh, ch, n, lo, hi, lane, 2, ids);
                    r.h[h].w = eval_range<16, uint32_t>(bh, ch, n, lo, hi, lane, 3, ids);
                }
                if (flags & kFlagStore) {
#pragma unroll
                    for (uint32_t h = 0; h < kRowHalves; ++h) st_cg(base + (size_t)(n_ev + g) * kRowVec + h * 32u + lane, r.h[h]);
                    stored = true;
                }
                if (flags & kFlagTally) {
                    const uint32_t c = __reduce_add_sync(0xFFFFFFFFu, popc_masked(r, mask));
                    if (lane == 0 && c) atomicAdd(tallies + g, (unsigned long long)c);
                }
            }
            // rows stored through the generic proxy are read by later layers through the async proxy (bulk copies)
            if (stored) fence_proxy_async();
            __syncthreads();
        }
    }
    // every row this CTA had to read has been read (each layer ends in a CTA barrier after its last ring pop)
    if (threadIdx.x == 0) { __threadfence(); atomicAdd(walk_done + parity, 1u); }
}

constexpr auto gate_ring_kernel = gate_ring_kernel_t<false>;     // bulk copy per row
constexpr auto gate_ring_kernel_g4 = gate_ring_kernel_t<true>;  // tile::gather4, four rows per copy

__global__ void finalize_kernel(unsigned long long* __restrict__ out, const unsigned long long* __restrict__ gate_tallies,
                                const uint8_t* __restrict__ kind, const uint32_t* __restrict__ gate, uint32_t n,
                                unsigned long long n_samples) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint8_t k = kind[i];
    unsigned long long v = 0;
    if (k == kTallyTrue) v = n_samples;
    else if (k == kTallyGate) v = gate_tallies[gate[i]];
    else if (k == kTallyNotGate) v = n_samples - gate_tallies[gate[i]];
    out[i] = v;
}

// Repack rows of one batch into [event][word] order for the sampler parity hook: out[ev][w] = word (word_off + w) of
// the batch's row of event ev (a row holds kSlabWords consecutive words per slab).
__global__ void gather_words_kernel(const uint4* __restrict__ vals, uint32_t n_ev, uint32_t n_nodes, uint64_t word_off,
                                    uint64_t n_words, unsigned long long* __restrict__ out /* [n_ev][n_words] */) {
    const uint32_t ev = blockIdx.x;
    const uint64_t w = (uint64_t)blockIdx.y * blockDim.x + threadIdx.x;
    if (w >= n_words) return;
    const uint64_t gw = word_off + w, slab = gw / kSlabWords;
    const unsigned long long* row = (const unsigned long long*)(vals + ((size_t)slab * n_nodes + ev) * kRowVec);
    out[(size_t)ev * n_words + w] = row[gw % kSlabWords];
}

// ================================================================================================
// Integer-pipe microbenchmarks (roofline denominators for the sampler; MEASURED_PEAKS.json has none).
// ================================================================================================
template <int KIND>
__global__ void __launch_bounds__(256) intpeak_kernel(uint32_t* out, uint32_t iters, uint32_t seed) {
    uint32_t a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = seed * (threadIdx.x + 1) + i * 0x9E3779B9u + blockIdx.x;
    const uint32_t b = seed ^ 0x5bd1e995u, c = seed + 77u;
    uint32_t u[8];  // kinds 5, 6: chains on warp-uniform values (block index and parameters only) -> uniform datapath
#pragma unroll
    for (int i = 0; i < 8; ++i) u[i] = seed * (blockIdx.x + 1) + i * 0x85EBCA6Bu;
    for (uint32_t it = 0; it < iters; ++it) {
#pragma unroll
        for (int rep = 0; rep < 4; ++rep) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (KIND == 5 || KIND == 6) u[i] = (u[i] + b) ^ (c + i);  // UIADD3 + ULOP3
                if (KIND == 6) {  // ... beside two vector LOP3
                    asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(b), "r"(c + i));
                    asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(c), "r"(b + i));
                }
                if (KIND == 0) {  // LOP3 with a loop-carried dependency per chain
                    asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(b), "r"(c + i));
                } else if (KIND == 1) {  // IMAD.WIDE.U32: 32x32 -> 64, consume hi ^ lo through the next multiply
                    uint32_t lo, hi;
                    asm volatile("{ .reg .u64 t; mul.wide.u32 t, %2, %3; mov.b64 {%0, %1}, t; }" : "=r"(lo), "=r"(hi) : "r"(a[i]), "r"(kPhiloxM0));
                    a[i] = hi ^ lo;  // +1 LOP3 on the other pipe; counted as one IMAD.WIDE
                } else if (KIND == 3) {
                    asm volatile("popc.b32 %0, %0;" : "+r"(a[i]));
                } else if (KIND == 4) {  // 1 IMAD.WIDE.U32 : 2 LOP3 per chain step
                    uint32_t lo, hi;
                    asm volatile("{ .reg .u64 t; mul.wide.u32 t, %2, %3; mov.b64 {%0, %1}, t; }" : "=r"(lo), "=r"(hi) : "r"(a[i]), "r"(kPhiloxM0));
                    asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(lo) : "r"(lo), "r"(b), "r"(c + i));
                    asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(a[i]) : "r"(lo), "r"(hi), "r"(c));
                }
            }
            if (KIND == 2) {  // two interleaved Philox rounds-chains: 4 x (2 IMAD.WIDE + 2 LOP3)
                philox_round(a[0], a[1], a[2], a[3], b, c);
                philox_round(a[4], a[5], a[6], a[7], c, b);
                philox_round(a[0], a[1], a[2], a[3], c, b);
                philox_round(a[4], a[5], a[6], a[7], b, c);
            }
        }
    }
    uint32_t x = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) x ^= a[i] ^ u[i];
    if (x == 0x12345678u) out[blockIdx.x * blockDim.x + threadIdx.x] = x;  // keep the chains alive
}

void microbench_int(int device, int kind, double* lane_ops_per_s, double* ms_out) {
    CUDA_CHECK(cudaSetDevice(device));
    cudaDeviceProp prop; CUDA_CHECK(cudaGetDeviceProperties(&prop, device));
    const int blocks = prop.multiProcessorCount * 8, threads = 256;
    const uint32_t iters = 2048;
    uint32_t* d = nullptr; CUDA_CHECK(cudaMalloc(&d, (size_t)blocks * threads * 4));
    cudaEvent_t e0, e1; CUDA_CHECK(cudaEventCreate(&e0)); CUDA_CHECK(cudaEventCreate(&e1));
    auto launch = [&]() {
        switch (kind) {
            case 0: intpeak_kernel<0><<<blocks, threads>>>(d, iters, 12345u); break;
            case 1: intpeak_kernel<1><<<blocks, threads>>>(d, iters, 12345u); break;
            case 2: intpeak_kernel<2><<<blocks, threads>>>(d, iters, 12345u); break;
            case 4: intpeak_kernel<4><<<blocks, threads>>>(d, iters, 12345u); break;
            case 5: intpeak_kernel<5><<<blocks, threads>>>(d, iters, 12345u); break;
            case 6: intpeak_kernel<6><<<blocks, threads>>>(d, iters, 12345u); break;
            default: intpeak_kernel<3><<<blocks, threads>>>(d, iters, 12345u); break;
        }
    };
    for (int w = 0; w < 3; ++w) launch();
    CUDA_CHECK(cudaDeviceSynchronize());
    const int reps = 5;
    CUDA_CHECK(cudaEventRecord(e0));
    for (int r = 0; r < reps; ++r) launch();
    CUDA_CHECK(cudaEventRecord(e1));
    CUDA_CHECK(cudaEventSynchronize(e1));
    float ms = 0; CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
    ms /= reps;
    // lane-instructions per launch: kinds 0,1,3: iters*4*8 ; kind 2: iters*4*4 rounds * 4 instr
    // kind 4 counts all three instructions of a step (1 IMAD.WIDE + 2 LOP3)
    // kind 5: 2 uniform instructions per step; kind 6: 2 uniform + 2 vector per step (uniform ones counted per warp x 32,
    // so that "lane-instr/s" / 32 is warp instructions per second for every kind)
    const double per_thread = kind == 2 ? (double)iters * 4 * 4 * 4 : kind == 4 ? (double)iters * 4 * 8 * 3
                            : kind == 5 ? (double)iters * 4 * 8 * 2 : kind == 6 ? (double)iters * 4 * 8 * 4 : (double)iters * 4 * 8;
    *lane_ops_per_s = per_thread * blocks * threads / (ms * 1e-3);
    *ms_out = ms;
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d);
}

// ================================================================================================
// Per-device state
// ================================================================================================
struct DeviceState {
    int device = -1;
    int sm_count = 0;
    uint32_t *ev_orig = nullptr, *thr32 = nullptr, *child = nullptr, *layer_off = nullptr, *tally_gate = nullptr;
    uint32_t *wide_off = nullptr, *split = nullptr, *split_child = nullptr;
    int ring_ctas_per_sm = 0;   // occupancy of gate_ring_kernel on this device
    CUtensorMap vmap[2];        // tile::gather4 view of vals[b]: tensor row = node row of kRowBytes; valid iff vmap_ok[b]
    bool vmap_ok[2] = {false, false};
    uint64_t* thr64 = nullptr;
    uint32_t* ev_perm = nullptr;   // spec 3 only
    uint64_t* ev_cdf = nullptr;    // spec 3 only
    uint4* grec = nullptr;
    uint8_t* tally_kind = nullptr;
    unsigned long long* gate_tallies = nullptr;
    unsigned long long* out_tallies = nullptr;   // used when the caller wants host output
    void* arena = nullptr;                       // one allocation holding every plan array above
    uint4* vals[2] = {nullptr, nullptr};
    size_t vals_bytes[2] = {0, 0};  // capacity of each value buffer
    size_t vals_slabs = 0;  // capacity of each buffer in slabs
    int n_bufs = 0;
    cudaStream_t main = nullptr, s_sample = nullptr, s_gate = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_t0 = nullptr, ev_t1 = nullptr;
    cudaEvent_t ev_sampled[2] = {nullptr, nullptr}, ev_consumed[2] = {nullptr, nullptr};
    uint64_t h2d_bytes = 0;
    std::mutex run_mu;  // one run at a time per device (the value buffers and streams are per-device scratch)
    // Completion of the last run that used this state's scratch (value buffers, gate tallies, thresholds).  A run may
    // return with its kernels still in flight (canopy_mc_run_device without a report): whatever touches the scratch
    // next from ANOTHER stream, or from the host, waits on this event first.
    cudaEvent_t ev_done = nullptr;
    cudaStream_t last_stream = nullptr;
    bool ran = false;
    uint32_t* walk_done = nullptr;   // 2 counters: walk CTAs of even / odd batches that have finished (pipeline hazard guard)
};

// Order `st` after the last run on this state, if that run was on another stream.
static void order_after_last_run(DeviceState* s, cudaStream_t st) {
    if (s->ran && s->last_stream != st) CUDA_CHECK(cudaStreamWaitEvent(st, s->ev_done, 0));
}
static void mark_run(DeviceState* s, cudaStream_t st) {
    CUDA_CHECK(cudaEventRecord(s->ev_done, st));
    s->ran = true; s->last_stream = st;
}
// The host is about to touch (or give away) scratch that the last run may still be using.
static void wait_last_run(DeviceState* s) {
    if (s->ran) CUDA_CHECK(cudaEventSynchronize(s->ev_done));
}

// ------------------------------------------------------------------------------------------------
// Value-buffer pool.  A value buffer is GBs (c2: 4 GB, c4: 48 GB per buffer); cudaMalloc + cudaFree of that size
// cost tens of milliseconds, which a one-shot quantification (plan -> run -> destroy) or an uncertainty loop that
// lowers a new plan per trial would pay every time.  Buffers of destroyed plans are parked here, per device, and
// handed to the next plan that fits; canopy_trim_device_cache() returns them to the driver.
// ------------------------------------------------------------------------------------------------
struct ValPool {
    struct Entry { int device; void* ptr; size_t bytes; };
    std::mutex mu;
    std::vector<Entry> parked;
    static constexpr size_t kMaxParkedPerDevice = 2;

    void* acquire(int device, size_t bytes, size_t* got) {
        {
            std::lock_guard<std::mutex> lock(mu);
            int best = -1;
            for (size_t i = 0; i < parked.size(); ++i)
                if (parked[i].device == device && parked[i].bytes >= bytes && (best < 0 || parked[i].bytes < parked[best].bytes)) best = (int)i;
            if (best >= 0) {
                Entry e = parked[best];
                parked.erase(parked.begin() + best);
                *got = e.bytes;
                return e.ptr;
            }
        }
        void* p = nullptr;
        if (cudaMalloc(&p, bytes) != cudaSuccess) {
            cudaGetLastError();
            trim(device);  // parked buffers that were too small still hold memory: give it back and retry once
            if (cudaMalloc(&p, bytes) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        }
        *got = bytes;
        return p;
    }
    void release(int device, void* ptr, size_t bytes) {
        if (!ptr) return;
        std::lock_guard<std::mutex> lock(mu);
        parked.push_back({device, ptr, bytes});
        size_t n = 0;
        for (auto& e : parked) n += e.device == device;
        while (n > kMaxParkedPerDevice) {  // evict the smallest buffer of this device
            int worst = -1;
            for (size_t i = 0; i < parked.size(); ++i)
                if (parked[i].device == device && (worst < 0 || parked[i].bytes < parked[worst].bytes)) worst = (int)i;
            cudaFree(parked[worst].ptr);
            parked.erase(parked.begin() + worst);
            --n;
        }
    }
    int count_at_least(int device, size_t bytes) {
        std::lock_guard<std::mutex> lock(mu);
        int n = 0;
        for (auto& e : parked) n += e.device == device && e.bytes >= bytes;
        return n;
    }
    size_t parked_bytes(int device) {
        std::lock_guard<std::mutex> lock(mu);
        size_t b = 0;
        for (auto& e : parked) if (e.device == device) b += e.bytes;
        return b;
    }
    void trim(int device) {  // device < 0: every device
        std::lock_guard<std::mutex> lock(mu);
        int cur = -1; cudaGetDevice(&cur);
        for (size_t i = 0; i < parked.size();) {
            if (device < 0 || parked[i].device == device) {
                cudaSetDevice(parked[i].device);
                cudaFree(parked[i].ptr);
                parked.erase(parked.begin() + i);
            } else ++i;
        }
        if (cur >= 0) cudaSetDevice(cur);
    }
};
static ValPool g_val_pool;

void trim_device_cache(int device) { g_val_pool.trim(device); }

void destroy_device_state(DeviceState* s) {
    if (!s) return;
    cudaSetDevice(s->device);
    cudaDeviceSynchronize();  // nothing of this plan (on its own or on a caller's stream) may still run on a parked buffer
    cudaFree(s->arena);
    for (int b = 0; b < 2; ++b) g_val_pool.release(s->device, s->vals[b], s->vals_bytes[b]);
    if (s->main) cudaStreamDestroy(s->main);
    if (s->s_sample) cudaStreamDestroy(s->s_sample);
    if (s->s_gate) cudaStreamDestroy(s->s_gate);
    for (cudaEvent_t e : {s->ev_fork, s->ev_join, s->ev_t0, s->ev_t1, s->ev_sampled[0], s->ev_sampled[1],
                          s->ev_consumed[0], s->ev_consumed[1], s->ev_done})
        if (e) cudaEventDestroy(e);
    delete s;
}

// Plan arrays go into ONE device allocation (256-byte aligned pieces): a dozen cudaMalloc calls cost more than the
// upload itself for the graphs of this path (c2: 0.3 MB, c4: 3.4 MB).
struct ArenaPlan {
    struct Piece { void** dst; const void* src; size_t bytes, off; };
    std::vector<Piece> pieces;
    size_t total = 0;
    template <class P, class T>
    void add(P*& dst, const std::vector<T>& v, size_t min_elems = 1) {
        const size_t bytes = std::max(v.size(), min_elems) * sizeof(T);
        pieces.push_back({reinterpret_cast<void**>(&dst), v.empty() ? nullptr : v.data(), bytes, total});
        pieces.back().bytes = v.size() * sizeof(T);  // bytes to copy; the reservation below may be larger
        total += (bytes + 255) & ~size_t(255);
    }
    template <class P>
    void reserve(P*& dst, size_t bytes) {
        pieces.push_back({reinterpret_cast<void**>(&dst), nullptr, 0, total});
        total += (std::max<size_t>(bytes, 8) + 255) & ~size_t(255);
    }
};

// What the engine needs to know about a device, queried once per process: cudaGetDeviceProperties costs milliseconds
// per call, which a one-shot plan -> run -> destroy would pay every time.
struct DeviceInfo { bool ready = false; int major = 0, minor = 0, sm_count = 0, ring_ctas_per_sm = 0, hi_prio = 0; };
static std::mutex g_info_mu;
static DeviceInfo g_info[64];

static DeviceInfo device_info(int device) {
    std::lock_guard<std::mutex> lock(g_info_mu);
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0)
        throw Error(CANOPY_E_CUDA, "no CUDA device available: libcanopy has no CPU fallback");
    if (device < 0 || device >= count || device >= 64) throw Error(CANOPY_E_INVALID, "device ordinal out of range");
    DeviceInfo& I = g_info[device];
    if (I.ready) return I;
    CUDA_CHECK(cudaDeviceGetAttribute(&I.major, cudaDevAttrComputeCapabilityMajor, device));
    CUDA_CHECK(cudaDeviceGetAttribute(&I.minor, cudaDevAttrComputeCapabilityMinor, device));
    CUDA_CHECK(cudaDeviceGetAttribute(&I.sm_count, cudaDevAttrMultiProcessorCount, device));
    if (I.major != 10) {
        cudaDeviceProp prop; CUDA_CHECK(cudaGetDeviceProperties(&prop, device));
        throw Error(CANOPY_E_CUDA, std::string("device '") + prop.name + "' is sm_" + std::to_string(I.major) +
                                       std::to_string(I.minor) + "; libcanopy is built for sm_100a only");
    }
    CUDA_CHECK(cudaSetDevice(device));
    // function attributes are per device (context): set once
    CUDA_CHECK(cudaFuncSetAttribute(gate_ring_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kRingSmemBytes));
    CUDA_CHECK(cudaFuncSetAttribute(gate_ring_kernel_g4, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kRingSmemBytes));
    CUDA_CHECK(cudaFuncSetAttribute(sampler_sliced_kernel<32, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    CUDA_CHECK(cudaFuncSetAttribute(sampler_sliced_kernel<32, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    CUDA_CHECK(cudaFuncSetAttribute(sampler_sliced_kernel<64, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&I.ring_ctas_per_sm, gate_ring_kernel, kGateThreads, kRingSmemBytes));
    if (I.ring_ctas_per_sm < 1) throw Error(CANOPY_E_CUDA, "gate_ring_kernel does not fit on an SM");
    int lo_prio = 0;
    CUDA_CHECK(cudaDeviceGetStreamPriorityRange(&lo_prio, &I.hi_prio));
    I.ready = true;
    return I;
}

static DeviceState* get_state(canopy_plan* plan, int device) {
    for (DeviceState* s : plan->dev) if (s->device == device) { CUDA_CHECK(cudaSetDevice(device)); return s; }
    const DeviceInfo info = device_info(device);
    CUDA_CHECK(cudaSetDevice(device));
    const Lowered& L = plan->low;
    DeviceState* s = new DeviceState();
    s->device = device; s->sm_count = info.sm_count; s->ring_ctas_per_sm = info.ring_ctas_per_sm;
    try {
        static_assert(sizeof(GateRec) == sizeof(uint4), "GateRec must be 16 bytes");
        ArenaPlan A;
        A.add(s->ev_orig, L.event_orig); A.add(s->thr32, L.event_thr32); A.add(s->thr64, L.event_thr64);
        A.add(s->child, L.child); A.add(s->layer_off, L.layer_off); A.add(s->tally_gate, L.tally_gate);
        A.add(s->wide_off, L.wide_off); A.add(s->split, L.split); A.add(s->split_child, L.split_child);
        A.add(s->tally_kind, L.tally_kind); A.add(s->grec, L.gate_rec);
        A.add(s->ev_cdf, L.event_cdf); A.add(s->ev_perm, L.event_perm);
        A.reserve(s->gate_tallies, (std::max<size_t>(L.n_gates(), 1) + 1) * 8);  // + the two walk_done counters
        A.reserve(s->out_tallies, std::max<size_t>(L.tally_kind.size(), 1) * 8);
        if (cudaMalloc(&s->arena, A.total) != cudaSuccess) { cudaGetLastError(); throw Error(CANOPY_E_OOM, "cudaMalloc failed for plan arrays"); }
        {   // one staged upload of all plan arrays instead of a dozen small synchronous copies
            std::vector<char> stage(A.total);
            for (auto& p : A.pieces) {
                *p.dst = static_cast<char*>(s->arena) + p.off;
                if (p.bytes) std::memcpy(stage.data() + p.off, p.src, p.bytes);
                s->h2d_bytes += p.bytes;
            }
            CUDA_CHECK(cudaMemcpy(s->arena, stage.data(), A.total, cudaMemcpyHostToDevice));
        }
        CUDA_CHECK(cudaStreamCreateWithFlags(&s->main, cudaStreamNonBlocking));
        CUDA_CHECK(cudaStreamCreateWithFlags(&s->s_sample, cudaStreamNonBlocking));
        // the gate walk of batch b competes with the sampler of batch b+1 for SM slots: its (few, persistent)
        // CTAs must win, so its stream gets the highest priority
        CUDA_CHECK(cudaStreamCreateWithPriority(&s->s_gate, cudaStreamNonBlocking, info.hi_prio));
        for (cudaEvent_t* e : {&s->ev_fork, &s->ev_join, &s->ev_sampled[0], &s->ev_sampled[1], &s->ev_consumed[0], &s->ev_consumed[1], &s->ev_done})
            CUDA_CHECK(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
        CUDA_CHECK(cudaEventCreate(&s->ev_t0)); CUDA_CHECK(cudaEventCreate(&s->ev_t1));
        s->walk_done = reinterpret_cast<uint32_t*>(s->gate_tallies + std::max<size_t>(L.n_gates(), 1));
    } catch (...) { destroy_device_state(s); throw; }
    plan->dev.push_back(s);
    return s;
}

// 2-D tensor map over a value buffer for the gather4 copies of the gate walk: [rows][kRowBytes / 4] of uint32, box = one
// row (tile::gather4 takes four row coordinates per copy).  Row coordinates are int32: larger buffers fall back to the
// per-row bulk copies.
static bool encode_vals_map(CUtensorMap* map, void* base, size_t bytes) {
    static const auto encode = []() {
        void* fn = nullptr; cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr) != cudaSuccess || qr != cudaDriverEntryPointSuccess) fn = nullptr;
        return reinterpret_cast<CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill)>(fn);
    }();
    static const bool off = getenv("CANOPY_RING_GATHER4") && !strcmp(getenv("CANOPY_RING_GATHER4"), "0");  // A/B knob
    const size_t rows = bytes / kRowBytes;
    if (off || !encode || rows == 0 || rows > 0x7FFFFFFFull) return false;
    const cuuint64_t gdim[2] = {kRowBytes / 4, (cuuint64_t)rows};
    const cuuint64_t gstr[1] = {kRowBytes};
    const cuuint32_t box[2] = {kRowBytes / 4, 1}, estr[2] = {1, 1};
    return encode(map, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, base, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

static void ensure_vals(DeviceState* s, const Lowered& L, size_t slabs, int n_bufs) {
    const size_t per_slab = (size_t)L.n_nodes() * kRowBytes;
    const size_t bytes = std::max<size_t>(slabs * per_slab, 16);
    bool ok = s->n_bufs >= n_bufs;
    for (int b = 0; ok && b < n_bufs; ++b) ok = s->vals_bytes[b] >= bytes;
    if (ok) { s->vals_slabs = std::max(s->vals_slabs, slabs); return; }
    wait_last_run(s);  // the buffers go back to the pool, where another plan may pick them up
    for (int b = 0; b < 2; ++b) { g_val_pool.release(s->device, s->vals[b], s->vals_bytes[b]); s->vals[b] = nullptr; s->vals_bytes[b] = 0; s->vmap_ok[b] = false; }
    s->vals_slabs = 0; s->n_bufs = 0;
    for (int b = 0; b < n_bufs; ++b) {
        s->vals[b] = static_cast<uint4*>(g_val_pool.acquire(s->device, bytes, &s->vals_bytes[b]));
        if (!s->vals[b]) throw Error(CANOPY_E_OOM, "cudaMalloc failed for the value buffer (" + std::to_string(bytes >> 20) + " MiB)");
        s->vmap_ok[b] = encode_vals_map(&s->vmap[b], s->vals[b], s->vals_bytes[b]);
    }
    s->vals_slabs = slabs; s->n_bufs = n_bufs;
}

static PhiloxKeys make_keys(uint64_t seed) {
    PhiloxKeys K;
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
    for (int r = 0; r < 10; ++r) { K.k0[r] = k0; K.k1[r] = k1; k0 += kPhiloxW0; k1 += kPhiloxW1; }
    return K;
}

// `shared_sm`: the launch runs beside the other kernel of the pipeline (sampler of batch b+1 beside the gate walk of
// batch b).  Measured on B200 (gpurun_out/rows_sweep.txt, ovl1.txt): the two co-reside best with the sampler capped at
// 4 CTAs per SM (by a dummy dynamic-shared-memory request that leaves room for the gate rings) and 64 KB of gate rings
// per SM (1 CTA with 1 KB rows, 2 CTAs with 512 B rows) on the high-priority stream.
constexpr int kSharedSamplerSmem = kRowHalves == 2 ? 38000 : 47000;
constexpr int kSharedRingCtas = 2 / kRowHalves;

// `pdl`: launch as a programmatic dependent of the kernel before it in the stream (the gate walk of the previous
// batch): the sampler starts once that walk's CTAs are resident instead of when the walk has finished.
static void launch_sampler(const DeviceState* s, const Lowered& L, uint4* vals, uint64_t first_slab, uint32_t n_slabs,
                           const PhiloxKeys& K, cudaStream_t st, bool shared_sm = false, bool pdl = false) {
    if (L.n_events() == 0 || n_slabs == 0) return;
    const uint32_t n_blk = n_slabs * kRowHalves;  // 4096-sample blocks
    const dim3 grid((L.n_events() + 7) / 8, n_blk), block(256);
    // sliced sampler: each warp loops over up to CANOPY_SLICED_SLABS_PER_WARP blocks of its event
    const dim3 sgrid(grid.x, (n_blk + CANOPY_SLICED_SLABS_PER_WARP - 1) / CANOPY_SLICED_SLABS_PER_WARP);
#define CANOPY_SAMPLER_ARGS vals, s->ev_orig, s->thr32, s->thr64, L.n_events(), L.n_nodes(), first_slab, K
#define CANOPY_SLICED_ARGS vals, s->ev_orig, s->thr32, s->thr64, L.n_events(), L.n_nodes(), first_slab, n_slabs, K
    static const int pack = getenv("CANOPY_SAMPLER_PACK") ? atoi(getenv("CANOPY_SAMPLER_PACK")) : 1;  // tuning knob
    if (L.opts.sampler == CANOPY_SAMPLER_SLICED || L.opts.sampler == CANOPY_SAMPLER_SPARSE) {
        // tuning knob: dynamic shared memory requested only to cap the sampler's CTAs per SM (room for gate CTAs)
        static const int env_smem = getenv("CANOPY_SAMPLER_SMEM") ? atoi(getenv("CANOPY_SAMPLER_SMEM")) : -1;
        // (a programmatic dependent starts after the walk's CTAs are resident: no room has to be kept for them)
        const int cap_smem = !shared_sm ? 0 : env_smem >= 0 ? env_smem : pdl ? 0 : kSharedSamplerSmem;
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = sgrid; cfg.blockDim = block; cfg.dynamicSmemBytes = (size_t)cap_smem; cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
        const uint32_t n_ev = L.n_events(), n_nodes = L.n_nodes();
        // spec 3 with 64-bit thresholds is spec 2 (no event takes the sparse path): the plain kernel serves it
        const bool sparse = L.opts.sampler == CANOPY_SAMPLER_SPARSE && L.opts.threshold_bits == 32;
        auto kern = L.opts.threshold_bits == 64 ? sampler_sliced_kernel<64, false>
                    : sparse                    ? sampler_sliced_kernel<32, true>
                                                : sampler_sliced_kernel<32, false>;
        CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, vals, (const uint32_t*)s->ev_orig, (const uint32_t*)s->thr32,
                                      (const uint64_t*)s->thr64, n_ev, n_nodes, first_slab, n_slabs, K,
                                      (const uint32_t*)s->ev_perm, (const uint64_t*)s->ev_cdf));
    }
    else if (L.opts.threshold_bits == 64) sampler_kernel<64, 0><<<grid, block, 0, st>>>(CANOPY_SAMPLER_ARGS);
    else if (pack == 0) sampler_kernel<32, 0><<<grid, block, 0, st>>>(CANOPY_SAMPLER_ARGS);
    else sampler_kernel<32, 1><<<grid, block, 0, st>>>(CANOPY_SAMPLER_ARGS);
#undef CANOPY_SAMPLER_ARGS
#undef CANOPY_SLICED_ARGS
    CUDA_CHECK(cudaGetLastError());
}

// Returns the number of CTAs launched (what the next walk waits for in walk_done[parity]).
static uint32_t launch_gates(const DeviceState* s, const Lowered& L, uint4* vals, uint64_t first_slab, uint32_t n_slabs,
                             uint64_t begin, uint64_t end, cudaStream_t st, bool shared_sm = false, uint32_t parity = 0,
                             uint32_t wait_val = 0) {
    if (L.n_gates() == 0 || n_slabs == 0) return 0;
#if CANOPY_ROW_HALVES == 1
    static const bool direct = getenv("CANOPY_GATE_KERNEL") && !strcmp(getenv("CANOPY_GATE_KERNEL"), "direct");  // A/B knob
    if (direct) {
        gate_kernel<<<n_slabs, kGateThreads, 0, st>>>(vals, s->grec, s->child, s->layer_off, L.n_layers(), L.n_nodes(),
                                                      L.n_events(), s->gate_tallies, first_slab, begin, end);
    } else
#endif
    {
        // persistent: one CTA per resident slot, each walks slabs blockIdx.x, +gridDim.x, ...
        static const int cap = getenv("CANOPY_RING_CTAS") ? atoi(getenv("CANOPY_RING_CTAS")) : 0;  // tuning knob
        const int want = cap > 0 ? cap : shared_sm ? kSharedRingCtas : s->ring_ctas_per_sm;
        const int per_sm = std::min(want, s->ring_ctas_per_sm);
        const uint32_t grid = std::min<uint32_t>(n_slabs, (uint32_t)(s->sm_count * per_sm));
        const int b = vals == s->vals[1] ? 1 : 0;
        if (vals == s->vals[b] && s->vmap_ok[b] && kGroupRows == 4)
            gate_ring_kernel_g4<<<grid, kGateThreads, kRingSmemBytes, st>>>(
                s->vmap[b], vals, s->grec, s->child, s->layer_off, s->wide_off, s->split, s->split_child, L.n_layers(), L.n_nodes(),
                L.n_events(), s->gate_tallies, first_slab, n_slabs, begin, end, s->walk_done, parity, wait_val);
        else
            gate_ring_kernel<<<grid, kGateThreads, kRingSmemBytes, st>>>(
                s->vmap[0], vals, s->grec, s->child, s->layer_off, s->wide_off, s->split, s->split_child, L.n_layers(), L.n_nodes(),
                L.n_events(), s->gate_tallies, first_slab, n_slabs, begin, end, s->walk_done, parity, wait_val);
        CUDA_CHECK(cudaGetLastError());
        return grid;
    }
    CUDA_CHECK(cudaGetLastError());
    return 0;
}

static uint32_t pick_batch_slabs(const DeviceState* s, const Lowered& L, const canopy_mc_opts& o, uint64_t total_slabs, int n_bufs) {
    uint64_t want = o.batch_slabs > 0 ? (uint64_t)o.batch_slabs : (uint64_t)s->sm_count * (4 / kRowHalves);  // 16384 samples per SM
    want = std::min<uint64_t>(want, 65535 / kRowHalves);  // gridDim.y of the direct sampler
    want = std::min(want, std::max<uint64_t>(total_slabs, 1));
    {   // buffers on hand (own or parked by an earlier plan) that already hold `want` slabs: no need to ask the driver
        // how much memory is free (cudaMemGetInfo is slow)
        const size_t need = (size_t)want * L.n_nodes() * kRowBytes;
        int have = 0;
        for (int b = 0; b < 2; ++b) have += s->vals_bytes[b] >= need;
        if (have >= n_bufs || have + g_val_pool.count_at_least(s->device, need) >= n_bufs) return (uint32_t)want;
    }
    size_t free_b = 0, total_b = 0;
    CUDA_CHECK(cudaMemGetInfo(&free_b, &total_b));
    free_b += s->vals_bytes[0] + s->vals_bytes[1] + g_val_pool.parked_bytes(s->device);  // what a regrow can reuse or release
    const size_t per_slab = (size_t)L.n_nodes() * kRowBytes;
    const uint64_t fit = per_slab ? (uint64_t)((free_b / 2) / (per_slab * n_bufs)) : want;
    if (fit == 0) throw Error(CANOPY_E_OOM, "one slab of node rows does not fit in device memory");
    want = std::min(want, fit);
    want = std::min<uint64_t>(want, 65535 / kRowHalves);  // gridDim.y of the direct sampler
    want = std::min(want, std::max<uint64_t>(total_slabs, 1));
    return (uint32_t)want;
}

// Slabs per batch.  Uniform batches of `batch` slabs; with `taper` (experiment, CANOPY_TAPER=1) a pipelined run makes
// its first and last batch a quarter of a full one, because nothing overlaps the first sampler launch and the last gate
// walk.  Measured neutral (gpurun_out/pdl_taper_*.json): the pipelined step costs sampler + stolen issue slots, not
// fill/drain.  Tallies do not depend on the schedule (tests: batching invariance).
static std::vector<uint32_t> batch_schedule(uint64_t total_slabs, uint32_t batch, bool taper, uint32_t sm_count) {
    std::vector<uint32_t> v;
    const uint32_t edge = std::max<uint32_t>(std::min<uint32_t>(batch, sm_count), batch / 4);
    if (!taper || total_slabs < 2ull * edge + batch) {
        for (uint64_t done = 0; done < total_slabs; done += batch) v.push_back((uint32_t)std::min<uint64_t>(batch, total_slabs - done));
        return v;
    }
    const uint64_t mid = total_slabs - 2ull * edge, n_mid = (mid + batch - 1) / batch;
    v.push_back(edge);
    for (uint64_t i = 0; i < n_mid; ++i) v.push_back((uint32_t)(mid / n_mid + (i < mid % n_mid ? 1 : 0)));
    v.push_back(edge);
    return v;
}

void mc_run(canopy_plan* plan, const canopy_mc_opts& o, uint64_t* tallies_host, uint64_t* tallies_dev, canopy_mc_report* rep) {
    const Lowered& L = plan->low;
    DeviceState* s = nullptr;
    { std::lock_guard<std::mutex> lock(plan->mu); s = get_state(plan, o.device); }
    std::lock_guard<std::mutex> run_lock(s->run_mu);
    if (o.sample_begin + o.n_samples < o.sample_begin) throw Error(CANOPY_E_INVALID, "sample range overflows 64 bits");
    const uint64_t begin = o.sample_begin, end = o.sample_begin + o.n_samples;
    const uint64_t slab0 = begin / kSlabSamples;
    const uint64_t total_slabs = o.n_samples ? (end - 1) / kSlabSamples - slab0 + 1 : 0;
    const bool profile = o.profile != 0;
    bool overlap = o.overlap != 0 && !profile;
    cudaStream_t user = o.stream ? (cudaStream_t)o.stream : s->main;
    const uint32_t n_tallies = (uint32_t)L.tally_kind.size();
    unsigned long long* d_out = tallies_dev ? (unsigned long long*)tallies_dev : s->out_tallies;

    uint32_t batch = 1; uint64_t n_batches = 0;
    std::vector<uint32_t> sched;  // slabs per batch, in order
    if (total_slabs && L.n_gates()) {
        batch = pick_batch_slabs(s, L, o, total_slabs, overlap ? 2 : 1);
        if ((total_slabs + batch - 1) / batch < 2) overlap = false;
        static const bool taper = getenv("CANOPY_TAPER") != nullptr;  // A/B knob: measured neutral on B200, off by default
        sched = batch_schedule(total_slabs, batch, overlap && o.batch_slabs <= 0 && taper, (uint32_t)s->sm_count);
        n_batches = sched.size();
        ensure_vals(s, L, batch, overlap ? 2 : 1);
    }
    const PhiloxKeys K = make_keys(o.seed);
    canopy_mc_report R{};
    R.slab_samples = kSlabSamples; R.batch_slabs = batch; R.n_batches = n_batches;

    order_after_last_run(s, user);
    CUDA_CHECK(cudaEventRecord(s->ev_t0, user));
    CUDA_CHECK(cudaMemsetAsync(s->gate_tallies, 0, (std::max<size_t>(L.n_gates(), 1) + 1) * 8, user));  // tallies + walk_done
    R.other_launches += 1;
    struct EventList {  // profile events, 3 per batch; destroyed on every path out of this function
        std::vector<cudaEvent_t> v;
        ~EventList() { for (cudaEvent_t e : v) cudaEventDestroy(e); }
        size_t size() const { return v.size(); }
        cudaEvent_t operator[](size_t i) const { return v[i]; }
        void push_back(cudaEvent_t e) { v.push_back(e); }
    } pe;
    if (!overlap) {
        uint64_t fs = slab0;
        for (uint64_t b = 0; b < n_batches; fs += sched[b], ++b) {
            const uint32_t ns = sched[b];
            if (profile) { cudaEvent_t e[3]; for (auto& x : e) { CUDA_CHECK(cudaEventCreate(&x)); pe.push_back(x); } CUDA_CHECK(cudaEventRecord(e[0], user)); }
            launch_sampler(s, L, s->vals[0], fs, ns, K, user);
            if (profile) CUDA_CHECK(cudaEventRecord(pe[pe.size() - 2], user));
            launch_gates(s, L, s->vals[0], fs, ns, begin, end, user);
            if (profile) CUDA_CHECK(cudaEventRecord(pe[pe.size() - 1], user));
            R.sampler_launches += L.n_events() ? 1 : 0; R.gate_launches += 1;
        }
    } else {
        // sampler of batch b+1 (integer-issue bound) runs beside the gate walk of batch b (HBM bound)
        static const bool use_streams = getenv("CANOPY_PIPELINE") && !strcmp(getenv("CANOPY_PIPELINE"), "streams");  // A/B knob
        if (!use_streams) {
            // One stream, programmatic dependent launch: S0, G0, S1*, G1, S2*, ...  (* = dependent of the walk before
            // it).  G_b is an ordinary launch, so it starts when S_b and everything before it have finished; S_{b+1}
            // starts as soon as the CTAs of G_b (one per SM, all resident at once on the idle GPU) have signalled, fills
            // the rest of every SM and keeps it full when the walk ends.  The order in which the two kernels get their
            // SM slots is fixed by construction; with two streams it depended on how the driver mapped the streams
            // to hardware queues (40 % of fresh plans ran without any overlap, gpurun_out/var.txt).
            // Hazards: S_{b+2} rewrites the buffer G_b read; it starts after G_{b+1} started, i.e. after G_b finished.
            uint64_t fs = slab0;
            uint32_t done[2] = {0, 0};  // walk CTAs launched so far on even / odd batches
            for (uint64_t b = 0; b < n_batches; fs += sched[b], ++b) {
                const int buf = (int)(b & 1);
                launch_sampler(s, L, s->vals[buf], fs, sched[b], K, user, b > 0, b > 0);
                done[buf] += launch_gates(s, L, s->vals[buf], fs, sched[b], begin, end, user, b + 1 < n_batches, (uint32_t)buf,
                                          done[buf ^ 1]);
                R.sampler_launches += L.n_events() ? 1 : 0; R.gate_launches += 1;
            }
        } else {
        CUDA_CHECK(cudaEventRecord(s->ev_fork, user));
        CUDA_CHECK(cudaStreamWaitEvent(s->s_sample, s->ev_fork, 0));
        CUDA_CHECK(cudaStreamWaitEvent(s->s_gate, s->ev_fork, 0));
        uint64_t fs = slab0;
        for (uint64_t b = 0; b < n_batches; fs += sched[b], ++b) {
            const int buf = (int)(b & 1);
            const uint32_t ns = sched[b];
            if (b >= 2) CUDA_CHECK(cudaStreamWaitEvent(s->s_sample, s->ev_consumed[buf], 0));
            launch_sampler(s, L, s->vals[buf], fs, ns, K, s->s_sample, b > 0);               // batch 0 has the GPU to itself
            CUDA_CHECK(cudaEventRecord(s->ev_sampled[buf], s->s_sample));
            CUDA_CHECK(cudaStreamWaitEvent(s->s_gate, s->ev_sampled[buf], 0));
            launch_gates(s, L, s->vals[buf], fs, ns, begin, end, s->s_gate, b + 1 < n_batches);  // so has the last gate walk
            CUDA_CHECK(cudaEventRecord(s->ev_consumed[buf], s->s_gate));
            R.sampler_launches += L.n_events() ? 1 : 0; R.gate_launches += 1;
        }
        CUDA_CHECK(cudaEventRecord(s->ev_join, s->s_gate));
        CUDA_CHECK(cudaStreamWaitEvent(user, s->ev_join, 0));
        }
    }
    if (n_tallies) {
        finalize_kernel<<<(n_tallies + 255) / 256, 256, 0, user>>>(d_out, s->gate_tallies, s->tally_kind, s->tally_gate,
                                                                  n_tallies, (unsigned long long)o.n_samples);
        CUDA_CHECK(cudaGetLastError());
        R.other_launches += 1;
    }
    CUDA_CHECK(cudaEventRecord(s->ev_t1, user));
    if (tallies_host && n_tallies) {
        CUDA_CHECK(cudaMemcpyAsync(tallies_host, d_out, (size_t)n_tallies * 8, cudaMemcpyDeviceToHost, user));
        R.d2h_bytes = (uint64_t)n_tallies * 8;
    }
    if (tallies_host || rep || profile) {
        CUDA_CHECK(cudaStreamSynchronize(user));
        float ms = 0; CUDA_CHECK(cudaEventElapsedTime(&ms, s->ev_t0, s->ev_t1)); R.total_ms = ms;
        for (size_t i = 0; i + 3 <= pe.size(); i += 3) {
            float a = 0, b2 = 0;
            CUDA_CHECK(cudaEventElapsedTime(&a, pe[i], pe[i + 1]));
            CUDA_CHECK(cudaEventElapsedTime(&b2, pe[i + 1], pe[i + 2]));
            R.sampler_ms += a; R.gate_ms += b2;
        }
    }
    mark_run(s, user);
    R.h2d_bytes = s->h2d_bytes;
    if (rep) *rep = R;
}

void mc_sample_words(canopy_plan* plan, const canopy_mc_opts& o, uint64_t first_word, uint64_t n_words, uint64_t* words_host) {
    const Lowered& L = plan->low;
    DeviceState* s = nullptr;
    { std::lock_guard<std::mutex> lock(plan->mu); s = get_state(plan, o.device); }
    std::lock_guard<std::mutex> run_lock(s->run_mu);
    if (first_word % 64 || n_words % 64 || n_words == 0)
        throw Error(CANOPY_E_INVALID, "first_word and n_words must be multiples of 64 (4096 samples)");
    if (L.n_events() == 0) return;
    const uint64_t slab0 = first_word / kSlabWords, n_slabs = (first_word + n_words - 1) / kSlabWords - slab0 + 1;
    if (n_slabs > 65535 / kRowHalves || n_words / 64 > 65535) throw Error(CANOPY_E_INVALID, "too many words requested");
    ensure_vals(s, L, n_slabs, 1);
    const PhiloxKeys K = make_keys(o.seed);
    order_after_last_run(s, s->main);
    launch_sampler(s, L, s->vals[0], slab0, (uint32_t)n_slabs, K, s->main);
    struct DevBuf {  // test hook only: a scratch buffer per call, freed on every path
        unsigned long long* p = nullptr;
        ~DevBuf() { if (p) cudaFree(p); }
    } d;
    const size_t bytes = (size_t)L.n_events() * n_words * 8;
    if (cudaMalloc(&d.p, bytes) != cudaSuccess) { cudaGetLastError(); throw Error(CANOPY_E_OOM, "cudaMalloc failed"); }
    gather_words_kernel<<<dim3(L.n_events(), (uint32_t)(n_words / 64)), 64, 0, s->main>>>(
        s->vals[0], L.n_events(), L.n_nodes(), first_word - slab0 * kSlabWords, n_words, d.p);
    CUDA_CHECK(cudaGetLastError());
    CUDA_CHECK(cudaMemcpyAsync(words_host, d.p, bytes, cudaMemcpyDeviceToHost, s->main));
    mark_run(s, s->main);
    CUDA_CHECK(cudaStreamSynchronize(s->main));
}

void refresh_thresholds(canopy_plan* plan) {
    const Lowered& L = plan->low;
    for (DeviceState* s : plan->dev) {
        std::lock_guard<std::mutex> run_lock(s->run_mu);  // never while a run is being enqueued ...
        CUDA_CHECK(cudaSetDevice(s->device));
        wait_last_run(s);  // ... nor under one still in flight on a user stream (blocking copies below do not order against it)
        if (!L.event_thr32.empty()) {
            CUDA_CHECK(cudaMemcpy(s->thr32, L.event_thr32.data(), L.event_thr32.size() * 4, cudaMemcpyHostToDevice));
            CUDA_CHECK(cudaMemcpy(s->thr64, L.event_thr64.data(), L.event_thr64.size() * 8, cudaMemcpyHostToDevice));
        }
        if (!L.event_perm.empty()) {   // spec 3: tables and warp order follow the thresholds (sizes do not change)
            CUDA_CHECK(cudaMemcpy(s->ev_cdf, L.event_cdf.data(), L.event_cdf.size() * 8, cudaMemcpyHostToDevice));
            CUDA_CHECK(cudaMemcpy(s->ev_perm, L.event_perm.data(), L.event_perm.size() * 4, cudaMemcpyHostToDevice));
        }
    }
}

}  // namespace canopy
