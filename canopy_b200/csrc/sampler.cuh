// Basic-event sampler kernels (north-star subsystem 2; SURVEY.md 8a row a12, ABSENT in the reference).
//
//   sampler_direct_kernel<BITS>    spec 1: one whole variate per (event, sample), borrow-chain compare + pack.
//   sampler_sliced_kernel<BITS>    spec 2: MSB-first comparison on Philox bit planes (H head planes + sparse tail).
//   sampler_classed_kernel<BITS>   spec 4: dense events (p >= 2^-Z) take the planes of spec 2; a rarer event draws, per group
//                                  of 2^k samples (k = 7 .. 13 by probability class), the NUMBER of ones from a cumulative
//                                  binomial table and their positions from the same Philox call.  The rows of a unit of four
//                                  events are assembled as a 4-KB image in shared memory and leave through TMA bulk stores.
//   sparse_table_kernel            the count tables of spec 4, built on the device from the 64-bit thresholds.
//
// Every kernel writes the value buffer vals[slab][node][half][lane] (uint4; plan.h): sample s of a slab's row is bit
// s & 31 of 32-bit word s >> 5 of the 1-KB row, so a row is also a plain little-endian bitmap of its 8192 samples.
#pragma once
#include "kernels.cuh"
#include "sparse_table.h"

namespace canopy {

// ================================================================================================
// Spec 1 ("direct").  One warp per (basic event, 4096-sample block); lane r owns words 2r and 2r+1 of the block, i.e.
// 128 consecutive samples = 32 Philox calls (call index q = s >> 2, output word s & 3).  All 32 calls are independent.
// ================================================================================================
// bits = 2*bits + (u < thr) in two instructions: `sub.cc (thr-1) - u` leaves the hardware carry set exactly
// when u <= thr-1 (the subtraction is thr-1 + ~u + 1, carry = no borrow), and addc shifts it in.
// thr >= 1 always holds: events whose threshold quantises to 0 are folded to constants by the lowering.
__device__ __forceinline__ uint32_t shift_in_lt(uint32_t bits, uint32_t u, uint32_t thr_m1) {
    asm("{ .reg .u32 t; sub.cc.u32 t, %2, %1; addc.u32 %0, %0, %0; }" : "+r"(bits) : "r"(u), "r"(thr_m1));
    return bits;
}

template <int BITS>
__global__ void __launch_bounds__(256)
sampler_direct_kernel(uint4* __restrict__ vals, const uint32_t* __restrict__ ev_orig,
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
            uint32_t bits = 0;  // samples are visited last-to-first so sample i lands on bit i
#pragma unroll
            for (int c = 7; c >= 0; --c) {
                const uint4 r = philox4x32_10(q_lo + h * 8 + c, q_hi, e, 0u, K);
                bits = shift_in_lt(bits, r.w, thr - 1); bits = shift_in_lt(bits, r.z, thr - 1);
                bits = shift_in_lt(bits, r.y, thr - 1); bits = shift_in_lt(bits, r.x, thr - 1);
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
// Spec 2 ("sliced") planes: one warp per event, lane r owns the 128 samples of group g = block * 32 + r.
//   head: H = CANOPY_SLICED_HEAD_BITS Philox calls ctr = (g, e, 0x100+k); word j of call k is bit (H-1-k)
//         of the variates' top H bits for samples 32j..32j+31 of the group.  Per plane and word: a lane
//         still undecided (all higher bits equal to the threshold's) becomes "less" where thr has 1 and u
//         has 0, "greater" where thr has 0 and u has 1.
//   tail: the samples of the group still undecided after H planes (probability 2^-H each) take their low
//         32-H (64-H) bits, in ascending sample order, from the output words of the group's tail calls
//         ctr = (g, e, 0x200+n) (resp. 0x300+n): four (two) undecided samples per call, so one tail call
//         per lane covers all but 3.5e-3 of the groups (H = 7) and is issued without divergence.
// ================================================================================================
constexpr int kHeadBits = CANOPY_SLICED_HEAD_BITS;
constexpr int kBlocksPerWarp = 8;   // 4096-sample blocks a plane warp walks per launch row (gridDim.y = blocks / 8)

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

// The threshold of a plane event as the group code needs it: everything here is warp-uniform (loaded once and broadcast),
// so the compiler keeps it in uniform registers.
struct PlaneThr {
    uint32_t head;                 // top H bits of the threshold, left-aligned in a 32-bit word
    uint32_t trans;                // bit k (MSB first) set iff threshold bits k and k-1 differ
    uint32_t tail32; uint64_t tail64;  // (thr mod 2^(BITS-H)) << H: compared with the raw tail variate
};
template <int BITS>
__device__ __forceinline__ PlaneThr plane_thr(uint32_t t32, uint64_t t64) {
    PlaneThr T; T.tail32 = 0; T.tail64 = 0;
    if (BITS == 32) { T.head = t32; T.tail32 = t32 << kHeadBits; }
    else {
        const uint32_t lo = bcast0((uint32_t)t64), hi = bcast0((uint32_t)(t64 >> 32));
        T.head = hi; T.tail64 = (((uint64_t)hi << 32) | lo) << kHeadBits;
    }
    T.trans = T.head ^ (T.head >> 1);
    return T;
}

// The 128 samples of group g of event e (original index): four 32-sample words.  Called by a whole warp (the tail call is
// issued warp-wide), each lane with its own group.
template <int BITS>
__device__ __forceinline__ uint4 plane_group(uint64_t g, uint32_t e, const PlaneThr& T, const PhiloxKeys& K) {
    const uint32_t g_lo = (uint32_t)g, g_hi = (uint32_t)(g >> 32);
    const uint32_t head = T.head, trans = T.trans;
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
                if (Q.pop_lt(T.tail32, T.tail64)) r[w] |= b;
            }
        }
    }
    return make_uint4(r[0], r[1], r[2], r[3]);
}

// All 4096-sample blocks blockIdx.y, +gridDim.y, ... of the batch for one event; lane r owns group block * 32 + r.
template <int BITS>
__device__ __forceinline__ void plane_rows(uint4* __restrict__ vals, uint32_t ev, uint32_t e, uint32_t t32, uint64_t t64,
                                           uint32_t n_nodes, uint64_t first_slab, uint32_t n_slabs, const PhiloxKeys& K) {
    const uint32_t lane = threadIdx.x & 31;
    const PlaneThr T = plane_thr<BITS>(t32, t64);
    const uint32_t n_blk = n_slabs * kRowHalves;
    for (uint32_t blk = blockIdx.y; blk < n_blk; blk += gridDim.y) {
        uint4* row = vals + ((size_t)(blk / kRowHalves) * n_nodes + ev) * kRowVec + (blk % kRowHalves) * 32u + lane;
        st_cg(row, plane_group<BITS>((first_slab * kRowHalves + blk) * 32u + lane, e, T, K));
    }
}

template <int BITS>
__global__ void __launch_bounds__(256, 6)
sampler_sliced_kernel(uint4* __restrict__ vals, const uint32_t* __restrict__ ev_orig,
                      const uint32_t* __restrict__ thr32, const uint64_t* __restrict__ thr64,
                      uint32_t n_ev, uint32_t n_nodes, uint64_t first_slab, uint32_t n_slabs, PhiloxKeys K) {
    const uint32_t ev = blockIdx.x * (blockDim.x >> 5) + bcast0(threadIdx.x >> 5);
    if (ev >= n_ev) return;
    const uint32_t e = bcast0(__ldg(ev_orig + ev));
    const uint32_t t32 = BITS == 32 ? bcast0(__ldg(thr32 + ev)) : 0u;
    const uint64_t t64 = BITS == 64 ? __ldg(thr64 + ev) : 0ull;
    plane_rows<BITS>(vals, ev, e, t32, t64, n_nodes, first_slab, n_slabs, K);
}

// ================================================================================================
// Spec 4 ("classed sparse").
// ================================================================================================
constexpr uint32_t kClassedWarps = 4;                       // warps per CTA of sampler_classed_kernel
constexpr uint32_t kUnitRows = 4;                           // events per sparse unit: a 4-KB row image per warp
constexpr uint32_t kUnitBytes = kUnitRows * kRowBytes;
constexpr uint32_t kClassedSmemBytes = kClassedWarps * kUnitBytes;
static_assert(kRowHalves == 2, "the sparse units of spec 4 are written for 8192-sample slabs (1-KB rows)");
static_assert((1u << CANOPY_SPARSE_MAX_LOG2) <= kSlabSamples, "a sparse group must not span slabs");

struct SparseClasses {            // kernel parameter, by value
    uint32_t n_dense;             // ev_perm[0, n_dense): dense events (planes)
    uint32_t ev_off[8];           // class c (k = 7 + c): ev_perm[ev_off[c], ev_off[c+1])
    uint32_t unit_off[8];         // its units of kUnitRows events: [unit_off[c], unit_off[c+1])
};

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void sts128_zero(uint32_t addr) {
    asm volatile("st.shared.v4.u32 [%0], {%1, %1, %1, %1};" ::"r"(addr), "r"(0u) : "memory");
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
    uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory"); return v;
}
__device__ __forceinline__ void sts32(uint32_t addr, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }
__device__ __forceinline__ void bulk_s2g(void* dst, uint32_t src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read_all() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ uint32_t ge64(uint64_t a, uint64_t b) { return a >= b ? 1u : 0u; }

// One unit = up to four events of one class; lane = one 2^k-sample group of the unit's 4-KB image (k = 7: eight groups per
// lane, one after the other; k >= 11: some lanes idle).
struct SparseUnit {
    uint32_t k, n_rows, rows_log, n_groups;
    uint32_t first;         // position in ev_perm of the unit's first event
    uint32_t my_ev, my_e;   // lane r < n_rows: device id and original index of event r of the unit
};
__device__ __forceinline__ SparseUnit sparse_unit_setup(const uint32_t* __restrict__ ev_orig, const uint32_t* __restrict__ ev_perm,
                                                        uint32_t unit, const SparseClasses& C) {
    const uint32_t lane = threadIdx.x & 31;
    // class of this unit (warp-uniform); constant indices only, so that the parameter struct stays in the constant bank
    uint32_t c = 0, u_lo = C.unit_off[0], e_lo = C.ev_off[0], e_hi = C.ev_off[1];
#pragma unroll
    for (uint32_t i = 1; i < 7; ++i)
        if (unit >= C.unit_off[i]) { c = i; u_lo = C.unit_off[i]; e_lo = C.ev_off[i]; e_hi = C.ev_off[i + 1]; }
    SparseUnit U;
    U.k = 7u + c;
    const uint32_t first = e_lo + (unit - u_lo) * kUnitRows;
    U.first = first;
    U.n_rows = min(kUnitRows, e_hi - first);
    U.my_ev = 0; U.my_e = 0;
    if (lane < U.n_rows) { U.my_ev = __ldg(ev_perm + first + lane); U.my_e = __ldg(ev_orig + U.my_ev); }
    U.rows_log = 13u - U.k;                       // log2(groups per row)
    U.n_groups = U.n_rows << U.rows_log;
    return U;
}
// The unit's rows of one slab (global slab index `gslab`) as bitmaps in the shared-memory image `img` (n_rows x 1 KB,
// cleared here).  Ends with the image complete and visible to the whole warp.
__device__ __forceinline__ void sparse_unit_fill(const SparseUnit& U, const uint64_t* __restrict__ cdf, uint64_t gslab,
                                                 const PhiloxKeys& K, uint32_t img) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t k = U.k, rows_log = U.rows_log;
    const uint32_t fpw = 32u / k, kmask = (1u << k) - 1u;
    for (uint32_t i = lane; i < U.n_rows * (kRowBytes / 16u); i += 32u) sts128_zero(img + i * 16u);
    __syncwarp();
    const uint64_t gbase = gslab << rows_log;
    for (uint32_t u0 = 0; u0 < U.n_groups; u0 += 32u) {
        const uint32_t uu = u0 + lane;
        const bool active = uu < U.n_groups;
        const uint32_t r = active ? uu >> rows_log : 0u, gr = uu & ((1u << rows_log) - 1u);
        const uint32_t ev = __shfl_sync(0xFFFFFFFFu, U.my_ev, r), e = __shfl_sync(0xFFFFFFFFu, U.my_e, r);
        const uint64_t g = gbase + gr;
        const uint32_t g_lo = (uint32_t)g, g_hi = (uint32_t)(g >> 32);
        const uint4 x = philox4x32_10(g_lo, g_hi, e, 0x400u, K);
        const uint64_t Uv = active ? ((uint64_t)x.y << 32) | x.x : 0ull;   // an idle lane draws no ones
        // count: N = #{j : U >= cdf_j}, four table entries per trip, while some lane still exceeds the last one
        const ulonglong2* tab = reinterpret_cast<const ulonglong2*>(cdf + (size_t)ev * CANOPY_SPARSE_CDF_LEN);
        uint32_t n;
        {
            const ulonglong2 a = __ldg(tab), b = __ldg(tab + 1);
            n = ge64(Uv, a.x) + ge64(Uv, a.y) + ge64(Uv, b.x) + ge64(Uv, b.y);
            if (__any_sync(0xFFFFFFFFu, Uv >= b.y)) {
#pragma unroll 1
                for (uint32_t j = 2; j < CANOPY_SPARSE_CDF_LEN / 2; j += 2) {
                    const ulonglong2 p = __ldg(tab + j), q = __ldg(tab + j + 1);
                    n += ge64(Uv, p.x) + ge64(Uv, p.y) + ge64(Uv, q.x) + ge64(Uv, q.y);
                    if (!__any_sync(0xFFFFFFFFu, Uv >= q.y)) break;
                }
            }
        }
        n = active ? n : 0u;   // Uv = 0 keeps an idle lane at zero only while cdf_0 > 0 (every real table; not the test hook's all-zero one)
        // placement: the first n distinct k-bit fields.  The lane owns the group's words of the image outright, so a
        // plain read-modify-write does; the old value tells whether the position was taken already.
        const uint32_t wbase = img + r * kRowBytes + (gr << (k - 3u));
        uint32_t f = x.z, avail = fpw, fw = 0;
        uint4 ext = x;
#pragma unroll 1
        while (n) {
            if (avail == 0) {   // field word fw: 0, 1 = words 2, 3 of call 0x400; then word (fw-2)&3 of call 0x401 + ((fw-2)>>2)
                avail = fpw; ++fw;
                if (fw == 1u) f = x.w;
                else {
                    const uint32_t kx = fw - 2u;
                    if ((kx & 3u) == 0u) ext = philox4x32_10(g_lo, g_hi, e, 0x401u + (kx >> 2), K);
                    f = (kx & 3u) == 0u ? ext.x : (kx & 3u) == 1u ? ext.y : (kx & 3u) == 2u ? ext.z : ext.w;
                }
            }
            const uint32_t pos = f & kmask;
            f >>= k; --avail;
            const uint32_t addr = wbase + ((pos >> 5) << 2), bit = 1u << (pos & 31u);
            const uint32_t old = lds32(addr);
            sts32(addr, old | bit);
            n -= (old & bit) ? 0u : 1u;
        }
    }
    __syncwarp();
}

// The unit x the slabs blockIdx.y, +gridDim.y, ... of the batch, into the value buffer.
__device__ __forceinline__ void sparse_unit(uint4* __restrict__ vals, const uint32_t* __restrict__ ev_orig,
                                            const uint32_t* __restrict__ ev_perm, const uint64_t* __restrict__ cdf,
                                            uint32_t unit, const SparseClasses& C, uint32_t n_nodes, uint64_t first_slab,
                                            uint32_t n_slabs, const PhiloxKeys& K, uint32_t img) {
    const uint32_t lane = threadIdx.x & 31;
    const SparseUnit U = sparse_unit_setup(ev_orig, ev_perm, unit, C);
    uint4* row_ptr[kUnitRows];
#pragma unroll
    for (uint32_t r = 0; r < kUnitRows; ++r) row_ptr[r] = vals + (size_t)__shfl_sync(0xFFFFFFFFu, U.my_ev, r) * kRowVec;
    for (uint32_t slab = blockIdx.y; slab < n_slabs; slab += gridDim.y) {
        if (lane == 0) bulk_wait_read_all();                 // the stores of the previous slab (lane 0's bulk group) have read the image
        __syncwarp();
        sparse_unit_fill(U, cdf, first_slab + slab, K, img);
        // image -> global: one bulk store per row (the rows of a unit's events are not adjacent in the value buffer)
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) {   // always the same lane: bulk groups are tracked per thread
            const size_t slab_off = (size_t)slab * n_nodes * kRowVec;
#pragma unroll
            for (uint32_t r = 0; r < kUnitRows; ++r)
                if (r < U.n_rows) bulk_s2g(row_ptr[r] + slab_off, img + r * kRowBytes, kRowBytes);
            bulk_commit();
        }
        __syncwarp();
    }
    if (lane == 0) bulk_wait_read_all();   // shared memory must outlive the reads of the last stores
    __syncwarp();
}

// grid.x = dense CTAs (one warp per dense event) ++ sparse CTAs (one warp per unit); the dense ones come first: they are
// the expensive ones, so the CTAs still running when a launch drains are the cheap units.
template <int BITS>
#ifndef CANOPY_CLASSED_MIN_BLOCKS
#define CANOPY_CLASSED_MIN_BLOCKS 8
#endif
__global__ void __launch_bounds__(kClassedWarps * 32, CANOPY_CLASSED_MIN_BLOCKS)
sampler_classed_kernel(uint4* __restrict__ vals, const uint32_t* __restrict__ ev_orig,
                       const uint32_t* __restrict__ thr32, const uint64_t* __restrict__ thr64,
                       const uint32_t* __restrict__ ev_perm, const uint64_t* __restrict__ cdf, const SparseClasses C,
                       uint32_t dense_ctas, uint32_t n_units, uint32_t n_nodes, uint64_t first_slab, uint32_t n_slabs,
                       const PhiloxKeys K) {
    extern __shared__ __align__(128) unsigned char classed_smem[];
    const uint32_t warp = bcast0(threadIdx.x >> 5);
    if (blockIdx.x < dense_ctas) {
        const uint32_t i = blockIdx.x * kClassedWarps + warp;
        if (i >= C.n_dense) return;
        const uint32_t ev = bcast0(__ldg(ev_perm + i));
        const uint32_t e = bcast0(__ldg(ev_orig + ev));
        const uint32_t t32 = BITS == 32 ? bcast0(__ldg(thr32 + ev)) : 0u;
        const uint64_t t64 = BITS == 64 ? __ldg(thr64 + ev) : 0ull;
        plane_rows<BITS>(vals, ev, e, t32, t64, n_nodes, first_slab, n_slabs, K);
    } else {
        const uint32_t unit = (blockIdx.x - dense_ctas) * kClassedWarps + warp;
        if (unit >= n_units) return;
        sparse_unit(vals, ev_orig, ev_perm, cdf, unit, C, n_nodes, first_slab, n_slabs, K, smem_addr(classed_smem) + warp * kUnitBytes);
    }
}

// Count tables of spec 4, one thread per live event (a dense event's row is left untouched).
__global__ void sparse_table_kernel(const uint64_t* __restrict__ thr64, uint32_t n_ev, uint64_t* __restrict__ cdf) {
    const uint32_t ev = blockIdx.x * blockDim.x + threadIdx.x;
    if (ev >= n_ev) return;
    const uint64_t t = thr64[ev];
    const uint32_t k = sparse_class_k(t);
    if (k) sparse_cdf(t, k, cdf + (size_t)ev * CANOPY_SPARSE_CDF_LEN);
}

}  // namespace canopy
