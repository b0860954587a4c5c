// Layered gate evaluation with fused tallies (north-star subsystem 3; SURVEY.md 8a row a13, ABSENT in the reference).
//
//   gate_ring_kernel_t<G4>   one CTA walks every layer of the DAG for one slab; child rows stream through per-warp
//                            shared-memory rings filled by TMA (tile::gather4 tensor copies, or one bulk copy per row),
//                            bit-sliced adders for k-of-n, popcount tallies reduced with warp REDUX and one 64-bit RED
//                            per (gate, slab).
//   finalize_kernel          maps device gate tallies to the caller-visible outputs, in place in the buffer that the
//                            NCCL all-reduce consumes.
//   gather_words_kernel      parity hook: repacks sampler rows as [event][word].
#pragma once
#include <cuda.h>  // CUtensorMap (types only; the encoder is fetched through cudaGetDriverEntryPoint)

#include "kernels.cuh"

namespace canopy {

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

// count-in-range with a bit-sliced ripple counter of P planes (P bits hold counts < 2^P).
// V is the per-lane vector: uint4 (one pass, 16 B per lane per child) for the common fan-ins, or
// uint32_t (four passes, one 32-bit component each) for fan-in >= 512 where 16 planes x 4 registers
// would not fit the register budget of the common path.
__device__ __forceinline__ uint32_t and4(uint32_t a, uint32_t b) { return a & b; }
__device__ __forceinline__ uint32_t or4(uint32_t a, uint32_t b) { return a | b; }
__device__ __forceinline__ uint32_t xor4(uint32_t a, uint32_t b) { return a ^ b; }
__device__ __forceinline__ uint32_t andn4(uint32_t a, uint32_t b) { return a & ~b; }
template <class V> __device__ __forceinline__ V splat(uint32_t v);
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

constexpr int kGateThreads = kGateWarps * 32;

// ================================================================================================
// Ring gate kernel (slab walker, TMA staging).
//
// One CTA walks every layer of one slab, __syncthreads() between layers; child rows are never loaded into registers by
// address.  Each warp owns a ring of kRingSlots 1-KB row slots in shared memory (kRingGroups groups of kGroupRows rows)
// and streams the child rows of its share of the layer through it, one group per copy: a tile::gather4 tensor copy
// (SASS UTMALDG.2D.GATHER4: four row coordinates, rows land back to back) or, for buffers of more than 2^31 rows, one
// cp.async.bulk per row (UBLKCP), completing on the group's mbarrier.  The lowering cuts every layer into kGateWarps
// contiguous runs of equal cost and lays their child codes out run by run, every run padded to whole groups
// (Lowered::split / split_child / dev_child), so the producer side is a linear stream of 128-bit code loads independent
// of gate boundaries, and the consumer side walks the gates and pops rows.  A group is re-requested one group after it
// was read, so kRingSlots - kGroupRows rows per warp (64 KB per CTA in all) are in flight without holding a register.
// Arguments of a gate come plain-first, complemented-last (GateRec::n_children >> 16 = #complemented), so the consumer
// needs no per-row flag.  Count gates keep uint4 bit planes in registers (3/6/9 planes, carry-save adders); count gates
// with fan-in >= kWideFanin sit at the end of the layer and take a multi-pass register path (eval_range).
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

// Per-warp stream of child rows through the shared-memory ring.  Rows travel in groups of kGroupRows consecutive rows of
// the run: one mbarrier phase (one expect_tx, one try_wait) and one copy instruction per group, so the mbarrier round trip
// and the copy-issue sequence are paid once per 4 KB.  Everything but `lane_off` is warp-uniform (kernel parameters,
// blockIdx, loop counters, broadcast loads), so the copy operands live in uniform registers; one elected lane issues.
template <bool G4>
struct RowRing {
    const uint4* base;          // this slab's rows
    const CUtensorMap* tmap;    // G4: tensor map of the value buffer, rows of kRowBytes
    uint32_t row0;              // G4: tensor row of node 0 of this slab
    const uint4* ids;           // the warp's run of child codes for this layer, four per group (the lowering pads every run)
    uint32_t ring;              // shared address of slot 0 of this warp's ring
    uint32_t bars;              // shared address of mbarrier 0 of this warp's ring
    uint32_t lane_off;          // lane * 16
    uint32_t n_groups;          // groups in the run
    uint32_t k_cur;             // groups of the run handed to the consumer so far
    uint32_t n_used;            // ring groups consumed since the kernel started: group gi = n_used & (kRingGroups-1), phase parity from the rest
    uint32_t cur;               // shared address of this lane's 16 bytes of the next row to read
    uint32_t in_group;          // rows of the current group not read yet
    uint4 codes;                // child codes of the group that will be requested next (loaded one request ahead)

    __device__ __forceinline__ void issue_group(uint32_t k, uint32_t gi) {  // request rows [4k, 4k+4) of the run into ring group gi
        static_assert(kGroupRows == 4, "a group is the four codes of one 128-bit load");
        const uint4 c = codes;
        if (k + 1 < n_groups) codes = __ldg(ids + k + 1);   // warp-uniform address: one broadcast load, needed one request later
        const uint32_t bar = bars + gi * 8u;
        const uint32_t dst = ring + gi * (kGroupRows * kRowBytes);
        if (G4) {
            // one tensor copy gathers the four rows of the group (tile::gather4: a column offset and four row
            // coordinates; rows land back to back)
            if (elect_one()) {
                mbar_expect_tx(bar, kGroupRows * kRowBytes);
                asm volatile("cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];"
                             ::"r"(dst), "l"(tmap), "r"(0), "r"(row0 + (c.x & 0x7FFFFFFFu)), "r"(row0 + (c.y & 0x7FFFFFFFu)),
                               "r"(row0 + (c.z & 0x7FFFFFFFu)), "r"(row0 + (c.w & 0x7FFFFFFFu)), "r"(bar) : "memory");
            }
        } else if (elect_one()) {
            mbar_expect_tx(bar, kGroupRows * kRowBytes);
            bulk_g2s(dst, base + (size_t)(c.x & 0x7FFFFFFFu) * kRowVec, kRowBytes, bar);
            bulk_g2s(dst + kRowBytes, base + (size_t)(c.y & 0x7FFFFFFFu) * kRowVec, kRowBytes, bar);
            bulk_g2s(dst + 2 * kRowBytes, base + (size_t)(c.z & 0x7FFFFFFFu) * kRowVec, kRowBytes, bar);
            bulk_g2s(dst + 3 * kRowBytes, base + (size_t)(c.w & 0x7FFFFFFFu) * kRowVec, kRowBytes, bar);
        }
    }
    __device__ __forceinline__ void begin(const uint4* base_, const uint32_t* ids_, uint32_t total_) {   // total_: multiple of 4
        base = base_; ids = reinterpret_cast<const uint4*>(ids_); k_cur = 0; in_group = 0;
        n_groups = total_ / kGroupRows;
        if (n_groups) codes = __ldg(ids);
        const uint32_t ng = min(n_groups, kRingGroups);
        for (uint32_t k = 0; k < ng; ++k) issue_group(k, (n_used + k) & (kRingGroups - 1));
    }
    __device__ __forceinline__ void advance() {  // current group exhausted: recycle it, wait for the next one
        const uint32_t gi = n_used & (kRingGroups - 1);
        // the group read before this one is free now (its values have been consumed by the ALU): refill it
        if (k_cur && k_cur - 1 + kRingGroups < n_groups) issue_group(k_cur - 1 + kRingGroups, (gi - 1) & (kRingGroups - 1));
        static_assert((kRingGroups & (kRingGroups - 1)) == 0, "ring groups: a power of two");
        mbar_wait(bars + gi * 8u, (n_used / kRingGroups) & 1u);   // ring group gi is on its (n_used / kRingGroups)-th use
        ++n_used;
        cur = ring + gi * (kGroupRows * kRowBytes) + lane_off;
        in_group = kGroupRows;
        ++k_cur;
    }
    __device__ __forceinline__ Row pop() {
        if (in_group == 0) advance();
        Row v;
#pragma unroll
        for (uint32_t h = 0; h < kRowHalves; ++h) v.h[h] = lds128(cur + h * 512u);
        cur += kRowBytes; --in_group;
        return v;
    }
    __device__ __forceinline__ void pop2(Row& a, Row& b) {   // two rows of the current group (in_group >= 2): one bookkeeping step
#pragma unroll
        for (uint32_t h = 0; h < kRowHalves; ++h) { a.h[h] = lds128(cur + h * 512u); b.h[h] = lds128(cur + kRowBytes + h * 512u); }
        cur += 2 * kRowBytes; in_group -= 2;
    }
};

template <int OP, class RR>
__device__ __forceinline__ Row ring_simple(RR& R, uint32_t n_plain, uint32_t n_comp) {
    Row acc = splat<Row>(OP == kDevAnd ? 0xFFFFFFFFu : 0u);
    uint32_t j = 0;
    while (j < n_plain) {
        if (R.in_group >= 2u && j + 2u <= n_plain) {   // two rows per trip while the group and the gate both have them
            Row a, b;
            R.pop2(a, b);
            acc = OP == kDevAnd ? and4(acc, and4(a, b)) : OP == kDevOr ? or4(acc, or4(a, b)) : xor4(acc, xor4(a, b));
            j += 2u;
        } else {
            const Row v = R.pop();
            acc = OP == kDevAnd ? and4(acc, v) : OP == kDevOr ? or4(acc, v) : xor4(acc, v);
            ++j;
        }
    }
    for (uint32_t j = 0; j < n_comp; ++j) {
        const Row v = xors4(R.pop(), 0xFFFFFFFFu);
        acc = OP == kDevAnd ? and4(acc, v) : OP == kDevOr ? or4(acc, v) : xor4(acc, v);
    }
    return acc;
}

// sum and carry of a full adder on every bit lane: one LOP3 each (0x96, 0xE8)
__device__ __forceinline__ uint4 xor3(uint4 a, uint4 b, uint4 c) { return make_uint4(a.x ^ b.x ^ c.x, a.y ^ b.y ^ c.y, a.z ^ b.z ^ c.z, a.w ^ b.w ^ c.w); }
__device__ __forceinline__ uint32_t maj1(uint32_t a, uint32_t b, uint32_t c) { return (a & b) | (c & (a | b)); }
__device__ __forceinline__ uint4 maj3(uint4 a, uint4 b, uint4 c) { return make_uint4(maj1(a.x, b.x, c.x), maj1(a.y, b.y, c.y), maj1(a.z, b.z, c.z), maj1(a.w, b.w, c.w)); }
__device__ __forceinline__ Row xor3(Row a, Row b, Row c) { Row r; _Pragma("unroll") for (uint32_t i = 0; i < kRowHalves; ++i) r.h[i] = xor3(a.h[i], b.h[i], c.h[i]); return r; }
__device__ __forceinline__ Row maj3(Row a, Row b, Row c) { Row r; _Pragma("unroll") for (uint32_t i = 0; i < kRowHalves; ++i) r.h[i] = maj3(a.h[i], b.h[i], c.h[i]); return r; }

// count-in-range over the next n rows of the ring with a bit-sliced counter of P planes.  Inputs are absorbed four at a time
// through carry-save adders -- two full adders into plane 0, one into plane 1, and only their carry ripples through planes
// 2 .. P-1 -- which costs 6 + 2 (P - 2) LOP3 per word for four inputs instead of 8 P with a ripple counter per input
// (P = 9: 5 instead of 18 per input); the up to three inputs left over ripple one by one.
template <int P, class RR>
__device__ __forceinline__ Row ring_range(RR& R, uint32_t n_plain, uint32_t n_comp, uint32_t lo, uint32_t hi) {
    Row pl[P];
#pragma unroll
    for (int p = 0; p < P; ++p) pl[p] = splat<Row>(0u);
    const uint32_t n = n_plain + n_comp;
    uint32_t j = 0;
    auto next = [&]() { const Row v = xors4(R.pop(), j < n_plain ? 0u : 0xFFFFFFFFu); ++j; return v; };
    while (j + 4u <= n) {
        Row a = next(), b = next();
        const Row tA = maj3(pl[0], a, b);
        pl[0] = xor3(pl[0], a, b);
        a = next(); b = next();
        const Row tB = maj3(pl[0], a, b);
        pl[0] = xor3(pl[0], a, b);
        Row carry = maj3(pl[1], tA, tB);
        pl[1] = xor3(pl[1], tA, tB);
#pragma unroll
        for (int p = 2; p < P; ++p) {
            const Row t = and4(pl[p], carry);
            pl[p] = xor4(pl[p], carry);
            carry = t;
        }
    }
    while (j < n) {
        Row carry = next();
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
__global__ void __launch_bounds__(kGateThreads, CANOPY_RING_MIN_BLOCKS)
gate_ring_kernel_t(const __grid_constant__ CUtensorMap vmap, uint4* __restrict__ vals, const uint4* __restrict__ grec,
                   const uint32_t* __restrict__ child,
                   const uint32_t* __restrict__ layer_off, const uint32_t* __restrict__ wide_off,
                   const uint32_t* __restrict__ split, const uint32_t* __restrict__ split_child,
                   uint32_t n_layers, uint32_t n_nodes, uint32_t n_ev, unsigned long long* __restrict__ tallies,
                   uint64_t first_slab, uint32_t n_slabs, uint64_t begin, uint64_t end,
                   uint32_t* __restrict__ walk_done, uint32_t parity, uint32_t wait_val, uint32_t qsel) {
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
    // Slabs are dealt from a queue (walk_done[2 + qsel]: the next slab of this batch; qsel = batch index & 1): CTAs that draw
    // cheap slabs or start early take more of them.  This walk clears the other queue for the batch after it: the walk
    // that used it has finished (this launch is an ordinary one), and its next user starts after this kernel.
    uint32_t* queue = walk_done + 2 + (qsel & 1u);
    __shared__ uint32_t s_slab;
    if (threadIdx.x == 0 && blockIdx.x == 0) walk_done[2 + ((qsel & 1u) ^ 1u)] = 0u;
    __syncthreads();
    // Pipelined run: the sampler of the next batch is launched as a programmatic dependent of this walk and needs
    // nothing from it, so it may start as soon as every CTA of this (persistent, single-wave) grid is resident.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    RowRing<G4> R;
    R.tmap = &vmap; R.row0 = 0;
    R.lane_off = lane * 16u; R.n_used = 0;
    R.ring = smem_u32(smem) + warp * (kRingSlots * kRowBytes);
    R.bars = smem_u32(bars) + warp * (kRingGroups * 8u);

    for (;;) {
        if (threadIdx.x == 0) s_slab = atomicAdd(queue, 1u);
        __syncthreads();
        const uint32_t slab = s_slab;
        __syncthreads();
        if (slab >= n_slabs) break;
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


}  // namespace canopy
