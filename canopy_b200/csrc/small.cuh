// Fused path for small models (north-star subsystems 2 + 3 in one kernel): when the node rows of one slab fit in shared
// memory (n_nodes x 1 KB <= kSmallMaxBytes; config c1 is 66 rows), one CTA samples every basic event of a slab straight
// into shared-memory rows, walks all layers there and tallies -- no value buffer, no second launch, and literally "no
// intermediate bitvector written back to HBM" (BASELINE.json north_star).  The bits are those of the streamed path: the same
// plane_group / sparse_unit_fill code, the same counters; only the destination of a row differs.
//
// Shared-memory rows: events first, in the SAMPLER's order (spec 4: dense events, then the classes k = 7 .. 13, so that the
// up-to-four rows of a sparse unit are adjacent and its image is filled in place), then the gates in layer order.  The
// host translates the child codes to these row indices once per plan (DeviceState::child_small).
#pragma once
#include "sampler.cuh"
#include "walk.cuh"

namespace canopy {

constexpr uint32_t kSmallWarps = 16;
constexpr uint32_t kSmallThreads = kSmallWarps * 32;
constexpr uint32_t kSmallMaxBytes = 200u * 1024u;   // rows of one slab; leaves room for static shared memory and the runtime's 1 KB

__device__ __forceinline__ Row small_row(uint32_t rows, uint32_t code, uint32_t lane) {   // child row with its edge complement applied
    const uint32_t addr = rows + (code & 0x7FFFFFFFu) * kRowBytes + lane * 16u;
    Row v;
#pragma unroll
    for (uint32_t h = 0; h < kRowHalves; ++h) v.h[h] = lds128(addr + h * 512u);
    return xors4(v, (uint32_t)((int32_t)code >> 31));
}

template <int P>
__device__ __forceinline__ Row small_range(uint32_t rows, const uint32_t* __restrict__ ch, uint32_t n, uint32_t lo, uint32_t hi, uint32_t lane) {
    Row pl[P];
#pragma unroll
    for (int p = 0; p < P; ++p) pl[p] = splat<Row>(0u);
    for (uint32_t j0 = 0; j0 < n; j0 += 32) {
        const uint32_t cnt = min(32u, n - j0);
        const uint32_t my = lane < cnt ? __ldg(ch + j0 + lane) : 0u;
        for (uint32_t j = 0; j < cnt; ++j) {
            Row carry = small_row(rows, __shfl_sync(0xFFFFFFFFu, my, j), lane);
#pragma unroll
            for (int p = 0; p < P; ++p) {
                const Row t = and4(pl[p], carry);
                pl[p] = xor4(pl[p], carry);
                carry = t;
            }
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

// CLASSED: spec 4 (ev_perm / cdf / C are valid); otherwise spec 2 (every event takes the planes, rows in device order).
template <int BITS, bool CLASSED>
__global__ void __launch_bounds__(kSmallThreads, 1)
small_model_kernel(const uint4* __restrict__ grec, const uint32_t* __restrict__ child_small, const uint32_t* __restrict__ layer_off,
                   const uint32_t* __restrict__ ev_orig, const uint32_t* __restrict__ thr32, const uint64_t* __restrict__ thr64,
                   const uint32_t* __restrict__ ev_perm, const uint64_t* __restrict__ cdf, const SparseClasses C, uint32_t n_units,
                   uint32_t n_ev, uint32_t n_layers, unsigned long long* __restrict__ tallies,
                   uint64_t first_slab, uint32_t n_slabs, uint64_t begin, uint64_t end, const PhiloxKeys K) {
    extern __shared__ __align__(128) unsigned char small_smem[];
    const uint32_t rows = smem_u32(small_smem);
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = bcast0(threadIdx.x >> 5);
    const uint32_t n_dense = CLASSED ? C.n_dense : n_ev;
    for (uint32_t slab = blockIdx.x; slab < n_slabs; slab += gridDim.x) {
        const uint64_t gslab = first_slab + slab;
        // ---- subsystem 2: every live basic event of this slab, into its shared-memory row --------------------------
        for (uint32_t i = warp; i < n_dense + n_units; i += kSmallWarps) {
            if (i < n_dense) {   // planes: this warp's event, lane = one 128-sample group of each 4096-sample half
                const uint32_t ev = CLASSED ? bcast0(__ldg(ev_perm + i)) : i;
                const uint32_t e = bcast0(__ldg(ev_orig + ev));
                const PlaneThr T = plane_thr<BITS>(BITS == 32 ? bcast0(__ldg(thr32 + ev)) : 0u, BITS == 64 ? __ldg(thr64 + ev) : 0ull);
#pragma unroll 1
                for (uint32_t h = 0; h < kRowHalves; ++h) {
                    const uint4 v = plane_group<BITS>((gslab * kRowHalves + h) * 32u + lane, e, T, K);
                    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(rows + i * kRowBytes + h * 512u + lane * 16u),
                                 "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
                }
            } else if (CLASSED) {   // a sparse unit: its rows are adjacent, the image is filled in place
                const SparseUnit U = sparse_unit_setup(ev_orig, ev_perm, i - n_dense, C);
                sparse_unit_fill(U, cdf, gslab, K, rows + U.first * kRowBytes);
            }
        }
        __syncthreads();
        // ---- subsystem 3: all layers over the shared-memory rows, fused tallies -------------------------------------
        Row mask;
#pragma unroll
        for (uint32_t h = 0; h < kRowHalves; ++h) {
            const uint64_t s0 = gslab * (uint64_t)kSlabSamples + h * 4096u + lane * 128u;
            mask.h[h] = make_uint4(valid_mask32(s0, begin, end), valid_mask32(s0 + 32, begin, end),
                                   valid_mask32(s0 + 64, begin, end), valid_mask32(s0 + 96, begin, end));
        }
        for (uint32_t l = 0; l < n_layers; ++l) {
            const uint32_t g1 = __ldg(layer_off + l + 1);
            for (uint32_t g = __ldg(layer_off + l) + warp; g < g1; g += kSmallWarps) {
                const uint4 rec = __ldg(grec + g);
                const uint32_t n = rec.y & 0xFFFFu, op = rec.z & 0xFFu, flags = rec.z >> 8;
                const uint32_t* ch = child_small + rec.x;
                Row r;
                if (op == kDevRange) {
                    const uint32_t lo = rec.w & 0xFFFFu, hi = rec.w >> 16;
                    if (n < 8) r = small_range<3>(rows, ch, n, lo, hi, lane);
                    else if (n < 64) r = small_range<6>(rows, ch, n, lo, hi, lane);
                    else r = small_range<9>(rows, ch, n, lo, hi, lane);   // n < 512: a slab of a small model has fewer rows than that
                } else {
                    r = splat<Row>(op == kDevAnd ? 0xFFFFFFFFu : 0u);
                    for (uint32_t j0 = 0; j0 < n; j0 += 32) {
                        const uint32_t cnt = min(32u, n - j0);
                        const uint32_t my = lane < cnt ? __ldg(ch + j0 + lane) : 0u;
                        for (uint32_t j = 0; j < cnt; ++j) {
                            const Row v = small_row(rows, __shfl_sync(0xFFFFFFFFu, my, j), lane);
                            r = op == kDevAnd ? and4(r, v) : op == kDevOr ? or4(r, v) : xor4(r, v);
                        }
                    }
                }
                if (flags & kFlagStore) {
#pragma unroll
                    for (uint32_t h = 0; h < kRowHalves; ++h)
                        asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(rows + (n_ev + g) * kRowBytes + h * 512u + lane * 16u),
                                     "r"(r.h[h].x), "r"(r.h[h].y), "r"(r.h[h].z), "r"(r.h[h].w) : "memory");
                }
                if (flags & kFlagTally) {
                    const uint32_t c = __reduce_add_sync(0xFFFFFFFFu, popc_masked(r, mask));
                    if (lane == 0 && c) atomicAdd(tallies + g, (unsigned long long)c);
                }
            }
            __syncthreads();   // the layer's rows are complete; after the last layer: every read of this slab's rows is over
        }
        if (n_layers == 0) __syncthreads();
    }
}

}  // namespace canopy
