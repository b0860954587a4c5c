// Monte Carlo engine, host side: per-device state, the value-buffer pool and the batch loop.  sm_100a only, no CPU
// fallback.  Kernels: sampler.cuh (subsystem 2), walk.cuh (subsystem 3); layout, spec and roofline accounting: DESIGN.md.
#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "sampler.cuh"
#include "small.cuh"
#include "walk.cuh"

namespace canopy {

#define CUDA_CHECK(expr)                                                                              \
    do {                                                                                              \
        cudaError_t _e = (expr);                                                                      \
        if (_e != cudaSuccess)                                                                        \
            throw Error(CANOPY_E_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e));           \
    } while (0)

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
    uint32_t* ev_perm = nullptr;   // spec 4 only: sampler order (dense events, then the classes k = 7 .. 13)
    uint64_t* ev_cdf = nullptr;    // spec 4 only: [n_events][CANOPY_SPARSE_CDF_LEN] count tables, built on the device
    uint4* grec = nullptr;
    uint8_t* tally_kind = nullptr;
    unsigned long long* gate_tallies = nullptr;
    unsigned long long* out_tallies = nullptr;   // used when the caller wants host output
    void* arena = nullptr;                       // one allocation holding every plan array above
    size_t arena_bytes = 0;
    uint4* vals[2] = {nullptr, nullptr};
    size_t vals_bytes[2] = {0, 0};  // capacity of each value buffer
    size_t vals_slabs = 0;  // capacity of each buffer in slabs
    int n_bufs = 0;
    cudaStream_t main = nullptr;
    cudaEvent_t ev_t0 = nullptr, ev_t1 = nullptr;
    uint64_t h2d_bytes = 0;
    std::mutex run_mu;  // one run at a time per device (the value buffers and streams are per-device scratch)
    // Completion of the last run that used this state's scratch (value buffers, gate tallies, thresholds).  A run may
    // return with its kernels still in flight (canopy_mc_run_device without a report): whatever touches the scratch
    // next from ANOTHER stream, or from the host, waits on this event first.
    cudaEvent_t ev_done = nullptr;
    cudaStream_t last_stream = nullptr;
    bool ran = false;
    uint32_t* child_small = nullptr; // fused path only: Lowered::child with node ids replaced by shared-memory row indices (small.cuh)
    uint32_t* walk_done = nullptr;   // [0..1] walk CTAs of even / odd batches that have finished (pipeline hazard guard); [2..3] slab queues
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
    size_t kMaxParkedPerDevice;
    explicit ValPool(size_t max_parked) : kMaxParkedPerDevice(max_parked) {}

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
static ValPool g_val_pool(2);
// The plan arrays of destroyed plans are parked the same way: cudaFree synchronises the device and was measured at anything
// between 0.6 and 250 ms per call next to 90 GB of live value buffers (it dominated the end-to-end time of a one-shot
// quantification), cudaMalloc at a few ms.
static ValPool g_arena_pool(4);

void trim_device_cache(int device) { g_val_pool.trim(device); g_arena_pool.trim(device); }

void destroy_device_state(DeviceState* s) {
    if (!s) return;
    cudaSetDevice(s->device);
    cudaDeviceSynchronize();  // nothing of this plan (on its own or on a caller's stream) may still run on a parked buffer
    g_arena_pool.release(s->device, s->arena, s->arena_bytes);
    for (int b = 0; b < 2; ++b) g_val_pool.release(s->device, s->vals[b], s->vals_bytes[b]);
    if (s->main) cudaStreamDestroy(s->main);
    for (cudaEvent_t e : {s->ev_t0, s->ev_t1, s->ev_done})
        if (e) cudaEventDestroy(e);
    delete s;
}

// Plan arrays go into ONE device allocation (256-byte aligned pieces): a dozen cudaMalloc calls cost more than the
// upload itself for the graphs of this path (c2: 0.3 MB, c4: 3.4 MB).
struct ArenaPlan {
    struct Piece { void** dst; const void* src; size_t bytes, off; };
    std::vector<Piece> pieces;
    size_t total = 0;
    size_t upload = 0;   // the arrays with host data come first: only [0, upload) is staged and copied
    template <class P, class T>
    void add(P*& dst, const std::vector<T>& v, size_t min_elems = 1) {
        const size_t bytes = std::max(v.size(), min_elems) * sizeof(T);
        pieces.push_back({reinterpret_cast<void**>(&dst), v.empty() ? nullptr : v.data(), bytes, total});
        pieces.back().bytes = v.size() * sizeof(T);  // bytes to copy; the reservation below may be larger
        total += (bytes + 255) & ~size_t(255);
        upload = total;
    }
    template <class P>
    void reserve(P*& dst, size_t bytes) {
        pieces.push_back({reinterpret_cast<void**>(&dst), nullptr, 0, total});
        total += (std::max<size_t>(bytes, 8) + 255) & ~size_t(255);
    }
};

// What the engine needs to know about a device, queried once per process: cudaGetDeviceProperties costs milliseconds
// per call, which a one-shot plan -> run -> destroy would pay every time.
struct DeviceInfo { bool ready = false; int major = 0, minor = 0, sm_count = 0, ring_ctas_per_sm = 0; };
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
#ifndef CANOPY_CARVEOUT
#define CANOPY_CARVEOUT 100
#endif
    // Kernels that are to share an SM (the walk of batch b and the sampler of batch b+1) must ask for the same shared-memory
    // carve-out: an SM cannot change its L1 / shared split while a CTA of the other kernel is resident on it.
    if (CANOPY_CARVEOUT >= 0) {
        CUDA_CHECK(cudaFuncSetAttribute(gate_ring_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, CANOPY_CARVEOUT));
        CUDA_CHECK(cudaFuncSetAttribute(gate_ring_kernel_g4, cudaFuncAttributePreferredSharedMemoryCarveout, CANOPY_CARVEOUT));
        CUDA_CHECK(cudaFuncSetAttribute(sampler_classed_kernel<32>, cudaFuncAttributePreferredSharedMemoryCarveout, CANOPY_CARVEOUT));
        CUDA_CHECK(cudaFuncSetAttribute(sampler_classed_kernel<64>, cudaFuncAttributePreferredSharedMemoryCarveout, CANOPY_CARVEOUT));
        CUDA_CHECK(cudaFuncSetAttribute(sampler_sliced_kernel<32>, cudaFuncAttributePreferredSharedMemoryCarveout, CANOPY_CARVEOUT));
        CUDA_CHECK(cudaFuncSetAttribute(sampler_sliced_kernel<64>, cudaFuncAttributePreferredSharedMemoryCarveout, CANOPY_CARVEOUT));
    }
    CUDA_CHECK(cudaFuncSetAttribute(small_model_kernel<32, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmallMaxBytes));
    CUDA_CHECK(cudaFuncSetAttribute(small_model_kernel<64, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmallMaxBytes));
    CUDA_CHECK(cudaFuncSetAttribute(small_model_kernel<32, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmallMaxBytes));
    CUDA_CHECK(cudaFuncSetAttribute(small_model_kernel<64, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmallMaxBytes));
    CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&I.ring_ctas_per_sm, gate_ring_kernel, kGateThreads, kRingSmemBytes));
    if (I.ring_ctas_per_sm < 1) throw Error(CANOPY_E_CUDA, "gate_ring_kernel does not fit on an SM");
    I.ready = true;
    return I;
}

// Fused path (small.cuh): may runs of this plan use it, and do they?
bool small_eligible(const Lowered& L) {
    return L.n_gates() > 0 && (size_t)L.n_nodes() * kRowBytes <= kSmallMaxBytes && L.opts.sampler != CANOPY_SAMPLER_DIRECT;
}
static bool use_small(const Lowered& L) { return small_eligible(L) && L.opts.engine != CANOPY_ENGINE_STREAMED; }
// Child codes of the fused kernel: an event's row index is its position in the sampler's order (spec 4), a gate keeps its
// device node id.  Follows the thresholds (the order does), so it is rebuilt with them.
static std::vector<uint32_t> small_children(const Lowered& L) {
    const uint32_t n_ev = L.n_events();
    std::vector<uint32_t> pos(n_ev);
    for (uint32_t i = 0; i < n_ev; ++i) pos[L.event_perm.empty() ? i : L.event_perm[i]] = i;
    std::vector<uint32_t> out(L.dev_child.size());   // same offsets as dev_rec.child_off
    for (size_t c = 0; c < L.dev_child.size(); ++c) {
        const uint32_t code = L.dev_child[c], node = code & 0x7FFFFFFFu;
        out[c] = (node < n_ev ? pos[node] : node) | (code & 0x80000000u);
    }
    return out;
}

// Spec 4: (re)build the count tables of the sparse events on the device from the thresholds already uploaded.
static void build_tables(DeviceState* s, const Lowered& L) {
    if (L.event_perm.empty() || L.n_events() == 0) return;
    sparse_table_kernel<<<(L.n_events() + 127) / 128, 128, 0, s->main>>>(s->thr64, L.n_events(), s->ev_cdf);
    CUDA_CHECK(cudaGetLastError());
    CUDA_CHECK(cudaStreamSynchronize(s->main));
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
        A.add(s->child, L.dev_child); A.add(s->layer_off, L.layer_off); A.add(s->tally_gate, L.tally_gate);
        A.add(s->wide_off, L.wide_off); A.add(s->split, L.split); A.add(s->split_child, L.split_child);
        A.add(s->tally_kind, L.tally_kind); A.add(s->grec, L.dev_rec);
        A.add(s->ev_perm, L.event_perm);
        const std::vector<uint32_t> csmall = small_eligible(L) ? small_children(L) : std::vector<uint32_t>();
        A.add(s->child_small, csmall);
        if (!L.event_perm.empty()) A.reserve(s->ev_cdf, (size_t)L.n_events() * CANOPY_SPARSE_CDF_LEN * 8);
        A.reserve(s->gate_tallies, (std::max<size_t>(L.n_gates(), 1) + 2) * 8);  // + walk_done[4]
        A.reserve(s->out_tallies, std::max<size_t>(L.tally_kind.size(), 1) * 8);
        s->arena = g_arena_pool.acquire(device, A.total, &s->arena_bytes);
        if (!s->arena) throw Error(CANOPY_E_OOM, "cudaMalloc failed for plan arrays");
        {   // one staged upload of all plan arrays instead of a dozen small synchronous copies
            std::vector<char> stage(A.upload);
            for (auto& p : A.pieces) {
                *p.dst = static_cast<char*>(s->arena) + p.off;
                if (p.bytes) std::memcpy(stage.data() + p.off, p.src, p.bytes);
                s->h2d_bytes += p.bytes;
            }
            CUDA_CHECK(cudaMemcpy(s->arena, stage.data(), A.upload, cudaMemcpyHostToDevice));
        }
        CUDA_CHECK(cudaStreamCreateWithFlags(&s->main, cudaStreamNonBlocking));
        CUDA_CHECK(cudaEventCreateWithFlags(&s->ev_done, cudaEventDisableTiming));
        CUDA_CHECK(cudaEventCreate(&s->ev_t0)); CUDA_CHECK(cudaEventCreate(&s->ev_t1));
        s->walk_done = reinterpret_cast<uint32_t*>(s->gate_tallies + std::max<size_t>(L.n_gates(), 1));
        build_tables(s, L);
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
    const size_t rows = bytes / kRowBytes;
    if (!encode || rows == 0 || rows > 0x7FFFFFFFull) return false;
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

// `pdl`: launch as a programmatic dependent of the kernel before it in the stream (the gate walk of the previous
// batch): the sampler starts once that walk's CTAs are resident instead of when the walk has finished.
static void launch_sampler(const DeviceState* s, const Lowered& L, uint4* vals, uint64_t first_slab, uint32_t n_slabs,
                           const PhiloxKeys& K, cudaStream_t st, bool pdl = false) {
    const uint32_t n_ev = L.n_events(), n_nodes = L.n_nodes();
    if (n_ev == 0 || n_slabs == 0) return;
    const uint32_t n_blk = n_slabs * kRowHalves;  // 4096-sample blocks
    const bool wide = L.opts.threshold_bits == 64;
    if (L.opts.sampler == CANOPY_SAMPLER_DIRECT) {
        const dim3 grid((n_ev + 7) / 8, n_blk);
        if (wide) sampler_direct_kernel<64><<<grid, 256, 0, st>>>(vals, s->ev_orig, s->thr32, s->thr64, n_ev, n_nodes, first_slab, K);
        else sampler_direct_kernel<32><<<grid, 256, 0, st>>>(vals, s->ev_orig, s->thr32, s->thr64, n_ev, n_nodes, first_slab, K);
        CUDA_CHECK(cudaGetLastError());
        return;
    }
    cudaLaunchConfig_t cfg{};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
    cfg.stream = st;
    const uint32_t grid_y = (n_blk + kBlocksPerWarp - 1) / kBlocksPerWarp;  // each warp walks up to 8 blocks (4 slabs) of its work item
    if (L.opts.sampler == CANOPY_SAMPLER_SLICED) {
        cfg.gridDim = dim3((n_ev + 7) / 8, grid_y); cfg.blockDim = dim3(256);
        auto kern = wide ? sampler_sliced_kernel<64> : sampler_sliced_kernel<32>;
        CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, vals, (const uint32_t*)s->ev_orig, (const uint32_t*)s->thr32,
                                      (const uint64_t*)s->thr64, n_ev, n_nodes, first_slab, n_slabs, K));
        return;
    }
    // spec 4: dense CTAs (one warp per dense event), then sparse CTAs (one warp per unit of kUnitRows events of one class)
    SparseClasses C{};
    C.n_dense = L.n_dense;
    uint32_t units = 0;
    for (int c = 0; c < 8; ++c) {
        C.ev_off[c] = L.class_off[c];
        C.unit_off[c] = units;
        if (c < 7) units += (L.class_off[c + 1] - L.class_off[c] + kUnitRows - 1) / kUnitRows;
    }
    const uint32_t dense_ctas = (L.n_dense + kClassedWarps - 1) / kClassedWarps;
    cfg.gridDim = dim3(dense_ctas + (units + kClassedWarps - 1) / kClassedWarps, grid_y);
    cfg.blockDim = dim3(kClassedWarps * 32);
    cfg.dynamicSmemBytes = kClassedSmemBytes;
    auto kern = wide ? sampler_classed_kernel<64> : sampler_classed_kernel<32>;
    CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, vals, (const uint32_t*)s->ev_orig, (const uint32_t*)s->thr32, (const uint64_t*)s->thr64,
                                  (const uint32_t*)s->ev_perm, (const uint64_t*)s->ev_cdf, C, dense_ctas, units, n_nodes,
                                  first_slab, n_slabs, K));
}

// Test hook (canopy-testing.h, exported by libcanopy_testing.so only): take the fallback the walk uses when no tensor map
// can be had (value buffers of more than 2^31 rows, or a driver without cuTensorMapEncodeTiled) on buffers of any size, so
// that the parity tests reach gate_ring_kernel_t<false>.
static std::atomic<bool> g_force_row_copies{false};
void debug_force_row_copies(bool on) { g_force_row_copies.store(on); }

#ifndef CANOPY_WALK_SHARED_CTAS
#define CANOPY_WALK_SHARED_CTAS 1   // walk CTAs per SM while the sampler of the next batch shares the GPU
#endif
#ifndef CANOPY_WALK_ALONE_CTAS
#define CANOPY_WALK_ALONE_CTAS 8    // ... and when the walk has the GPU to itself (capped by its occupancy)
#endif
// Returns the number of CTAs launched (what the next walk waits for in walk_done[parity]).  `qsel`: slab queue of this
// batch (batch index & 1).
static uint32_t launch_gates(const DeviceState* s, const Lowered& L, uint4* vals, uint64_t first_slab, uint32_t n_slabs,
                             uint64_t begin, uint64_t end, cudaStream_t st, bool shared_sm = false, uint32_t parity = 0,
                             uint32_t wait_val = 0, uint32_t qsel = 0) {
    if (L.n_gates() == 0 || n_slabs == 0) return 0;
    // persistent: one CTA per resident slot, each takes slabs from the batch's queue; beside the sampler of the
    // next batch the walk keeps one CTA per SM (64 KB of row rings in flight), which measured best on B200
    const int per_sm = std::min(shared_sm ? CANOPY_WALK_SHARED_CTAS : CANOPY_WALK_ALONE_CTAS, s->ring_ctas_per_sm);
    const uint32_t grid = std::min<uint32_t>(n_slabs, (uint32_t)(s->sm_count * per_sm));
    const int b = vals == s->vals[1] ? 1 : 0;
    if (vals == s->vals[b] && s->vmap_ok[b] && kGroupRows == 4 && !g_force_row_copies.load(std::memory_order_relaxed))
        gate_ring_kernel_g4<<<grid, kGateThreads, kRingSmemBytes, st>>>(
            s->vmap[b], vals, s->grec, s->child, s->layer_off, s->wide_off, s->split, s->split_child, L.n_layers(), L.n_nodes(),
            L.n_events(), s->gate_tallies, first_slab, n_slabs, begin, end, s->walk_done, parity, wait_val, qsel);
    else   // value buffers of more than 2^31 rows: one bulk copy per row instead of tile::gather4
        gate_ring_kernel<<<grid, kGateThreads, kRingSmemBytes, st>>>(
            s->vmap[0], vals, s->grec, s->child, s->layer_off, s->wide_off, s->split, s->split_child, L.n_layers(), L.n_nodes(),
            L.n_events(), s->gate_tallies, first_slab, n_slabs, begin, end, s->walk_done, parity, wait_val, qsel);
    CUDA_CHECK(cudaGetLastError());
    return grid;
}

// Fused path: the whole sample range in one launch, one CTA per resident slot striding over the slabs.
static void launch_small(const DeviceState* s, const Lowered& L, uint64_t first_slab, uint64_t n_slabs, uint64_t begin, uint64_t end,
                         const PhiloxKeys& K, cudaStream_t st) {
    if (n_slabs > 0xFFFFFFFFull) throw Error(CANOPY_E_INVALID, "too many slabs for one run; split the sample range");
    const bool classed = L.opts.sampler == CANOPY_SAMPLER_SPARSE, wide = L.opts.threshold_bits == 64;
    SparseClasses C{};
    uint32_t units = 0;
    if (classed) {
        C.n_dense = L.n_dense;
        for (int c = 0; c < 8; ++c) {
            C.ev_off[c] = L.class_off[c];
            C.unit_off[c] = units;
            if (c < 7) units += (L.class_off[c + 1] - L.class_off[c] + kUnitRows - 1) / kUnitRows;
        }
    }
    const uint32_t smem = L.n_nodes() * kRowBytes;
    auto kern = classed ? (wide ? small_model_kernel<64, true> : small_model_kernel<32, true>)
                        : (wide ? small_model_kernel<64, false> : small_model_kernel<32, false>);
    int per_sm = 1;
    CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, (int)kSmallThreads, smem));
    const uint32_t grid = (uint32_t)std::min<uint64_t>(n_slabs, (uint64_t)s->sm_count * std::max(per_sm, 1));
    kern<<<grid, kSmallThreads, smem, st>>>(s->grec, s->child_small, s->layer_off, s->ev_orig, s->thr32, s->thr64, s->ev_perm, s->ev_cdf,
                                           C, units, L.n_events(), L.n_layers(), s->gate_tallies, first_slab, (uint32_t)n_slabs,
                                           begin, end, K);
    CUDA_CHECK(cudaGetLastError());
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

void mc_run(canopy_plan* plan, const canopy_mc_opts& o, uint64_t* tallies_host, uint64_t* tallies_dev, canopy_mc_report* rep) {
    std::shared_lock<std::shared_mutex> low_lock(plan->low_mu);   // no re-quantising while this run reads the plan (plan.h)
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

    const bool fused = use_small(L);
    uint32_t batch = 1; uint64_t n_batches = 0;
    std::vector<uint32_t> sched;  // slabs per batch, in order
    if (fused) { overlap = false; batch = (uint32_t)std::min<uint64_t>(total_slabs, 0xFFFFFFFFull); n_batches = total_slabs ? 1 : 0; }
    else if (total_slabs && L.n_gates()) {
        batch = pick_batch_slabs(s, L, o, total_slabs, overlap ? 2 : 1);
        if ((total_slabs + batch - 1) / batch < 2) overlap = false;
        for (uint64_t done = 0; done < total_slabs; done += batch) sched.push_back((uint32_t)std::min<uint64_t>(batch, total_slabs - done));
        n_batches = sched.size();
        ensure_vals(s, L, batch, overlap ? 2 : 1);
    }
    const PhiloxKeys K = make_keys(o.seed);
    canopy_mc_report R{};
    R.slab_samples = kSlabSamples; R.batch_slabs = batch; R.n_batches = n_batches;

    order_after_last_run(s, user);
    CUDA_CHECK(cudaEventRecord(s->ev_t0, user));
    CUDA_CHECK(cudaMemsetAsync(s->gate_tallies, 0, (std::max<size_t>(L.n_gates(), 1) + 2) * 8, user));  // tallies + walk_done[4]
    R.other_launches += 1;
    struct EventList {  // profile events, 3 per batch; destroyed on every path out of this function
        std::vector<cudaEvent_t> v;
        ~EventList() { for (cudaEvent_t e : v) cudaEventDestroy(e); }
        size_t size() const { return v.size(); }
        cudaEvent_t operator[](size_t i) const { return v[i]; }
        void push_back(cudaEvent_t e) { v.push_back(e); }
    } pe;
    if (fused) {   // sampler, walk and tallies of the whole range in one kernel over shared-memory rows (small.cuh)
        cudaEvent_t e[2] = {nullptr, nullptr};
        if (profile) { for (auto& x : e) { CUDA_CHECK(cudaEventCreate(&x)); pe.push_back(x); } CUDA_CHECK(cudaEventRecord(e[0], user)); }
        if (total_slabs) { launch_small(s, L, slab0, total_slabs, begin, end, K, user); R.gate_launches += 1; }
        if (profile) CUDA_CHECK(cudaEventRecord(e[1], user));
    } else if (!overlap) {
        uint64_t fs = slab0;
        for (uint64_t b = 0; b < n_batches; fs += sched[b], ++b) {
            const uint32_t ns = sched[b];
            if (profile) { cudaEvent_t e[3]; for (auto& x : e) { CUDA_CHECK(cudaEventCreate(&x)); pe.push_back(x); } CUDA_CHECK(cudaEventRecord(e[0], user)); }
            launch_sampler(s, L, s->vals[0], fs, ns, K, user);
            if (profile) CUDA_CHECK(cudaEventRecord(pe[pe.size() - 2], user));
            launch_gates(s, L, s->vals[0], fs, ns, begin, end, user, profile && o.profile_walk_shape == 1, 0, 0, (uint32_t)(b & 1));
            if (profile) CUDA_CHECK(cudaEventRecord(pe[pe.size() - 1], user));
            R.sampler_launches += L.n_events() ? 1 : 0; R.gate_launches += 1;
        }
    } else {
        // The sampler of batch b+1 runs beside the gate walk of batch b, on double-buffered value rows.
        // One stream, programmatic dependent launch: S0, G0, S1*, G1, S2*, ...  (* = dependent of the walk before it).
        // G_b is an ordinary launch, so it starts when S_b and everything before it have finished; S_{b+1} starts as
        // soon as the CTAs of G_b (one per SM, all resident at once on the idle GPU) have signalled, fills the rest of
        // every SM and keeps it full when the walk ends.  The order in which the two kernels get their SM slots is fixed
        // by construction (with two streams it depended on how the driver mapped the streams to hardware queues).
        // Hazard: S_{b+2} rewrites the buffer G_b read; it starts after G_{b+1} released it, which G_{b+1} does only
        // once G_b has counted all its CTAs out (walk_done; see the kernel).
        uint64_t fs = slab0;
        uint32_t done[2] = {0, 0};  // walk CTAs launched so far on even / odd batches
        for (uint64_t b = 0; b < n_batches; fs += sched[b], ++b) {
            const int buf = (int)(b & 1);
            launch_sampler(s, L, s->vals[buf], fs, sched[b], K, user, b > 0);
            done[buf] += launch_gates(s, L, s->vals[buf], fs, sched[b], begin, end, user, b + 1 < n_batches, (uint32_t)buf,
                                      done[buf ^ 1], (uint32_t)buf);
            R.sampler_launches += L.n_events() ? 1 : 0; R.gate_launches += 1;
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
        if (fused && pe.size() == 2) { float a = 0; CUDA_CHECK(cudaEventElapsedTime(&a, pe[0], pe[1])); R.gate_ms = a; }   // one fused kernel: reported as the walk
        for (size_t i = 0; !fused && i + 3 <= pe.size(); i += 3) {
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
    std::shared_lock<std::shared_mutex> low_lock(plan->low_mu);
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
        if (!L.event_perm.empty()) {   // spec 4: sampler order and tables follow the thresholds (sizes do not change)
            CUDA_CHECK(cudaMemcpy(s->ev_perm, L.event_perm.data(), L.event_perm.size() * 4, cudaMemcpyHostToDevice));
            if (small_eligible(L) && !L.dev_child.empty()) {   // the fused path's row indices follow the sampler's order
                const std::vector<uint32_t> cs = small_children(L);
                CUDA_CHECK(cudaMemcpy(s->child_small, cs.data(), cs.size() * 4, cudaMemcpyHostToDevice));
            }
            build_tables(s, L);
        }
    }
}

// canopy-testing.h: overwrite the count table of one sparse event on every device the plan is resident on (and make it
// resident on device 0 if it is on none yet); the next refresh_thresholds() restores the real tables.
void debug_set_sparse_table(canopy_plan* plan, uint32_t live_event, const uint64_t* cdf) {
    if (plan->dev.empty()) get_state(plan, 0);
    for (DeviceState* s : plan->dev) {
        std::lock_guard<std::mutex> run_lock(s->run_mu);
        CUDA_CHECK(cudaSetDevice(s->device));
        wait_last_run(s);
        CUDA_CHECK(cudaMemcpy(s->ev_cdf + (size_t)live_event * CANOPY_SPARSE_CDF_LEN, cdf, CANOPY_SPARSE_CDF_LEN * 8, cudaMemcpyHostToDevice));
    }
}

// ================================================================================================
// canopy_mc_run_static (include/canopy.h): the fused path for callers that must not allocate device memory.
//
// Who that is: a library like this one is often loaded INTO somebody else's process -- a plant-monitoring service, a
// solver, a framework with its own caching allocator -- that claimed the GPU's memory for its own pool at start-up, or
// that runs under a memory quota (MIG slice, container limit) it has already committed.  A cudaMalloc from a plug-in then
// fails or eats into a budget it was never given.  Others simply forbid cudaMalloc / cudaFree in their steady-state loop,
// because both synchronise the whole device and stall every other stream for the duration.  canopy_mc_run() is no use to
// them on its first run of a plan: that is when it makes the plan resident (cudaMalloc of the plan arena, stream and events).
//
// What this path does instead: everything the fused kernel needs in global memory beyond the caller's output -- the device
// copy of the plan arrays, the count tables of spec 4, the per-gate tally counters, the mapped tallies -- lives in ONE
// file-scope array that is part of the library's device image.  Every GPU that loads libcanopy holds those
// CANOPY_STATIC_SCRATCH_BYTES until the process exits, whether or not this entry point is ever called.  Nothing here calls
// cudaMalloc, cudaFree, cudaStreamCreate or cudaEventCreate.
//
// Consequences the caller has to know (repeated in canopy.h):
//   - one array per device, shared by every call and every plan: a call overwrites what the previous one left, so the plan
//     image is uploaded and the counters are cleared on EVERY call, and two calls whose GPU work overlaps (different
//     streams, or device-output calls from two threads) corrupt each other.  The host side is serialised by a mutex per
//     device; ordering the GPU work is the caller's job.
//   - the first call on a device resolves the array's device address (below) and sets the kernels' attributes; make that
//     call before starting any stream capture.
// ================================================================================================
static __device__ __align__(256) unsigned char g_static_scratch[CANOPY_STATIC_SCRATCH_BYTES];

struct StaticDev {
    std::mutex mu;                    // serialises the host side of canopy_mc_run_static per device
    bool ready = false;
    unsigned char* base = nullptr;    // DEVICE address of g_static_scratch on this device
};
static StaticDev g_static_dev[64];

void mc_run_static(canopy_plan* plan, const canopy_mc_opts& o, uint64_t* tallies_host, uint64_t* tallies_dev) {
    const DeviceInfo info = device_info(o.device);   // ordinal and sm_100 checks; kernel attributes, once per device
    CUDA_CHECK(cudaSetDevice(o.device));
    StaticDev& D = g_static_dev[o.device];
    std::lock_guard<std::mutex> dev_lock(D.mu);
    if (!D.ready) {
        // In host code `g_static_scratch` is only a handle for the symbol, not somewhere a kernel can write: ask the
        // runtime where the array lives on the device that is current now.
        void* p = nullptr;
        CUDA_CHECK(cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p || (reinterpret_cast<uintptr_t>(p) & 255u)) throw Error(CANOPY_E_CUDA, "static scratch array has no usable device address");
        D.base = static_cast<unsigned char*>(p);
        D.ready = true;
    }
    if (o.sample_begin + o.n_samples < o.sample_begin) throw Error(CANOPY_E_INVALID, "sample range overflows 64 bits");
    const uint64_t begin = o.sample_begin, end = o.sample_begin + o.n_samples;
    const uint64_t slab0 = begin / kSlabSamples;
    const uint64_t total_slabs = o.n_samples ? (end - 1) / kSlabSamples - slab0 + 1 : 0;
    cudaStream_t st = o.stream ? (cudaStream_t)o.stream : cudaStreamPerThread;   // no library-owned stream: that would be a create

    std::shared_lock<std::shared_mutex> low_lock(plan->low_mu);   // the thresholds may be re-quantised by canopy_plan_update_p
    const Lowered& L = plan->low;
    if (!small_eligible(L))
        throw Error(CANOPY_E_UNSUPPORTED, "canopy_mc_run_static runs the fused kernel only: the plan needs at least one live gate, at most " +
                                              std::to_string(kSmallMaxBytes / kRowBytes) + " live nodes (it has " + std::to_string(L.n_nodes()) +
                                              ") and a sliced or classed sampler");
    // Same pieces, same order and alignment as the arena of get_state(), laid over the static array.
    DeviceState v;   // a view: pointers into the static array, nothing owned
    v.device = o.device; v.sm_count = info.sm_count;
    ArenaPlan A;
    A.add(v.ev_orig, L.event_orig); A.add(v.thr32, L.event_thr32); A.add(v.thr64, L.event_thr64);
    A.add(v.layer_off, L.layer_off); A.add(v.tally_gate, L.tally_gate); A.add(v.tally_kind, L.tally_kind);
    A.add(v.grec, L.dev_rec); A.add(v.ev_perm, L.event_perm);
    const std::vector<uint32_t> csmall = small_children(L);
    A.add(v.child_small, csmall);
    if (!L.event_perm.empty()) A.reserve(v.ev_cdf, (size_t)L.n_events() * CANOPY_SPARSE_CDF_LEN * 8);
    const size_t tally_bytes = (size_t)L.n_gates() * 8;
    A.reserve(v.gate_tallies, tally_bytes);
    A.reserve(v.out_tallies, std::max<size_t>(L.tally_kind.size(), 1) * 8);
    if (A.total > (size_t)CANOPY_STATIC_SCRATCH_BYTES)
        throw Error(CANOPY_E_UNSUPPORTED, "canopy_mc_run_static: this plan needs " + std::to_string(A.total) + " bytes of device scratch, the static "
                                              "array holds CANOPY_STATIC_SCRATCH_BYTES = " + std::to_string((size_t)CANOPY_STATIC_SCRATCH_BYTES) +
                                              "; use canopy_mc_run");
    std::vector<char> stage(A.upload);
    for (auto& p : A.pieces) {
        *p.dst = D.base + p.off;
        if (p.bytes) std::memcpy(stage.data() + p.off, p.src, p.bytes);
    }
    // Pageable source: the call returns once the bytes are in the driver's staging buffer, so `stage` may go out of scope.
    CUDA_CHECK(cudaMemcpyAsync(D.base, stage.data(), A.upload, cudaMemcpyHostToDevice, st));
    CUDA_CHECK(cudaMemsetAsync(v.gate_tallies, 0, tally_bytes, st));   // the kernel adds into these; the last call's counts are still there
    if (!L.event_perm.empty() && L.n_events()) {
        sparse_table_kernel<<<(L.n_events() + 127) / 128, 128, 0, st>>>(v.thr64, L.n_events(), v.ev_cdf);
        CUDA_CHECK(cudaGetLastError());
    }
    if (total_slabs) launch_small(&v, L, slab0, total_slabs, begin, end, make_keys(o.seed), st);
    const uint32_t n_tallies = (uint32_t)L.tally_kind.size();
    unsigned long long* d_out = tallies_dev ? (unsigned long long*)tallies_dev : v.out_tallies;
    if (n_tallies) {
        finalize_kernel<<<(n_tallies + 255) / 256, 256, 0, st>>>(d_out, v.gate_tallies, v.tally_kind, v.tally_gate, n_tallies,
                                                                (unsigned long long)o.n_samples);
        CUDA_CHECK(cudaGetLastError());
    }
    if (tallies_host) {
        if (n_tallies) CUDA_CHECK(cudaMemcpyAsync(tallies_host, d_out, (size_t)n_tallies * 8, cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaStreamSynchronize(st));
    }
}

}  // namespace canopy
