// Internal structures of libcanopy: the lowered (device-ready) form of a MEF gate DAG.
// Public boundary: include/canopy.h.  Nothing here is visible to callers.
#pragma once
#include <cstdint>
#include <mutex>
#include <shared_mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "canopy.h"

namespace canopy {

// Error carrying a status code; translated to canopy_status + thread-local message at the C boundary.
// Mirrors the reference's exception hierarchy only by code (reference src/mef/openpsa/error.h:23-96).
struct Error : std::runtime_error {
    canopy_status code;
    Error(canopy_status c, const std::string& m) : std::runtime_error(m), code(c) {}
};

// Device connectives after canonicalisation. Every MEF connective (event.h:143-157) lowers to one of
// these plus edge complements: nand/nor/iff become a complemented reference to and/or/xor,
// imply(a,b) = or(!a,b), not/null become aliases, atleast/cardinality become kRange[lo,hi].
enum DevOp : uint8_t { kDevAnd = 0, kDevOr = 1, kDevXor = 2, kDevRange = 3 };
enum GateFlags : uint8_t { kFlagInvert = 1, kFlagStore = 2, kFlagTally = 4 };
enum TallyKind : uint8_t { kTallyFalse = 0, kTallyTrue = 1, kTallyGate = 2, kTallyNotGate = 3 };

// Geometry of the value buffer. A slab is a tile of kSlabSamples consecutive samples; one node row inside a slab is
// kRowHalves blocks of 512 bytes = 32 lanes x 16 bytes (4096 samples each), so one warp moves one block with one
// 128-bit access per lane.  kRowHalves = 2 (8192-sample slabs, 1 KB rows) halves the per-row bookkeeping of the gate
// walk per byte moved: one bulk copy, one ring pop and one trip through the gate loop per KB.
#ifndef CANOPY_ROW_HALVES
#define CANOPY_ROW_HALVES 2
#endif
constexpr uint32_t kRowHalves = CANOPY_ROW_HALVES;   // 1 or 2
constexpr uint32_t kSlabWords = 64 * kRowHalves;
constexpr uint32_t kSlabSamples = kSlabWords * 64;   // 4096 * kRowHalves
constexpr uint32_t kRowBytes = kSlabWords * 8;       // 512 * kRowHalves
constexpr uint32_t kRowVec = 32 * kRowHalves;        // uint4 per row
constexpr uint32_t kShardAlign = 4096;               // canopy_shard_range cuts on multiples of this (API contract)
static_assert(kRowHalves == 1 || kRowHalves == 2, "CANOPY_ROW_HALVES must be 1 or 2");

// Gate kernel geometry fixed at lowering time.
#ifndef CANOPY_GATE_WARPS
#define CANOPY_GATE_WARPS 8
#endif
constexpr uint32_t kGateWarps = CANOPY_GATE_WARPS;     // warps of the CTA that walks one slab; the per-layer split is cut for this
constexpr uint32_t kWideFanin = 512;   // count gates with at least this fan-in take the multi-pass register path

struct GateRec {           // 16 bytes, loaded as one uint4 by the gate kernel
    uint32_t child_off;    // into Lowered::child
    uint32_t n_children;   // fan-in | (number of complemented arguments, which come last) << 16
    uint32_t op_flags;     // op | flags << 8
    uint32_t krange;       // lo | hi << 16   (kDevRange only)
};

struct Lowered {
    canopy_plan_opts opts{};
    canopy_plan_stats stats{};
    // live basic events, ascending original index; device node id == position
    std::vector<uint32_t> event_orig;
    std::vector<uint32_t> event_thr32;
    std::vector<uint64_t> event_thr64;
    // spec 4: group exponent k of each live event (0 = dense, spec 2 planes; 7..13 = sparse with 2^k-sample groups), the
    // order in which the sampler takes the events (dense first, then k = 7, 8, ... 13: the expensive work is issued first
    // and the events of one unit share a class), and the class boundaries inside that order.  The count tables are
    // built on the device from thr64 (sparse_table.h).  All empty / zero under specs 1 and 2.
    std::vector<uint8_t> event_k;
    std::vector<uint32_t> event_perm;
    uint32_t n_dense = 0;                  // event_perm[0, n_dense) are the dense events
    uint32_t class_off[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // class c (k = 7 + c) is event_perm[class_off[c], class_off[c+1])
    std::vector<uint8_t> event_class;      // [n_basic_in] how lowering classed each ORIGINAL event: 0 false, 1 true, 2 sampled/dead
    // live gates in layer order; device node id == n_events + position
    std::vector<GateRec> gate_rec;
    std::vector<uint32_t> gate_child_off;  // [n_gates+1] (same offsets as GateRec, kept for the view)
    std::vector<uint32_t> child;           // device node id | complement << 31
    std::vector<uint8_t> gate_op, gate_flags;
    std::vector<uint16_t> gate_kmin, gate_kmax;
    std::vector<uint32_t> layer_off;       // [n_layers+1]
    std::vector<uint32_t> wide_off;        // [n_layers] first wide count gate of the layer (== layer end if none)
    std::vector<uint32_t> split;           // [n_layers][kGateWarps+1] equal-cost chunks of the ring-path gates
    std::vector<uint32_t> split_child;     // [n_layers][kGateWarps+1] run w of the layer is dev_child[split_child[w], split_child[w+1])
    // What the kernels read: gate_rec with child_off pointing into dev_child, and the child codes laid out per layer and
    // per warp run, every run padded to a multiple of four codes (lowering.cpp).  `child` / `gate_rec` above stay the
    // canonical form (plan view, fingerprint).
    std::vector<GateRec> dev_rec;
    std::vector<uint32_t> dev_child;
    // output mapping: user-visible tallies -> device gate tallies
    std::vector<uint8_t> tally_kind;
    std::vector<uint32_t> tally_gate;
    uint32_t n_events() const { return (uint32_t)event_orig.size(); }
    uint32_t n_gates() const { return (uint32_t)gate_rec.size(); }
    uint32_t n_nodes() const { return n_events() + n_gates(); }
    uint32_t n_layers() const { return layer_off.empty() ? 0 : (uint32_t)layer_off.size() - 1; }
};

void lower_graph(const canopy_graph_desc& d, const canopy_plan_opts& opts, Lowered& out);
void lower_graph_cached(const canopy_graph_desc& d, const canopy_plan_opts& opts, Lowered& out);  // through the plan cache
void plan_cache_clear();
// Re-quantise the thresholds of the live events from new probabilities (canopy_plan_update_p); throws if the
// constant folding of the existing lowering would no longer hold.
void requantise(Lowered& L, const double* p, uint32_t n_basic);
void build_sparse_classes(Lowered& L);   // spec 4: event_k, event_perm, n_dense, class_off from event_thr64
uint64_t fingerprint(const Lowered& L);

struct DeviceState;  // engine.cu
void destroy_device_state(DeviceState*);

}  // namespace canopy

struct canopy_plan {
    canopy::Lowered low;
    std::mutex mu;  // serialises runs on one plan (per-device scratch is owned by the plan)
    // Guards the parts of `low` that canopy_plan_update_p / canopy_uncertainty_run rewrite (thresholds, classes, sampler
    // order): re-quantising holds it exclusively, a run holds it shared for as long as its host side reads `low`, so a run
    // sees the old thresholds or the new ones, never a vector in mid-resize.  Lock order: low_mu, then mu, then a device's run_mu.
    std::shared_mutex low_mu;
    std::vector<canopy::DeviceState*> dev;  // lazily created, one per CUDA ordinal used
};

namespace canopy {
// engine.cu
void mc_run(canopy_plan* plan, const canopy_mc_opts& o, uint64_t* tallies_host, uint64_t* tallies_dev,
            canopy_mc_report* rep);
void mc_run_static(canopy_plan* plan, const canopy_mc_opts& o, uint64_t* tallies_host, uint64_t* tallies_dev);  // no device allocation
void mc_sample_words(canopy_plan* plan, const canopy_mc_opts& o, uint64_t first_word, uint64_t n_words,
                     uint64_t* words_host);
void mc_run_multi(canopy_plan* plan, const canopy_mc_opts& o, const int32_t* devices, int32_t n_devices,
                  uint64_t* tallies_host, canopy_mc_report* rep);
void microbench_int(int device, int kind, double* lane_ops_per_s, double* ms);
bool small_eligible(const Lowered& L);        // may runs of this plan take the fused shared-memory path (small.cuh)
void trim_device_cache(int device);           // free the value buffers parked by destroyed plans (-1: all devices)
void refresh_thresholds(canopy_plan* plan);  // re-upload thresholds and sampler order to every device the plan is resident on; rebuild the tables
void debug_set_sparse_table(canopy_plan* plan, uint32_t live_event, const uint64_t* cdf);  // canopy-testing.h hook
void debug_force_row_copies(bool on);   // canopy-testing.h hook: the gate walk fills its ring with one bulk copy per row instead of tile::gather4
void shard_range(uint64_t begin, uint64_t n, int rank, int world, uint64_t* ob, uint64_t* on);
}  // namespace canopy
