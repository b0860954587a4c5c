/*
 * canopy-testing.h -- test-only entry points.  NOT part of the public boundary (include/canopy.h): nothing a caller of
 * the engine needs is declared here, the installed package does not ship this header, and libcanopy.so does not export
 * these symbols.  They exist in libcanopy_testing.so only -- the same objects as libcanopy.so plus csrc/capi_testing.cpp
 * (csrc/Makefile) -- which the test suite loads, so that the parity tests can drive paths that real inputs never reach.
 */
#ifndef CANOPY_TESTING_H
#define CANOPY_TESTING_H
#include "canopy.h"
#ifdef __cplusplus
extern "C" {
#endif

/* Spec 4 test hook: replace the count table of one sparse event of a plan (live_event = position in
 * canopy_plan_view.event_orig) by cdf[CANOPY_SPARSE_CDF_LEN] (non-decreasing).  Lets the parity tests drive the count, and
 * with it the number of position fields and Philox calls a group consumes, far beyond what a real probability reaches
 * (an all-zero table gives 32 ones in every group).  The next canopy_plan_update_p restores the real table.
 * CANOPY_E_INVALID unless the plan samples with spec 4 and the event takes the sparse path. */
CANOPY_API canopy_status canopy_plan_debug_set_sparse_table(canopy_plan* plan, uint32_t live_event, const uint64_t* cdf);

/* The child stream the gate walk reads (host only; pointers stay valid until canopy_plan_destroy, gate_child_off until the
 * next call of this function on the same thread): per layer the runs of the
 * walk's warps one after the other, each padded to a multiple of four codes, then the arguments of the layer's wide gates.
 * gate_child_off[g] = where gate g's arguments start in dev_child (same order as canopy_plan_view.gate_child);
 * split / split_child: [n_layers][warps + 1] gate and dev_child offsets of the runs.  Lets the host-logic tests check the
 * layout without a GPU. */
typedef struct canopy_plan_streams {
    uint32_t n_dev_child, warps, n_layers, reserved;
    const uint32_t* dev_child;
    const uint32_t* gate_child_off;   /* [n_gates] */
    const uint32_t* split;
    const uint32_t* split_child;
} canopy_plan_streams;
CANOPY_API canopy_status canopy_plan_debug_streams(const canopy_plan* plan, canopy_plan_streams* out);

/* Process-wide switch: with on != 0 every later gate walk fills its shared-memory ring with one bulk copy per row
 * (gate_ring_kernel_t<false>) instead of four-row tile::gather4 tensor copies -- the path the engine otherwise takes only
 * for value buffers of more than 2^31 rows or when the driver offers no tensor-map encoder, neither of which a 180 GB part
 * can produce.  Tallies must not change. */
CANOPY_API canopy_status canopy_debug_force_row_copies(int32_t on);

#ifdef __cplusplus
}
#endif
#endif /* CANOPY_TESTING_H */
