/*
 * canopy-testing.h -- test-only entry points of libcanopy.  NOT part of the public boundary (include/canopy.h): nothing a
 * caller of the engine needs is declared here, and the installed package does not ship this header.  The symbols are
 * exported so that the parity tests can drive paths a real probability never reaches.
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

#ifdef __cplusplus
}
#endif
#endif /* CANOPY_TESTING_H */
