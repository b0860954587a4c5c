// Test-only entry points (include/canopy-testing.h).  This translation unit is linked into libcanopy_testing.so only: the
// product library, libcanopy.so, exports nothing the public header does not declare.
#include <mutex>
#include <string>
#include <vector>

#include "plan.h"
#include "canopy-testing.h"

namespace canopy { void set_last_error(const std::string& m); }   // capi.cpp

namespace {
template <class F>
canopy_status guard(F&& f) {
    try { canopy::set_last_error(""); f(); return CANOPY_OK; }
    catch (const canopy::Error& e) { canopy::set_last_error(e.what()); return e.code; }
    catch (const std::bad_alloc&) { canopy::set_last_error("out of host memory"); return CANOPY_E_OOM; }
    catch (const std::exception& e) { canopy::set_last_error(e.what()); return CANOPY_E_INVALID; }
}
void require(bool ok, const char* what) {
    if (!ok) throw canopy::Error(CANOPY_E_INVALID, what);
}
}  // namespace

extern "C" {

canopy_status canopy_plan_debug_set_sparse_table(canopy_plan* plan, uint32_t live_event, const uint64_t* cdf) {
    return guard([&] {
        require(plan && cdf, "canopy_plan_debug_set_sparse_table: null argument");
        std::lock_guard<std::mutex> lock(plan->mu);
        canopy::Lowered& L = plan->low;
        require(L.opts.sampler == CANOPY_SAMPLER_SPARSE && live_event < L.n_events() && L.event_k[live_event] != 0,
                "canopy_plan_debug_set_sparse_table: not a sparse event of a spec 4 plan");
        for (uint32_t j = 0; j + 1 < CANOPY_SPARSE_CDF_LEN; ++j) require(cdf[j] <= cdf[j + 1], "canopy_plan_debug_set_sparse_table: table must be non-decreasing");
        canopy::debug_set_sparse_table(plan, live_event, cdf);
    });
}

canopy_status canopy_plan_debug_streams(const canopy_plan* plan, canopy_plan_streams* out) {
    return guard([&] {
        require(plan && out, "canopy_plan_debug_streams: null argument");
        const canopy::Lowered& L = plan->low;
        static thread_local std::vector<uint32_t> off;   // valid until the next call on this thread
        off.resize(L.dev_rec.size());
        for (size_t g = 0; g < L.dev_rec.size(); ++g) off[g] = L.dev_rec[g].child_off;
        out->n_dev_child = (uint32_t)L.dev_child.size(); out->warps = canopy::kGateWarps; out->n_layers = L.n_layers(); out->reserved = 0;
        out->dev_child = L.dev_child.data(); out->gate_child_off = off.data();
        out->split = L.split.data(); out->split_child = L.split_child.data();
    });
}

canopy_status canopy_debug_force_row_copies(int32_t on) {
    return guard([&] { canopy::debug_force_row_copies(on != 0); });
}

}  // extern "C"
