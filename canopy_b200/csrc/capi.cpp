// C ABI glue: exception -> status translation, thread-local error text, small host helpers.
// Boundary declared in include/canopy.h; replaces the empty reference TU src/canopy.cpp.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <new>
#include <shared_mutex>
#include <string>
#include <vector>

#include "plan.h"
#include "sparse_table.h"

namespace {
thread_local std::string g_last_error;

template <class F>
canopy_status guard(F&& f) {
    try {
        g_last_error.clear();
        f();
        return CANOPY_OK;
    } catch (const canopy::Error& e) {
        g_last_error = e.what();
        return e.code;
    } catch (const std::bad_alloc&) {
        g_last_error = "out of host memory";
        return CANOPY_E_OOM;
    } catch (const std::exception& e) {
        g_last_error = e.what();
        return CANOPY_E_INVALID;
    } catch (...) {
        g_last_error = "unknown error";
        return CANOPY_E_INVALID;
    }
}
void require(bool ok, const char* what) {
    if (!ok) throw canopy::Error(CANOPY_E_INVALID, what);
}
}  // namespace

extern "C" {

canopy_status canopy_plan_create(const canopy_graph_desc* desc, const canopy_plan_opts* opts, canopy_plan** out_plan) {
    return guard([&] {
        require(desc && out_plan, "canopy_plan_create: null argument");
        *out_plan = nullptr;
        canopy_plan_opts o{};
        if (opts) o = *opts;
        if (o.threshold_bits == 0) o.threshold_bits = 32;
        if (o.sampler == CANOPY_SAMPLER_DEFAULT) o.sampler = CANOPY_SPEC_VERSION;
        auto* p = new canopy_plan();
        try {
            require(o.engine >= CANOPY_ENGINE_AUTO && o.engine <= CANOPY_ENGINE_FUSED, "canopy_plan_create: bad engine option");
            canopy::lower_graph_cached(*desc, o, p->low);
            if (o.engine == CANOPY_ENGINE_FUSED && !canopy::small_eligible(p->low))
                throw canopy::Error(CANOPY_E_UNSUPPORTED, "the fused engine needs a model whose node rows of one slab fit in shared memory "
                                                          "(about 200 live nodes) and a sliced or classed sampler");
        } catch (...) { delete p; throw; }
        *out_plan = p;
    });
}

void canopy_plan_destroy(canopy_plan* plan) {
    if (!plan) return;
    for (auto* s : plan->dev) canopy::destroy_device_state(s);
    delete plan;
}

canopy_status canopy_plan_info(const canopy_plan* plan, canopy_plan_stats* out) {
    return guard([&] {
        require(plan && out, "canopy_plan_info: null argument");
        *out = plan->low.stats;
    });
}

canopy_status canopy_plan_view_get(const canopy_plan* plan, canopy_plan_view* v) {
    return guard([&] {
        require(plan && v, "canopy_plan_view_get: null argument");
        const canopy::Lowered& L = plan->low;
        v->n_events = L.n_events(); v->n_gates = L.n_gates(); v->n_layers = L.n_layers();
        v->n_tallies = (uint32_t)L.tally_kind.size(); v->n_edges = L.child.size();
        v->event_orig = L.event_orig.data(); v->event_thr32 = L.event_thr32.data(); v->event_thr64 = L.event_thr64.data();
        v->gate_child_off = L.gate_child_off.data(); v->gate_child = L.child.data();
        v->gate_op = L.gate_op.data(); v->gate_flags = L.gate_flags.data();
        v->gate_kmin = L.gate_kmin.data(); v->gate_kmax = L.gate_kmax.data();
        v->layer_off = L.layer_off.data(); v->tally_kind = L.tally_kind.data(); v->tally_gate = L.tally_gate.data();
        v->event_k = L.event_k.empty() ? nullptr : L.event_k.data();
    });
}

void canopy_mc_opts_default(canopy_mc_opts* o) {
    if (!o) return;
    std::memset(o, 0, sizeof *o);
    o->overlap = 1;
}

canopy_status canopy_mc_run(canopy_plan* plan, const canopy_mc_opts* opts, uint64_t* tallies_host, canopy_mc_report* report) {
    return guard([&] {
        require(plan && opts, "canopy_mc_run: null argument");
        require(tallies_host || plan->low.tally_kind.empty(), "canopy_mc_run: tallies_host is null");
        canopy_mc_report dummy;
        canopy::mc_run(plan, *opts, tallies_host, nullptr, report ? report : &dummy);
    });
}

canopy_status canopy_mc_run_device(canopy_plan* plan, const canopy_mc_opts* opts, uint64_t* tallies_device, canopy_mc_report* report) {
    return guard([&] {
        require(plan && opts, "canopy_mc_run_device: null argument");
        require(tallies_device || plan->low.tally_kind.empty(), "canopy_mc_run_device: tallies_device is null");
        canopy::mc_run(plan, *opts, nullptr, tallies_device, report);
    });
}

canopy_status canopy_mc_run_static(canopy_plan* plan, const canopy_mc_opts* opts, uint64_t* tallies_host, uint64_t* tallies_device) {
    return guard([&] {
        require(plan && opts, "canopy_mc_run_static: null argument");
        require(tallies_host || tallies_device || plan->low.tally_kind.empty(), "canopy_mc_run_static: both output pointers are null");
        canopy::mc_run_static(plan, *opts, tallies_host, tallies_device);
    });
}

canopy_status canopy_mc_run_multi(canopy_plan* plan, const canopy_mc_opts* opts, const int32_t* devices, int32_t n_devices,
                                  uint64_t* tallies_host, canopy_mc_report* report) {
    return guard([&] {
        require(plan && opts && devices && n_devices > 0, "canopy_mc_run_multi: bad argument");
        require(tallies_host || plan->low.tally_kind.empty(), "canopy_mc_run_multi: tallies_host is null");
        canopy::mc_run_multi(plan, *opts, devices, n_devices, tallies_host, report);
    });
}

canopy_status canopy_trim_device_cache(int32_t device) {
    return guard([&] { canopy::trim_device_cache(device); });
}

canopy_status canopy_plan_cache_clear(void) {
    return guard([&] { canopy::plan_cache_clear(); });
}

canopy_status canopy_plan_fingerprint(const canopy_plan* plan, uint64_t* out) {
    return guard([&] {
        require(plan && out, "canopy_plan_fingerprint: null argument");
        *out = canopy::fingerprint(plan->low);
    });
}

canopy_status canopy_checkpoint_init(const canopy_plan* plan, uint64_t seed, uint64_t sample_begin, canopy_checkpoint* out) {
    return guard([&] {
        require(plan && out, "canopy_checkpoint_init: null argument");
        std::memset(out, 0, sizeof *out);
        out->magic = CANOPY_CHECKPOINT_MAGIC;
        out->spec_version = plan->low.stats.spec_version;
        out->threshold_bits = (uint32_t)plan->low.opts.threshold_bits;
        out->n_tallies = (uint32_t)plan->low.tally_kind.size();
        out->plan_fingerprint = canopy::fingerprint(plan->low);
        out->seed = seed; out->sample_begin = sample_begin; out->next_sample = sample_begin;
    });
}

canopy_status canopy_mc_run_resume(canopy_plan* plan, const canopy_mc_opts* opts, canopy_checkpoint* ck,
                                   uint64_t* tallies_host_accum, canopy_mc_report* report) {
    return guard([&] {
        require(plan && opts && ck, "canopy_mc_run_resume: null argument");
        require(tallies_host_accum || plan->low.tally_kind.empty(), "canopy_mc_run_resume: tallies_host_accum is null");
        require(ck->magic == CANOPY_CHECKPOINT_MAGIC, "canopy_mc_run_resume: not a checkpoint record (bad magic)");
        require(ck->spec_version == plan->low.stats.spec_version, "canopy_mc_run_resume: checkpoint was taken under another sampling spec");
        require(ck->threshold_bits == (uint32_t)plan->low.opts.threshold_bits && ck->n_tallies == plan->low.tally_kind.size() &&
                    ck->plan_fingerprint == canopy::fingerprint(plan->low),
                "canopy_mc_run_resume: checkpoint belongs to another plan (graph, options or probabilities differ)");
        require(ck->next_sample >= ck->sample_begin && ck->next_sample + opts->n_samples >= ck->next_sample,
                "canopy_mc_run_resume: sample range overflows 64 bits");
        canopy_mc_opts o = *opts;
        o.seed = ck->seed; o.sample_begin = ck->next_sample;
        std::vector<uint64_t> part(plan->low.tally_kind.size());
        canopy_mc_report dummy;
        canopy::mc_run(plan, o, part.data(), nullptr, report ? report : &dummy);
        for (size_t i = 0; i < part.size(); ++i) tallies_host_accum[i] += part[i];
        ck->next_sample += o.n_samples;
    });
}

canopy_status canopy_plan_update_p(canopy_plan* plan, const double* p, uint32_t n_basic) {
    return guard([&] {
        require(plan && (p || n_basic == 0), "canopy_plan_update_p: null argument");
        std::unique_lock<std::shared_mutex> low_lock(plan->low_mu);   // waits for the host side of runs in progress
        std::lock_guard<std::mutex> lock(plan->mu);
        canopy::requantise(plan->low, p, n_basic);
        canopy::refresh_thresholds(plan);
    });
}

canopy_status canopy_mc_sample_words(canopy_plan* plan, const canopy_mc_opts* opts, uint64_t first_word, uint64_t n_words,
                                     uint64_t* words_host) {
    return guard([&] {
        require(plan && opts && words_host, "canopy_mc_sample_words: null argument");
        canopy::mc_sample_words(plan, *opts, first_word, n_words, words_host);
    });
}

canopy_status canopy_shard_range(uint64_t sample_begin, uint64_t n_samples, int32_t rank, int32_t world,
                                 uint64_t* out_begin, uint64_t* out_n) {
    return guard([&] {
        require(out_begin && out_n, "canopy_shard_range: null argument");
        canopy::shard_range(sample_begin, n_samples, rank, world, out_begin, out_n);
    });
}

canopy_status canopy_estimate_from_tally(uint64_t tally, uint64_t n_samples, canopy_estimate* out) {
    return guard([&] {
        require(out != nullptr, "canopy_estimate_from_tally: null argument");
        require(n_samples > 0 && tally <= n_samples, "canopy_estimate_from_tally: need 0 <= tally <= n_samples, n_samples > 0");
        const double n = (double)n_samples, p = (double)tally / n;
        const double sigma = std::sqrt(p * (1.0 - p) / n);
        out->p = p; out->sigma = sigma;
        out->ci_lo = std::fmax(0.0, p - 1.959963984540054 * sigma);
        out->ci_hi = std::fmin(1.0, p + 1.959963984540054 * sigma);
    });
}

canopy_status canopy_microbench_int(int32_t device, int32_t kind, double* out_lane_ops_per_s, double* out_ms) {
    return guard([&] {
        require(out_lane_ops_per_s && out_ms && kind >= 0 && kind <= 6, "canopy_microbench_int: bad argument");
        canopy::microbench_int(device, kind, out_lane_ops_per_s, out_ms);
    });
}

const char* canopy_last_error(void) { return g_last_error.c_str(); }

const char* canopy_status_string(canopy_status s) {
    switch (s) {
        case CANOPY_OK: return "ok";
        case CANOPY_E_INVALID: return "invalid argument";
        case CANOPY_E_VALIDITY: return "validity error";
        case CANOPY_E_CYCLE: return "cycle error";
        case CANOPY_E_UNSUPPORTED: return "unsupported";
        case CANOPY_E_CUDA: return "CUDA error";
        case CANOPY_E_NCCL: return "NCCL error";
        case CANOPY_E_OOM: return "out of memory";
        case CANOPY_E_DOMAIN: return "domain error";
        case CANOPY_E_IO: return "I/O error";
    }
    return "unknown";
}

uint32_t canopy_version(void) { return (CANOPY_VERSION_MAJOR << 16) | CANOPY_VERSION_MINOR; }
uint32_t canopy_spec_version(void) { return CANOPY_SPEC_VERSION; }

canopy_status canopy_sparse_class(uint64_t thr64, uint32_t* out_k) {
    return guard([&] {
        require(out_k != nullptr, "canopy_sparse_class: null output");
        *out_k = canopy::sparse_class_k(thr64);
    });
}

canopy_status canopy_sparse_cdf(uint64_t thr64, uint32_t k, uint64_t* out_cdf) {
    return guard([&] {
        require(out_cdf != nullptr, "canopy_sparse_cdf: null output");
        require(thr64 != 0 && k >= 7 && k <= CANOPY_SPARSE_MAX_LOG2 && canopy::clz64(thr64) >= CANOPY_SPARSE_MIN_ZEROS,
                "canopy_sparse_cdf: threshold is not in the sparse range, or k is not in 7..CANOPY_SPARSE_MAX_LOG2");
        canopy::sparse_cdf(thr64, k, out_cdf);
    });
}

}  // extern "C"

namespace canopy {
// shared by mef_xml.cpp
void set_last_error(const std::string& m) { g_last_error = m; }
}
