// Parameter-uncertainty outer loop (reference Settings::uncertainty_analysis settings.h:281-295,
// Expression::Sample()/Reset() expression.h:98-117, the deviates of expr/random_deviate.h:25-264).
// Boundary: the canopy_psource_* / canopy_uncertainty_run entries of include/canopy.h.
#include <algorithm>
#include <cmath>
#include <future>
#include <memory>
#include <new>
#include <string>
#include <vector>

#include "model.h"
#include "plan.h"

namespace canopy {
void set_last_error(const std::string& m);  // capi.cpp
namespace {

template <class F>
canopy_status guard(F&& f) {
    try { set_last_error(""); f(); return CANOPY_OK; }
    catch (const Error& e) { set_last_error(e.what()); return e.code; }
    catch (const std::bad_alloc&) { set_last_error("out of host memory"); return CANOPY_E_OOM; }
    catch (const std::exception& e) { set_last_error(e.what()); return CANOPY_E_INVALID; }
    catch (...) { set_last_error("unknown error"); return CANOPY_E_INVALID; }
}
void require(bool ok, const char* what) { if (!ok) throw Error(CANOPY_E_INVALID, what); }
[[noreturn]] void bad_dist(canopy_status c, uint32_t i, const char* msg) {
    throw Error(c, "basic event " + std::to_string(i) + ": " + msg);
}

// Same validation rules and messages as the XML path (mef_model.cpp), which restates random_deviate.h's Validate().
uint32_t add_deviate(ExprTape& T, const canopy_deviate& d, uint32_t i) {
    auto K = [&](double c) { return T.constant(c); };
    auto finite = [&](double x) { if (!(x == x) || std::isinf(x)) bad_dist(CANOPY_E_DOMAIN, i, "distribution parameter is not finite"); };
    switch (d.kind) {
        case CANOPY_DEV_CONSTANT: finite(d.a); return K(d.a);
        case CANOPY_DEV_UNIFORM:
            finite(d.a); finite(d.b);
            if (d.a >= d.b) bad_dist(CANOPY_E_VALIDITY, i, "Min value is more than max for Uniform distribution.");
            return T.add(X_UNIFORM, {K(d.a), K(d.b)});
        case CANOPY_DEV_NORMAL:
            finite(d.a); finite(d.b);
            if (d.b <= 0) bad_dist(CANOPY_E_DOMAIN, i, "Standard deviation cannot be negative or zero.");
            return T.add(X_NORMAL, {K(d.a), K(d.b)});
        case CANOPY_DEV_LOGNORMAL:
            finite(d.a); finite(d.b);
            if (d.b <= 0) bad_dist(CANOPY_E_DOMAIN, i, "Standard deviation cannot be negative or zero.");
            return T.add(X_LOGNORMAL2, {K(d.a), K(d.b)});
        case CANOPY_DEV_LOGNORMAL_EF:
            finite(d.a); finite(d.b); finite(d.c);
            if (d.a <= 0) bad_dist(CANOPY_E_DOMAIN, i, "The mean of Log-Normal distribution cannot be negative or zero.");
            if (d.b <= 1) bad_dist(CANOPY_E_DOMAIN, i, "The Error Factor for Log-Normal distribution cannot be less than 1.");
            if (d.c <= 0 || d.c >= 1) bad_dist(CANOPY_E_DOMAIN, i, "The confidence level is not within (0, 1).");
            if (d.c == 0.5) bad_dist(CANOPY_E_DOMAIN, i, "The confidence level 0.5 leaves the Log-Normal scale undefined.");
            return T.add(X_LOGNORMAL3, {K(d.a), K(d.b), K(d.c)});
        case CANOPY_DEV_GAMMA:
            finite(d.a); finite(d.b);
            if (d.a <= 0 || d.b <= 0) bad_dist(CANOPY_E_DOMAIN, i, "The k and theta parameters of Gamma distribution must be positive.");
            return T.add(X_GAMMA, {K(d.a), K(d.b)});
        case CANOPY_DEV_BETA:
            finite(d.a); finite(d.b);
            if (d.a <= 0 || d.b <= 0) bad_dist(CANOPY_E_DOMAIN, i, "The alpha and beta parameters of Beta distribution must be positive.");
            return T.add(X_BETA, {K(d.a), K(d.b)});
        case CANOPY_DEV_HISTOGRAM: {
            if (d.n_bins == 0 || !d.bounds || !d.weights) bad_dist(CANOPY_E_VALIDITY, i, "<histogram> needs a lower boundary and at least one <bin>");
            std::vector<uint32_t> a{K(d.bounds[0])};
            double sum = 0;
            for (uint32_t j = 0; j < d.n_bins; ++j) {
                finite(d.bounds[j + 1]); finite(d.weights[j]);
                if (!(d.bounds[j + 1] > d.bounds[j])) bad_dist(CANOPY_E_VALIDITY, i, "Histogram upper boundaries are not strictly increasing.");
                if (d.weights[j] < 0) bad_dist(CANOPY_E_VALIDITY, i, "Histogram weights cannot be negative.");
                sum += d.weights[j];
                a.push_back(K(d.bounds[j + 1])); a.push_back(K(d.weights[j]));
            }
            if (sum <= 0) bad_dist(CANOPY_E_VALIDITY, i, "Histogram weights sum to zero.");
            return T.add(X_HISTOGRAM, a);
        }
        default: bad_dist(CANOPY_E_INVALID, i, "unknown canopy_deviate_kind");
    }
}

void source_p(const canopy_psource& S, bool sample, uint64_t seed, uint64_t trial, std::vector<double>& scratch, double* p_out) {
    S.tape.evaluate(sample, seed, trial, scratch);
    for (size_t i = 0; i < S.p_node.size(); ++i) {
        const double x = scratch[S.p_node[i]];
        if (x != x) throw Error(CANOPY_E_DOMAIN, "probability of basic event " + std::to_string(i) + " is not a number in trial " + std::to_string(trial));
        p_out[i] = sample ? std::fmin(1.0, std::fmax(0.0, x)) : x;
    }
}

}  // namespace
}  // namespace canopy

using namespace canopy;

extern "C" {

canopy_status canopy_psource_from_model(const canopy_model* model, canopy_psource** out) {
    return guard([&] {
        require(model && out, "canopy_psource_from_model: null argument");
        auto s = std::make_unique<canopy_psource>();
        s->tape = model->tape; s->p_node = model->p_node;
        *out = s.release();
    });
}

canopy_status canopy_psource_from_deviates(const canopy_deviate* dist, uint32_t n_basic, canopy_psource** out) {
    return guard([&] {
        require((dist || n_basic == 0) && out, "canopy_psource_from_deviates: null argument");
        auto s = std::make_unique<canopy_psource>();
        for (uint32_t i = 0; i < n_basic; ++i) s->p_node.push_back(add_deviate(s->tape, dist[i], i));
        *out = s.release();
    });
}

canopy_status canopy_psource_info(const canopy_psource* src, uint32_t* n_basic, uint32_t* n_deviates) {
    return guard([&] {
        require(src != nullptr, "canopy_psource_info: null argument");
        if (n_basic) *n_basic = (uint32_t)src->p_node.size();
        if (n_deviates) *n_deviates = src->tape.n_deviates;
    });
}

canopy_status canopy_psource_mean(const canopy_psource* src, double* p_out) {
    return guard([&] {
        require(src && (p_out || src->p_node.empty()), "canopy_psource_mean: null argument");
        std::vector<double> v;
        source_p(*src, false, 0, 0, v, p_out);
    });
}

canopy_status canopy_psource_sample(const canopy_psource* src, uint64_t seed, uint64_t trial, double* p_out) {
    return guard([&] {
        require(src && (p_out || src->p_node.empty()), "canopy_psource_sample: null argument");
        std::vector<double> v;
        source_p(*src, true, seed, trial, v, p_out);
    });
}

void canopy_psource_destroy(canopy_psource* src) { delete src; }

canopy_status canopy_uncertainty_run(canopy_plan* plan, const canopy_psource* src, const canopy_mc_opts* mc,
                                     const canopy_uncertainty_opts* opts, canopy_uncertainty_stats* stats, double* quantiles,
                                     double* bin_edges, uint64_t* bin_counts, double* p_hat) {
    return guard([&] {
        require(plan && src && mc && opts, "canopy_uncertainty_run: null argument");
        Lowered& L = plan->low;
        const uint32_t nb = (uint32_t)src->p_node.size(), nt = (uint32_t)L.tally_kind.size();
        require(nb == L.event_class.size(), "canopy_uncertainty_run: the probability source and the plan differ in the number of basic events");
        require(stats || nt == 0, "canopy_uncertainty_run: stats is null");
        require(opts->n_trials > 0 && mc->n_samples > 0, "canopy_uncertainty_run: n_trials and n_samples must be positive");
        require(opts->n_trials <= (1ull << 32), "canopy_uncertainty_run: too many trials");
        const uint32_t nq = opts->num_quantiles ? opts->num_quantiles : 20, nbins = opts->num_bins ? opts->num_bins : 20;
        const uint64_t T = opts->n_trials;
        const double p_lo = std::ldexp(1.0, -L.opts.threshold_bits), p_hi = 1.0 - std::ldexp(1.0, -53);
        std::vector<double> mean_p(nb), p(nb), scratch;
        source_p(*src, false, 0, 0, scratch, mean_p.data());
        std::vector<double> ph((size_t)T * nt);
        std::vector<uint64_t> tallies(nt);
        struct Restore {  // leave the plan at the mean probabilities on every path out
            canopy_plan* plan; const std::vector<double>& p;
            ~Restore() { try { std::unique_lock<std::shared_mutex> low_lock(plan->low_mu); std::lock_guard<std::mutex> lock(plan->mu); requantise(plan->low, p.data(), (uint32_t)p.size()); refresh_thresholds(plan); } catch (...) {} }
        } restore{plan, mean_p};
        // the draws of trial t + 1 are made on a helper thread while the device runs trial t (100k lognormal draws cost 9 ms
        // of host time, an inner run of 2^22 samples of c4 50 ms)
        std::vector<double> p_next(nb), scratch_next;
        auto draw = [&](uint64_t t, std::vector<double>& sc, std::vector<double>& out) { source_p(*src, true, opts->seed, opts->first_trial + t, sc, out.data()); };
        std::future<void> pending = std::async(std::launch::async, draw, (uint64_t)0, std::ref(scratch_next), std::ref(p_next));
        for (uint64_t t = 0; t < T; ++t) {
            pending.get();   // rethrows what the draw threw
            p.swap(p_next);
            if (t + 1 < T) pending = std::async(std::launch::async, draw, t + 1, std::ref(scratch_next), std::ref(p_next));
            for (uint32_t i = 0; i < nb; ++i)
                p[i] = L.event_class[i] == 2 ? std::fmin(p_hi, std::fmax(p_lo, p[i])) : mean_p[i];
            {
                std::unique_lock<std::shared_mutex> low_lock(plan->low_mu);
                std::lock_guard<std::mutex> lock(plan->mu);
                requantise(L, p.data(), nb);
                refresh_thresholds(plan);
            }
            canopy_mc_opts o = *mc;
            o.sample_begin = mc->sample_begin + t * mc->n_samples;
            mc_run(plan, o, tallies.data(), nullptr, nullptr);
            for (uint32_t i = 0; i < nt; ++i) ph[(size_t)t * nt + i] = (double)tallies[i] / (double)mc->n_samples;
        }
        if (p_hat) std::copy(ph.begin(), ph.end(), p_hat);
        std::vector<double> col(T);
        for (uint32_t i = 0; i < nt; ++i) {
            double s = 0;
            for (uint64_t t = 0; t < T; ++t) { col[t] = ph[(size_t)t * nt + i]; s += col[t]; }
            const double mean = s / (double)T;
            double ss = 0;
            for (double x : col) ss += (x - mean) * (x - mean);
            const double sigma = T > 1 ? std::sqrt(ss / (double)(T - 1)) : 0.0;
            std::sort(col.begin(), col.end());
            auto q = [&](double level) {  // order statistic ceil(level T), 1-based
                const uint64_t r = (uint64_t)std::ceil(level * (double)T);
                return col[std::min<uint64_t>(std::max<uint64_t>(r, 1), T) - 1];
            };
            canopy_uncertainty_stats& S = stats[i];
            S.mean = mean; S.sigma = sigma;
            S.ci_lo = mean - 1.959963984540054 * sigma / std::sqrt((double)T);
            S.ci_hi = mean + 1.959963984540054 * sigma / std::sqrt((double)T);
            S.min = col.front(); S.max = col.back();
            const double q50 = q(0.5), q95 = q(0.95);
            S.error_factor = q50 > 0 ? q95 / q50 : 0.0;
            if (quantiles) for (uint32_t j = 0; j < nq; ++j) quantiles[(size_t)i * nq + j] = q((double)(j + 1) / (double)nq);
            if (bin_edges || bin_counts) {
                const double lo = S.min, w = (S.max - S.min) / (double)nbins;
                if (bin_edges) for (uint32_t j = 0; j <= nbins; ++j) bin_edges[(size_t)i * (nbins + 1) + j] = j == nbins ? S.max : lo + w * j;
                if (bin_counts) {
                    uint64_t* c = bin_counts + (size_t)i * nbins;
                    std::fill(c, c + nbins, 0);
                    for (double x : col) { uint32_t j = w > 0 ? (uint32_t)std::fmin((double)(nbins - 1), std::floor((x - lo) / w)) : 0u; ++c[j]; }
                }
            }
        }
    });
}

}  // extern "C"
