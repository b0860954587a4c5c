/*
 * canopy-cpp.h -- C++ layer over the C ABI of canopy.h.
 *
 * Keeps the shape of the reference header (reference include/canopy-cpp.h:32-40: C++ only, <memory>, then
 * "canopy.h"); everything below it is new, because the reference header declares nothing.
 *
 *   canopy::GraphBuilder   incremental construction with the reference's vocabulary: basic events, house
 *                          events, gates with Formula(connective, args, min, max)
 *                          (reference src/mef/openpsa/event/formula.h:72-80, event/event.h:143-157)
 *   canopy::Plan           RAII owner of a lowered graph (canopy_plan)
 *   canopy::McSettings     Monte Carlo subset of reference Settings (settings.h:118-187,324-327)
 *   canopy::lower_model()  adapter from reference mef::openpsa::Model objects; only compiled when the caller
 *                          defines CANOPY_WITH_MEF_OPENPSA and has the reference headers (+ Boost) on its include
 *                          path.  This repository compiles and runs it against tests/cpp/mef_shim/ (stand-ins with
 *                          the reference's signatures; Boost is not in this image): tests/test_mef_adapter.py.
 */
#pragma once

#ifndef __cplusplus
#error "This header is for C++ only"
#endif

#include <cstddef>
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

#include "canopy.h"

namespace canopy {

/// Thrown by this layer; carries the canopy_status of the failed C call.
/// Status -> reference exception (src/mef/openpsa/error.h): E_VALIDITY ValidityError, E_CYCLE CycleError,
/// E_DOMAIN DomainError, E_INVALID LogicError, E_IO IOError.
class Error : public std::runtime_error {
  public:
    Error(canopy_status s, const std::string& msg) : std::runtime_error(msg), status_(s) {}
    canopy_status status() const noexcept { return status_; }

  private:
    canopy_status status_;
};

inline void check(canopy_status s) {
    if (s != CANOPY_OK) throw Error(s, std::string(canopy_status_string(s)) + ": " + canopy_last_error());
}

/// Same numbering as mef::openpsa::Connective.
enum Connective : std::uint8_t {
    kAnd = CANOPY_AND, kOr = CANOPY_OR, kAtleast = CANOPY_ATLEAST, kXor = CANOPY_XOR, kNot = CANOPY_NOT,
    kNand = CANOPY_NAND, kNor = CANOPY_NOR, kNull = CANOPY_NULL, kIff = CANOPY_IFF, kImply = CANOPY_IMPLY,
    kCardinality = CANOPY_CARDINALITY
};

/// A formula argument: node handle + complement flag (reference Formula::Arg, formula.h:12-15).
struct Arg {
    std::uint32_t node;
    bool complement = false;
};

class GraphBuilder {
  public:
    /// Gate handles carry this tag; basic-event handles are plain indices, so events and gates may be added in any order
    /// (the adapter meets CCF events only while it walks the gates).  desc() resolves handles to descriptor node ids.
    static constexpr std::uint32_t kGateTag = 0x40000000u;

    /// @returns handle of a new basic event with mean probability p (BasicEvent::p(), event.h:93-96).
    std::uint32_t basic_event(double p) {
        if (p_.size() >= kGateTag) throw Error(CANOPY_E_INVALID, "too many basic events");
        p_.push_back(p);
        return static_cast<std::uint32_t>(p_.size() - 1);
    }
    /// Handles of the two Boolean constants (HouseEvent::kTrue / kFalse, event.h:37-38).
    static constexpr std::uint32_t constant(bool state) { return state ? CANOPY_CONST_TRUE : CANOPY_CONST_FALSE; }
    /// A house event is its state (event.h:51).
    static constexpr std::uint32_t house_event(bool state) { return constant(state); }

    /// @returns handle of a new gate with formula (connective, args, min_number, max_number).
    std::uint32_t gate(Connective c, const std::vector<Arg>& args, int min_number = 0, int max_number = 0) {
        if (gates_op_.size() >= kGateTag - 2u) throw Error(CANOPY_E_INVALID, "too many gates");
        gates_op_.push_back(c);
        kmin_.push_back(static_cast<std::uint16_t>(min_number));
        kmax_.push_back(static_cast<std::uint16_t>(max_number));
        for (const Arg& a : args) child_.push_back(a.node | (a.complement ? CANOPY_CHILD_COMPLEMENT : 0u));
        child_off_.push_back(child_.size());
        return kGateTag | static_cast<std::uint32_t>(gates_op_.size() - 1);
    }
    void root(std::uint32_t node) { roots_.push_back(node); }
    std::size_t n_basic() const noexcept { return p_.size(); }
    std::size_t n_gates() const noexcept { return gates_op_.size(); }

    /// The flat descriptor; valid until the builder is modified or destroyed.
    canopy_graph_desc desc() const {
        const std::uint32_t nb = static_cast<std::uint32_t>(p_.size());
        auto resolve = [nb](std::uint32_t h) {
            const std::uint32_t id = h & ~CANOPY_CHILD_COMPLEMENT, c = h & CANOPY_CHILD_COMPLEMENT;
            if (id == CANOPY_CONST_TRUE || id == CANOPY_CONST_FALSE || !(id & kGateTag)) return h;
            return (nb + (id & ~kGateTag)) | c;
        };
        child_out_.resize(child_.size());
        for (std::size_t i = 0; i < child_.size(); ++i) child_out_[i] = resolve(child_[i]);
        roots_out_.resize(roots_.size());
        for (std::size_t i = 0; i < roots_.size(); ++i) roots_out_[i] = resolve(roots_[i]);
        canopy_graph_desc d{};
        d.n_basic = nb;
        d.n_gates = static_cast<std::uint32_t>(gates_op_.size());
        d.n_roots = static_cast<std::uint32_t>(roots_.size());
        d.n_edges = child_.size();
        d.p = p_.data(); d.op = gates_op_.data(); d.kmin = kmin_.data(); d.kmax = kmax_.data();
        d.child_off = child_off_.data(); d.child = child_out_.data(); d.roots = roots_out_.data();
        return d;
    }

  private:
    std::vector<double> p_;
    std::vector<std::uint8_t> gates_op_;
    std::vector<std::uint16_t> kmin_, kmax_;
    std::vector<std::uint64_t> child_off_{0};
    std::vector<std::uint32_t> child_, roots_;                 // handles
    mutable std::vector<std::uint32_t> child_out_, roots_out_; // descriptor node ids, filled by desc()
};

/// Monte Carlo knobs of reference Settings, widened from int to 64 bit.
struct McSettings {
    std::uint64_t num_trials = 1000;  ///< settings.h:325
    std::uint64_t seed = 0;           ///< settings.h:324
    std::uint64_t sample_begin = 0;   ///< restart / shard by counter range
    int batch_size = 0;               ///< slabs per batch (settings.h:326 counts trials); 0 = engine default
    int device = 0;
    bool tally_all_gates = false;
    int threshold_bits = 32;
    int sampler = CANOPY_SAMPLER_DEFAULT;  ///< sampling spec (canopy.h CANOPY_SAMPLER_*)
};

struct Result {
    std::vector<std::uint64_t> tallies;
    canopy_mc_report report{};
    canopy_estimate estimate(std::size_t i, std::uint64_t n) const {
        canopy_estimate e{};
        check(canopy_estimate_from_tally(tallies.at(i), n, &e));
        return e;
    }
};

class Plan {
  public:
    Plan(const canopy_graph_desc& d, bool tally_all_gates = false, int threshold_bits = 32,
         int sampler = CANOPY_SAMPLER_DEFAULT) {
        canopy_plan_opts o{};
        o.tally_all_gates = tally_all_gates;
        o.threshold_bits = threshold_bits;
        o.sampler = sampler;
        canopy_plan* p = nullptr;
        check(canopy_plan_create(&d, &o, &p));
        plan_.reset(p);
        check(canopy_plan_info(p, &stats_));
    }
    const canopy_plan_stats& stats() const noexcept { return stats_; }
    canopy_plan* get() const noexcept { return plan_.get(); }

    Result run(const McSettings& s, void* stream = nullptr) const {
        canopy_mc_opts o;
        canopy_mc_opts_default(&o);
        o.seed = s.seed; o.sample_begin = s.sample_begin; o.n_samples = s.num_trials;
        o.device = s.device; o.batch_slabs = s.batch_size; o.stream = stream;
        Result r;
        r.tallies.resize(stats_.n_tallies);
        check(canopy_mc_run(plan_.get(), &o, r.tallies.data(), &r.report));
        return r;
    }

    /// New basic-event probabilities on the same lowered graph (uncertainty outer loop; canopy_plan_update_p).
    void update_p(const std::vector<double>& p) { check(canopy_plan_update_p(plan_.get(), p.data(), (std::uint32_t)p.size())); }
    std::uint64_t fingerprint() const {
        std::uint64_t f = 0;
        check(canopy_plan_fingerprint(plan_.get(), &f));
        return f;
    }

  private:
    struct Deleter { void operator()(canopy_plan* p) const noexcept { canopy_plan_destroy(p); } };
    std::unique_ptr<canopy_plan, Deleter> plan_;
    canopy_plan_stats stats_{};
};

/// Probability source of the parameter-uncertainty loop (canopy_psource): BasicEvent::p() of every event as the mean
/// (Expression::value()) or as one joint draw of all random deviates (Expression::Sample()/Reset(), reference
/// expression.h:98-117; the deviates of expr/random_deviate.h).
class ProbabilitySource {
  public:
    explicit ProbabilitySource(const std::vector<canopy_deviate>& dist) {
        canopy_psource* s = nullptr;
        check(canopy_psource_from_deviates(dist.data(), (std::uint32_t)dist.size(), &s));
        src_.reset(s);
    }
    explicit ProbabilitySource(const canopy_model* model) {
        canopy_psource* s = nullptr;
        check(canopy_psource_from_model(model, &s));
        src_.reset(s);
    }
    std::uint32_t n_basic() const { std::uint32_t n = 0; check(canopy_psource_info(src_.get(), &n, nullptr)); return n; }
    std::uint32_t n_deviates() const { std::uint32_t n = 0; check(canopy_psource_info(src_.get(), nullptr, &n)); return n; }
    std::vector<double> mean() const { std::vector<double> p(n_basic()); check(canopy_psource_mean(src_.get(), p.data())); return p; }
    std::vector<double> sample(std::uint64_t seed, std::uint64_t trial) const {
        std::vector<double> p(n_basic());
        check(canopy_psource_sample(src_.get(), seed, trial, p.data()));
        return p;
    }
    const canopy_psource* get() const noexcept { return src_.get(); }

  private:
    struct Deleter { void operator()(canopy_psource* p) const noexcept { canopy_psource_destroy(p); } };
    std::unique_ptr<canopy_psource, Deleter> src_;
};

/// Result of the uncertainty analysis (reference Settings::uncertainty_analysis / num_quantiles / num_bins,
/// settings.h:154-175,281-295), per output.
struct UncertaintyResult {
    std::uint32_t num_quantiles = 0, num_bins = 0;
    std::vector<canopy_uncertainty_stats> stats;   ///< [n_tallies]
    std::vector<double> quantiles;                 ///< [n_tallies][num_quantiles]
    std::vector<double> bin_edges;                 ///< [n_tallies][num_bins + 1]
    std::vector<std::uint64_t> bin_counts;         ///< [n_tallies][num_bins]
    std::vector<double> p_hat;                     ///< [n_trials][n_tallies]
};
/// `n_outer` epistemic trials of `s.num_trials` samples each on one lowered plan (canopy_uncertainty_run).
inline UncertaintyResult uncertainty(Plan& plan, const ProbabilitySource& src, const McSettings& s, std::uint64_t n_outer,
                                     std::uint64_t deviate_seed, std::uint32_t num_quantiles = 20, std::uint32_t num_bins = 20) {
    canopy_mc_opts o;
    canopy_mc_opts_default(&o);
    o.seed = s.seed; o.sample_begin = s.sample_begin; o.n_samples = s.num_trials; o.device = s.device; o.batch_slabs = s.batch_size;
    canopy_uncertainty_opts u{};
    u.n_trials = n_outer; u.seed = deviate_seed; u.num_quantiles = num_quantiles; u.num_bins = num_bins;
    const std::size_t nt = plan.stats().n_tallies;
    UncertaintyResult r;
    r.num_quantiles = num_quantiles; r.num_bins = num_bins;
    r.stats.resize(nt); r.quantiles.resize(nt * num_quantiles); r.bin_edges.resize(nt * (num_bins + 1));
    r.bin_counts.resize(nt * num_bins); r.p_hat.resize(n_outer * nt);
    check(canopy_uncertainty_run(plan.get(), src.get(), &o, &u, r.stats.data(), r.quantiles.data(), r.bin_edges.data(),
                                 r.bin_counts.data(), r.p_hat.data()));
    return r;
}

/// Restartable run: (seed, next sample index, tallies) -- canopy_checkpoint + canopy_mc_run_resume.
/// `record` and `tallies` are plain data and may be written to disk between calls.
class Session {
  public:
    Session(const Plan& plan, std::uint64_t seed, std::uint64_t sample_begin = 0) : plan_(&plan), tallies(plan.stats().n_tallies, 0) {
        check(canopy_checkpoint_init(plan.get(), seed, sample_begin, &record));
    }
    Session(const Plan& plan, const canopy_checkpoint& rec, std::vector<std::uint64_t> t) : plan_(&plan), record(rec), tallies(std::move(t)) {}
    std::uint64_t done() const noexcept { return record.next_sample - record.sample_begin; }
    canopy_mc_report advance(std::uint64_t n_samples, int device = 0) {
        canopy_mc_opts o;
        canopy_mc_opts_default(&o);
        o.n_samples = n_samples; o.device = device;
        canopy_mc_report rep{};
        check(canopy_mc_run_resume(plan_->get(), &o, &record, tallies.data(), &rep));
        return rep;
    }

  private:
    const Plan* plan_;

  public:
    canopy_checkpoint record{};
    std::vector<std::uint64_t> tallies;
};

inline Result quantify(const canopy_graph_desc& d, const McSettings& s) {
    return Plan(d, s.tally_all_gates, s.threshold_bits, s.sampler).run(s);
}

}  // namespace canopy

#ifdef CANOPY_WITH_MEF_OPENPSA
// Adapter from the reference object model. Needs the reference headers (and Boost) on the include path.
#include <variant>

#include "mef/openpsa/event/event.h"
#include "mef/openpsa/event/formula.h"
#include "mef/openpsa/event/gate.h"
#include "mef/openpsa/fault_tree.h"
#include "mef/openpsa/model.h"
#include "mef/openpsa/settings.h"

namespace canopy {

/// Flattens Model -> FaultTree::top_events() -> Gate::formula().args() into a GraphBuilder.
/// Basic events: the rows of Model::basic_events() first, in table order (model.h:89), so that an event's Philox stream
/// (keyed by its descriptor index) does not depend on which gates use it.  With settings.ccf_analysis() (settings.h:298) a
/// member of a CCF group is replaced by its proxy gate BasicEvent::ccf_gate() (event.h:108-114), whose arguments are
/// CcfEvent objects owned by the group (ccf_group.h:215-260), NOT rows of Model::basic_events(): those are added when the
/// walk first meets them, in walk order (top events in fault-tree order, arguments left to right, depth first).
/// Must run on the thread that owns the model: MEF objects are not thread-safe (model.h:74-76).
inline GraphBuilder lower_model(const mef::openpsa::Model& model, const mef::openpsa::Settings& settings) {
    using namespace mef::openpsa;
    GraphBuilder b;
    using EventIds = std::unordered_map<const BasicEvent*, std::uint32_t>;
    using GateIds = std::unordered_map<const Gate*, std::uint32_t>;
    EventIds events;
    for (const BasicEvent& e : model.basic_events()) events.emplace(&e, b.basic_event(e.p()));
    GateIds gates;
    struct Walk {
        GraphBuilder& b; EventIds& ev; GateIds& done; bool ccf;
        std::uint32_t basic(const BasicEvent* e) {
            if (ccf && e->HasCcf()) return gate(e->ccf_gate());
            auto it = ev.find(e);
            if (it != ev.end()) return it->second;
            return ev.emplace(e, b.basic_event(e->p())).first->second;   // a CcfEvent (or an event outside the model's table)
        }
        std::uint32_t gate(const Gate& g) {
            auto it = done.find(&g);
            if (it != done.end()) return it->second;
            const Formula& f = g.formula();
            std::vector<Arg> args;
            for (const Formula::Arg& a : f.args()) {
                std::uint32_t node = std::visit([&](auto* e) -> std::uint32_t {
                    using T = std::decay_t<decltype(*e)>;
                    if constexpr (std::is_same_v<T, Gate>) return gate(*e);
                    else if constexpr (std::is_same_v<T, BasicEvent>) return basic(e);
                    else return GraphBuilder::house_event(e->state());
                }, a.event);
                args.push_back({node, a.complement});
            }
            std::uint32_t h = b.gate(static_cast<Connective>(f.connective()), args, f.min_number().value_or(0),
                                     f.max_number().value_or(0));
            done.emplace(&g, h);
            return h;
        }
    } walk{b, events, gates, settings.ccf_analysis()};
    for (const FaultTree& ft : model.fault_trees())
        for (const Gate* top : ft.top_events()) b.root(walk.gate(*top));
    return b;
}

inline McSettings mc_settings(const mef::openpsa::Settings& s) {
    McSettings m;
    m.num_trials = static_cast<std::uint64_t>(s.num_trials());
    m.seed = static_cast<std::uint64_t>(s.seed());
    return m;
}

}  // namespace canopy
#endif  // CANOPY_WITH_MEF_OPENPSA
