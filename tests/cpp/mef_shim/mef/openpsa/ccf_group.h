// TEST SHIM -- not the reference (see event/event.h).  Stand-in for the part of src/mef/openpsa/ccf_group.h the adapter meets:
// CcfEvent objects are owned by their group (ccf_group.h:240 `ccf_events_`), NOT by Model::basic_events().
#pragma once
#include "mef/openpsa/event/gate.h"

namespace mef::openpsa {

class CcfEvent : public BasicEvent {                 // ccf_group.h:30-72
  public:
    explicit CcfEvent(std::vector<Gate*> members) : BasicEvent(MakeName(members)), members_(std::move(members)) {}
    const std::vector<Gate*>& members() const { return members_; }               // ccf_group.h:48
  private:
    static std::string MakeName(const std::vector<Gate*>& m) {                    // ccf_group.h:57-66
        std::string s = "[";
        for (std::size_t i = 0; i < m.size(); ++i) s += (i ? " " : "") + m[i]->name();
        return s + "]";
    }
    std::vector<Gate*> members_;
};

/// The proxy-gate construction of CcfGroup::ApplyModel (ccf_group.h:215-260) for given per-level probabilities
/// (level k = k members fail together): every member gets a proxy OR gate over the CCF events it takes part in.
class CcfGroup {
  public:
    void AddMember(BasicEvent* e) { members_.push_back(e); }
    void ApplyModel(const std::vector<std::pair<int, double>>& level_probabilities) {
        std::vector<std::pair<Gate*, Formula::ArgSet>> proxy;
        for (BasicEvent* m : members_) {
            auto g = std::make_unique<Gate>(m->name(), m->base_path());
            proxy.push_back({g.get(), {}});
            m->ccf_gate(std::move(g));
        }
        const std::size_t n = proxy.size();
        for (const auto& [level, prob] : level_probabilities) {
            exprs_.push_back(std::make_unique<ConstantExpression>(prob));
            std::vector<std::size_t> idx(level);
            for (int i = 0; i < level; ++i) idx[i] = i;
            for (;;) {                                // all combinations of `level` members, lexicographic
                std::vector<Gate*> comb;
                for (std::size_t i : idx) comb.push_back(proxy[i].first);
                auto ev = std::make_unique<CcfEvent>(std::move(comb));
                for (std::size_t i : idx) proxy[i].second.Add(ev.get());
                ev->expression(exprs_.back().get());
                ccf_events_.push_back(std::move(ev));
                int k = level - 1;
                while (k >= 0 && idx[k] == n - level + k) --k;
                if (k < 0) break;
                ++idx[k];
                for (int j = k + 1; j < level; ++j) idx[j] = idx[j - 1] + 1;
            }
        }
        for (auto& [gate, args] : proxy) gate->formula(std::make_unique<Formula>(kOr, std::move(args)));
    }
    const std::vector<std::unique_ptr<CcfEvent>>& ccf_events() const { return ccf_events_; }
  private:
    std::vector<BasicEvent*> members_;
    std::vector<std::unique_ptr<CcfEvent>> ccf_events_;
    std::vector<std::unique_ptr<ConstantExpression>> exprs_;
};

}  // namespace mef::openpsa
