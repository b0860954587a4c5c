// TEST SHIM -- not the reference (see event/event.h).  Stand-in for src/mef/openpsa/fault_tree.h.
#pragma once
#include "mef/openpsa/event/gate.h"

namespace mef::openpsa {

class FaultTree {                                    // fault_tree.h:128-190
  public:
    explicit FaultTree(std::string name) : name_(std::move(name)) {}
    const std::string& name() const { return name_; }
    const std::vector<const Gate*>& top_events() const { return top_events_; }   // fault_tree.h:142
    void add_top_event(const Gate* g) { top_events_.push_back(g); }              // shim: stands for CollectTopEvents() :151-186
  private:
    std::string name_;
    std::vector<const Gate*> top_events_;
};

}  // namespace mef::openpsa
