// TEST SHIM -- not the reference (see event/event.h).  Stand-in for src/mef/openpsa/event/gate.h.
#pragma once
#include "mef/openpsa/event/formula.h"

namespace mef::openpsa {

class Gate : public Event {                          // gate.h:31-65
  public:
    using Event::Event;
    [[nodiscard]] bool HasFormula() const { return formula_ != nullptr; }       // gate.h:38
    [[nodiscard]] const Formula& formula() const { assert(formula_); return *formula_; }  // gate.h:45-48
    std::unique_ptr<Formula> formula(std::unique_ptr<Formula> f) { formula_.swap(f); return f; }  // gate.h:57-61
  private:
    std::unique_ptr<Formula> formula_;
};

inline void BasicEvent::ccf_gate(std::unique_ptr<Gate> gate) { assert(!ccf_gate_); ccf_gate_ = std::move(gate); }
inline BasicEvent::~BasicEvent() = default;

}  // namespace mef::openpsa
