// TEST SHIM -- not the reference (see event/event.h).  Stand-in for src/mef/openpsa/event/formula.h.
#pragma once
#include "mef/openpsa/event/event.h"

namespace mef::openpsa {

class Formula {                                      // formula.h:6-148
  public:
    using ArgEvent = std::variant<Gate*, BasicEvent*, HouseEvent*>;   // formula.h:9
    struct Arg {                                     // formula.h:12-15
        bool complement;
        ArgEvent event;
    };
    class ArgSet {                                   // formula.h:18-69 (no duplicate check in the shim)
      public:
        ArgSet() = default;
        ArgSet(std::initializer_list<Arg> l) : args_(l) {}
        void Add(ArgEvent event, bool complement = false) { args_.push_back({complement, event}); }
        const std::vector<Arg>& data() const { return args_; }
        std::size_t size() const { return args_.size(); }
      private:
        std::vector<Arg> args_;
    };
    Formula(Connective connective, ArgSet args, std::optional<int> min_number = {}, std::optional<int> max_number = {})  // formula.h:80-82
        : connective_(connective), min_(min_number), max_(max_number), args_(std::move(args)) {}
    [[nodiscard]] Connective connective() const { return connective_; }        // formula.h:91
    [[nodiscard]] std::optional<int> min_number() const { return min_; }        // formula.h:94
    std::optional<int> max_number() const { return max_; }                      // formula.h:97
    [[nodiscard]] const std::vector<Arg>& args() const { return args_.data(); } // formula.h:100
  private:
    Connective connective_;
    std::optional<int> min_, max_;
    ArgSet args_;
};

}  // namespace mef::openpsa
