// TEST SHIM -- not the reference.  Minimal stand-ins with exactly the signatures canopy::lower_model (include/canopy-cpp.h)
// uses from the reference's src/mef/openpsa/event/event.h, so that the adapter is compiled and run in this repository
// (the reference headers themselves need Boost, which this image does not have).  Reference file:line beside each item.
#pragma once
#include <cassert>
#include <cstdint>
#include <memory>
#include <optional>
#include <string>
#include <utility>
#include <variant>
#include <vector>

namespace mef::openpsa {

class Expression {                                   // expression.h:52-117 (value() only: the mean)
  public:
    virtual ~Expression() = default;
    virtual double value() noexcept = 0;             // expression.h:74
};
class ConstantExpression : public Expression {       // expr/constant.h (a float literal)
  public:
    explicit ConstantExpression(double v) : v_(v) {}
    double value() noexcept override { return v_; }
  private:
    double v_;
};

class Event {                                        // event.h:21-28 (Id: name / base_path / role, element.h)
  public:
    explicit Event(std::string name, std::string base_path = "") : name_(std::move(name)), base_path_(std::move(base_path)) {}
    virtual ~Event() = default;
    const std::string& name() const { return name_; }
    const std::string& base_path() const { return base_path_; }
  private:
    std::string name_, base_path_;
};

class HouseEvent : public Event {                    // event.h:33-57
  public:
    using Event::Event;
    void state(bool constant) { state_ = constant; } // event.h:48
    [[nodiscard]] bool state() const { return state_; }  // event.h:51
  private:
    bool state_ = false;                             // event.h:56: unset house events are false
};

class Gate;

class BasicEvent : public Event {                    // event.h:62-135
  public:
    using Event::Event;
    void expression(Expression* expression) { expression_ = expression; }   // event.h:76
    [[nodiscard]] double p() const noexcept { assert(expression_); return expression_->value(); }  // event.h:93-96
    [[nodiscard]] bool HasCcf() const { return ccf_gate_ != nullptr; }      // event.h:108
    [[nodiscard]] const Gate& ccf_gate() const { assert(ccf_gate_); return *ccf_gate_; }  // event.h:111-114
    void ccf_gate(std::unique_ptr<Gate> gate);       // event.h:123-126
    ~BasicEvent() override;
  private:
    Expression* expression_ = nullptr;
    std::unique_ptr<Gate> ccf_gate_;
};

enum Connective : std::uint8_t {                     // event.h:143-157, same numbering
    kAnd = 0, kOr, kAtleast, kXor, kNot, kNand, kNor, kNull, kIff, kImply, kCardinality
};

}  // namespace mef::openpsa
