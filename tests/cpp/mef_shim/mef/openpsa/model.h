// TEST SHIM -- not the reference (see event/event.h).  Stand-in for src/mef/openpsa/model.h: tables of elements that
// iterate as references (the reference's boost::multi_index tables of unique_ptr do, model.h:80-93).
#pragma once
#include "mef/openpsa/ccf_group.h"
#include "mef/openpsa/fault_tree.h"

namespace mef::openpsa {

template <class T>
class Table {
  public:
    class iterator {
      public:
        explicit iterator(typename std::vector<std::unique_ptr<T>>::const_iterator it) : it_(it) {}
        const T& operator*() const { return **it_; }
        iterator& operator++() { ++it_; return *this; }
        bool operator!=(const iterator& o) const { return it_ != o.it_; }
      private:
        typename std::vector<std::unique_ptr<T>>::const_iterator it_;
    };
    explicit Table(const std::vector<std::unique_ptr<T>>& v) : v_(&v) {}
    iterator begin() const { return iterator(v_->begin()); }
    iterator end() const { return iterator(v_->end()); }
    std::size_t size() const { return v_->size(); }
  private:
    const std::vector<std::unique_ptr<T>>* v_;
};

class Model {                                        // model.h:23-140
  public:
    auto fault_trees() const { return Table<FaultTree>(fault_trees_); }    // model.h:84
    auto house_events() const { return Table<HouseEvent>(house_events_); } // model.h:88
    auto basic_events() const { return Table<BasicEvent>(basic_events_); } // model.h:89
    auto gates() const { return Table<Gate>(gates_); }                     // model.h:90
    template <class T> T* Add(std::unique_ptr<T> e) {                      // model.h:105-125
        T* raw = e.get();
        if constexpr (std::is_same_v<T, FaultTree>) fault_trees_.push_back(std::move(e));
        else if constexpr (std::is_same_v<T, HouseEvent>) house_events_.push_back(std::move(e));
        else if constexpr (std::is_same_v<T, BasicEvent>) basic_events_.push_back(std::move(e));
        else gates_.push_back(std::move(e));
        return raw;
    }
  private:
    std::vector<std::unique_ptr<FaultTree>> fault_trees_;
    std::vector<std::unique_ptr<HouseEvent>> house_events_;
    std::vector<std::unique_ptr<BasicEvent>> basic_events_;
    std::vector<std::unique_ptr<Gate>> gates_;
};

}  // namespace mef::openpsa
