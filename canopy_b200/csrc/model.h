// The flat model a MEF reader produces (csrc/mef_model.cpp) and the probability source of the uncertainty loop
// (csrc/uncertainty.cpp).  Internal; the boundary is include/canopy.h.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "expr_tape.h"

struct canopy_model {
    std::vector<double> p;
    std::vector<uint8_t> op;
    std::vector<uint16_t> kmin, kmax;
    std::vector<uint64_t> child_off{0};
    std::vector<uint32_t> child;
    std::vector<uint32_t> roots;
    std::vector<std::string> basic_names, gate_names, root_names;
    canopy::ExprTape tape;          // every probability expression of the model (expr_tape.h)
    std::vector<uint32_t> p_node;   // [n_basic] tape node whose value is p[i]
};

struct canopy_psource {
    canopy::ExprTape tape;
    std::vector<uint32_t> p_node;   // [n_basic]
};
