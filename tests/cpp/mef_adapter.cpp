// Compiles include/canopy-cpp.h with -DCANOPY_WITH_MEF_OPENPSA against tests/cpp/mef_shim/ (stand-ins with the reference's
// signatures), builds a small model the way the reference's Initializer would leave it -- two fault trees, a house event,
// nested gates, complemented arguments, an atleast gate and a CCF group whose CcfEvents are NOT rows of
// Model::basic_events() -- lowers it with canopy::lower_model, prints the descriptor as JSON and, with `run`, the tallies.
//   mef_adapter <ccf 0|1> [run <seed> <n>]
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "canopy-cpp.h"

using namespace mef::openpsa;

int main(int argc, char** argv) {
    const bool ccf = argc > 1 && std::atoi(argv[1]) != 0;
    Model model;
    std::vector<std::unique_ptr<ConstantExpression>> exprs;
    auto basic = [&](const char* name, double p) {
        exprs.push_back(std::make_unique<ConstantExpression>(p));
        BasicEvent* e = model.Add(std::make_unique<BasicEvent>(name));
        e->expression(exprs.back().get());
        return e;
    };
    BasicEvent* pump_a = basic("pump-a", 0.05);
    BasicEvent* pump_b = basic("pump-b", 0.05);
    BasicEvent* pump_c = basic("pump-c", 0.05);
    BasicEvent* valve = basic("valve", 0.02);
    BasicEvent* power = basic("power", 0.003);
    BasicEvent* unused = basic("unused", 0.5);   // in the model's table, referenced by no gate
    (void)unused;
    HouseEvent* maintenance = model.Add(std::make_unique<HouseEvent>("maintenance"));
    maintenance->state(true);
    HouseEvent* bypass = model.Add(std::make_unique<HouseEvent>("bypass"));   // unset: false (event.h:56)

    CcfGroup pumps;                               // three members; 1-, 2- and 3-fold failures
    pumps.AddMember(pump_a); pumps.AddMember(pump_b); pumps.AddMember(pump_c);
    pumps.ApplyModel({{1, 0.04}, {2, 0.006}, {3, 0.004}});

    auto gate = [&](const char* name, Connective c, Formula::ArgSet args, std::optional<int> lo = {}, std::optional<int> hi = {}) {
        Gate* g = model.Add(std::make_unique<Gate>(name));
        g->formula(std::make_unique<Formula>(c, std::move(args), lo, hi));
        return g;
    };
    Gate* two_pumps = gate("two-pumps", kAtleast, {{false, pump_a}, {false, pump_b}, {false, pump_c}}, 2);
    Gate* isolation = gate("isolation", kOr, {{false, valve}, {false, bypass}});
    Gate* cooling = gate("cooling", kOr, {{false, two_pumps}, {false, isolation}});
    Gate* supply = gate("supply", kAnd, {{false, maintenance}, {false, power}});
    Gate* top1 = gate("loss-of-cooling", kOr, {{false, cooling}, {false, supply}});
    Gate* top2 = gate("pump-a-only", kAnd, {{false, pump_a}, {true, pump_b}, {true, two_pumps}});
    FaultTree* ft1 = model.Add(std::make_unique<FaultTree>("cooling-system"));
    ft1->add_top_event(top1);
    FaultTree* ft2 = model.Add(std::make_unique<FaultTree>("diagnostics"));
    ft2->add_top_event(top2);

    Settings settings;
    settings.ccf_analysis(ccf).seed(argc > 3 ? std::atoi(argv[3]) : 7).num_trials(argc > 4 ? std::atoi(argv[4]) : 1000);
    canopy::GraphBuilder b = canopy::lower_model(model, settings);
    const canopy_graph_desc d = b.desc();
    std::printf("{\"ccf\": %d, \"n_basic\": %u, \"n_gates\": %u, \"p\": [", (int)ccf, d.n_basic, d.n_gates);
    for (uint32_t i = 0; i < d.n_basic; ++i) std::printf("%s%.17g", i ? ", " : "", d.p[i]);
    std::printf("], \"op\": [");
    for (uint32_t i = 0; i < d.n_gates; ++i) std::printf("%s%u", i ? ", " : "", d.op[i]);
    std::printf("], \"kmin\": [");
    for (uint32_t i = 0; i < d.n_gates; ++i) std::printf("%s%u", i ? ", " : "", d.kmin[i]);
    std::printf("], \"kmax\": [");
    for (uint32_t i = 0; i < d.n_gates; ++i) std::printf("%s%u", i ? ", " : "", d.kmax[i]);
    std::printf("], \"child_off\": [");
    for (uint32_t i = 0; i <= d.n_gates; ++i) std::printf("%s%llu", i ? ", " : "", (unsigned long long)d.child_off[i]);
    std::printf("], \"child\": [");
    for (uint64_t i = 0; i < d.n_edges; ++i) std::printf("%s%u", i ? ", " : "", d.child[i]);
    std::printf("], \"roots\": [");
    for (uint32_t i = 0; i < d.n_roots; ++i) std::printf("%s%u", i ? ", " : "", d.roots[i]);
    std::printf("]");
    if (argc > 2 && !std::strcmp(argv[2], "run")) {
        canopy::McSettings mc = canopy::mc_settings(settings);
        mc.tally_all_gates = true;
        try {
            canopy::Result r = canopy::quantify(d, mc);
            std::printf(", \"seed\": %llu, \"n\": %llu, \"tallies\": [", (unsigned long long)mc.seed, (unsigned long long)mc.num_trials);
            for (std::size_t i = 0; i < r.tallies.size(); ++i) std::printf("%s%llu", i ? ", " : "", (unsigned long long)r.tallies[i]);
            std::printf("]");
        } catch (const canopy::Error& e) {
            std::printf(", \"run_status\": %d", (int)e.status());
        }
    }
    std::printf("}\n");
    return 0;
}
