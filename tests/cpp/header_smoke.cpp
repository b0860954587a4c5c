#include "canopy-cpp.h"
#include <cstdio>
int main() {
    canopy::GraphBuilder b;
    auto e0 = b.basic_event(0.1), e1 = b.basic_event(0.2), e2 = b.basic_event(0.3);
    auto g = b.gate(canopy::kAtleast, {{e0}, {e1}, {e2, true}}, 2);
    b.root(g);
    canopy::Plan plan(b.desc(), true);
    std::printf("layers=%u gates=%u tallies=%u\n", plan.stats().n_layers, plan.stats().n_gates_live, plan.stats().n_tallies);
    try { canopy::McSettings s; s.num_trials = 1000; plan.run(s); } catch (const canopy::Error& e) { std::printf("run: status %d (%s)\n", e.status(), e.what()); }
    try { canopy::GraphBuilder c; auto x = c.basic_event(0.5); c.root(c.gate(canopy::kAnd, {{x}})); canopy::Plan p(c.desc()); }
    catch (const canopy::Error& e) { std::printf("validity: status %d\n", e.status()); }
    canopy::Session sess(plan, 7, 4096);
    std::printf("session: fingerprint %d done=%llu tallies=%zu\n", sess.record.plan_fingerprint == plan.fingerprint(),
                (unsigned long long)sess.done(), sess.tallies.size());
    try { plan.update_p({0.1, 0.2, 1.0}); } catch (const canopy::Error& e) { std::printf("update_p: status %d\n", e.status()); }
    return 0;
}
