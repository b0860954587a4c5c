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
    {   // probability source of the uncertainty loop: host only up to the inner run
        std::vector<canopy_deviate> d(3);
        d[0].kind = CANOPY_DEV_UNIFORM; d[0].a = 0.05; d[0].b = 0.15;
        d[1].kind = CANOPY_DEV_LOGNORMAL_EF; d[1].a = 0.2; d[1].b = 2.0; d[1].c = 0.95;
        d[2].kind = CANOPY_DEV_CONSTANT; d[2].a = 0.3;
        canopy::ProbabilitySource src(d);
        const auto m = src.mean(), x = src.sample(1, 0), y = src.sample(1, 0), z = src.sample(1, 1);
        std::printf("psource: n=%u deviates=%u mean_ok=%d repeat=%d fresh=%d\n", src.n_basic(), src.n_deviates(),
                    m[0] == 0.1 && m[1] == 0.2 && m[2] == 0.3, x == y, x != z && z[2] == 0.3);
        try { canopy::McSettings s; s.num_trials = 1000; canopy::uncertainty(plan, src, s, 2, 9); }
        catch (const canopy::Error& e) { std::printf("uncertainty: status %d\n", e.status()); }
    }
    return 0;
}
