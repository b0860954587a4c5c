// Host lowering pass: MEF gate DAG (canopy_graph_desc) -> topologically layered CSR arrays.
//
// The reference has no lowering pass (ABSENT); what it does define, and what is
// honoured here, is the input contract:
//   - connective set and numbering            reference src/mef/openpsa/event/event.h:143-157
//   - arity / min / max rules per connective  reference src/mef/openpsa/event/event.cpp:140-191
//   - min > max is a validity error           reference src/mef/openpsa/event/event.cpp:133-136
//   - no complemented arg under not/null, no constant under not   event.cpp:193-204
//   - formula arguments are unique by event   reference src/mef/openpsa/event/event.cpp:35-45
//   - per-argument complement flag            reference src/mef/openpsa/event/formula.h:12-15
//   - gate graph must be acyclic              reference src/mef/openpsa/cycle.h:114-131
//   - house events are Boolean constants      reference src/mef/openpsa/event/event.h:34-57
//
// Passes: validate -> topological order (cycle check) -> resolve each gate to a canonical reference
// (constant | possibly complemented basic event | possibly complemented materialised gate) with
// constant folding and not/null aliasing -> dead-code elimination from the tallied outputs ->
// levelisation -> per-layer ordering by descending fan-in -> CSR emission with device node ids.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <deque>
#include <memory>
#include <numeric>
#include <thread>

#include "plan.h"
#include "sparse_table.h"

namespace canopy {
namespace {

enum RefKind : uint8_t { kConst = 0, kEvent = 1, kGate = 2 };
struct Ref {
    uint32_t id;   // kConst: 0/1; kEvent: original event index; kGate: materialised gate index
    uint8_t kind;
    uint8_t comp;
};
inline Ref make_const(bool v) { return Ref{v ? 1u : 0u, kConst, 0}; }
inline Ref negate(Ref r) {
    if (r.kind == kConst) r.id ^= 1u; else r.comp ^= 1u;
    return r;
}

struct MGate {          // materialised (device) gate before renumbering
    uint32_t first, count;  // slice of mchild
    uint8_t op;
    uint16_t lo, hi;
    uint32_t level;
    uint32_t orig;      // an original gate index it came from (for debugging), or ~0u for buffers
};

[[noreturn]] void fail(canopy_status c, const std::string& m) { throw Error(c, m); }

void validate(const canopy_graph_desc& d) {
    const uint32_t B = d.n_basic, G = d.n_gates;
    if ((uint64_t)B + G >= CANOPY_CONST_FALSE) fail(CANOPY_E_INVALID, "too many nodes");
    if (B && !d.p) fail(CANOPY_E_INVALID, "p is null");
    if (G && (!d.op || !d.child_off)) fail(CANOPY_E_INVALID, "op/child_off is null");
    if (d.n_edges && !d.child) fail(CANOPY_E_INVALID, "child is null");
    if (d.n_roots && !d.roots) fail(CANOPY_E_INVALID, "roots is null");
    if (d.n_edges >= 0xFFFFFFFFull) fail(CANOPY_E_UNSUPPORTED, "more than 2^32-1 edges");
    for (uint32_t e = 0; e < B; ++e) {
        double p = d.p[e];
        if (!(p >= 0.0) || !(p <= 1.0))
            fail(CANOPY_E_DOMAIN, "basic event " + std::to_string(e) + ": probability outside [0,1]");
    }
    if (G) {
        if (d.child_off[0] != 0) fail(CANOPY_E_INVALID, "child_off[0] != 0");
        if (d.child_off[G] != d.n_edges) fail(CANOPY_E_INVALID, "child_off[n_gates] != n_edges");
    } else if (d.n_edges) {
        fail(CANOPY_E_INVALID, "edges without gates");
    }
    std::vector<uint32_t> ids;
    for (uint32_t g = 0; g < G; ++g) {
        if (d.child_off[g + 1] < d.child_off[g]) fail(CANOPY_E_INVALID, "child_off not monotone");
        const uint64_t o = d.child_off[g], n = d.child_off[g + 1] - o;
        if (d.child_off[g + 1] > d.n_edges) fail(CANOPY_E_INVALID, "child_off beyond n_edges");
        const uint8_t op = d.op[g];
        auto where = [g] { return "gate " + std::to_string(g) + ": "; };   // built only when a rule fails
        if (op >= CANOPY_NUM_CONNECTIVES) fail(CANOPY_E_INVALID, where() + "unknown connective");
        const unsigned lo = d.kmin ? d.kmin[g] : 0, hi = d.kmax ? d.kmax[g] : 0;
        switch (op) {  // reference event.cpp:140-191
            case CANOPY_AND: case CANOPY_OR: case CANOPY_NAND: case CANOPY_NOR:
                if (n < 2) fail(CANOPY_E_VALIDITY, where() + "The connective must have 2 or more arguments.");
                break;
            case CANOPY_NOT: case CANOPY_NULL:
                if (n != 1) fail(CANOPY_E_VALIDITY, where() + "The connective must have only one argument.");
                break;
            case CANOPY_XOR: case CANOPY_IFF: case CANOPY_IMPLY:
                if (n != 2) fail(CANOPY_E_VALIDITY, where() + "The connective must have exactly 2 arguments.");
                break;
            case CANOPY_ATLEAST:
                if (!d.kmin) fail(CANOPY_E_VALIDITY, where() + "The connective requires min number for its args.");
                if (lo < 2) fail(CANOPY_E_VALIDITY, where() + "Min number cannot be less than 2: " + std::to_string(lo));
                if (n <= lo) fail(CANOPY_E_VALIDITY, where() + "The connective must have more arguments than its min number:" +
                                                        std::to_string(n) + " <= " + std::to_string(lo));
                break;
            case CANOPY_CARDINALITY:
                if (!d.kmin || !d.kmax) fail(CANOPY_E_VALIDITY, where() + "The connective requires min and max numbers for args.");
                if (lo > hi) fail(CANOPY_E_VALIDITY, where() + "The connective min number cannot be greater than max number:" +
                                                     std::to_string(lo) + " > " + std::to_string(hi));
                if (n == 0) fail(CANOPY_E_VALIDITY, where() + "The connective requires one or more arguments.");
                if (n < hi) fail(CANOPY_E_VALIDITY, where() + "The connective max number cannot be greater than the number of arguments: " +
                                                    std::to_string(hi) + " > " + std::to_string(n));
                break;
        }
        ids.clear();
        for (uint64_t j = 0; j < n; ++j) {
            const uint32_t c = d.child[o + j], id = c & CANOPY_CHILD_ID_MASK;
            const bool is_const = id == CANOPY_CONST_FALSE || id == CANOPY_CONST_TRUE;
            if (!is_const && id >= B + G) fail(CANOPY_E_INVALID, where() + "argument id out of range");
            if (!is_const && id == B + g) fail(CANOPY_E_CYCLE, where() + "gate is its own argument");
            if ((c & CANOPY_CHILD_COMPLEMENT) && (op == CANOPY_NOT || op == CANOPY_NULL))
                fail(CANOPY_E_INVALID, where() + "Invalid nesting of a complement arg.");   // event.cpp:194-197
            if (is_const && op == CANOPY_NOT)
                fail(CANOPY_E_INVALID, where() + "Invalid nesting of a constant arg.");     // event.cpp:198-203
            ids.push_back(id);
        }
        std::sort(ids.begin(), ids.end());
        if (std::adjacent_find(ids.begin(), ids.end()) != ids.end())
            fail(CANOPY_E_VALIDITY, where() + "Duplicate Element Error: duplicate argument");  // event.cpp:35-45
    }
    for (uint32_t r = 0; r < d.n_roots; ++r)
        if (d.roots[r] >= B + G) fail(CANOPY_E_INVALID, "root id out of range");
}

// Children-first order of all gates; throws on a cycle (reference cycle.h:114-131 does the same DFS
// with NodeMark temporary/permanent marks).
std::vector<uint32_t> topo_order(const canopy_graph_desc& d) {
    const uint32_t B = d.n_basic, G = d.n_gates;
    std::vector<uint8_t> mark(G, 0);
    std::vector<uint32_t> order; order.reserve(G);
    std::vector<std::pair<uint32_t, uint64_t>> st;
    for (uint32_t g0 = 0; g0 < G; ++g0) {
        if (mark[g0]) continue;
        mark[g0] = 1; st.push_back({g0, d.child_off[g0]});
        while (!st.empty()) {
            auto& top = st.back();
            const uint32_t g = top.first;
            if (top.second == d.child_off[g + 1]) { mark[g] = 2; order.push_back(g); st.pop_back(); continue; }
            const uint32_t id = d.child[top.second++] & CANOPY_CHILD_ID_MASK;
            if (id == CANOPY_CONST_FALSE || id == CANOPY_CONST_TRUE || id < B) continue;
            const uint32_t c = id - B;
            if (mark[c] == 1) fail(CANOPY_E_CYCLE, "Cycle Error: gate " + std::to_string(c) + " is reachable from itself");
            if (mark[c] == 0) { mark[c] = 1; st.push_back({c, d.child_off[c]}); }
        }
    }
    return order;
}

}  // namespace

void lower_graph(const canopy_graph_desc& d, const canopy_plan_opts& opts, Lowered& L) {
    if (opts.threshold_bits != 32 && opts.threshold_bits != 64)
        fail(CANOPY_E_INVALID, "threshold_bits must be 32 or 64");
    if (opts.sampler == 3)
        fail(CANOPY_E_UNSUPPORTED, "sampling spec 3 was retired in favour of spec 4 (CANOPY_SAMPLER_SPARSE): tallies taken under it cannot be continued");
    if (opts.sampler != CANOPY_SAMPLER_DIRECT && opts.sampler != CANOPY_SAMPLER_SLICED && opts.sampler != CANOPY_SAMPLER_SPARSE)
        fail(CANOPY_E_INVALID, "sampler must be CANOPY_SAMPLER_DIRECT (1), CANOPY_SAMPLER_SLICED (2) or CANOPY_SAMPLER_SPARSE (4)");
    validate(d);
    const uint32_t B = d.n_basic, G = d.n_gates;
    const bool tally_all = opts.tally_all_gates != 0;
    const bool wide = opts.threshold_bits == 64;
    L = Lowered{};
    L.opts = opts;

    // ---- basic events: quantise (DESIGN.md SPEC v1 "Threshold quantisation") --------------------
    std::vector<Ref> ev_ref(B);
    std::vector<uint64_t> thr(B, 0);
    L.event_class.assign(B, 0);
    for (uint32_t e = 0; e < B; ++e) {
        const double p = d.p[e];
        if (p <= 0.0) { ev_ref[e] = make_const(false); continue; }
        if (p >= 1.0) { ev_ref[e] = make_const(true); L.event_class[e] = 1; continue; }
        thr[e] = wide ? (uint64_t)(p * 18446744073709551616.0) : (uint64_t)(p * 4294967296.0);
        if (thr[e] == 0) { ev_ref[e] = make_const(false); continue; }  // u < 0 never holds
        ev_ref[e] = Ref{e, kEvent, 0};
        L.event_class[e] = 2;
    }

    // ---- resolve gates in children-first order --------------------------------------------------
    const std::vector<uint32_t> order = topo_order(d);
    std::vector<Ref> gate_ref(G);
    std::vector<MGate> mg; mg.reserve(G);
    std::vector<Ref> mchild; mchild.reserve(d.n_edges);
    std::vector<Ref> args;
    uint32_t n_const = 0, n_alias = 0;

    auto resolve = [&](uint32_t code) -> Ref {
        const uint32_t id = code & CANOPY_CHILD_ID_MASK;
        Ref r = id == CANOPY_CONST_TRUE ? make_const(true)
              : id == CANOPY_CONST_FALSE ? make_const(false)
              : id < B ? ev_ref[id] : gate_ref[id - B];
        return (code & CANOPY_CHILD_COMPLEMENT) ? negate(r) : r;
    };
    auto materialise = [&](uint8_t op, uint16_t lo, uint16_t hi, uint32_t orig) -> Ref {
        MGate m{(uint32_t)mchild.size(), (uint32_t)args.size(), op, lo, hi, 0, orig};
        mchild.insert(mchild.end(), args.begin(), args.end());
        mg.push_back(m);
        return Ref{(uint32_t)mg.size() - 1, kGate, 0};
    };

    for (uint32_t g : order) {
        const uint64_t o = d.child_off[g];
        const uint32_t n = (uint32_t)(d.child_off[g + 1] - o);
        uint8_t op = d.op[g];
        bool inv = false;
        args.clear();
        for (uint32_t j = 0; j < n; ++j) args.push_back(resolve(d.child[o + j]));
        int64_t lo = 0, hi = 0;
        switch (op) {
            case CANOPY_NAND: op = CANOPY_AND; inv = true; break;
            case CANOPY_NOR: op = CANOPY_OR; inv = true; break;
            case CANOPY_IFF: op = CANOPY_XOR; inv = true; break;
            case CANOPY_IMPLY: op = CANOPY_OR; args[0] = negate(args[0]); break;
            case CANOPY_NOT: op = CANOPY_NULL; inv = true; break;
            case CANOPY_ATLEAST: lo = d.kmin[g]; hi = n; op = CANOPY_CARDINALITY; break;
            case CANOPY_CARDINALITY: lo = d.kmin[g]; hi = d.kmax[g]; break;
            default: break;
        }
        Ref res{};
        bool done = false;
        if (op == CANOPY_CARDINALITY) {
            // drop constants, shifting the window by the number of true constants
            int64_t t = 0; size_t w = 0;
            for (const Ref& a : args) { if (a.kind == kConst) t += a.id; else args[w++] = a; }
            args.resize(w);
            const int64_t n2 = (int64_t)args.size();
            lo = std::max<int64_t>(0, lo - t); hi = std::min<int64_t>(hi - t, n2);
            if (hi < 0 || lo > hi) { res = make_const(false); done = true; }
            else if (lo == 0 && hi == n2) { res = make_const(true); done = true; }
            else if (lo == 1 && hi == n2) op = CANOPY_OR;
            else if (lo == n2) op = CANOPY_AND;                                   // hi == n2 too
            else if (hi == 0) { op = CANOPY_OR; inv = !inv; }                     // none true
            else if (lo == 0 && hi == n2 - 1) { op = CANOPY_AND; inv = !inv; }    // not all true
        }
        if (!done && (op == CANOPY_AND || op == CANOPY_OR)) {
            const bool absorbing = op == CANOPY_OR;  // constant that decides the result
            size_t w = 0; bool decided = false;
            for (const Ref& a : args) {
                if (a.kind == kConst) { if ((a.id != 0) == absorbing) decided = true; }
                else args[w++] = a;
            }
            args.resize(w);
            if (decided) { res = make_const(absorbing); done = true; }
            else if (args.empty()) { res = make_const(!absorbing); done = true; }
            else if (args.size() == 1) { res = args[0]; done = true; ++n_alias; }
            else res = materialise(op == CANOPY_AND ? kDevAnd : kDevOr, 0, 0, g), done = true;
        }
        if (!done && op == CANOPY_XOR) {
            size_t w = 0;
            for (const Ref& a : args) { if (a.kind == kConst) inv ^= (a.id != 0); else args[w++] = a; }
            args.resize(w);
            if (args.empty()) res = make_const(false);
            else if (args.size() == 1) { res = args[0]; ++n_alias; }
            else {
                // move complements out of the operands: (!a ^ b) == !(a ^ b)
                for (Ref& a : args) if (a.comp) { a.comp = 0; inv = !inv; }
                res = materialise(kDevXor, 0, 0, g);
            }
            done = true;
        }
        if (!done && op == CANOPY_NULL) { res = args[0]; ++n_alias; done = true; }
        if (!done) res = materialise(kDevRange, (uint16_t)lo, (uint16_t)hi, g);
        if (inv) res = negate(res);
        if (res.kind == kConst) ++n_const;
        gate_ref[g] = res;
    }

    // ---- outputs ---------------------------------------------------------------------------------
    const uint32_t n_tallies = d.n_roots + (tally_all ? G : 0);
    std::vector<Ref> out_ref(n_tallies);
    for (uint32_t r = 0; r < d.n_roots; ++r) {
        const uint32_t id = d.roots[r];
        out_ref[r] = id < B ? ev_ref[id] : gate_ref[id - B];
    }
    if (tally_all) for (uint32_t g = 0; g < G; ++g) out_ref[d.n_roots + g] = gate_ref[g];
    // an output that resolved to a bare basic event needs a device gate to carry its tally: a 1-input OR
    std::vector<uint32_t> buffer_of(B, ~0u);
    for (Ref& r : out_ref) {
        if (r.kind != kEvent) continue;
        if (buffer_of[r.id] == ~0u) {
            args.assign(1, Ref{r.id, kEvent, 0});
            buffer_of[r.id] = materialise(kDevOr, 0, 0, ~0u).id;
        }
        r = Ref{buffer_of[r.id], kGate, r.comp};
    }

    // ---- dead-code elimination: materialised gates are in children-first creation order ----------
    const uint32_t M = (uint32_t)mg.size();
    std::vector<uint8_t> live(M, 0), stored(M, 0), tallied(M, 0), ev_live(B, 0);
    for (const Ref& r : out_ref) if (r.kind == kGate) { live[r.id] = 1; tallied[r.id] = 1; }
    for (uint32_t m = M; m-- > 0;) {
        if (!live[m]) continue;
        for (uint32_t j = 0; j < mg[m].count; ++j) {
            const Ref& a = mchild[mg[m].first + j];
            if (a.kind == kGate) { live[a.id] = 1; stored[a.id] = 1; }
            else if (a.kind == kEvent) ev_live[a.id] = 1;
        }
    }
    // ---- levels ----------------------------------------------------------------------------------
    uint32_t max_level = 0;
    for (uint32_t m = 0; m < M; ++m) {
        if (!live[m]) continue;
        uint32_t lv = 0;
        for (uint32_t j = 0; j < mg[m].count; ++j) {
            const Ref& a = mchild[mg[m].first + j];
            if (a.kind == kGate) lv = std::max(lv, mg[a.id].level);
        }
        mg[m].level = lv + 1;
        max_level = std::max(max_level, lv + 1);
    }
    // ---- renumber ----------------------------------------------------------------------------------
    std::vector<uint32_t> ev_dev(B, ~0u);
    for (uint32_t e = 0; e < B; ++e) if (ev_live[e]) {
        ev_dev[e] = (uint32_t)L.event_orig.size();
        L.event_orig.push_back(e);
        L.event_thr32.push_back(wide ? (uint32_t)(thr[e] >> 32) : (uint32_t)thr[e]);
        L.event_thr64.push_back(wide ? thr[e] : (thr[e] << 32));
    }
    const uint32_t n_ev = L.n_events();
    std::vector<uint32_t> live_ids;
    for (uint32_t m = 0; m < M; ++m) if (live[m]) live_ids.push_back(m);
    // inside a layer: ring-path gates by descending fan-in, then the "wide" count gates (fan-in >= kWideFanin,
    // evaluated in several register passes outside the shared-memory ring) at the end of the layer
    auto is_wide = [&](uint32_t m) { return mg[m].op == kDevRange && mg[m].count >= kWideFanin; };
    std::stable_sort(live_ids.begin(), live_ids.end(), [&](uint32_t a, uint32_t b) {
        if (mg[a].level != mg[b].level) return mg[a].level < mg[b].level;
        if (is_wide(a) != is_wide(b)) return is_wide(b);
        return mg[a].count > mg[b].count;
    });
    std::vector<uint32_t> gate_dev(M, ~0u);
    for (uint32_t i = 0; i < live_ids.size(); ++i) gate_dev[live_ids[i]] = i;
    L.layer_off.assign(max_level + 1, 0);
    {
        const size_t ng = live_ids.size();
        L.child.reserve(mchild.size()); L.gate_rec.reserve(ng); L.gate_child_off.reserve(ng + 1);
        L.gate_op.reserve(ng); L.gate_flags.reserve(ng); L.gate_kmin.reserve(ng); L.gate_kmax.reserve(ng);
    }
    L.gate_child_off.push_back(0);
    uint32_t max_fanin = 0, n_stored = 0, n_tallied = 0;
    for (uint32_t i = 0; i < live_ids.size(); ++i) {
        const MGate& m = mg[live_ids[i]];
        L.layer_off[m.level] = i + 1;  // running end of this level
        const uint8_t flags = (stored[live_ids[i]] ? kFlagStore : 0) | (tallied[live_ids[i]] ? kFlagTally : 0);
        n_stored += stored[live_ids[i]]; n_tallied += tallied[live_ids[i]];
        if (m.count > 0xFFFFu)
            fail(CANOPY_E_UNSUPPORTED, "a gate with more than 65535 arguments after folding is not supported");
        // every device op is symmetric in its inputs, so the arguments are emitted plain-first,
        // complemented-last: the kernel then needs one count instead of one flag per row
        uint32_t n_comp = 0;
        for (uint32_t j = 0; j < m.count; ++j) n_comp += mchild[m.first + j].comp ? 1u : 0u;
        GateRec rec{(uint32_t)L.child.size(), m.count | (n_comp << 16), (uint32_t)m.op | ((uint32_t)flags << 8),
                    (uint32_t)m.lo | ((uint32_t)m.hi << 16)};
        for (int pass = 0; pass < 2; ++pass)
            for (uint32_t j = 0; j < m.count; ++j) {
                const Ref& a = mchild[m.first + j];
                if ((a.comp != 0) != (pass == 1)) continue;
                const uint32_t node = a.kind == kEvent ? ev_dev[a.id] : n_ev + gate_dev[a.id];
                L.child.push_back(node | (a.comp ? 0x80000000u : 0u));
            }
        L.gate_rec.push_back(rec);
        L.gate_child_off.push_back((uint32_t)L.child.size());
        L.gate_op.push_back(m.op); L.gate_flags.push_back(flags);
        L.gate_kmin.push_back(m.lo); L.gate_kmax.push_back(m.hi);
        max_fanin = std::max(max_fanin, m.count);
    }
    // layer_off[l] currently holds the end of level l (0 for empty slots); make it a proper prefix array
    for (uint32_t l = 1; l <= max_level; ++l) if (L.layer_off[l] < L.layer_off[l - 1]) L.layer_off[l] = L.layer_off[l - 1];
    // layer_off has max_level+1 entries: [0]=0 start of level 1, [l]=end of level l
    // ---- per-layer work split for the gate kernel ------------------------------------------------------
    // wide_off[l]: first wide gate of layer l.  split[l*(kGateWarps+1) + w]: the ring-path gates of layer l are
    // cut into kGateWarps contiguous chunks of equal cost (rows read + rows written + a per-gate constant),
    // so each warp of the slab's CTA streams one contiguous run of the child array.
    {
        const uint32_t nl = max_level;
        L.wide_off.resize(nl); L.split.assign((size_t)nl * (kGateWarps + 1), 0);
        L.split_child.assign((size_t)nl * (kGateWarps + 1), 0);
        L.dev_rec = L.gate_rec;
        L.dev_child.clear(); L.dev_child.reserve(L.child.size() + (size_t)nl * 4 * (kGateWarps + 1));
        for (uint32_t l = 0; l < nl; ++l) {
            const uint32_t a = L.layer_off[l], b = L.layer_off[l + 1];
            uint32_t wv = b;
            while (wv > a && L.gate_op[wv - 1] == kDevRange && (L.gate_rec[wv - 1].n_children & 0xFFFFu) >= kWideFanin) --wv;
            L.wide_off[l] = wv;
            std::vector<uint64_t> cost(wv - a + 1, 0);
            for (uint32_t g = a; g < wv; ++g) cost[g - a + 1] = cost[g - a] + (L.gate_rec[g].n_children & 0xFFFFu) + 2;
            uint32_t* sp = &L.split[(size_t)l * (kGateWarps + 1)];
            for (uint32_t w = 0; w <= kGateWarps; ++w) {
                const uint64_t target = cost[wv - a] * w / kGateWarps;
                sp[w] = a + (uint32_t)(std::lower_bound(cost.begin(), cost.end(), target) - cost.begin());
            }
            sp[0] = a; sp[kGateWarps] = wv;
            // Device child stream of the layer: the runs of the kGateWarps warps one after the other, each padded to a whole
            // number of four-row groups (the pad repeats the run's first row: fetched with its group, never read), then
            // the arguments of the layer's wide gates.  Every run therefore starts on a 16-byte boundary of dev_child[]
            // and the walk reads the four codes of a group with one 128-bit load.
            uint32_t* sc = &L.split_child[(size_t)l * (kGateWarps + 1)];
            for (uint32_t w = 0; w < kGateWarps; ++w) {
                sc[w] = (uint32_t)L.dev_child.size();
                for (uint32_t g = sp[w]; g < sp[w + 1]; ++g) {
                    L.dev_rec[g].child_off = (uint32_t)L.dev_child.size();
                    L.dev_child.insert(L.dev_child.end(), L.child.begin() + L.gate_child_off[g], L.child.begin() + L.gate_child_off[g + 1]);
                }
                while ((L.dev_child.size() - sc[w]) % 4u) L.dev_child.push_back(L.dev_child[sc[w]] & 0x7FFFFFFFu);
            }
            sc[kGateWarps] = (uint32_t)L.dev_child.size();
            for (uint32_t g = wv; g < b; ++g) {
                L.dev_rec[g].child_off = (uint32_t)L.dev_child.size();
                L.dev_child.insert(L.dev_child.end(), L.child.begin() + L.gate_child_off[g], L.child.begin() + L.gate_child_off[g + 1]);
            }
            while (L.dev_child.size() % 4u) L.dev_child.push_back(0u);
        }
    }
    // ---- tally mapping -------------------------------------------------------------------------------
    L.tally_kind.resize(n_tallies); L.tally_gate.assign(n_tallies, 0);
    for (uint32_t i = 0; i < n_tallies; ++i) {
        const Ref& r = out_ref[i];
        if (r.kind == kConst) L.tally_kind[i] = r.id ? kTallyTrue : kTallyFalse;
        else { L.tally_kind[i] = r.comp ? kTallyNotGate : kTallyGate; L.tally_gate[i] = gate_dev[r.id]; }
    }
    // ---- stats ---------------------------------------------------------------------------------------
    canopy_plan_stats& s = L.stats;
    s.n_basic_in = B; s.n_gates_in = G; s.n_edges_in = d.n_edges;
    s.n_basic_live = n_ev; s.n_gates_live = L.n_gates(); s.n_edges_live = L.child.size();
    s.n_gates_stored = n_stored; s.n_gates_tallied = n_tallied; s.n_layers = L.n_layers();
    s.max_fanin = max_fanin; s.mean_fanin = L.n_gates() ? (double)L.child.size() / L.n_gates() : 0.0;
    s.n_tallies = n_tallies; s.n_const_nodes = n_const; s.n_alias_nodes = n_alias;
    s.spec_version = (uint32_t)opts.sampler;  // spec number == sampler id
    build_sparse_classes(L);
}

// Spec 4: class of every live event and the sampler's order.  Pure functions of thr64.
void build_sparse_classes(Lowered& L) {
    const uint32_t n = L.n_events();
    L.event_k.clear(); L.event_perm.clear(); L.n_dense = 0;
    for (uint32_t& c : L.class_off) c = 0;
    if (L.opts.sampler != CANOPY_SAMPLER_SPARSE) return;
    L.event_k.resize(n);
    uint32_t count[8] = {0, 0, 0, 0, 0, 0, 0, 0};   // [0] dense, [1 + c] class k = 7 + c
    for (uint32_t i = 0; i < n; ++i) {
        const uint32_t k = sparse_class_k(L.event_thr64[i]);
        L.event_k[i] = (uint8_t)k;
        ++count[k ? k - 6 : 0];
    }
    uint32_t start[8], acc = 0;
    for (int c = 0; c < 8; ++c) { start[c] = acc; acc += count[c]; }
    L.n_dense = count[0];
    for (int c = 0; c < 7; ++c) L.class_off[c] = start[c + 1];
    L.class_off[7] = n;
    L.event_perm.resize(n);
    for (uint32_t i = 0; i < n; ++i) { const uint32_t k = L.event_k[i]; L.event_perm[start[k ? k - 6 : 0]++] = i; }
}

void shard_range(uint64_t begin, uint64_t n, int rank, int world, uint64_t* ob, uint64_t* on) {
    if (world < 1 || rank < 0 || rank >= world) fail(CANOPY_E_INVALID, "bad rank/world");
    // split on 4096-sample boundaries of the GLOBAL sample index (the kernels mask ragged slab ends, so any cut is
    // exact; aligned cuts keep every rank's interior blocks whole); the first and last rank absorb the ragged ends.
    const uint64_t end = begin + n;
    const uint64_t s0 = begin / kShardAlign, s1 = (end + kShardAlign - 1) / kShardAlign;  // 4096-sample block index range
    const uint64_t slabs = s1 - s0;
    auto cut = [&](int r) -> uint64_t {
        if (r <= 0) return begin;
        if (r >= world) return end;
        const uint64_t c = (s0 + (slabs * (uint64_t)r) / (uint64_t)world) * kShardAlign;
        return std::min(std::max(c, begin), end);
    };
    const uint64_t b = cut(rank), e = cut(rank + 1);
    *ob = b; *on = e - b;
}

void requantise(Lowered& L, const double* p, uint32_t n_basic) {
    if (n_basic != L.event_class.size()) throw Error(CANOPY_E_INVALID, "canopy_plan_update_p: n_basic does not match the plan");
    const bool wide = L.opts.threshold_bits == 64;
    std::vector<uint64_t> thr(n_basic, 0);
    for (uint32_t e = 0; e < n_basic; ++e) {
        const double x = p[e];
        if (!(x >= 0.0 && x <= 1.0)) throw Error(CANOPY_E_DOMAIN, "probability of basic event " + std::to_string(e) + " is outside [0,1]");
        uint8_t cls = 2;
        if (x <= 0.0) cls = 0;
        else if (x >= 1.0) cls = 1;
        else {
            thr[e] = wide ? (uint64_t)(x * 18446744073709551616.0) : (uint64_t)(x * 4294967296.0);
            if (thr[e] == 0) cls = 0;
        }
        if (cls != L.event_class[e])
            throw Error(CANOPY_E_INVALID, "canopy_plan_update_p: basic event " + std::to_string(e) +
                                              " changes between constant and sampled; lower the graph again");
    }
    for (size_t i = 0; i < L.event_orig.size(); ++i) {
        const uint64_t t = thr[L.event_orig[i]];
        L.event_thr32[i] = wide ? (uint32_t)(t >> 32) : (uint32_t)t;
        L.event_thr64[i] = wide ? t : (t << 32);
    }
    build_sparse_classes(L);
}

// ------------------------------------------------------------------------------------------------
// Plan cache.  Lowering a large graph costs tens of milliseconds of host time (c4: 42 ms), which a caller who
// quantifies the same model again and again (one plan per request, per uncertainty trial, per rank of a job) would pay
// every time.  canopy_plan_create keeps the last few lowered forms keyed by a 64-bit hash of the DESCRIPTOR (arrays
// and options, hashed before any lowering work) and verifies a hit by comparing the arrays themselves, so a collision
// can never hand out the wrong plan.  A hit costs one pass over the descriptor and one copy of the lowered arrays.
// ------------------------------------------------------------------------------------------------
namespace {
struct DescCopy {
    uint32_t n_basic = 0, n_gates = 0, n_roots = 0;
    uint64_t n_edges = 0;
    std::vector<double> p;
    std::vector<uint8_t> op;
    std::vector<uint16_t> kmin, kmax;
    std::vector<uint64_t> child_off;
    std::vector<uint32_t> child, roots;
    explicit DescCopy(const canopy_graph_desc& d)
        : n_basic(d.n_basic), n_gates(d.n_gates), n_roots(d.n_roots), n_edges(d.n_edges), p(d.p, d.p + d.n_basic),
          op(d.op, d.op + d.n_gates), kmin(d.n_gates, 0), kmax(d.n_gates, 0), child_off(d.child_off, d.child_off + d.n_gates + 1),
          child(d.child, d.child + d.n_edges), roots(d.roots, d.roots + d.n_roots) {
        if (d.kmin) kmin.assign(d.kmin, d.kmin + d.n_gates);
        if (d.kmax) kmax.assign(d.kmax, d.kmax + d.n_gates);
    }
    template <class T> static bool same(const std::vector<T>& v, const T* q) { return v.empty() || std::memcmp(v.data(), q, v.size() * sizeof(T)) == 0; }
    bool equals(const canopy_graph_desc& d) const {
        if (n_basic != d.n_basic || n_gates != d.n_gates || n_roots != d.n_roots || n_edges != d.n_edges) return false;
        if (!same(p, d.p) || !same(op, d.op) || !same(child_off, d.child_off) || !same(child, d.child) || !same(roots, d.roots)) return false;
        if (d.kmin ? !same(kmin, d.kmin) : std::any_of(kmin.begin(), kmin.end(), [](uint16_t v) { return v != 0; })) return false;
        if (d.kmax ? !same(kmax, d.kmax) : std::any_of(kmax.begin(), kmax.end(), [](uint16_t v) { return v != 0; })) return false;
        return true;
    }
};
struct CacheEntry { uint64_t hash; canopy_plan_opts opts; DescCopy desc; std::shared_ptr<const Lowered> low; };
std::mutex g_cache_mu;
std::deque<CacheEntry> g_cache;
constexpr size_t kCacheEntries = 4;

uint64_t mix64(uint64_t h, uint64_t v) {
    h ^= v; h *= 0x9E3779B97F4A7C15ull; h ^= h >> 29;
    return h;
}
uint64_t hash_bytes(uint64_t h, const void* data, size_t n) {
    const unsigned char* b = static_cast<const unsigned char*>(data);
    size_t i = 0;
    for (; i + 8 <= n; i += 8) { uint64_t v; std::memcpy(&v, b + i, 8); h = mix64(h, v); }
    uint64_t tail = 0;
    if (i < n) std::memcpy(&tail, b + i, n - i);
    return mix64(mix64(h, tail), n);
}
uint64_t hash_desc(const canopy_graph_desc& d, const canopy_plan_opts& o) {
    uint64_t h = 0x243F6A8885A308D3ull;
    const uint64_t head[7] = {d.n_basic, d.n_gates, d.n_roots, d.n_edges, (uint64_t)o.tally_all_gates, (uint64_t)o.threshold_bits, (uint64_t)o.sampler};
    h = hash_bytes(h, head, sizeof head);
    h = hash_bytes(h, d.p, (size_t)d.n_basic * 8);
    h = hash_bytes(h, d.op, d.n_gates);
    if (d.kmin) h = hash_bytes(h, d.kmin, (size_t)d.n_gates * 2);
    if (d.kmax) h = hash_bytes(h, d.kmax, (size_t)d.n_gates * 2);
    h = hash_bytes(h, d.child_off, ((size_t)d.n_gates + 1) * 8);
    h = hash_bytes(h, d.child, (size_t)d.n_edges * 4);
    h = hash_bytes(h, d.roots, (size_t)d.n_roots * 4);
    return h;
}
}  // namespace

void lower_graph_cached(const canopy_graph_desc& d, const canopy_plan_opts& opts, Lowered& out) {
    if (opts.no_plan_cache) { lower_graph(d, opts, out); return; }
    if (!d.p && d.n_basic) fail(CANOPY_E_INVALID, "descriptor array p is null");
    if ((!d.op || !d.child_off) && d.n_gates) fail(CANOPY_E_INVALID, "descriptor arrays op / child_off are null");
    if (!d.child_off || (!d.child && d.n_edges) || (!d.roots && d.n_roots)) { lower_graph(d, opts, out); return; }  // let validate() report it
    if (d.child_off[d.n_gates] != d.n_edges) { lower_graph(d, opts, out); return; }
    const uint64_t h = hash_desc(d, opts);
    std::shared_ptr<const Lowered> hit;
    {
        std::lock_guard<std::mutex> lock(g_cache_mu);
        for (auto it = g_cache.begin(); it != g_cache.end(); ++it)
            if (it->hash == h && it->opts.tally_all_gates == opts.tally_all_gates && it->opts.threshold_bits == opts.threshold_bits &&
                it->opts.sampler == opts.sampler && it->desc.equals(d)) {
                hit = it->low;
                if (it != g_cache.begin()) { CacheEntry e = std::move(*it); g_cache.erase(it); g_cache.push_front(std::move(e)); }
                break;
            }
    }
    if (hit) { out = *hit; out.opts = opts; out.stats.plan_cache_hit = 1; return; }
    lower_graph(d, opts, out);
    auto keep = std::make_shared<Lowered>(out);
    std::lock_guard<std::mutex> lock(g_cache_mu);
    g_cache.push_front(CacheEntry{h, opts, DescCopy(d), std::move(keep)});
    while (g_cache.size() > kCacheEntries) g_cache.pop_back();
}

void plan_cache_clear() {
    std::lock_guard<std::mutex> lock(g_cache_mu);
    g_cache.clear();
}

uint64_t fingerprint(const Lowered& L) {
    uint64_t h = 0xcbf29ce484222325ull;
    auto mix = [&](const void* data, size_t n) {
        const unsigned char* b = static_cast<const unsigned char*>(data);
        for (size_t i = 0; i < n; ++i) { h ^= b[i]; h *= 0x100000001b3ull; }
        const uint64_t len = n;  // length-delimited so adjacent arrays cannot trade bytes
        for (int i = 0; i < 8; ++i) { h ^= (len >> (8 * i)) & 0xFF; h *= 0x100000001b3ull; }
    };
    const int32_t o[3] = {L.opts.tally_all_gates, L.opts.threshold_bits, L.opts.sampler};
    mix(o, sizeof o);
    if (L.opts.sampler == CANOPY_SAMPLER_SPARSE) {  // the constants of spec 4 are part of what a restart record must match
        const int32_t c[4] = {CANOPY_SLICED_HEAD_BITS, CANOPY_SPARSE_MIN_ZEROS, CANOPY_SPARSE_CDF_LEN, CANOPY_SPARSE_MAX_LOG2};
        mix(c, sizeof c);
    }
    mix(L.event_orig.data(), L.event_orig.size() * 4);
    mix(L.event_thr32.data(), L.event_thr32.size() * 4);
    mix(L.event_thr64.data(), L.event_thr64.size() * 8);
    mix(L.gate_rec.data(), L.gate_rec.size() * sizeof(GateRec));
    mix(L.child.data(), L.child.size() * 4);
    mix(L.layer_off.data(), L.layer_off.size() * 4);
    mix(L.tally_kind.data(), L.tally_kind.size());
    mix(L.tally_gate.data(), L.tally_gate.size() * 4);
    return h;
}

}  // namespace canopy
