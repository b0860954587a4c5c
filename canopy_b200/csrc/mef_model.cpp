// Open-PSA MEF XML -> flat gate DAG (canopy_graph_desc), host only, no third-party dependency.
//
// Stands in for the reference loader, which needs Boost + libxml2 and does not compile as shipped
// (SURVEY.md section 0).  What it follows, by reference file:line (src/mef/openpsa/):
//   - two passes: register every name, then define bodies                       initializer.cpp:252-297, 415-473, 655-682
//   - fault trees / components / model-data containers                          initializer.cpp:712-770
//   - formula extraction: connective names (event.h:164-166), a bare event reference or <constant> is a NULL
//     pass-through, <not> around an argument sets its complement flag, <constant> arguments are
//     HouseEvent::kTrue/kFalse                                                   initializer.cpp:772-851
//   - duplicate ids across gates / basic events / house events                   model.h:151-155
//   - house event without a body is false                                        event.h:33
//   - top events = gates not used as an argument inside their own tree           fault_tree.h:151-186
//   - BasicEvent::p() = Expression::value(), the MEAN of the expression          event.h:93-96, expression.h:74
//     (constants, parameters, arithmetic, built-ins, mean of the random deviates: expr/*.h)
//   - CCF groups: beta-factor / MGL / alpha-factor / phi-factor probabilities and the proxy OR gate over the
//     generated CCF events that replaces each member when Settings::ccf_analysis is on
//                                                                                ccf_group.h:84-175, 215-260, 330-440
//   - event trees: the data model of event_tree.h:18-224 and instruction.h:21-215.  The reference has NO walker
//     (SURVEY.md 8a row a8); the walk below follows the Open-PSA MEF semantics: a sequence is reached by a set of
//     paths; along a path collect-formula instructions are conjoined, collect-expression instructions multiply
//     into one extra basic event, set-house-event overrides apply to the formulas collected after them.
// Not covered (CANOPY_E_UNSUPPORTED): nested formulas (the reference's Formula::Arg cannot hold them either,
// formula.h:9-15), the 11-argument periodic-test, extern functions, alignments/substitutions, XInclude.
// Roles: public / private elements, base paths of fault trees and nested components, and the reference's lookup order
// (local scope, public id, full path) are followed (element.h:300-379, initializer.cpp:42-57,1378-1437).
#include <algorithm>
#include <cmath>
#include <fstream>
#include <map>
#include <set>
#include <sstream>

#include "mef_xml.h"
#include "model.h"


namespace canopy {
namespace {

struct MefOpts { bool ccf_analysis = false; double mission_time = 8760.0; };  // settings.h:320,330

struct ArgRef {
    enum Kind : uint8_t { kGate, kBasic, kHouse, kConst } kind;
    bool complement;
    uint32_t idx;  // kConst: 0 false / 1 true
};
struct GateDef {
    std::string name;
    uint8_t conn = CANOPY_NULL;
    uint16_t kmin = 0, kmax = 0;
    std::vector<ArgRef> args;
    int tree = -1;               // fault-tree ordinal, -1 for synthetic gates
    const XNode* body = nullptr; // <define-gate>, null for synthetic gates
    std::string base_path;       // scope its formula's references are resolved in (initializer.cpp:483)
};
struct BasicDef { std::string name, base_path; const XNode* body = nullptr; double p = 0.0; bool has_p = false; int proxy_gate = -1; uint32_t node = 0; };  // node: tape node of p
struct HouseDef { std::string name; bool state = false; };
struct ParamDef { std::string name, base_path; const XNode* body = nullptr; double value = 0.0; int mark = 0; uint32_t node = 0; };
struct CcfDef { std::string name, model, base_path; const XNode* node = nullptr; std::vector<uint32_t> members; };
struct EventTreeDef {
    std::string name;
    const XNode* node = nullptr;
    std::set<std::string> functional_events;
    std::map<std::string, const XNode*> branches;
    std::set<std::string> sequences;
};
enum class Kind { kGate, kBasic, kHouse };
struct Sym { Kind kind; uint32_t idx; };

// What an expression may look at while an event tree is walked (test-initiating-event, test-functional-event).
struct EvalCtx {
    const std::map<std::string, std::string>* functional_states = nullptr;
    const std::string* initiating_event = nullptr;
};

[[noreturn]] void bad(canopy_status c, const XNode& n, const std::string& msg) {
    throw Error(c, "line " + std::to_string(n.line) + ": " + msg);
}
const std::string& need_attr(const XNode& n, const char* k) {
    const std::string* s = n.attr(k);
    if (!s || s->empty()) bad(CANOPY_E_VALIDITY, n, "<" + n.name + "> needs attribute '" + k + "'");
    return *s;
}
const std::string& need_name(const XNode& n) { return need_attr(n, "name"); }
bool is_meta(const XNode& n) { return n.name == "label" || n.name == "attributes"; }
std::vector<const XNode*> kids_of(const XNode& n) {  // GetNonAttributeElements, initializer.cpp:107-120
    std::vector<const XNode*> v;
    for (auto& k : n.kids) if (!is_meta(*k)) v.push_back(k.get());
    return v;
}
const XNode* first_kid(const XNode& n) {
    for (auto& k : n.kids) if (!is_meta(*k)) return k.get();
    return nullptr;
}
bool parse_bool(const XNode& n, const std::string& v) {
    if (v == "true" || v == "1") return true;
    if (v == "false" || v == "0") return false;
    bad(CANOPY_E_VALIDITY, n, "bad boolean '" + v + "'");
}
double parse_number(const XNode& n, const std::string& v) {
    char* end = nullptr;
    const double x = std::strtod(v.c_str(), &end);
    if (end == v.c_str() || *end) bad(CANOPY_E_VALIDITY, n, "bad number '" + v + "'");
    return x;
}
double combination_reciprocal(int n, int k) {  // 1 / C(n, k), algorithm.h:230-241
    if (n - k > k) k = n - k;
    double r = 1;
    for (int i = 1; i <= n - k; ++i) r *= (double)i / (double)(k + i);
    return r;
}

struct Builder {
    canopy_model& m;
    MefOpts opts;
    std::map<std::string, Sym> sym;
    std::vector<GateDef> gates;
    std::vector<BasicDef> basics;
    std::vector<HouseDef> houses;
    std::vector<ParamDef> params;
    std::map<std::string, uint32_t> param_idx;
    // Roles (element.h:300-379, initializer.cpp:42-57,98-106): an element's id is its name when public and
    // base_path.name when private; `sym` / `param_idx` are the id tables (Model::table<T>), `path_sym` / `param_path`
    // the full-path tables (Initializer::path_gates_ ...), which hold public and private elements alike.
    std::map<std::string, Sym> path_sym[3];   // one full-path table per Kind (path_gates_, path_basic_events_, path_house_events_)
    std::map<std::string, uint32_t> param_path;
    std::string cur_base;   // base path of the element whose body is being read (scope of its references)
    std::vector<CcfDef> ccfs;
    std::vector<EventTreeDef> trees;
    std::map<std::string, uint32_t> tree_idx;
    std::map<std::string, const XNode*> sequences, rules;
    std::vector<std::pair<std::string, std::string>> initiating;  // (initiating event, event tree)
    int n_fault_trees = 0;
    uint32_t n_xml_gates = 0;  // gates defined in the XML come first; synthetic gates are appended

    explicit Builder(canopy_model& model, const MefOpts& o) : m(model), opts(o) {}

    // ---- pass 1: register names ---------------------------------------------------------------------
    static bool role_of(const XNode& n, bool container_public) {  // GetRole, initializer.cpp:42-57
        const std::string* r = n.attr("role");
        if (!r || r->empty()) return container_public;
        if (*r != "public" && *r != "private") bad(CANOPY_E_VALIDITY, n, "bad role '" + *r + "'");
        return *r == "public";
    }
    static std::string make_id(const std::string& nm, const std::string& base, bool is_public) {  // Id::id(), element.h:341-344
        return is_public ? nm : base + "." + nm;
    }
    // Returns the id.  `in_path_table`: CCF members are added to the model only (initializer.cpp:1324-1339).
    std::string reg(const XNode& n, const std::string& nm, Kind k, uint32_t idx, const std::string& base, bool is_public,
                    bool in_path_table = true) {
        if (!is_public && base.empty()) bad(CANOPY_E_VALIDITY, n, "Elements cannot be private at model scope: " + nm);  // element.h: Role ctor
        const std::string id = make_id(nm, base, is_public);
        if (!sym.emplace(id, Sym{k, idx}).second)
            bad(CANOPY_E_VALIDITY, n, "Duplicate Element Error: Redefinition of event: " + id);  // model.h:151-155
        if (in_path_table && !base.empty()) path_sym[(int)k].emplace(base + "." + nm, Sym{k, idx});
        return id;
    }
    uint32_t add_basic(const XNode& n, const std::string& nm, const XNode* body, const std::string& base, bool is_public,
                       bool in_path_table = true) {
        BasicDef b; b.body = body; b.base_path = base;
        b.name = reg(n, nm, Kind::kBasic, (uint32_t)basics.size(), base, is_public, in_path_table);
        basics.push_back(b);
        return (uint32_t)basics.size() - 1;
    }
    void register_ccf(const XNode& n, const std::string& base, bool container_public) {  // initializer.cpp:375-401, 1324-1339
        const bool is_public = role_of(n, container_public);
        CcfDef c; c.name = make_id(need_name(n), base, is_public); c.model = need_attr(n, "model"); c.node = &n; c.base_path = base;
        if (c.model != "beta-factor" && c.model != "MGL" && c.model != "alpha-factor" && c.model != "phi-factor")
            bad(CANOPY_E_VALIDITY, n, "unknown CCF model '" + c.model + "'");
        for (auto& g : ccfs) if (g.name == c.name) bad(CANOPY_E_VALIDITY, n, "Duplicate Element Error: Redefinition of CCF group: " + c.name);
        const XNode* mem = nullptr;
        for (auto& k : n.kids) if (k->name == "members") mem = k.get();
        if (!mem) bad(CANOPY_E_VALIDITY, n, "CCF group " + c.name + " has no <members>");
        for (auto& e : mem->kids) {
            if (e->name != "basic-event") bad(CANOPY_E_VALIDITY, *e, "CCF members must be <basic-event>");
            c.members.push_back(add_basic(*e, need_name(*e), nullptr, base, is_public, false));  // the group's base path and role
        }
        ccfs.push_back(c);
    }
    void register_container(const XNode& c, int tree, const std::string& base, bool container_public) {  // initializer.cpp:712-770
        for (auto& k : c.kids) {
            const std::string& t = k->name;
            if (t == "define-gate") {
                GateDef g; g.body = k.get(); g.tree = tree; g.base_path = base;
                g.name = reg(*k, need_name(*k), Kind::kGate, (uint32_t)gates.size(), base, role_of(*k, container_public));
                gates.push_back(g);
            } else if (t == "define-basic-event") {
                add_basic(*k, need_name(*k), first_kid(*k), base, role_of(*k, container_public));
            } else if (t == "define-house-event") {
                HouseDef h;
                h.name = reg(*k, need_name(*k), Kind::kHouse, (uint32_t)houses.size(), base, role_of(*k, container_public));
                if (const XNode* b = first_kid(*k)) {
                    if (b->name != "constant" || !b->attr("value")) bad(CANOPY_E_VALIDITY, *b, "house event body must be <constant value=.../>");
                    h.state = parse_bool(*b, *b->attr("value"));
                }
                houses.push_back(h);
            } else if (t == "define-parameter") {
                const bool is_public = role_of(*k, container_public);
                if (!is_public && base.empty()) bad(CANOPY_E_VALIDITY, *k, "Elements cannot be private at model scope: " + need_name(*k));
                const std::string nm = make_id(need_name(*k), base, is_public);
                if (!param_idx.emplace(nm, (uint32_t)params.size()).second)
                    bad(CANOPY_E_VALIDITY, *k, "Duplicate Element Error: Redefinition of parameter: " + nm);
                if (!base.empty()) param_path.emplace(base + "." + need_name(*k), (uint32_t)params.size());
                ParamDef p; p.name = nm; p.base_path = base; p.body = first_kid(*k);
                if (!p.body) bad(CANOPY_E_VALIDITY, *k, "parameter " + nm + " has no expression");
                params.push_back(p);
            } else if (t == "define-CCF-group") {
                register_ccf(*k, base, container_public);
            } else if (t == "define-component") {  // DefineComponent, initializer.cpp:718-726: nested scope, inherited role
                if (base.empty()) bad(CANOPY_E_VALIDITY, *k, "<define-component> outside a fault tree");
                register_container(*k, tree, base + "." + need_name(*k), role_of(*k, container_public));
            } else if (is_meta(*k)) {
            } else {
                bad(CANOPY_E_UNSUPPORTED, *k, "unexpected element <" + t + ">");
            }
        }
    }
    void register_event_tree(const XNode& n) {  // initializer.cpp:684-710
        EventTreeDef et; et.name = need_name(n); et.node = &n;
        if (tree_idx.count(et.name)) bad(CANOPY_E_VALIDITY, n, "Duplicate Element Error: Redefinition of event tree: " + et.name);
        for (auto& k : n.kids) {
            if (k->name == "define-sequence") {
                const std::string& nm = need_name(*k);
                if (!sequences.emplace(nm, k.get()).second) bad(CANOPY_E_VALIDITY, *k, "Duplicate Element Error: Redefinition of sequence: " + nm);
                et.sequences.insert(nm);
            } else if (k->name == "define-branch") {
                if (!et.branches.emplace(need_name(*k), k.get()).second) bad(CANOPY_E_VALIDITY, *k, "Duplicate Element Error: Redefinition of branch: " + need_name(*k));
            } else if (k->name == "define-functional-event") {
                if (!et.functional_events.insert(need_name(*k)).second) bad(CANOPY_E_VALIDITY, *k, "Duplicate Element Error: Redefinition of functional event: " + need_name(*k));
            } else if (k->name == "initial-state" || is_meta(*k)) {
            } else bad(CANOPY_E_UNSUPPORTED, *k, "unexpected element <" + k->name + "> in an event tree");
        }
        tree_idx[et.name] = (uint32_t)trees.size();
        trees.push_back(et);
    }
    void load(const XNode& root) {  // initializer.cpp:415-473
        if (root.name != "opsa-mef") bad(CANOPY_E_IO, root, "document element must be <opsa-mef>");
        for (auto& k : root.kids) {
            const std::string& t = k->name;
            if (t == "define-fault-tree") register_container(*k, n_fault_trees++, need_name(*k), true);  // a fault tree is public; its name opens the scope
            else if (t == "model-data") register_container(*k, -1, "", true);
            else if (t == "define-CCF-group") register_ccf(*k, "", true);
            else if (t == "define-event-tree") register_event_tree(*k);
            else if (t == "define-initiating-event") {
                const std::string& nm = need_name(*k);
                for (auto& ie : initiating) if (ie.first == nm) bad(CANOPY_E_VALIDITY, *k, "Duplicate Element Error: Redefinition of initiating event: " + nm);
                const std::string* et = k->attr("event-tree");
                initiating.emplace_back(nm, et ? *et : std::string());
            } else if (t == "define-rule") {
                if (!rules.emplace(need_name(*k), k.get()).second) bad(CANOPY_E_VALIDITY, *k, "Duplicate Element Error: Redefinition of rule: " + need_name(*k));
            } else if (is_meta(*k)) {
            } else bad(CANOPY_E_UNSUPPORTED, *k, "<" + t + "> is not supported by this loader");
        }
    }

    // ---- expressions (expression.h:74-117; expr/*.h) -----------------------------------------------------
    // Every expression is compiled onto the model's tape (expr_tape.h) while its mean value -- Expression::value() -- is
    // computed for validation, so that the same probabilities can later be re-evaluated with sampled deviates
    // (Expression::Sample(), canopy_model_sample_p).  Ex = (mean value, tape node).
    struct Ex { double v; uint32_t id; };
    Ex emit(uint16_t op, const std::vector<Ex>& a, double v) {
        std::vector<uint32_t> ids;
        for (const Ex& x : a) ids.push_back(x.id);
        return Ex{v, m.tape.add(op, ids)};
    }
    Ex konst(double c) { return Ex{c, m.tape.constant(c)}; }
    // GetEntity, initializer.cpp:1378-1403: the local scope first, then the public id, or the full path as written.
    const uint32_t* find_parameter(const std::string& name) const {
        if (!cur_base.empty()) {
            auto it = param_path.find(cur_base + "." + name);
            if (it != param_path.end()) return &it->second;
        }
        if (name.find('.') == std::string::npos) {
            auto it = param_idx.find(name);
            return it == param_idx.end() ? nullptr : &it->second;
        }
        auto it = param_path.find(name);
        return it == param_path.end() ? nullptr : &it->second;
    }
    Ex define_parameter(uint32_t idx, const XNode& at) {
        ParamDef& p = params[idx];
        if (p.mark == 2) return Ex{p.value, p.node};   // one node per parameter: sampled once per trial, shared by its users
        if (p.mark == 1) bad(CANOPY_E_CYCLE, at, "Cycle Error: parameter " + p.name + " refers to itself");  // cycle.h:219-229
        p.mark = 1;
        const std::string saved = cur_base;
        cur_base = p.base_path;   // a parameter's own references resolve in ITS scope (initializer.cpp:499-503)
        const Ex x = ev(*p.body, EvalCtx{});
        cur_base = saved;
        params[idx].value = x.v; params[idx].node = x.id;
        params[idx].mark = 2;
        return x;
    }
    Ex parameter_value(const XNode& at, const std::string& name) {
        const uint32_t* idx = find_parameter(name);
        if (!idx) bad(CANOPY_E_VALIDITY, at, "Undefined Element Error: parameter " + name);
        return define_parameter(*idx, at);
    }
    // The reference defines EVERY registered parameter (tbd_, initializer.cpp:668-681), cycle-checks the whole parameter
    // table and validates every expression (initializer.cpp:1860-1874), used or not.  Those no event reached are compiled
    // here, after everything else, so that the tape nodes (and with them the deviate streams) of the used ones keep their
    // numbers.
    void define_unused_parameters() {
        for (uint32_t i = 0; i < params.size(); ++i)
            if (params[i].mark != 2) define_parameter(i, *params[i].body);
        cur_base.clear();
    }
    std::vector<Ex> eval_args(const XNode& e, const EvalCtx& cx, size_t lo, size_t hi) {
        std::vector<Ex> v;
        for (const XNode* k : kids_of(e)) v.push_back(ev(*k, cx));
        if (v.size() < lo || v.size() > hi)
            bad(CANOPY_E_VALIDITY, e, "<" + e.name + "> takes " + (lo == hi ? std::to_string(lo) : std::to_string(lo) + "+") +
                                          " argument(s), got " + std::to_string(v.size()));
        return v;
    }
    double eval(const XNode& e, const EvalCtx& cx) { return ev(e, cx).v; }
    Ex ev(const XNode& e, const EvalCtx& cx) {
        const std::string& t = e.name;
        constexpr size_t kMany = 1u << 20;
        if (t == "float" || t == "int") return konst(parse_number(e, need_attr(e, "value")));
        if (t == "bool") return konst(parse_bool(e, need_attr(e, "value")) ? 1.0 : 0.0);
        if (t == "parameter") return parameter_value(e, need_name(e));
        if (t == "system-mission-time") return konst(opts.mission_time);
        if (t == "pi") return konst(3.14159265358979323846);
        if (t == "test-initiating-event") {  // test_event.h: true iff the walk started from this initiating event
            if (!cx.initiating_event) bad(CANOPY_E_VALIDITY, e, "<test-initiating-event> outside an event-tree walk");
            return konst(*cx.initiating_event == need_name(e) ? 1.0 : 0.0);
        }
        if (t == "test-functional-event") {  // test_event.h: true iff the functional event is in the given state on this path
            if (!cx.functional_states) bad(CANOPY_E_VALIDITY, e, "<test-functional-event> outside an event-tree walk");
            auto it = cx.functional_states->find(need_name(e));
            return konst(it != cx.functional_states->end() && it->second == need_attr(e, "state") ? 1.0 : 0.0);
        }
        if (t == "neg") { const auto v = eval_args(e, cx, 1, 1); return emit(X_NEG, v, -v[0].v); }
        if (t == "add" || t == "sub" || t == "mul" || t == "div") {
            const auto v = eval_args(e, cx, t == "add" || t == "mul" ? 1 : 2, kMany);
            double r = v[0].v;
            for (size_t i = 1; i < v.size(); ++i) {
                if (t == "add") r += v[i].v; else if (t == "sub") r -= v[i].v; else if (t == "mul") r *= v[i].v;
                else { if (v[i].v == 0.0) bad(CANOPY_E_DOMAIN, e, "Division by 0."); r /= v[i].v; }  // numerical.h: Div::Validate
            }
            return emit(t == "add" ? X_ADD : t == "sub" ? X_SUB : t == "mul" ? X_MUL : X_DIV, v, r);
        }
        struct Unary { uint16_t op; double (*f)(double); };
        static const std::map<std::string, Unary> kUnary = {
            {"abs", {X_ABS, [](double x) { return std::fabs(x); }}}, {"acos", {X_ACOS, [](double x) { return std::acos(x); }}},
            {"asin", {X_ASIN, [](double x) { return std::asin(x); }}}, {"atan", {X_ATAN, [](double x) { return std::atan(x); }}},
            {"cos", {X_COS, [](double x) { return std::cos(x); }}}, {"sin", {X_SIN, [](double x) { return std::sin(x); }}},
            {"tan", {X_TAN, [](double x) { return std::tan(x); }}}, {"cosh", {X_COSH, [](double x) { return std::cosh(x); }}},
            {"sinh", {X_SINH, [](double x) { return std::sinh(x); }}}, {"tanh", {X_TANH, [](double x) { return std::tanh(x); }}},
            {"exp", {X_EXP, [](double x) { return std::exp(x); }}}, {"log", {X_LOG, [](double x) { return std::log(x); }}},
            {"log10", {X_LOG10, [](double x) { return std::log10(x); }}}, {"sqrt", {X_SQRT, [](double x) { return std::sqrt(x); }}},
            {"ceil", {X_CEIL, [](double x) { return std::ceil(x); }}}, {"floor", {X_FLOOR, [](double x) { return std::floor(x); }}}};
        if (auto it = kUnary.find(t); it != kUnary.end()) {
            const auto v = eval_args(e, cx, 1, 1);
            const double x = v[0].v;
            if ((t == "acos" || t == "asin") && (x < -1 || x > 1)) bad(CANOPY_E_DOMAIN, e, "argument of " + t + " outside [-1, 1]");
            if ((t == "log" || t == "log10") && x <= 0) bad(CANOPY_E_DOMAIN, e, "argument of " + t + " is not positive");
            if (t == "sqrt" && x < 0) bad(CANOPY_E_DOMAIN, e, "argument of sqrt is negative");
            return emit(it->second.op, v, it->second.f(x));
        }
        if (t == "mod") { const auto v = eval_args(e, cx, 2, 2); if ((long long)v[1].v == 0) bad(CANOPY_E_DOMAIN, e, "Modulo second operand must not be 0."); return emit(X_MOD, v, (double)((long long)v[0].v % (long long)v[1].v)); }
        if (t == "pow") { const auto v = eval_args(e, cx, 2, 2); return emit(X_POW, v, std::pow(v[0].v, v[1].v)); }
        if (t == "min" || t == "max" || t == "mean") {
            const auto v = eval_args(e, cx, 1, kMany);
            double r = v[0].v, s = 0;
            for (const Ex& x : v) { r = t == "min" ? std::fmin(r, x.v) : std::fmax(r, x.v); s += x.v; }
            return t == "mean" ? emit(X_MEAN, v, s / (double)v.size()) : emit(t == "min" ? X_MIN : X_MAX, v, r);
        }
        if (t == "not") { const auto v = eval_args(e, cx, 1, 1); return emit(X_NOT, v, v[0].v != 0.0 ? 0.0 : 1.0); }
        if (t == "and") { const auto v = eval_args(e, cx, 2, kMany); double r = 1.0; for (const Ex& x : v) if (x.v == 0.0) r = 0.0; return emit(X_AND, v, r); }
        if (t == "or") { const auto v = eval_args(e, cx, 2, kMany); double r = 0.0; for (const Ex& x : v) if (x.v != 0.0) r = 1.0; return emit(X_OR, v, r); }
        if (t == "eq" || t == "df" || t == "lt" || t == "gt" || t == "leq" || t == "geq") {
            const auto v = eval_args(e, cx, 2, 2);
            const double a = v[0].v, b = v[1].v;
            const bool r = t == "eq" ? a == b : t == "df" ? a != b : t == "lt" ? a < b : t == "gt" ? a > b : t == "leq" ? a <= b : a >= b;
            return emit(t == "eq" ? X_EQ : t == "df" ? X_DF : t == "lt" ? X_LT : t == "gt" ? X_GT : t == "leq" ? X_LEQ : X_GEQ, v, r ? 1.0 : 0.0);
        }
        if (t == "ite") { const auto v = eval_args(e, cx, 3, 3); return emit(X_ITE, v, v[0].v != 0.0 ? v[1].v : v[2].v); }
        if (t == "switch") {  // conditional.h: the first <case> whose condition holds, else the trailing default; = nested ite
            const auto ks = kids_of(e);
            if (ks.empty()) bad(CANOPY_E_VALIDITY, e, "<switch> needs a default value");
            Ex r = ev(*ks.back(), cx);
            for (size_t i = ks.size() - 1; i-- > 0;) {
                const auto cs = kids_of(*ks[i]);
                if (ks[i]->name != "case" || cs.size() != 2) bad(CANOPY_E_VALIDITY, *ks[i], "<switch> takes <case> elements with 2 arguments, then a default");
                const Ex c = ev(*cs[0], cx), x = ev(*cs[1], cx);
                r = emit(X_ITE, {c, x, r}, c.v != 0.0 ? x.v : r.v);
            }
            return r;
        }
        if (t == "exponential") {  // exponential.h:15-41: 1 - exp(-lambda t)
            const auto v = eval_args(e, cx, 2, 2);
            if (v[0].v < 0) bad(CANOPY_E_DOMAIN, e, "The rate of failure cannot be negative.");
            if (v[1].v < 0) bad(CANOPY_E_DOMAIN, e, "The mission time cannot be negative.");
            return emit(X_EXPONENTIAL, v, 1.0 - std::exp(-v[0].v * v[1].v));
        }
        if (t == "GLM") {  // exponential.h:43-72: (lambda - (lambda - gamma r) exp(-r t)) / r, r = lambda + mu
            const auto v = eval_args(e, cx, 4, 4);
            const double gamma = v[0].v, lambda = v[1].v, mu = v[2].v, time = v[3].v, r = lambda + mu;
            if (gamma < 0 || gamma > 1) bad(CANOPY_E_DOMAIN, e, "Invalid value for probability on demand.");
            if (lambda < 0 || mu < 0) bad(CANOPY_E_DOMAIN, e, "The rate of failure or repair cannot be negative.");
            if (time < 0) bad(CANOPY_E_DOMAIN, e, "The mission time cannot be negative.");
            return emit(X_GLM, v, r == 0 ? gamma : (lambda - (lambda - gamma * r) * std::exp(-r * time)) / r);
        }
        if (t == "Weibull") {  // exponential.h:74-103: 1 - exp(-((t - t0) / alpha)^beta)
            const auto v = eval_args(e, cx, 4, 4);
            if (v[0].v <= 0) bad(CANOPY_E_DOMAIN, e, "The scale parameter for Weibull distribution must be positive.");
            if (v[1].v <= 0) bad(CANOPY_E_DOMAIN, e, "The shape parameter for Weibull distribution must be positive.");
            if (v[2].v < 0) bad(CANOPY_E_DOMAIN, e, "Invalid value for time shift.");
            if (v[3].v < 0) bad(CANOPY_E_DOMAIN, e, "The mission time cannot be negative.");
            return emit(X_WEIBULL, v, v[3].v <= v[2].v ? 0.0 : 1.0 - std::exp(-std::pow((v[3].v - v[2].v) / v[0].v, v[1].v)));
        }
        if (t == "periodic-test") {  // exponential.h:105-245, initializer.cpp:1128-1141 (4, 5 or 11 arguments)
            const auto ks = kids_of(e);
            if (ks.size() != 4 && ks.size() != 5 && ks.size() != 11) bad(CANOPY_E_VALIDITY, e, "Invalid number of arguments for Periodic Test.");
            if (ks.size() == 11) bad(CANOPY_E_UNSUPPORTED, e, "the 11-argument <periodic-test> is not supported by this loader");
            const auto v = eval_args(e, cx, 4, 5);
            const bool five = v.size() == 5;
            const double lambda = v[0].v, tau = v[five ? 2 : 1].v, theta = v[five ? 3 : 2].v, time = v[five ? 4 : 3].v;
            if (lambda < 0) bad(CANOPY_E_DOMAIN, e, "The rate of failure cannot be negative.");
            if (five && v[1].v < 0) bad(CANOPY_E_DOMAIN, e, "The rate of repair cannot be negative.");
            if (tau <= 0) bad(CANOPY_E_DOMAIN, e, "The time between tests must be positive.");
            if (theta < 0) bad(CANOPY_E_DOMAIN, e, "The time before the first test cannot be negative.");
            if (time < 0) bad(CANOPY_E_DOMAIN, e, "The mission time cannot be negative.");
            if ((time - theta) / tau > 1e7) bad(CANOPY_E_DOMAIN, e, "periodic-test: more than 1e7 test intervals in the mission time");
            return five ? emit(X_PERIODIC5, v, periodic_test5(lambda, v[1].v, tau, theta, time)) : emit(X_PERIODIC4, v, periodic_test4(lambda, tau, theta, time));
        }
        // random deviates: value() is the mean of the distribution (random_deviate.h:44-264); sampling: expr_tape.h
        if (t == "uniform-deviate") { const auto v = eval_args(e, cx, 2, 2); if (v[0].v >= v[1].v) bad(CANOPY_E_VALIDITY, e, "Min value is more than max for Uniform distribution."); return emit(X_UNIFORM, v, (v[0].v + v[1].v) / 2); }
        if (t == "normal-deviate") { const auto v = eval_args(e, cx, 2, 2); if (v[1].v <= 0) bad(CANOPY_E_DOMAIN, e, "Standard deviation cannot be negative or zero."); return emit(X_NORMAL, v, v[0].v); }
        if (t == "lognormal-deviate") {
            const auto v = eval_args(e, cx, 2, 3);
            if (v.size() == 3) {  // (mean, error factor, confidence level)
                if (v[0].v <= 0) bad(CANOPY_E_DOMAIN, e, "The mean of Log-Normal distribution cannot be negative or zero.");
                if (v[1].v <= 1) bad(CANOPY_E_DOMAIN, e, "The Error Factor for Log-Normal distribution cannot be less than 1.");
                if (v[2].v <= 0 || v[2].v >= 1) bad(CANOPY_E_DOMAIN, e, "The confidence level is not within (0, 1).");
                if (v[2].v == 0.5) bad(CANOPY_E_DOMAIN, e, "The confidence level 0.5 leaves the Log-Normal scale undefined.");
                return emit(X_LOGNORMAL3, v, v[0].v);
            }
            if (v[1].v <= 0) bad(CANOPY_E_DOMAIN, e, "Standard deviation cannot be negative or zero.");
            return emit(X_LOGNORMAL2, v, std::exp(v[0].v + v[1].v * v[1].v / 2));  // (mu, sigma) of the underlying normal
        }
        if (t == "gamma-deviate") { const auto v = eval_args(e, cx, 2, 2); if (v[0].v <= 0 || v[1].v <= 0) bad(CANOPY_E_DOMAIN, e, "The k and theta parameters of Gamma distribution must be positive."); return emit(X_GAMMA, v, v[0].v * v[1].v); }
        if (t == "beta-deviate") { const auto v = eval_args(e, cx, 2, 2); if (v[0].v <= 0 || v[1].v <= 0) bad(CANOPY_E_DOMAIN, e, "The alpha and beta parameters of Beta distribution must be positive."); return emit(X_BETA, v, v[0].v / (v[0].v + v[1].v)); }
        if (t == "histogram") {
            // <histogram> lower-bound <bin> upper-bound weight </bin>... (initializer.cpp:1094-1112); validation as
            // random_deviate.h:240-247: boundaries strictly increasing, weights not negative.  value() has no body in
            // the reference; the mean of the piecewise-constant distribution whose bin i carries probability mass
            // w_i / sum(w) (random_deviate.h:234-238 "weights of intervals") is sum_i w_i (b_{i-1} + b_i) / (2 sum w).
            const auto kids = kids_of(e);
            if (kids.size() < 2) bad(CANOPY_E_VALIDITY, e, "<histogram> needs a lower boundary and at least one <bin>");
            std::vector<Ex> v{ev(*kids[0], cx)};
            double prev = v[0].v, sum_w = 0, sum_wm = 0;
            for (size_t i = 1; i < kids.size(); ++i) {
                const auto bin = kids_of(*kids[i]);
                if (kids[i]->name != "bin" || bin.size() != 2) bad(CANOPY_E_VALIDITY, *kids[i], "<bin> takes an upper boundary and a weight");
                const Ex b = ev(*bin[0], cx), w = ev(*bin[1], cx);
                if (!(b.v > prev)) bad(CANOPY_E_VALIDITY, e, "Histogram upper boundaries are not strictly increasing.");
                if (w.v < 0) bad(CANOPY_E_VALIDITY, e, "Histogram weights cannot be negative.");
                sum_w += w.v; sum_wm += w.v * (prev + b.v); prev = b.v;
                v.push_back(b); v.push_back(w);
            }
            if (sum_w <= 0) bad(CANOPY_E_VALIDITY, e, "Histogram weights sum to zero.");
            return emit(X_HISTOGRAM, v, sum_wm / (2 * sum_w));
        }
        bad(CANOPY_E_UNSUPPORTED, e, "expression <" + t + "> is not supported by this loader");
    }

    // ---- CCF groups (ccf_group.h) ----------------------------------------------------------------------
    void apply_ccf() {
        for (CcfDef& c : ccfs) {
            const XNode& n = *c.node;
            cur_base = c.base_path;   // distribution and factors resolve in the group's scope (initializer.cpp:1341-1352)
            const int n_mem = (int)c.members.size();
            if (n_mem < 2) bad(CANOPY_E_VALIDITY, n, "CCF group must have at least 2 members: " + c.name);  // ccf_group.h:123-125
            const XNode* distr = nullptr;
            std::vector<const XNode*> factor_nodes;
            for (auto& k : n.kids) {
                if (k->name == "distribution") distr = k.get();
                else if (k->name == "factor") factor_nodes.push_back(k.get());
                else if (k->name == "factors") { for (auto& f : k->kids) if (f->name == "factor") factor_nodes.push_back(f.get()); }
                else if (k->name != "members" && !is_meta(*k)) bad(CANOPY_E_UNSUPPORTED, *k, "unexpected element <" + k->name + "> in CCF group " + c.name);
            }
            if (!distr || !first_kid(*distr)) bad(CANOPY_E_VALIDITY, n, "CCF group " + c.name + " has no distribution");
            const Ex qx = ev(*first_kid(*distr), EvalCtx{});
            const double q = qx.v;
            if (!(q >= 0 && q <= 1)) bad(CANOPY_E_DOMAIN, *distr, "Invalid probability value for CCF group distribution " + c.name);
            for (uint32_t b : c.members) { basics[b].p = q; basics[b].node = qx.id; basics[b].has_p = true; }  // AddDistribution, ccf_group.h:127-130
            // AddFactor, ccf_group.h:144-175
            const int min_level = c.model == "beta-factor" ? n_mem : c.model == "MGL" ? 2 : 1;
            std::vector<std::pair<int, Ex>> factors;  // index = level - min_level; level 0 = missing
            int prev_level = 0;
            for (const XNode* f : factor_nodes) {
                int level = prev_level ? prev_level + 1 : min_level;
                if (const std::string* lv = f->attr("level")) level = (int)parse_number(*f, *lv);
                if (level <= 0) bad(CANOPY_E_INVALID, *f, "Invalid CCF group factor setup.");
                if (level < min_level) bad(CANOPY_E_VALIDITY, *f, "The CCF factor level (" + std::to_string(level) + ") is less than the minimum level (" + std::to_string(min_level) + ") in " + c.name);
                if (level > n_mem) bad(CANOPY_E_VALIDITY, *f, "The CCF factor level " + std::to_string(level) + " is more than the number of members (" + std::to_string(n_mem) + ") in " + c.name);
                const size_t idx = (size_t)(level - min_level);
                if (idx < factors.size() && factors[idx].first) bad(CANOPY_E_VALIDITY, *f, "Redefinition of CCF factor for level " + std::to_string(level) + " in " + c.name);
                if (idx >= factors.size()) factors.resize(idx + 1, {0, Ex{0.0, 0u}});
                if (!first_kid(*f)) bad(CANOPY_E_VALIDITY, *f, "CCF factor without an expression");
                const Ex v = ev(*first_kid(*f), EvalCtx{});
                if (!(v.v >= 0 && v.v <= 1)) bad(CANOPY_E_DOMAIN, *f, "Invalid probability value for CCF group factor in " + c.name);
                factors[idx] = {level, v};
                prev_level = level;
            }
            if (factors.empty()) bad(CANOPY_E_VALIDITY, n, "CCF group " + c.name + " has no factors");
            for (auto& f : factors) if (!f.first) bad(CANOPY_E_VALIDITY, n, "Missing some CCF factors in " + c.name);  // ccf_group.h:195-197
            // CalculateProbabilities of the four models: (level, probability of ONE event of that level), built as tape
            // expressions over the distribution and the factors so that sampled deviates flow through them
            auto mul = [&](Ex a, Ex b) { return emit(X_MUL, {a, b}, a.v * b.v); };
            auto one_minus = [&](Ex a) { return emit(X_SUB, {konst(1.0), a}, 1.0 - a.v); };
            std::vector<std::pair<int, Ex>> probs;
            const int max_level = factors.back().first;
            if (c.model == "beta-factor") {  // ccf_group.h:330-350
                probs.emplace_back(1, mul(one_minus(factors[0].second), qx));
                probs.emplace_back(factors[0].first, mul(factors[0].second, qx));
            } else if (c.model == "MGL") {  // ccf_group.h:357-381
                for (int i = 0; i < max_level; ++i) {
                    Ex v = konst(combination_reciprocal(n_mem - 1, i));
                    for (int j = 0; j < i; ++j) v = mul(v, factors[j].second);
                    if (i < max_level - 1) v = mul(v, one_minus(factors[i].second));
                    probs.emplace_back(i + 1, mul(v, qx));
                }
            } else if (c.model == "alpha-factor") {  // ccf_group.h:388-416
                Ex sum = konst(0.0);
                for (auto& f : factors) sum = emit(X_ADD, {sum, mul(konst((double)f.first), f.second)}, sum.v + f.first * f.second.v);
                if (sum.v == 0) bad(CANOPY_E_DOMAIN, n, "alpha factors of " + c.name + " sum to 0");
                for (int i = 0; i < max_level; ++i) {
                    const Ex frac = emit(X_DIV, {factors[i].second, sum}, factors[i].second.v / sum.v);
                    probs.emplace_back(i + 1, mul(mul(konst((i + 1) * combination_reciprocal(n_mem - 1, i)), frac), qx));
                }
            } else {  // phi-factor, ccf_group.h:422-455
                double sum = 0;
                for (auto& f : factors) sum += f.second.v;
                if (std::fabs(sum - 1.0) > 1e-4) bad(CANOPY_E_VALIDITY, n, "The factors for Phi model CCF must sum to 1.");
                for (auto& f : factors) probs.emplace_back(f.first, mul(f.second, qx));
            }
            cur_base.clear();
            if (!opts.ccf_analysis) continue;
            // ApplyModel, ccf_group.h:215-260: proxy OR gate per member over the CCF events that contain it
            std::vector<uint32_t> proxy(n_mem);
            for (int i = 0; i < n_mem; ++i) {
                GateDef g; g.name = basics[c.members[i]].name; g.conn = CANOPY_OR;
                proxy[i] = (uint32_t)gates.size();
                basics[c.members[i]].proxy_gate = (int)proxy[i];
                gates.push_back(g);
            }
            for (auto& lp : probs) {
                const int level = lp.first;
                if (!(lp.second.v >= 0 && lp.second.v <= 1)) bad(CANOPY_E_DOMAIN, n, "CCF event probability outside [0, 1] in " + c.name);
                std::vector<int> comb(level);
                for (int i = 0; i < level; ++i) comb[i] = i;
                for (;;) {  // combinations in lexicographic order (algorithm.h:214-221 for_each_combination)
                    std::string nm = "[";
                    for (int i = 0; i < level; ++i) nm += (i ? " " : "") + basics[c.members[comb[i]]].name;
                    nm += "]";
                    if (sym.count(nm)) bad(CANOPY_E_VALIDITY, n, "Duplicate Element Error: Redefinition of event: " + nm);
                    BasicDef b; b.name = nm; b.p = lp.second.v; b.node = lp.second.id; b.has_p = true;
                    const uint32_t bi = (uint32_t)basics.size();
                    basics.push_back(b);
                    for (int i = 0; i < level; ++i) gates[proxy[comb[i]]].args.push_back(ArgRef{ArgRef::kBasic, false, bi});
                    int i = level - 1;
                    while (i >= 0 && comb[i] == n_mem - level + i) --i;
                    if (i < 0) break;
                    ++comb[i];
                    for (int j = i + 1; j < level; ++j) comb[j] = comb[j - 1] + 1;
                }
            }
        }
    }

    // ---- formulas (initializer.cpp:772-851) --------------------------------------------------------------
    // An untyped reference (GetEvent, initializer.cpp:1416-1437) looks through the gates, the basic events and the house
    // events in that order; a typed one (GetGate / GetBasicEvent / GetHouseEvent -> GetEntity, :1372-1403) only through the
    // tables of its kind, so that an element of another kind with the same local name does not shadow it.  Either way:
    // base_path.reference in the full-path table(s); else a reference without a dot is a public id, one with dots a full path.
    const Sym* find_in_paths(const std::string& path, int kind) const {
        for (int k = 0; k < 3; ++k) {
            if (kind >= 0 && k != kind) continue;
            auto it = path_sym[k].find(path);
            if (it != path_sym[k].end()) return &it->second;
        }
        return nullptr;
    }
    const Sym* find_event(const std::string& ref, int kind = -1) const {
        if (!cur_base.empty())
            if (const Sym* s = find_in_paths(cur_base + "." + ref, kind)) return s;
        if (ref.find('.') == std::string::npos) {
            auto it = sym.find(ref);   // ids are unique across the three kinds (model.h:151-155)
            return it == sym.end() || (kind >= 0 && (int)it->second.kind != kind) ? nullptr : &it->second;
        }
        return find_in_paths(ref, kind);
    }
    ArgRef event_ref(const XNode& a, bool complement) {
        const std::string* type = a.attr("type");
        const std::string et = type && !type->empty() ? *type : a.name;
        if (et != "event" && et != "gate" && et != "basic-event" && et != "house-event")
            bad(CANOPY_E_UNSUPPORTED, a, "formula argument <" + a.name + "> is not an event reference (nested formulas are not supported)");
        const std::string& nm = need_name(a);
        const int want = et == "gate" ? (int)Kind::kGate : et == "basic-event" ? (int)Kind::kBasic : et == "house-event" ? (int)Kind::kHouse : -1;
        const Sym* found = find_event(nm, want);
        if (!found) bad(CANOPY_E_VALIDITY, a, "Undefined Element Error: " + nm + (want >= 0 && find_event(nm) ? " is not a " + et : std::string()));
        const Sym s = *found;
        if (s.kind == Kind::kGate) return ArgRef{ArgRef::kGate, complement, s.idx};
        if (s.kind == Kind::kHouse) return ArgRef{ArgRef::kHouse, complement, s.idx};
        if (basics[s.idx].proxy_gate >= 0) return ArgRef{ArgRef::kGate, complement, (uint32_t)basics[s.idx].proxy_gate};  // event.h:108-114
        return ArgRef{ArgRef::kBasic, complement, s.idx};
    }
    ArgRef formula_arg(const XNode& a) {
        if (a.name == "constant") return ArgRef{ArgRef::kConst, false, parse_bool(a, need_attr(a, "value")) ? 1u : 0u};
        if (a.name == "not") {
            const auto ks = kids_of(a);
            if (ks.size() != 1) bad(CANOPY_E_VALIDITY, a, "<not> as an argument takes exactly one event");
            return event_ref(*ks[0], true);
        }
        return event_ref(a, false);
    }
    void parse_formula(const XNode& f, GateDef& g) {
        static const char* kNames[] = {"and", "or", "atleast", "xor", "not", "nand", "nor", "null", "iff", "imply", "cardinality"};
        if (f.attr("name") || f.name == "constant") { g.conn = CANOPY_NULL; g.args.push_back(formula_arg(f)); return; }
        int pos = -1;
        for (int i = 0; i < CANOPY_NUM_CONNECTIVES; ++i) if (f.name == kNames[i]) pos = i;
        if (pos < 0) bad(CANOPY_E_UNSUPPORTED, f, "unknown connective <" + f.name + ">");
        g.conn = (uint8_t)pos;
        auto num = [&](const char* k) -> uint16_t {
            const std::string* v = f.attr(k);
            if (!v) return 0;
            char* end = nullptr; long x = std::strtol(v->c_str(), &end, 10);
            if (end == v->c_str() || *end) bad(CANOPY_E_VALIDITY, f, std::string("bad ") + k + " number");
            if (x < 0) bad(CANOPY_E_INVALID, f, std::string("The ") + k + " number cannot be negative: " + *v);  // event.cpp:118-119
            if (x > 65535) bad(CANOPY_E_UNSUPPORTED, f, std::string(k) + " number above 65535");
            return (uint16_t)x;
        };
        if (g.conn == CANOPY_ATLEAST && !f.attr("min")) bad(CANOPY_E_VALIDITY, f, "The connective requires min number for its args.");
        if (g.conn == CANOPY_CARDINALITY && (!f.attr("min") || !f.attr("max")))
            bad(CANOPY_E_VALIDITY, f, "The connective requires min and max numbers for args.");
        g.kmin = num("min"); g.kmax = num("max");
        for (const XNode* a : kids_of(f)) g.args.push_back(formula_arg(*a));
    }
    void define_gates() {
        for (uint32_t g = 0; g < n_xml_gates; ++g) {
            const XNode* f = first_kid(*gates[g].body);
            if (!f) bad(CANOPY_E_VALIDITY, *gates[g].body, "gate " + gates[g].name + " has no formula");
            cur_base = gates[g].base_path;
            parse_formula(*f, gates[g]);
            cur_base.clear();
        }
    }
    void check_cycles() {  // cycle.h:114-131 (DFS marks); the event-tree walk below recurses over the gate graph
        std::vector<uint8_t> mark(gates.size(), 0);
        std::vector<std::pair<uint32_t, size_t>> st;
        for (uint32_t r = 0; r < gates.size(); ++r) {
            if (mark[r]) continue;
            st.push_back({r, 0}); mark[r] = 1;
            while (!st.empty()) {
                auto& [g, i] = st.back();
                if (i == gates[g].args.size()) { mark[g] = 2; st.pop_back(); continue; }
                const ArgRef a = gates[g].args[i++];
                if (a.kind != ArgRef::kGate) continue;
                if (mark[a.idx] == 1) throw Error(CANOPY_E_CYCLE, "Cycle Error: detected a cycle through gate " + gates[a.idx].name);
                if (mark[a.idx] == 0) { mark[a.idx] = 1; st.push_back({a.idx, 0}); }
            }
        }
    }

    // ---- event trees (event_tree.h:18-224, instruction.h:21-215) -----------------------------------------
    using HouseCtx = std::map<uint32_t, bool>;  // set-house-event overrides in force on the path
    struct PathCollector {
        std::vector<ArgRef> formulas;  // gates collected along the path (conjoined)
        double expr = 1.0;             // product of the collected expressions ...
        uint32_t expr_node = 0;        // ... and its tape node (valid iff has_expr)
        bool has_expr = false;
        HouseCtx house;
    };
    struct Walk {
        std::string initiating_event;
        std::map<std::string, std::string> functional_states;
        std::vector<std::string> tree_stack, branch_stack;
        std::vector<std::string> seq_order;                          // sequences in the order first reached
        std::map<std::string, std::vector<PathCollector>> reached;   // sequence -> the paths that end in it
    };
    std::map<HouseCtx, int> ctx_ids;
    std::map<std::pair<int, uint32_t>, uint32_t> clone_memo;
    std::map<std::pair<int, uint32_t>, bool> touch_memo;

    int ctx_id(const HouseCtx& full) {  // only overrides that differ from the declared state matter
        HouseCtx eff;
        for (auto& kv : full) if (houses[kv.first].state != kv.second) eff.insert(kv);
        if (eff.empty()) return -1;
        return ctx_ids.emplace(eff, (int)ctx_ids.size()).first->second;
    }
    bool touches(uint32_t g, int cid, const HouseCtx& eff) {
        auto key = std::make_pair(cid, g);
        if (auto it = touch_memo.find(key); it != touch_memo.end()) return it->second;
        bool t = false;
        for (const ArgRef& a : gates[g].args) {
            if (a.kind == ArgRef::kHouse && eff.count(a.idx)) t = true;
            else if (a.kind == ArgRef::kGate && touches(a.idx, cid, eff)) t = true;
            if (t) break;
        }
        return touch_memo[key] = t;
    }
    ArgRef arg_under(ArgRef a, int cid, const HouseCtx& eff) {
        if (a.kind == ArgRef::kHouse) {
            auto it = eff.find(a.idx);
            if (it != eff.end()) return ArgRef{ArgRef::kConst, a.complement, it->second ? 1u : 0u};
        } else if (a.kind == ArgRef::kGate) a.idx = clone_under(a.idx, cid, eff);
        return a;
    }
    uint32_t clone_under(uint32_t g, int cid, const HouseCtx& eff) {  // the gate as it reads under the overrides
        if (!touches(g, cid, eff)) return g;
        auto key = std::make_pair(cid, g);
        if (auto it = clone_memo.find(key); it != clone_memo.end()) return it->second;
        GateDef c = gates[g];
        c.tree = -1; c.body = nullptr; c.name += "@" + std::to_string(cid);
        for (ArgRef& a : c.args) a = arg_under(a, cid, eff);
        gates.push_back(c);
        return clone_memo[key] = (uint32_t)gates.size() - 1;
    }
    ArgRef collect_formula(const XNode& f, const HouseCtx& house, const std::string& label) {
        GateDef g; g.name = label;
        parse_formula(f, g);
        const int cid = ctx_id(house);
        if (cid >= 0) {
            HouseCtx eff;
            for (auto& kv : house) if (houses[kv.first].state != kv.second) eff.insert(kv);
            for (ArgRef& a : g.args) a = arg_under(a, cid, eff);
        }
        gates.push_back(g);
        return ArgRef{ArgRef::kGate, false, (uint32_t)gates.size() - 1};
    }
    void exec(const XNode& ins, PathCollector& pc, Walk& w, bool in_sequence, bool* linked) {  // initializer.cpp:910-984
        const std::string& t = ins.name;
        const EvalCtx cx{&w.functional_states, &w.initiating_event};
        if (t == "rule") {
            auto it = rules.find(need_name(ins));
            if (it == rules.end()) bad(CANOPY_E_VALIDITY, ins, "Undefined Element Error: rule " + need_name(ins));
            if (std::count(w.branch_stack.begin(), w.branch_stack.end(), "rule:" + it->first)) bad(CANOPY_E_CYCLE, ins, "Cycle Error: rule " + it->first);
            w.branch_stack.push_back("rule:" + it->first);
            for (const XNode* k : kids_of(*it->second)) exec(*k, pc, w, in_sequence, linked);
            w.branch_stack.pop_back();
        } else if (t == "event-tree") {  // Link: only as an instruction of an end-state sequence (initializer.cpp:1659-1700)
            if (!in_sequence) bad(CANOPY_E_VALIDITY, ins, "Link " + need_name(ins) + " can only be used in end-state sequences.");
            auto it = tree_idx.find(need_name(ins));
            if (it == tree_idx.end()) bad(CANOPY_E_VALIDITY, ins, "Undefined Element Error: event tree " + need_name(ins));
            *linked = true;
            auto saved = w.functional_states;
            w.functional_states.clear();
            walk_tree(trees[it->second], pc, w, ins);
            w.functional_states = saved;
        } else if (t == "collect-expression") {
            const XNode* e = first_kid(ins);
            if (!e) bad(CANOPY_E_VALIDITY, ins, "<collect-expression> needs an expression");
            const Ex x = ev(*e, cx);
            if (pc.has_expr) { const Ex prod = emit(X_MUL, {Ex{pc.expr, pc.expr_node}, x}, pc.expr * x.v); pc.expr = prod.v; pc.expr_node = prod.id; }
            else { pc.expr = x.v; pc.expr_node = x.id; pc.has_expr = true; }
        } else if (t == "collect-formula") {
            const XNode* f = first_kid(ins);
            if (!f) bad(CANOPY_E_VALIDITY, ins, "<collect-formula> needs a formula");
            pc.formulas.push_back(collect_formula(*f, pc.house, "__collect_" + std::to_string(gates.size())));
        } else if (t == "if") {
            const auto ks = kids_of(ins);
            if (ks.size() < 2 || ks.size() > 3) bad(CANOPY_E_VALIDITY, ins, "<if> takes a condition, an instruction and an optional else instruction");
            if (eval(*ks[0], cx) != 0.0) exec(*ks[1], pc, w, in_sequence, linked);
            else if (ks.size() == 3) exec(*ks[2], pc, w, in_sequence, linked);
        } else if (t == "block") {
            for (const XNode* k : kids_of(ins)) exec(*k, pc, w, in_sequence, linked);
        } else if (t == "set-house-event") {
            const Sym* it = find_event(need_name(ins));   // event-tree instructions are read at model scope (initializer.cpp:950)
            if (!it || it->kind != Kind::kHouse) bad(CANOPY_E_VALIDITY, ins, "Undefined Element Error: house event " + need_name(ins));
            const XNode* c = first_kid(ins);
            if (!c || c->name != "constant") bad(CANOPY_E_VALIDITY, ins, "<set-house-event> needs <constant value=.../>");
            pc.house[it->idx] = parse_bool(*c, need_attr(*c, "value"));
        } else bad(CANOPY_E_UNSUPPORTED, ins, "unknown instruction <" + t + ">");
    }
    void walk_branch(const EventTreeDef& et, const std::vector<const XNode*>& nodes, PathCollector pc, Walk& w, const XNode& at) {
        if (nodes.empty()) bad(CANOPY_E_VALIDITY, at, "a branch needs at least its target");  // initializer.cpp:893
        bool linked = false;
        for (size_t i = 0; i + 1 < nodes.size(); ++i) exec(*nodes[i], pc, w, false, &linked);
        const XNode& tgt = *nodes.back();
        if (tgt.name == "fork") {  // initializer.cpp:856-875
            const std::string& fe = need_attr(tgt, "functional-event");
            if (!et.functional_events.count(fe)) bad(CANOPY_E_VALIDITY, tgt, "Undefined Element Error: functional event " + fe + " in event tree " + et.name);
            if (w.functional_states.count(fe)) bad(CANOPY_E_VALIDITY, tgt, "Functional event " + fe + " is forked twice on one path of event tree " + et.name);
            std::set<std::string> states;
            size_t n_paths = 0;
            for (const XNode* p : kids_of(tgt)) {
                if (p->name != "path") bad(CANOPY_E_VALIDITY, *p, "<fork> takes <path> elements");
                const std::string& st = need_attr(*p, "state");
                if (!states.insert(st).second) bad(CANOPY_E_VALIDITY, *p, "Duplicate state '" + st + "' path in a fork");  // event_tree.h:118-131
                w.functional_states[fe] = st;
                walk_branch(et, kids_of(*p), pc, w, *p);
                ++n_paths;
            }
            w.functional_states.erase(fe);
            if (!n_paths) bad(CANOPY_E_VALIDITY, tgt, "<fork> without paths");
        } else if (tgt.name == "sequence") {  // initializer.cpp:876-879
            auto it = sequences.find(need_name(tgt));
            if (it == sequences.end()) bad(CANOPY_E_VALIDITY, tgt, "Undefined Element Error: sequence " + need_name(tgt));
            for (const XNode* k : kids_of(*it->second)) exec(*k, pc, w, true, &linked);
            if (!linked) {
                if (!w.reached.count(it->first)) w.seq_order.push_back(it->first);
                w.reached[it->first].push_back(pc);
            }
        } else if (tgt.name == "branch") {  // initializer.cpp:880-885
            auto it = et.branches.find(need_name(tgt));
            if (it == et.branches.end()) bad(CANOPY_E_VALIDITY, tgt, "Undefined Element Error: branch " + need_name(tgt) + " in event tree " + et.name);
            const std::string key = et.name + "/" + it->first;
            if (std::count(w.branch_stack.begin(), w.branch_stack.end(), key)) bad(CANOPY_E_CYCLE, tgt, "Cycle Error: branch " + it->first);  // cycle.h:219-229 over NamedBranch
            w.branch_stack.push_back(key);
            walk_branch(et, kids_of(*it->second), pc, w, *it->second);
            w.branch_stack.pop_back();
        } else bad(CANOPY_E_VALIDITY, tgt, "a branch must end in <fork>, <sequence> or <branch>, not <" + tgt.name + ">");
    }
    void walk_tree(const EventTreeDef& et, const PathCollector& pc, Walk& w, const XNode& at) {
        if (std::count(w.tree_stack.begin(), w.tree_stack.end(), et.name)) bad(CANOPY_E_CYCLE, at, "Cycle Error: event tree " + et.name + " links to itself");
        const XNode* init = nullptr;
        for (auto& k : et.node->kids) if (k->name == "initial-state") init = k.get();
        if (!init) bad(CANOPY_E_VALIDITY, *et.node, "event tree " + et.name + " has no <initial-state>");
        w.tree_stack.push_back(et.name);
        walk_branch(et, kids_of(*init), pc, w, *init);
        w.tree_stack.pop_back();
    }
    uint32_t synth_gate(const std::string& name, uint8_t conn, std::vector<ArgRef> args) {
        GateDef g; g.name = name;
        if (args.empty()) { g.conn = CANOPY_NULL; g.args.push_back(ArgRef{ArgRef::kConst, false, 1u}); }  // nothing collected: certain
        else if (args.size() == 1) { g.conn = CANOPY_NULL; g.args = args; }
        else { g.conn = conn; g.args = std::move(args); }
        gates.push_back(g);
        return (uint32_t)gates.size() - 1;
    }
    std::vector<std::pair<std::string, uint32_t>> sequence_roots;  // ("initiating-event.sequence", gate)
    void walk_event_trees() {
        for (auto& ie : initiating) {
            if (ie.second.empty()) continue;  // an initiating event without an event tree is only a name
            auto it = tree_idx.find(ie.second);
            if (it == tree_idx.end()) throw Error(CANOPY_E_VALIDITY, "Undefined Element Error: event tree " + ie.second + " of initiating event " + ie.first);
            Walk w; w.initiating_event = ie.first;
            walk_tree(trees[it->second], PathCollector{}, w, *trees[it->second].node);
            for (const std::string& sq : w.seq_order) {
                std::vector<ArgRef> path_gates;
                size_t pi = 0;
                for (const PathCollector& pc : w.reached[sq]) {
                    std::vector<ArgRef> conj = pc.formulas;
                    const std::string tag = ie.first + "." + sq + "#" + std::to_string(pi++);
                    if (pc.has_expr) {  // the product of the collected expressions becomes one more basic event
                        if (!(pc.expr >= 0 && pc.expr <= 1)) throw Error(CANOPY_E_DOMAIN, "Invalid probability value " + std::to_string(pc.expr) + " collected on a path to sequence " + sq);
                        BasicDef b; b.name = "__expr_" + tag; b.p = pc.expr; b.node = pc.expr_node; b.has_p = true;
                        basics.push_back(b);
                        conj.push_back(ArgRef{ArgRef::kBasic, false, (uint32_t)basics.size() - 1});
                    }
                    path_gates.push_back(ArgRef{ArgRef::kGate, false, synth_gate("__path_" + tag, CANOPY_AND, conj)});
                }
                sequence_roots.emplace_back(ie.first + "." + sq, synth_gate(ie.first + "." + sq, CANOPY_OR, path_gates));
            }
        }
    }

    // ---- emission ----------------------------------------------------------------------------------------
    void finish() {
        n_xml_gates = (uint32_t)gates.size();
        apply_ccf();      // needs every parameter registered; allocates the proxy gates formulas may refer to
        define_gates();
        check_cycles();
        walk_event_trees();
        const uint32_t B = (uint32_t)basics.size(), G = (uint32_t)gates.size();

        m.p.resize(B); m.p_node.resize(B); m.basic_names.resize(B);
        for (uint32_t e = 0; e < B; ++e) {
            BasicDef& b = basics[e];
            m.basic_names[e] = b.name;
            if (!b.has_p) {
                // a basic event never given an expression cannot be quantified; the reference reports it when
                // the probability analysis asks for p() (event.h:93-96 asserts the expression)
                if (!b.body) throw Error(CANOPY_E_VALIDITY, "basic event " + b.name + " has no expression");
                cur_base = b.base_path;
                const Ex x = ev(*b.body, EvalCtx{});
                cur_base.clear();
                b.p = x.v; b.node = x.id;
            }
            if (!(b.p >= 0.0 && b.p <= 1.0)) throw Error(CANOPY_E_DOMAIN, "Invalid probability value for " + b.name);  // event.cpp:27-33
            m.p[e] = b.p; m.p_node[e] = b.node;
        }
        define_unused_parameters();
        m.op.resize(G); m.kmin.resize(G); m.kmax.resize(G); m.gate_names.resize(G);
        m.child_off.assign(1, 0);
        for (uint32_t g = 0; g < G; ++g) {
            const GateDef& d = gates[g];
            m.op[g] = d.conn; m.kmin[g] = d.kmin; m.kmax[g] = d.kmax; m.gate_names[g] = d.name;
            for (const ArgRef& a : d.args) {
                uint32_t id;
                if (a.kind == ArgRef::kBasic) id = a.idx;
                else if (a.kind == ArgRef::kGate) id = B + a.idx;
                else if (a.kind == ArgRef::kHouse) id = houses[a.idx].state ? CANOPY_CONST_TRUE : CANOPY_CONST_FALSE;
                else id = a.idx ? CANOPY_CONST_TRUE : CANOPY_CONST_FALSE;
                m.child.push_back(id | (a.complement ? CANOPY_CHILD_COMPLEMENT : 0u));
            }
            m.child_off.push_back(m.child.size());
        }
        // top events per fault tree (fault_tree.h:151-186), then one root per (initiating event, sequence)
        std::vector<bool> used(G, false);
        for (uint32_t g = 0; g < G; ++g)
            for (const ArgRef& a : gates[g].args)
                if (a.kind == ArgRef::kGate && gates[a.idx].tree == gates[g].tree) used[a.idx] = true;
        for (uint32_t g = 0; g < G; ++g)
            if (!used[g] && gates[g].tree >= 0) { m.roots.push_back(B + g); m.root_names.push_back(gates[g].name); }
        for (auto& sr : sequence_roots) { m.roots.push_back(B + sr.second); m.root_names.push_back(sr.first); }
    }
};

std::string slurp(const char* path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw Error(CANOPY_E_IO, std::string("File doesn't exist or inaccessible: ") + path);  // initializer.cpp:232-239
    std::ostringstream ss; ss << f.rdbuf();
    return ss.str();
}

MefOpts to_opts(const canopy_mef_opts* o) {
    MefOpts r;
    if (!o) return r;
    r.ccf_analysis = o->ccf_analysis != 0;
    if (o->mission_time < 0 || o->mission_time != o->mission_time) throw Error(CANOPY_E_INVALID, "The mission time cannot be negative.");  // settings.h:192-197
    r.mission_time = o->mission_time;
    return r;
}

canopy_model* load_docs(const std::vector<std::string>& texts, const MefOpts& opts) {
    auto model = std::make_unique<canopy_model>();
    Builder b(*model, opts);
    std::vector<std::unique_ptr<XNode>> docs;
    for (auto& t : texts) { XmlParser p(t); docs.push_back(p.parse()); }
    for (auto& d : docs) b.load(*d);
    b.finish();
    return model.release();
}

}  // namespace
void set_last_error(const std::string& m);  // capi.cpp
namespace {

template <class F>
canopy_status guard_io(F&& f) {
    try { set_last_error(""); f(); return CANOPY_OK; }
    catch (const Error& e) { set_last_error(e.what()); return e.code; }
    catch (const std::bad_alloc&) { set_last_error("out of host memory"); return CANOPY_E_OOM; }
    catch (const std::exception& e) { set_last_error(e.what()); return CANOPY_E_INVALID; }
}

}  // namespace
}  // namespace canopy

extern "C" {

void canopy_mef_opts_default(canopy_mef_opts* o) {
    if (!o) return;
    std::memset(o, 0, sizeof *o);
    o->mission_time = 8760.0;  // settings.h:330
}

canopy_status canopy_model_load_mef_ex(const char* const* xml_paths, int32_t n_paths, const canopy_mef_opts* opts, canopy_model** out) {
    return canopy::guard_io([&] {
        if (!xml_paths || n_paths <= 0 || !out) throw canopy::Error(CANOPY_E_INVALID, "canopy_model_load_mef: bad argument");
        *out = nullptr;
        std::vector<std::string> texts;
        for (int i = 0; i < n_paths; ++i) {
            if (!xml_paths[i]) throw canopy::Error(CANOPY_E_INVALID, "null path");
            for (int j = 0; j < i; ++j)
                if (std::strcmp(xml_paths[i], xml_paths[j]) == 0)
                    throw canopy::Error(CANOPY_E_VALIDITY, std::string("Duplicate input files: ") + xml_paths[i]);  // initializer.cpp:241-250
            texts.push_back(canopy::slurp(xml_paths[i]));
        }
        *out = canopy::load_docs(texts, canopy::to_opts(opts));
    });
}
canopy_status canopy_model_load_mef(const char* const* xml_paths, int32_t n_paths, canopy_model** out) {
    return canopy_model_load_mef_ex(xml_paths, n_paths, nullptr, out);
}

canopy_status canopy_model_load_mef_string_ex(const char* xml_text, const canopy_mef_opts* opts, canopy_model** out) {
    return canopy::guard_io([&] {
        if (!xml_text || !out) throw canopy::Error(CANOPY_E_INVALID, "canopy_model_load_mef_string: null argument");
        *out = nullptr;
        *out = canopy::load_docs({std::string(xml_text)}, canopy::to_opts(opts));
    });
}
canopy_status canopy_model_load_mef_string(const char* xml_text, canopy_model** out) {
    return canopy_model_load_mef_string_ex(xml_text, nullptr, out);
}

canopy_status canopy_model_desc(const canopy_model* m, canopy_graph_desc* d) {
    return canopy::guard_io([&] {
        if (!m || !d) throw canopy::Error(CANOPY_E_INVALID, "canopy_model_desc: null argument");
        std::memset(d, 0, sizeof *d);
        d->n_basic = (uint32_t)m->p.size(); d->n_gates = (uint32_t)m->op.size(); d->n_roots = (uint32_t)m->roots.size();
        d->n_edges = m->child.size();
        d->p = m->p.data(); d->op = m->op.data(); d->kmin = m->kmin.data(); d->kmax = m->kmax.data();
        d->child_off = m->child_off.data(); d->child = m->child.data(); d->roots = m->roots.data();
    });
}

const char* canopy_model_basic_name(const canopy_model* m, uint32_t i) { return m && i < m->basic_names.size() ? m->basic_names[i].c_str() : nullptr; }
const char* canopy_model_gate_name(const canopy_model* m, uint32_t i) { return m && i < m->gate_names.size() ? m->gate_names[i].c_str() : nullptr; }
const char* canopy_model_root_name(const canopy_model* m, uint32_t i) { return m && i < m->root_names.size() ? m->root_names[i].c_str() : nullptr; }
void canopy_model_destroy(canopy_model* m) { delete m; }

}  // extern "C"
