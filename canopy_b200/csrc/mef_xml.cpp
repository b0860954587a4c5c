// Open-PSA MEF XML (fault-tree subset) -> flat gate DAG, host only, no third-party dependency.
//
// Stands in for the reference loader, which needs Boost + libxml2 and does not compile as shipped
// (SURVEY.md section 0): Initializer::ProcessInputFiles / DefineFaultTree / GetFormula
// (reference src/mef/openpsa/initializer.cpp:252-297, 712-758, 772-851) and
// FaultTree::CollectTopEvents (reference src/mef/openpsa/fault_tree.h:151-186).
// Semantics kept from the reference:
//   - two passes: register every name, then define bodies (initializer.cpp:252-297)
//   - a formula element is a connective named as in kConnectiveToString (event.h:164-166); an element
//     with a `name` attribute or a <constant> directly under <define-gate> is a NULL pass-through
//     (initializer.cpp:774-782); <not> around an argument sets its complement flag (:826-829);
//     <constant value=.../> arguments are HouseEvent::kTrue/kFalse (:820-824)
//   - <event name=.. [type=..]/> resolves gate -> basic event -> house event (:787-812)
//   - duplicate ids across gates / basic events / house events are errors (model.h:151-155)
//   - a house event without a <constant> body is false (event.h:33)
//   - top events of a fault tree = its gates that are not an argument of another gate of the same tree
// Not covered (CANOPY_E_UNSUPPORTED): CCF groups, parameters and non-constant expressions, event trees,
// XInclude.  Those are SURVEY.md 8f rows 1-2, outside round 1.
#include <cctype>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

#include "plan.h"

namespace canopy {
namespace {

struct XNode {
    std::string name, text;
    std::vector<std::pair<std::string, std::string>> attrs;
    std::vector<std::unique_ptr<XNode>> kids;
    int line = 0;
    const std::string* attr(const char* k) const {
        for (auto& a : attrs) if (a.first == k) return &a.second;
        return nullptr;
    }
};

class XmlParser {
  public:
    explicit XmlParser(const std::string& s) : s_(s) {}
    std::unique_ptr<XNode> parse() {
        skip_misc();
        auto root = element();
        skip_misc();
        if (i_ != s_.size()) err("trailing content after the document element");
        return root;
    }

  private:
    const std::string& s_;
    size_t i_ = 0;
    int line_ = 1;
    [[noreturn]] void err(const std::string& m) { throw Error(CANOPY_E_IO, "XML line " + std::to_string(line_) + ": " + m); }
    bool starts(const char* t) const { return s_.compare(i_, std::strlen(t), t) == 0; }
    void adv(size_t n = 1) { for (size_t k = 0; k < n && i_ < s_.size(); ++k) { if (s_[i_] == '\n') ++line_; ++i_; } }
    void skip_ws() { while (i_ < s_.size() && std::isspace((unsigned char)s_[i_])) adv(); }
    void skip_until(const char* t) {
        size_t p = s_.find(t, i_);
        if (p == std::string::npos) err(std::string("unterminated construct, expected ") + t);
        adv(p - i_ + std::strlen(t));
    }
    void skip_misc() {
        for (;;) {
            skip_ws();
            if (starts("<?")) skip_until("?>");
            else if (starts("<!--")) skip_until("-->");
            else if (starts("<!DOCTYPE")) skip_until(">");
            else break;
        }
    }
    std::string name() {
        size_t b = i_;
        while (i_ < s_.size() && (std::isalnum((unsigned char)s_[i_]) || std::strchr("_-.:", s_[i_]))) ++i_;
        if (b == i_) err("expected a name");
        return s_.substr(b, i_ - b);
    }
    std::string unescape(const std::string& v) {
        std::string o; o.reserve(v.size());
        for (size_t k = 0; k < v.size(); ++k) {
            if (v[k] != '&') { o += v[k]; continue; }
            size_t e = v.find(';', k);
            if (e == std::string::npos) err("bad entity");
            std::string ent = v.substr(k + 1, e - k - 1);
            if (ent == "lt") o += '<'; else if (ent == "gt") o += '>'; else if (ent == "amp") o += '&';
            else if (ent == "quot") o += '"'; else if (ent == "apos") o += '\'';
            else if (!ent.empty() && ent[0] == '#') o += (char)std::strtol(ent.c_str() + (ent[1] == 'x' ? 2 : 1), nullptr, ent[1] == 'x' ? 16 : 10);
            else err("unknown entity &" + ent + ";");
            k = e;
        }
        return o;
    }
    std::unique_ptr<XNode> element() {
        if (i_ >= s_.size() || s_[i_] != '<') err("expected '<'");
        auto n = std::make_unique<XNode>();
        n->line = line_;
        adv();
        n->name = name();
        for (;;) {
            skip_ws();
            if (i_ >= s_.size()) err("unterminated start tag");
            if (s_[i_] == '/') { adv(); if (i_ >= s_.size() || s_[i_] != '>') err("expected '>'"); adv(); return n; }
            if (s_[i_] == '>') { adv(); break; }
            std::string k = name();
            skip_ws();
            if (i_ >= s_.size() || s_[i_] != '=') err("expected '=' after attribute name");
            adv(); skip_ws();
            if (i_ >= s_.size() || (s_[i_] != '"' && s_[i_] != '\'')) err("expected a quoted attribute value");
            const char q = s_[i_]; adv();
            size_t e = s_.find(q, i_);
            if (e == std::string::npos) err("unterminated attribute value");
            n->attrs.emplace_back(k, unescape(s_.substr(i_, e - i_)));
            adv(e - i_ + 1);
        }
        for (;;) {
            if (i_ >= s_.size()) err("unterminated element <" + n->name + ">");
            if (starts("<!--")) { skip_until("-->"); continue; }
            if (starts("<![CDATA[")) { size_t e = s_.find("]]>", i_); if (e == std::string::npos) err("unterminated CDATA"); n->text += s_.substr(i_ + 9, e - i_ - 9); adv(e - i_ + 3); continue; }
            if (starts("<?")) { skip_until("?>"); continue; }
            if (starts("</")) {
                adv(2);
                std::string c = name();
                if (c != n->name) err("mismatched end tag </" + c + "> for <" + n->name + ">");
                skip_ws();
                if (i_ >= s_.size() || s_[i_] != '>') err("expected '>'");
                adv();
                return n;
            }
            if (s_[i_] == '<') { n->kids.push_back(element()); continue; }
            size_t e = s_.find('<', i_);
            if (e == std::string::npos) e = s_.size();
            n->text += unescape(s_.substr(i_, e - i_));
            adv(e - i_);
        }
    }
};

enum class Kind { kGate, kBasic, kHouse };
struct Sym { Kind kind; uint32_t idx; };

}  // namespace
}  // namespace canopy

struct canopy_model {
    std::vector<double> p;
    std::vector<uint8_t> op;
    std::vector<uint16_t> kmin, kmax;
    std::vector<uint64_t> child_off{0};
    std::vector<uint32_t> child;
    std::vector<uint32_t> roots;
    std::vector<std::string> basic_names, gate_names;
};

namespace canopy {
namespace {

struct Builder {
    canopy_model& m;
    std::map<std::string, Sym> sym;
    std::vector<bool> house_state;
    std::vector<const XNode*> gate_body, basic_body;
    std::vector<int> gate_tree;  // fault tree ordinal of each gate
    std::vector<bool> basic_defined;
    int n_trees = 0;

    [[noreturn]] static void bad(canopy_status c, const XNode& n, const std::string& msg) {
        throw Error(c, "line " + std::to_string(n.line) + ": " + msg);
    }
    static const std::string& need_name(const XNode& n) {
        const std::string* s = n.attr("name");
        if (!s || s->empty()) bad(CANOPY_E_VALIDITY, n, "<" + n.name + "> needs a name");
        return *s;
    }
    void reg(const XNode& n, Kind k, uint32_t idx) {
        const std::string& nm = need_name(n);
        if (!sym.emplace(nm, Sym{k, idx}).second)
            bad(CANOPY_E_VALIDITY, n, "Duplicate Element Error: Redefinition of event: " + nm);
    }
    static const XNode* first_kid(const XNode& n) {
        for (auto& k : n.kids) if (k->name != "label" && k->name != "attributes") return k.get();
        return nullptr;
    }
    static bool parse_bool(const XNode& n, const std::string& v) {
        if (v == "true" || v == "1") return true;
        if (v == "false" || v == "0") return false;
        bad(CANOPY_E_VALIDITY, n, "bad boolean '" + v + "'");
    }
    // ---- pass 1 ----------------------------------------------------------------------------------
    void register_container(const XNode& c, int tree) {
        for (auto& k : c.kids) {
            const std::string& t = k->name;
            if (t == "define-gate") {
                reg(*k, Kind::kGate, (uint32_t)m.gate_names.size());
                m.gate_names.push_back(need_name(*k)); gate_body.push_back(k.get()); gate_tree.push_back(tree);
            } else if (t == "define-basic-event") {
                reg(*k, Kind::kBasic, (uint32_t)m.basic_names.size());
                m.basic_names.push_back(need_name(*k)); basic_body.push_back(k.get());
            } else if (t == "define-house-event") {
                reg(*k, Kind::kHouse, (uint32_t)house_state.size());
                bool st = false;
                if (const XNode* b = first_kid(*k)) {
                    if (b->name != "constant" || !b->attr("value")) bad(CANOPY_E_UNSUPPORTED, *b, "house event body must be <constant value=.../>");
                    st = parse_bool(*b, *b->attr("value"));
                }
                house_state.push_back(st);
            } else if (t == "define-component") {
                register_container(*k, tree);
            } else if (t == "define-CCF-group" || t == "define-parameter") {
                bad(CANOPY_E_UNSUPPORTED, *k, "<" + t + "> is not supported by this loader");
            } else if (t == "label" || t == "attributes") {
            } else {
                bad(CANOPY_E_UNSUPPORTED, *k, "unexpected element <" + t + ">");
            }
        }
    }
    // ---- pass 2 ----------------------------------------------------------------------------------
    double expression(const XNode& e) {
        if (e.name == "float" || e.name == "int") {
            const std::string* v = e.attr("value");
            if (!v) bad(CANOPY_E_VALIDITY, e, "<" + e.name + "> needs a value");
            char* end = nullptr;
            const double x = std::strtod(v->c_str(), &end);
            if (end == v->c_str() || *end) bad(CANOPY_E_VALIDITY, e, "bad number '" + *v + "'");
            return x;
        }
        if (e.name == "bool") { const std::string* v = e.attr("value"); if (!v) bad(CANOPY_E_VALIDITY, e, "<bool> needs a value"); return parse_bool(e, *v) ? 1.0 : 0.0; }
        bad(CANOPY_E_UNSUPPORTED, e, "expression <" + e.name + "> is not supported (constants only)");
    }
    uint32_t event_code(const XNode& a, bool complement) {
        const std::string* type = a.attr("type");
        const std::string et = type && !type->empty() ? *type : a.name;
        if (et != "event" && et != "gate" && et != "basic-event" && et != "house-event")
            bad(CANOPY_E_UNSUPPORTED, a, "formula argument <" + a.name + "> is not an event reference (nested formulas are not supported)");
        const std::string& nm = need_name(a);
        auto it = sym.find(nm);
        if (it == sym.end()) bad(CANOPY_E_VALIDITY, a, "Undefined Element Error: " + nm);
        const Sym s = it->second;
        if ((et == "gate" && s.kind != Kind::kGate) || (et == "basic-event" && s.kind != Kind::kBasic) ||
            (et == "house-event" && s.kind != Kind::kHouse))
            bad(CANOPY_E_VALIDITY, a, "Undefined Element Error: " + nm + " is not a " + et);
        uint32_t id;
        if (s.kind == Kind::kBasic) id = s.idx;
        else if (s.kind == Kind::kGate) id = (uint32_t)m.basic_names.size() + s.idx;
        else id = house_state[s.idx] ? CANOPY_CONST_TRUE : CANOPY_CONST_FALSE;
        return id | (complement ? CANOPY_CHILD_COMPLEMENT : 0u);
    }
    void add_arg(const XNode& a) {
        if (a.name == "constant") {
            const std::string* v = a.attr("value");
            if (!v) bad(CANOPY_E_VALIDITY, a, "<constant> needs a value");
            m.child.push_back(parse_bool(a, *v) ? CANOPY_CONST_TRUE : CANOPY_CONST_FALSE);
        } else if (a.name == "not") {
            const XNode* in = first_kid(a);
            if (!in || a.kids.size() != 1) bad(CANOPY_E_VALIDITY, a, "<not> as an argument takes exactly one event");
            m.child.push_back(event_code(*in, true));
        } else {
            m.child.push_back(event_code(a, false));
        }
    }
    void define_gate(uint32_t g) {
        const XNode& def = *gate_body[g];
        const XNode* f = first_kid(def);
        if (!f) bad(CANOPY_E_VALIDITY, def, "gate " + m.gate_names[g] + " has no formula");
        uint8_t conn;
        static const char* kNames[] = {"and", "or", "atleast", "xor", "not", "nand", "nor", "null", "iff", "imply", "cardinality"};
        if (f->attr("name") || f->name == "constant") conn = CANOPY_NULL;
        else {
            int pos = -1;
            for (int i = 0; i < CANOPY_NUM_CONNECTIVES; ++i) if (f->name == kNames[i]) pos = i;
            if (pos < 0) bad(CANOPY_E_UNSUPPORTED, *f, "unknown connective <" + f->name + ">");
            conn = (uint8_t)pos;
        }
        auto num = [&](const char* k) -> uint16_t {
            const std::string* v = f->attr(k);
            if (!v) return 0;
            char* end = nullptr; long x = std::strtol(v->c_str(), &end, 10);
            if (end == v->c_str() || *end) bad(CANOPY_E_VALIDITY, *f, std::string("bad ") + k + " number");
            if (x < 0) bad(CANOPY_E_INVALID, *f, std::string("The ") + k + " number cannot be negative: " + *v);  // event.cpp:118-119
            if (x > 65535) bad(CANOPY_E_UNSUPPORTED, *f, std::string(k) + " number above 65535");
            return (uint16_t)x;
        };
        if (conn == CANOPY_ATLEAST && !f->attr("min")) bad(CANOPY_E_VALIDITY, *f, "The connective requires min number for its args.");
        if (conn == CANOPY_CARDINALITY && (!f->attr("min") || !f->attr("max")))
            bad(CANOPY_E_VALIDITY, *f, "The connective requires min and max numbers for args.");
        m.op[g] = conn; m.kmin[g] = num("min"); m.kmax[g] = num("max");
        if (conn == CANOPY_NULL) add_arg(*f);
        else for (auto& a : f->kids) add_arg(*a);
    }
    void load(const XNode& root) {
        if (root.name != "opsa-mef") bad(CANOPY_E_IO, root, "document element must be <opsa-mef>");
        for (auto& k : root.kids) {
            if (k->name == "define-fault-tree") { need_name(*k); register_container(*k, n_trees++); }
            else if (k->name == "model-data") register_container(*k, -1);
            else if (k->name == "label" || k->name == "attributes") {}
            else bad(CANOPY_E_UNSUPPORTED, *k, "<" + k->name + "> is not supported by this loader");
        }
    }
    void finish() {
        const uint32_t B = (uint32_t)m.basic_names.size(), G = (uint32_t)m.gate_names.size();
        m.p.resize(B);
        for (uint32_t e = 0; e < B; ++e) {
            const XNode* b = first_kid(*basic_body[e]);
            if (!b) bad(CANOPY_E_VALIDITY, *basic_body[e], "basic event " + m.basic_names[e] + " has no expression");
            m.p[e] = expression(*b);
            if (!(m.p[e] >= 0.0 && m.p[e] <= 1.0)) bad(CANOPY_E_DOMAIN, *b, "Invalid probability value for " + m.basic_names[e]);
        }
        m.op.resize(G); m.kmin.resize(G); m.kmax.resize(G);
        m.child_off.assign(1, 0);
        for (uint32_t g = 0; g < G; ++g) { define_gate(g); m.child_off.push_back(m.child.size()); }
        // top events per fault tree (fault_tree.h:151-186)
        std::vector<bool> used(G, false);
        for (uint32_t g = 0; g < G; ++g)
            for (uint64_t j = m.child_off[g]; j < m.child_off[g + 1]; ++j) {
                const uint32_t id = m.child[j] & CANOPY_CHILD_ID_MASK;
                if (id >= B && id < B + G && gate_tree[id - B] == gate_tree[g]) used[id - B] = true;
            }
        for (uint32_t g = 0; g < G; ++g) if (!used[g] && gate_tree[g] >= 0) m.roots.push_back(B + g);
    }
};

std::string slurp(const char* path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw Error(CANOPY_E_IO, std::string("File doesn't exist or inaccessible: ") + path);
    std::ostringstream ss; ss << f.rdbuf();
    return ss.str();
}

canopy_model* load_docs(const std::vector<std::string>& texts) {
    auto model = std::make_unique<canopy_model>();
    Builder b{*model, {}, {}, {}, {}, {}, {}, 0};
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

canopy_status canopy_model_load_mef(const char* const* xml_paths, int32_t n_paths, canopy_model** out) {
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
        *out = canopy::load_docs(texts);
    });
}

canopy_status canopy_model_load_mef_string(const char* xml_text, canopy_model** out) {
    return canopy::guard_io([&] {
        if (!xml_text || !out) throw canopy::Error(CANOPY_E_INVALID, "canopy_model_load_mef_string: null argument");
        *out = nullptr;
        *out = canopy::load_docs({std::string(xml_text)});
    });
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
void canopy_model_destroy(canopy_model* m) { delete m; }

}  // extern "C"
