// Minimal XML DOM for the Open-PSA MEF reader (csrc/mef_model.cpp): elements, attributes, text, comments, CDATA,
// processing instructions and DOCTYPE are parsed or skipped; no namespaces, no external entities, no XInclude.
// Stands in for the reference's libxml2 wrapper (reference src/io/xml/document.h:15-33, element.h:18-96), which
// cannot be built here (no libxml2).  Malformed input -> CANOPY_E_IO with the line number.
#pragma once
#include <cctype>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <utility>
#include <vector>

#include "plan.h"

namespace canopy {

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
            else if (ent.size() > 1 && ent[0] == '#') {   // numeric character reference -> UTF-8
                char* endp = nullptr;
                const bool hex = ent[1] == 'x' || ent[1] == 'X';
                const long cp = std::strtol(ent.c_str() + (hex ? 2 : 1), &endp, hex ? 16 : 10);
                if (endp == ent.c_str() + (hex ? 2 : 1) || *endp || cp <= 0 || cp > 0x10FFFF || (cp >= 0xD800 && cp <= 0xDFFF))
                    err("bad character reference &" + ent + ";");
                if (cp < 0x80) o += (char)cp;
                else if (cp < 0x800) { o += (char)(0xC0 | (cp >> 6)); o += (char)(0x80 | (cp & 0x3F)); }
                else if (cp < 0x10000) { o += (char)(0xE0 | (cp >> 12)); o += (char)(0x80 | ((cp >> 6) & 0x3F)); o += (char)(0x80 | (cp & 0x3F)); }
                else { o += (char)(0xF0 | (cp >> 18)); o += (char)(0x80 | ((cp >> 12) & 0x3F)); o += (char)(0x80 | ((cp >> 6) & 0x3F)); o += (char)(0x80 | (cp & 0x3F)); }
            }
            else err("unknown entity &" + ent + ";");
            k = e;
        }
        return o;
    }
    // Recursion depth of element(): a model file nested deeper than this is refused (CANOPY_E_IO) instead of overflowing
    // the stack, which no exception handler could turn into a status code.
    static constexpr int kMaxDepth = 512;
    int depth_ = 0;
    struct DepthGuard { int& d; explicit DepthGuard(int& x) : d(x) { ++d; } ~DepthGuard() { --d; } };
    std::unique_ptr<XNode> element() {
        if (i_ >= s_.size() || s_[i_] != '<') err("expected '<'");
        DepthGuard dg(depth_);
        if (depth_ > kMaxDepth) err("elements nested deeper than " + std::to_string(kMaxDepth) + " levels");
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

}  // namespace canopy
