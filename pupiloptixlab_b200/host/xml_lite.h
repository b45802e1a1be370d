// xml_lite.h — a small XML DOM reader for mitsuba-style scene files: elements, attributes,
// comments, <?...?> / <!...> prolog items; text content is ignored (the scene format keeps all data
// in attributes).  Stands in for pugixml as used by framework/resource/xml/parser.cpp:53-59.
#pragma once
#include <memory>
#include <string>
#include <utility>
#include <vector>

namespace pupil::host::xml {

struct Node {
    std::string name;
    std::vector<std::pair<std::string, std::string>> attrs;
    std::vector<std::unique_ptr<Node>> children;
    const std::string* attr(const std::string& n) const {
        for (auto& a : attrs) if (a.first == n) return &a.second;
        return nullptr;
    }
    std::string attr_or(const std::string& n, const std::string& d = "") const { auto* a = attr(n); return a ? *a : d; }
    bool has(const std::string& n) const { return attr(n) != nullptr; }
};

class Reader {
public:
    explicit Reader(const std::string& text) : s_(text) {}
    // returns the document element or nullptr (error() says why)
    std::unique_ptr<Node> parse() {
        skip_misc();
        if (p_ >= s_.size() || s_[p_] != '<') { err_ = "no document element"; return nullptr; }
        auto root = element();
        if (!root && err_.empty()) err_ = "malformed document element";
        return root;
    }
    const std::string& error() const { return err_; }

private:
    const std::string& s_;
    size_t p_ = 0;
    std::string err_;

    static bool is_space(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r'; }
    static bool is_name(char c) { return (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || (c >= '0' && c <= '9') || c == '_' || c == '-' || c == ':' || c == '.'; }
    void skip_space() { while (p_ < s_.size() && is_space(s_[p_])) ++p_; }
    bool starts(const char* t) const { return s_.compare(p_, std::char_traits<char>::length(t), t) == 0; }
    void skip_until(const char* t) {
        size_t e = s_.find(t, p_);
        p_ = e == std::string::npos ? s_.size() : e + std::char_traits<char>::length(t);
    }
    void skip_misc() {  // whitespace, text, comments, processing instructions, doctype
        while (p_ < s_.size()) {
            if (starts("<!--")) skip_until("-->");
            else if (starts("<?")) skip_until("?>");
            else if (starts("<!")) skip_until(">");
            else if (s_[p_] == '<') return;
            else ++p_;
        }
    }
    static std::string decode(const std::string& v) {
        if (v.find('&') == std::string::npos) return v;
        std::string o;
        for (size_t i = 0; i < v.size(); ++i) {
            if (v[i] != '&') { o += v[i]; continue; }
            auto rep = [&](const char* ent, char c) { size_t n = std::char_traits<char>::length(ent); if (v.compare(i, n, ent) == 0) { o += c; i += n - 1; return true; } return false; };
            if (rep("&amp;", '&') || rep("&lt;", '<') || rep("&gt;", '>') || rep("&quot;", '"') || rep("&apos;", '\'')) continue;
            o += v[i];
        }
        return o;
    }
    std::unique_ptr<Node> element() {
        ++p_;  // '<'
        auto n = std::make_unique<Node>();
        while (p_ < s_.size() && is_name(s_[p_])) n->name += s_[p_++];
        if (n->name.empty()) { err_ = "element without a name at offset " + std::to_string(p_); return nullptr; }
        while (true) {
            skip_space();
            if (p_ >= s_.size()) { err_ = "unterminated element <" + n->name + ">"; return nullptr; }
            if (starts("/>")) { p_ += 2; return n; }
            if (s_[p_] == '>') { ++p_; break; }
            std::string an;
            while (p_ < s_.size() && is_name(s_[p_])) an += s_[p_++];
            skip_space();
            if (an.empty() || p_ >= s_.size() || s_[p_] != '=') { err_ = "malformed attribute in <" + n->name + ">"; return nullptr; }
            ++p_;
            skip_space();
            if (p_ >= s_.size() || (s_[p_] != '"' && s_[p_] != '\'')) { err_ = "attribute value must be quoted in <" + n->name + ">"; return nullptr; }
            const char q = s_[p_++];
            size_t e = s_.find(q, p_);
            if (e == std::string::npos) { err_ = "unterminated attribute value in <" + n->name + ">"; return nullptr; }
            n->attrs.emplace_back(an, decode(s_.substr(p_, e - p_)));
            p_ = e + 1;
        }
        while (true) {  // children until </name>
            skip_misc();
            if (p_ >= s_.size()) { err_ = "missing </" + n->name + ">"; return nullptr; }
            if (starts("</")) {
                p_ += 2;
                std::string cn;
                while (p_ < s_.size() && is_name(s_[p_])) cn += s_[p_++];
                skip_space();
                if (cn != n->name || p_ >= s_.size() || s_[p_] != '>') { err_ = "mismatched </" + cn + "> for <" + n->name + ">"; return nullptr; }
                ++p_;
                return n;
            }
            auto c = element();
            if (!c) return nullptr;
            n->children.push_back(std::move(c));
        }
    }
};

}  // namespace pupil::host::xml
