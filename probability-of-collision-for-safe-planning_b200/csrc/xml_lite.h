// xml_lite.h -- the subset of XML the OpenRAVE scene files use: elements, attributes, text, comments,
// processing instructions.  No entities beyond the five predefined ones, no DTD, no namespaces.
// Input side of the hot path: pr2custom.env.xml / pr2test2.env.xml and the kinbody / robot files they name.
#ifndef MCSIM_XML_LITE_H
#define MCSIM_XML_LITE_H

#include <cctype>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace xml_lite {

struct Node {
    std::string tag;                                          // as written
    std::string ltag;                                         // lower-cased (tag case varies in the scene files)
    std::vector<std::pair<std::string, std::string>> attrs;   // attribute names lower-cased
    std::string text;                                         // concatenated character data
    std::vector<std::unique_ptr<Node>> children;

    const std::string* attr(const std::string& lname) const {
        for (const auto& a : attrs) if (a.first == lname) return &a.second;
        return nullptr;
    }
    std::vector<const Node*> all(const std::string& lname) const {
        std::vector<const Node*> out;
        for (const auto& c : children) if (c->ltag == lname) out.push_back(c.get());
        return out;
    }
    const Node* first(const std::string& lname) const {
        for (const auto& c : children) if (c->ltag == lname) return c.get();
        return nullptr;
    }
};

inline std::string lower(std::string s) {
    for (auto& ch : s) ch = char(std::tolower(static_cast<unsigned char>(ch)));
    return s;
}

class Parser {
  public:
    explicit Parser(const std::string& src) : s_(src), i_(0) {}

    std::unique_ptr<Node> parse_document() {
        skip_misc();
        if (i_ >= s_.size() || s_[i_] != '<') fail("no root element");
        auto root = parse_element();
        skip_misc();
        return root;
    }

  private:
    const std::string& s_;
    size_t i_;

    [[noreturn]] void fail(const std::string& what) const {
        size_t line = 1;
        for (size_t k = 0; k < i_ && k < s_.size(); ++k) if (s_[k] == '\n') ++line;
        throw std::runtime_error("XML parse error at line " + std::to_string(line) + ": " + what);
    }
    bool starts(const char* lit) const { return s_.compare(i_, std::char_traits<char>::length(lit), lit) == 0; }
    void skip_ws() { while (i_ < s_.size() && std::isspace(static_cast<unsigned char>(s_[i_]))) ++i_; }
    void skip_until(const char* lit) {
        const size_t p = s_.find(lit, i_);
        if (p == std::string::npos) fail(std::string("unterminated construct, expected ") + lit);
        i_ = p + std::char_traits<char>::length(lit);
    }
    // whitespace, comments, <? ... ?>, <!DOCTYPE ...>
    void skip_misc() {
        for (;;) {
            skip_ws();
            if (starts("<!--")) { i_ += 4; skip_until("-->"); }
            else if (starts("<?")) { i_ += 2; skip_until("?>"); }
            else if (starts("<!")) { i_ += 2; skip_until(">"); }
            else return;
        }
    }
    static bool name_char(char c) {
        return std::isalnum(static_cast<unsigned char>(c)) || c == '_' || c == '-' || c == ':' || c == '.';
    }
    std::string parse_name() {
        const size_t b = i_;
        while (i_ < s_.size() && name_char(s_[i_])) ++i_;
        if (i_ == b) fail("expected a name");
        return s_.substr(b, i_ - b);
    }
    static std::string unescape(const std::string& in) {
        std::string out;
        out.reserve(in.size());
        for (size_t k = 0; k < in.size(); ++k) {
            if (in[k] != '&') { out.push_back(in[k]); continue; }
            static const std::pair<const char*, char> ents[] = {{"&lt;", '<'}, {"&gt;", '>'}, {"&amp;", '&'}, {"&quot;", '"'}, {"&apos;", '\''}};
            bool hit = false;
            for (const auto& e : ents) {
                const size_t n = std::char_traits<char>::length(e.first);
                if (in.compare(k, n, e.first) == 0) { out.push_back(e.second); k += n - 1; hit = true; break; }
            }
            if (!hit) out.push_back('&');
        }
        return out;
    }
    std::unique_ptr<Node> parse_element() {
        ++i_;  // '<'
        auto n = std::make_unique<Node>();
        n->tag = parse_name();
        n->ltag = lower(n->tag);
        for (;;) {
            skip_ws();
            if (i_ >= s_.size()) fail("unterminated start tag <" + n->tag);
            if (starts("/>")) { i_ += 2; return n; }
            if (s_[i_] == '>') { ++i_; break; }
            std::string an = lower(parse_name());
            skip_ws();
            if (i_ >= s_.size() || s_[i_] != '=') fail("attribute " + an + " has no value");
            ++i_;
            skip_ws();
            if (i_ >= s_.size() || (s_[i_] != '"' && s_[i_] != '\'')) fail("attribute value must be quoted");
            const char q = s_[i_++];
            const size_t e = s_.find(q, i_);
            if (e == std::string::npos) fail("unterminated attribute value");
            n->attrs.emplace_back(an, unescape(s_.substr(i_, e - i_)));
            i_ = e + 1;
        }
        // content
        for (;;) {
            if (i_ >= s_.size()) fail("unterminated element <" + n->tag + ">");
            if (s_[i_] != '<') {
                const size_t e = s_.find('<', i_);
                const size_t stop = e == std::string::npos ? s_.size() : e;
                n->text += unescape(s_.substr(i_, stop - i_));
                i_ = stop;
                continue;
            }
            if (starts("<!--")) { i_ += 4; skip_until("-->"); continue; }
            if (starts("<![CDATA[")) {
                i_ += 9;
                const size_t e = s_.find("]]>", i_);
                if (e == std::string::npos) fail("unterminated CDATA");
                n->text += s_.substr(i_, e - i_);
                i_ = e + 3;
                continue;
            }
            if (starts("<?")) { i_ += 2; skip_until("?>"); continue; }
            if (starts("</")) {
                i_ += 2;
                const std::string close = parse_name();
                if (lower(close) != n->ltag) fail("</" + close + "> closes <" + n->tag + ">");
                skip_ws();
                if (i_ >= s_.size() || s_[i_] != '>') fail("malformed end tag");
                ++i_;
                return n;
            }
            n->children.push_back(parse_element());
        }
    }
};

inline std::unique_ptr<Node> parse(const std::string& text) { return Parser(text).parse_document(); }

}  // namespace xml_lite
#endif
