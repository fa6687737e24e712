// xml.h — a small XML DOM reader, enough for the reference's scene files (inputs/*.xml; schema in
// src/parser/xmlload.cpp:71-630). Elements, attributes, comments, <? ?> and <! > declarations; text is skipped.
// Attribute queries follow the TinyXML calls the reference makes: Attribute(name) is case-sensitive,
// QueryDoubleAttribute parses a leading number (sscanf "%lf") and leaves the target untouched when the
// attribute is missing or not a number.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

namespace qb {

struct XmlElement {
  std::string name;
  std::vector<std::pair<std::string, std::string>> attrs;
  std::vector<std::unique_ptr<XmlElement>> children;

  const char *attribute(const char *key) const
  {
    for (auto &a : attrs)
      if (a.first == key) return a.second.c_str();
    return nullptr;
  }
  bool query_double(const char *key, double *out) const
  {
    const char *s = attribute(key);
    if (!s) return false;
    double d;
    if (sscanf(s, "%lf", &d) != 1) return false;
    *out = d;
    return true;
  }
  bool query_int(const char *key, int *out) const
  {
    const char *s = attribute(key);
    if (!s) return false;
    int d;
    if (sscanf(s, "%d", &d) != 1) return false;
    *out = d;
    return true;
  }
  const XmlElement *first_child(const char *tag) const
  {
    for (auto &c : children)
      if (c->name == tag) return c.get();
    return nullptr;
  }
};

class XmlParser {
 public:
  // returns a synthetic root whose children are the top-level elements; nullptr on failure (error() says why)
  std::unique_ptr<XmlElement> parse_file(const char *path)
  {
    FILE *fp = fopen(path, "rb");
    if (!fp) { err_ = std::string("cannot open ") + path; return nullptr; }
    std::string buf;
    char tmp[65536];
    size_t n;
    while ((n = fread(tmp, 1, sizeof tmp, fp)) > 0) buf.append(tmp, n);
    fclose(fp);
    return parse(buf);
  }
  std::unique_ptr<XmlElement> parse(const std::string &text)
  {
    s_ = text.c_str();
    e_ = s_ + text.size();
    err_.clear();
    std::unique_ptr<XmlElement> root(new XmlElement);
    root->name = "#document";
    if (!parse_children(root.get(), nullptr)) return nullptr;
    return root;
  }
  const std::string &error() const { return err_; }

 private:
  const char *s_ = nullptr, *e_ = nullptr;
  std::string err_;

  static bool is_space(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r'; }
  static bool is_name(char c)
  {
    return (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || (c >= '0' && c <= '9') || c == '_' || c == '-' || c == ':' || c == '.';
  }
  void skip_space() { while (s_ < e_ && is_space(*s_)) ++s_; }
  bool starts(const char *lit) const
  {
    size_t n = strlen(lit);
    return (size_t) (e_ - s_) >= n && memcmp(s_, lit, n) == 0;
  }
  bool skip_until(const char *lit)
  {
    size_t n = strlen(lit);
    while ((size_t) (e_ - s_) >= n) {
      if (memcmp(s_, lit, n) == 0) { s_ += n; return true; }
      ++s_;
    }
    err_ = std::string("unterminated construct, expected ") + lit;
    return false;
  }
  static std::string decode(const std::string &v)
  {
    if (v.find('&') == std::string::npos) return v;
    std::string o;
    for (size_t i = 0; i < v.size();) {
      if (v.compare(i, 5, "&amp;") == 0) { o += '&'; i += 5; }
      else if (v.compare(i, 4, "&lt;") == 0) { o += '<'; i += 4; }
      else if (v.compare(i, 4, "&gt;") == 0) { o += '>'; i += 4; }
      else if (v.compare(i, 6, "&quot;") == 0) { o += '"'; i += 6; }
      else if (v.compare(i, 6, "&apos;") == 0) { o += '\''; i += 6; }
      else o += v[i++];
    }
    return o;
  }
  // parses content until the closing tag of `parent` (or end of input for the document)
  bool parse_children(XmlElement *parent, const char *closing)
  {
    for (;;) {
      while (s_ < e_ && *s_ != '<') ++s_;  // text is ignored
      if (s_ >= e_) {
        if (closing) { err_ = std::string("missing </") + closing + ">"; return false; }
        return true;
      }
      if (starts("<!--")) { if (!skip_until("-->")) return false; continue; }
      if (starts("<?")) { if (!skip_until("?>")) return false; continue; }
      if (starts("<!")) { if (!skip_until(">")) return false; continue; }
      if (starts("</")) {
        s_ += 2;
        const char *b = s_;
        while (s_ < e_ && is_name(*s_)) ++s_;
        std::string nm(b, s_);
        skip_space();
        if (s_ >= e_ || *s_ != '>') { err_ = "malformed closing tag"; return false; }
        ++s_;
        if (!closing || nm != closing) { err_ = "mismatched closing tag </" + nm + ">"; return false; }
        return true;
      }
      ++s_;  // '<'
      const char *b = s_;
      while (s_ < e_ && is_name(*s_)) ++s_;
      if (s_ == b) { err_ = "malformed tag"; return false; }
      std::unique_ptr<XmlElement> el(new XmlElement);
      el->name.assign(b, s_);
      bool self_closed = false;
      for (;;) {
        skip_space();
        if (s_ >= e_) { err_ = "unterminated tag <" + el->name; return false; }
        if (*s_ == '>') { ++s_; break; }
        if (*s_ == '/' && s_ + 1 < e_ && s_[1] == '>') { s_ += 2; self_closed = true; break; }
        const char *kb = s_;
        while (s_ < e_ && is_name(*s_)) ++s_;
        if (s_ == kb) { err_ = "malformed attribute in <" + el->name; return false; }
        std::string key(kb, s_);
        skip_space();
        if (s_ >= e_ || *s_ != '=') { err_ = "attribute without value in <" + el->name; return false; }
        ++s_;
        skip_space();
        if (s_ >= e_ || (*s_ != '"' && *s_ != '\'')) { err_ = "unquoted attribute in <" + el->name; return false; }
        const char q = *s_++;
        const char *vb = s_;
        while (s_ < e_ && *s_ != q) ++s_;
        if (s_ >= e_) { err_ = "unterminated attribute value in <" + el->name; return false; }
        el->attrs.emplace_back(key, decode(std::string(vb, s_)));
        ++s_;
      }
      if (!self_closed) {
        const std::string nm = el->name;
        if (!parse_children(el.get(), nm.c_str())) return false;
      }
      parent->children.push_back(std::move(el));
    }
  }
};

}  // namespace qb
