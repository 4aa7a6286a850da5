// jinja.hpp -- the subset of Jinja2 the reference's chat templates use
// (templates/opt-chat-template.yaml:6-25, templates/phi-chat-template.yaml:6-25, key
// `template.jinja`): {% if/elif/else/endif %}, {% set %} (visible after the enclosing if, as in
// Jinja), {% for x in seq %}, {{ expr }}, subscripts a[0] / a['k'] / a[1:], attribute access, string
// and integer literals, + (concatenation / addition), == != and or not, and {%- -%} / {{- -}}
// whitespace control.  Values are b2json::Value so the OpenAI `messages` array plugs in directly.
#pragma once
#include <cstring>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../json.hpp"

namespace b2jinja {
using b2json::Value;

struct Error : std::runtime_error {
  using std::runtime_error::runtime_error;
};

// ------------------------------------------------------------------------------------------
// expressions
// ------------------------------------------------------------------------------------------
struct Scope {
  std::vector<std::map<std::string, Value>> frames;
  const Value* find(const std::string& n) const {
    for (size_t i = frames.size(); i-- > 0;) {
      auto it = frames[i].find(n);
      if (it != frames[i].end()) return &it->second;
    }
    return nullptr;
  }
  void set(const std::string& n, Value v) { frames.back()[n] = std::move(v); }
};

struct Expr {
  virtual ~Expr() {}
  virtual Value eval(const Scope& s) const = 0;
};
using ExprP = std::unique_ptr<Expr>;

inline bool truthy(const Value& v) {
  switch (v.type) {
    case Value::Null: return false;
    case Value::Bool: return v.b;
    case Value::Number: return v.num != 0;
    case Value::String: return !v.str.empty();
    default: return v.size() > 0;
  }
}
inline bool equal(const Value& a, const Value& b) {
  if (a.type != b.type) return false;
  switch (a.type) {
    case Value::Null: return true;
    case Value::Bool: return a.b == b.b;
    case Value::Number: return a.num == b.num;
    case Value::String: return a.str == b.str;
    default: return b2json::dump(a) == b2json::dump(b);
  }
}
inline std::string to_text(const Value& v) {
  switch (v.type) {
    case Value::Null: return "None";
    case Value::Bool: return v.b ? "True" : "False";
    case Value::String: return v.str;
    case Value::Number: {
      char b[40];
      if (v.is_int) snprintf(b, sizeof b, "%lld", (long long)v.inum);
      else snprintf(b, sizeof b, "%g", v.num);
      return b;
    }
    default: return b2json::dump(v);
  }
}

struct Lit : Expr {
  Value v;
  explicit Lit(Value x) : v(std::move(x)) {}
  Value eval(const Scope&) const override { return v; }
};
struct Var : Expr {
  std::string name;
  explicit Var(std::string n) : name(std::move(n)) {}
  Value eval(const Scope& s) const override {
    const Value* v = s.find(name);
    return v ? *v : Value();  // undefined renders as empty / falsy
  }
};
struct Index : Expr {
  ExprP base, idx;
  Value eval(const Scope& s) const override {
    Value b = base->eval(s), i = idx->eval(s);
    if (b.is_array() && i.is_number()) {
      long long k = i.as_int();
      long long n = (long long)b.size();
      if (k < 0) k += n;
      if (k < 0 || k >= n) throw Error("jinja: list index out of range");
      return b.at((size_t)k);
    }
    if (b.is_object() && i.is_string()) {
      const Value* v = b.find(i.str);
      return v ? *v : Value();
    }
    if (b.is_string() && i.is_number()) {
      long long k = i.as_int(), n = (long long)b.str.size();
      if (k < 0) k += n;
      if (k < 0 || k >= n) throw Error("jinja: string index out of range");
      return Value(std::string(1, b.str[(size_t)k]));
    }
    return Value();
  }
};
struct Slice : Expr {
  ExprP base, lo, hi;  // lo/hi may be null
  Value eval(const Scope& s) const override {
    Value b = base->eval(s);
    long long n = b.is_array() ? (long long)b.size() : b.is_string() ? (long long)b.str.size() : 0;
    auto clampi = [&](const ExprP& e, long long dflt) {
      if (!e) return dflt;
      long long k = e->eval(s).as_int();
      if (k < 0) k += n;
      return std::max(0ll, std::min(n, k));
    };
    long long a = clampi(lo, 0), z = clampi(hi, n);
    if (b.is_string()) return Value(a < z ? b.str.substr((size_t)a, (size_t)(z - a)) : std::string());
    Value out = Value::array();
    for (long long k = a; k < z; ++k) out.push(b.at((size_t)k));
    return out;
  }
};
struct Unary : Expr {
  ExprP x;
  Value eval(const Scope& s) const override { return Value(!truthy(x->eval(s))); }
};
struct Binary : Expr {
  std::string op;
  ExprP a, b;
  Value eval(const Scope& s) const override {
    if (op == "and") { Value l = a->eval(s); return truthy(l) ? b->eval(s) : l; }
    if (op == "or") { Value l = a->eval(s); return truthy(l) ? l : b->eval(s); }
    Value l = a->eval(s), r = b->eval(s);
    if (op == "==") return Value(equal(l, r));
    if (op == "!=") return Value(!equal(l, r));
    if (op == "+") {
      if (l.is_number() && r.is_number()) {
        if (l.is_int && r.is_int) return Value((int64_t)(l.inum + r.inum));
        return Value(l.num + r.num);
      }
      if (l.is_array() && r.is_array()) {
        Value out = Value::array();
        for (size_t i = 0; i < l.size(); ++i) out.push(l.at(i));
        for (size_t i = 0; i < r.size(); ++i) out.push(r.at(i));
        return out;
      }
      return Value(to_text(l) + to_text(r));
    }
    throw Error("jinja: unsupported operator " + op);
  }
};

class ExprParser {
 public:
  explicit ExprParser(const std::string& s) : s_(s) {}
  ExprP parse_all() {
    ExprP e = parse_or();
    ws();
    if (p_ != s_.size()) throw Error("jinja: unexpected '" + s_.substr(p_) + "' in expression");
    return e;
  }
  ExprP parse_or() {
    ExprP l = parse_and();
    while (kw("or")) { auto b = std::make_unique<Binary>(); b->op = "or"; b->a = std::move(l); b->b = parse_and(); l = std::move(b); }
    return l;
  }
  size_t pos() const { return p_; }
  void ws() { while (p_ < s_.size() && isspace((unsigned char)s_[p_])) ++p_; }
  bool kw(const char* k) {
    ws();
    size_t n = strlen(k);
    if (s_.compare(p_, n, k) == 0 && (p_ + n == s_.size() || !(isalnum((unsigned char)s_[p_ + n]) || s_[p_ + n] == '_'))) { p_ += n; return true; }
    return false;
  }
  std::string ident() {
    ws();
    size_t a = p_;
    while (p_ < s_.size() && (isalnum((unsigned char)s_[p_]) || s_[p_] == '_')) ++p_;
    if (a == p_) throw Error("jinja: identifier expected in '" + s_ + "'");
    return s_.substr(a, p_ - a);
  }
  bool sym(const char* k) {
    ws();
    size_t n = strlen(k);
    if (s_.compare(p_, n, k) == 0) { p_ += n; return true; }
    return false;
  }

 private:
  const std::string& s_;
  size_t p_ = 0;
  ExprP parse_and() {
    ExprP l = parse_not();
    while (kw("and")) { auto b = std::make_unique<Binary>(); b->op = "and"; b->a = std::move(l); b->b = parse_not(); l = std::move(b); }
    return l;
  }
  ExprP parse_not() {
    if (kw("not")) { auto u = std::make_unique<Unary>(); u->x = parse_not(); return u; }
    return parse_cmp();
  }
  ExprP parse_cmp() {
    ExprP l = parse_add();
    while (true) {
      std::string op;
      if (sym("==")) op = "=="; else if (sym("!=")) op = "!="; else break;
      auto b = std::make_unique<Binary>(); b->op = op; b->a = std::move(l); b->b = parse_add(); l = std::move(b);
    }
    return l;
  }
  ExprP parse_add() {
    ExprP l = parse_postfix();
    while (true) {
      ws();
      if (p_ < s_.size() && s_[p_] == '+') { ++p_; auto b = std::make_unique<Binary>(); b->op = "+"; b->a = std::move(l); b->b = parse_postfix(); l = std::move(b); }
      else break;
    }
    return l;
  }
  ExprP parse_postfix() {
    ExprP e = parse_primary();
    while (true) {
      ws();
      if (p_ < s_.size() && s_[p_] == '[') {
        ++p_;
        ws();
        ExprP lo, hi;
        bool slice = false;
        if (p_ < s_.size() && s_[p_] != ':') lo = parse_or();
        ws();
        if (p_ < s_.size() && s_[p_] == ':') {
          slice = true;
          ++p_;
          ws();
          if (p_ < s_.size() && s_[p_] != ']') hi = parse_or();
        }
        ws();
        if (p_ >= s_.size() || s_[p_] != ']') throw Error("jinja: ']' expected");
        ++p_;
        if (slice) { auto sl = std::make_unique<Slice>(); sl->base = std::move(e); sl->lo = std::move(lo); sl->hi = std::move(hi); e = std::move(sl); }
        else { auto ix = std::make_unique<Index>(); ix->base = std::move(e); ix->idx = std::move(lo); e = std::move(ix); }
      } else if (p_ < s_.size() && s_[p_] == '.' && p_ + 1 < s_.size() && (isalpha((unsigned char)s_[p_ + 1]) || s_[p_ + 1] == '_')) {
        ++p_;
        std::string name = ident();
        auto ix = std::make_unique<Index>(); ix->base = std::move(e); ix->idx = std::make_unique<Lit>(Value(name)); e = std::move(ix);
      } else break;
    }
    return e;
  }
  ExprP parse_primary() {
    ws();
    if (p_ >= s_.size()) throw Error("jinja: expression expected");
    char c = s_[p_];
    if (c == '(') { ++p_; ExprP e = parse_or(); ws(); if (p_ >= s_.size() || s_[p_] != ')') throw Error("jinja: ')' expected"); ++p_; return e; }
    if (c == '\'' || c == '"') {
      char q = c;
      ++p_;
      std::string out;
      while (p_ < s_.size() && s_[p_] != q) {
        if (s_[p_] == '\\' && p_ + 1 < s_.size()) {
          char e = s_[p_ + 1];
          p_ += 2;
          switch (e) { case 'n': out += '\n'; break; case 't': out += '\t'; break; case 'r': out += '\r'; break;
                       case '\\': out += '\\'; break; case '\'': out += '\''; break; case '"': out += '"'; break;
                       default: out += '\\'; out += e; }
        } else out += s_[p_++];
      }
      if (p_ >= s_.size()) throw Error("jinja: unterminated string literal");
      ++p_;
      return std::make_unique<Lit>(Value(out));
    }
    if (isdigit((unsigned char)c) || (c == '-' && p_ + 1 < s_.size() && isdigit((unsigned char)s_[p_ + 1]))) {
      size_t a = p_++;
      while (p_ < s_.size() && isdigit((unsigned char)s_[p_])) ++p_;
      return std::make_unique<Lit>(Value((int64_t)std::stoll(s_.substr(a, p_ - a))));
    }
    std::string id = ident();
    if (id == "true" || id == "True") return std::make_unique<Lit>(Value(true));
    if (id == "false" || id == "False") return std::make_unique<Lit>(Value(false));
    if (id == "none" || id == "None") return std::make_unique<Lit>(Value());
    return std::make_unique<Var>(id);
  }
};

// ------------------------------------------------------------------------------------------
// template nodes
// ------------------------------------------------------------------------------------------
struct Node {
  virtual ~Node() {}
  virtual void render(Scope& s, std::string& out) const = 0;
};
using NodeP = std::unique_ptr<Node>;
using Body = std::vector<NodeP>;
inline void render_body(const Body& b, Scope& s, std::string& out) { for (auto& n : b) n->render(s, out); }

struct TextNode : Node { std::string t; void render(Scope&, std::string& out) const override { out += t; } };
struct OutNode : Node { ExprP e; void render(Scope& s, std::string& out) const override { Value v = e->eval(s); if (!v.is_null()) out += to_text(v); } };
struct SetNode : Node { std::string name; ExprP e; void render(Scope& s, std::string&) const override { s.set(name, e->eval(s)); } };
struct IfNode : Node {
  std::vector<std::pair<ExprP, Body>> branches;  // null condition = else
  void render(Scope& s, std::string& out) const override {
    for (auto& br : branches)
      if (!br.first || truthy(br.first->eval(s))) { render_body(br.second, s, out); return; }  // no new scope: `set` escapes the if
  }
};
struct ForNode : Node {
  std::string var;
  ExprP iter;
  Body body;
  void render(Scope& s, std::string& out) const override {
    Value seq = iter->eval(s);
    if (!seq.is_array()) return;
    const size_t n = seq.size();
    for (size_t i = 0; i < n; ++i) {
      s.frames.emplace_back();  // the loop body has its own scope
      s.set(var, seq.at(i));
      Value loop = Value::object();
      loop.set("index0", Value((int64_t)i));
      loop.set("index", Value((int64_t)i + 1));
      loop.set("first", Value(i == 0));
      loop.set("last", Value(i + 1 == n));
      loop.set("length", Value((int64_t)n));
      s.set("loop", loop);
      render_body(body, s, out);
      s.frames.pop_back();
    }
  }
};

class Template {
 public:
  // trim_blocks / lstrip_blocks as HF's chat-template environment sets them (both true)
  explicit Template(const std::string& src, bool trim_blocks = true, bool lstrip_blocks = true) {
    lex(src, trim_blocks, lstrip_blocks);
    size_t i = 0;
    body_ = parse_until(i, {});
    if (i != toks_.size()) throw Error("jinja: unexpected '" + toks_[i].text + "'");
  }
  std::string render(const Value& context) const {
    Scope s;
    s.frames.emplace_back();
    if (context.is_object()) for (auto& kv : *context.obj) s.set(kv.first, kv.second);
    std::string out;
    render_body(body_, s, out);
    return out;
  }

 private:
  struct Tok { int kind; std::string text; };  // 0 text, 1 {{ expr }}, 2 {% stmt %}
  std::vector<Tok> toks_;
  Body body_;

  static void rstrip(std::string& s) { while (!s.empty() && isspace((unsigned char)s.back())) s.pop_back(); }
  void lex(const std::string& src, bool trim_blocks, bool lstrip_blocks) {
    size_t p = 0;
    bool strip_next = false, trim_nl_next = false;
    bool line_start = true;  // does the text chunk starting at p begin at the start of a line?
    while (p < src.size()) {
      size_t q = p;
      int kind = -1;
      while (q + 1 < src.size()) {
        if (src[q] == '{' && (src[q + 1] == '{' || src[q + 1] == '%' || src[q + 1] == '#')) { kind = src[q + 1] == '{' ? 1 : src[q + 1] == '%' ? 2 : 3; break; }
        ++q;
      }
      if (kind < 0) q = src.size();
      std::string text = src.substr(p, q - p);
      if (strip_next) { size_t a = 0; while (a < text.size() && isspace((unsigned char)text[a])) ++a; text.erase(0, a); }
      else if (trim_nl_next && !text.empty() && text[0] == '\n') { text.erase(0, 1); line_start = true; }
      strip_next = trim_nl_next = false;
      if (kind < 0) { if (!text.empty()) toks_.push_back({0, text}); break; }
      const char* close = kind == 1 ? "}}" : kind == 2 ? "%}" : "#}";
      size_t e = src.find(close, q + 2);
      if (e == std::string::npos) throw Error("jinja: unterminated tag");
      std::string inner = src.substr(q + 2, e - q - 2);
      bool lstrip = false, rstrip_after = false, no_lstrip = false;
      if (!inner.empty() && inner[0] == '-') { lstrip = true; inner.erase(0, 1); }
      else if (!inner.empty() && inner[0] == '+') { no_lstrip = true; inner.erase(0, 1); }
      if (!inner.empty() && inner.back() == '-') { rstrip_after = true; inner.pop_back(); }
      if (lstrip) rstrip(text);
      else if (lstrip_blocks && kind != 1 && !no_lstrip) {
        // block tag alone at the start of a line: drop the indentation in front of it
        size_t nl = text.rfind('\n');
        size_t a = nl == std::string::npos ? 0 : nl + 1;
        bool blank = true;
        for (size_t k = a; k < text.size(); ++k) if (text[k] != ' ' && text[k] != '\t') blank = false;
        if (blank && (nl != std::string::npos || line_start)) text.erase(a);
      }
      if (!text.empty()) toks_.push_back({0, text});
      if (kind != 3) toks_.push_back({kind, inner});
      strip_next = rstrip_after;
      trim_nl_next = !rstrip_after && trim_blocks && kind != 1;
      line_start = false;
      p = e + 2;
    }
  }
  static std::string first_word(const std::string& s, size_t* rest) {
    size_t a = 0;
    while (a < s.size() && isspace((unsigned char)s[a])) ++a;
    size_t b = a;
    while (b < s.size() && (isalnum((unsigned char)s[b]) || s[b] == '_')) ++b;
    *rest = b;
    return s.substr(a, b - a);
  }
  Body parse_until(size_t& i, const std::vector<std::string>& stops) {
    Body body;
    while (i < toks_.size()) {
      const Tok& t = toks_[i];
      if (t.kind == 0) { auto n = std::make_unique<TextNode>(); n->t = t.text; body.push_back(std::move(n)); ++i; continue; }
      if (t.kind == 1) { auto n = std::make_unique<OutNode>(); n->e = ExprParser(t.text).parse_all(); body.push_back(std::move(n)); ++i; continue; }
      size_t rest;
      std::string w = first_word(t.text, &rest);
      for (auto& s : stops) if (w == s) return body;
      std::string tail = t.text.substr(rest);
      if (w == "if") {
        auto n = std::make_unique<IfNode>();
        ExprP cond = ExprParser(tail).parse_all();
        ++i;
        while (true) {
          Body b = parse_until(i, {"elif", "else", "endif"});
          if (i >= toks_.size()) throw Error("jinja: missing endif");
          n->branches.emplace_back(std::move(cond), std::move(b));
          size_t r2;
          std::string w2 = first_word(toks_[i].text, &r2);
          std::string tail2 = toks_[i].text.substr(r2);
          ++i;
          if (w2 == "endif") break;
          if (w2 == "elif") { cond = ExprParser(tail2).parse_all(); continue; }
          Body eb = parse_until(i, {"endif"});  // else
          if (i >= toks_.size()) throw Error("jinja: missing endif");
          n->branches.emplace_back(nullptr, std::move(eb));
          ++i;
          break;
        }
        body.push_back(std::move(n));
      } else if (w == "for") {
        auto n = std::make_unique<ForNode>();
        ExprParser ep(tail);
        n->var = ep.ident();
        if (!ep.kw("in")) throw Error("jinja: 'in' expected in for");
        std::string it = tail.substr(ep.pos());
        n->iter = ExprParser(it).parse_all();
        ++i;
        n->body = parse_until(i, {"endfor"});
        if (i >= toks_.size()) throw Error("jinja: missing endfor");
        ++i;
        body.push_back(std::move(n));
      } else if (w == "set") {
        auto n = std::make_unique<SetNode>();
        ExprParser ep(tail);
        n->name = ep.ident();
        if (!ep.sym("=")) throw Error("jinja: '=' expected in set");
        std::string ex = tail.substr(ep.pos());
        n->e = ExprParser(ex).parse_all();
        body.push_back(std::move(n));
        ++i;
      } else {
        throw Error("jinja: unsupported statement '" + w + "'");
      }
    }
    return body;
  }
};

// Pulls the template text out of a ConfigMap YAML of the reference's shape (key `template.jinja: |-`
// followed by an indented block) or returns the input unchanged when it is already a bare template.
inline std::string extract_template(const std::string& file_text) {
  const std::string key = "template.jinja:";
  size_t k = file_text.find(key);
  if (k == std::string::npos) return file_text;
  size_t nl = file_text.find('\n', k);
  if (nl == std::string::npos) return std::string();
  std::string header = file_text.substr(k + key.size(), nl - k - key.size());
  const bool keep_last_newline = header.find('-') == std::string::npos;  // |- strips the final newline
  // indentation of the block = indentation of its first non-empty line
  size_t p = nl + 1;
  std::vector<std::string> lines;
  int indent = -1;
  while (p <= file_text.size()) {
    size_t e = file_text.find('\n', p);
    std::string line = file_text.substr(p, (e == std::string::npos ? file_text.size() : e) - p);
    size_t a = 0;
    while (a < line.size() && line[a] == ' ') ++a;
    const bool blank = a == line.size();
    if (indent < 0 && !blank) indent = (int)a;
    if (!blank && (int)a < indent) break;  // dedent: block ended
    lines.push_back(blank ? std::string() : line.substr(std::min((size_t)std::max(indent, 0), line.size())));
    if (e == std::string::npos) break;
    p = e + 1;
  }
  while (!lines.empty() && lines.back().empty()) lines.pop_back();
  std::string out;
  for (size_t i = 0; i < lines.size(); ++i) { out += lines[i]; if (i + 1 < lines.size()) out += '\n'; }
  if (keep_last_newline) out += '\n';
  return out;
}

}  // namespace b2jinja
