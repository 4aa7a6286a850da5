// json.hpp -- small self-contained JSON value / parser / writer used by the safetensors
// loader and by the HTTP server (OpenAI request/response bodies).  UTF-8 passes through
// untouched; \uXXXX escapes (including surrogate pairs) are decoded to UTF-8.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace b2json {

struct Value;
using Array = std::vector<Value>;
using Object = std::vector<std::pair<std::string, Value>>;  // insertion order preserved

struct Value {
  enum Type { Null, Bool, Number, String, Arr, Obj } type = Null;
  bool b = false;
  double num = 0;
  bool is_int = false;
  int64_t inum = 0;
  std::string str;
  std::shared_ptr<Array> arr;
  std::shared_ptr<Object> obj;

  Value() {}
  Value(bool v) : type(Bool), b(v) {}
  Value(double v) : type(Number), num(v) {}
  Value(int v) : type(Number), num(v), is_int(true), inum(v) {}
  Value(int64_t v) : type(Number), num((double)v), is_int(true), inum(v) {}
  Value(const char* s) : type(String), str(s) {}
  Value(const std::string& s) : type(String), str(s) {}
  static Value array() { Value v; v.type = Arr; v.arr = std::make_shared<Array>(); return v; }
  static Value object() { Value v; v.type = Obj; v.obj = std::make_shared<Object>(); return v; }

  bool is_null() const { return type == Null; }
  bool is_bool() const { return type == Bool; }
  bool is_number() const { return type == Number; }
  bool is_string() const { return type == String; }
  bool is_array() const { return type == Arr; }
  bool is_object() const { return type == Obj; }

  const Value* find(const std::string& k) const {
    if (type != Obj) return nullptr;
    for (auto& kv : *obj) if (kv.first == k) return &kv.second;
    return nullptr;
  }
  Value& set(const std::string& k, Value v) {
    if (type != Obj) { type = Obj; obj = std::make_shared<Object>(); }
    for (auto& kv : *obj) if (kv.first == k) { kv.second = std::move(v); return kv.second; }
    obj->emplace_back(k, std::move(v));
    return obj->back().second;
  }
  Value& push(Value v) {
    if (type != Arr) { type = Arr; arr = std::make_shared<Array>(); }
    arr->push_back(std::move(v));
    return arr->back();
  }
  size_t size() const { return type == Arr ? arr->size() : type == Obj ? obj->size() : 0; }
  const Value& at(size_t i) const { return (*arr)[i]; }
  int64_t as_int(int64_t dflt = 0) const { return type == Number ? (is_int ? inum : (int64_t)num) : dflt; }
  double as_double(double dflt = 0) const { return type == Number ? num : dflt; }
  bool as_bool(bool dflt = false) const { return type == Bool ? b : dflt; }
  const std::string& as_string() const { static const std::string e; return type == String ? str : e; }
};

class Parser {
 public:
  Parser(const char* p, size_t n) : p_(p), end_(p + n) {}
  Value parse() {
    ws();
    Value v = value(0);
    ws();
    if (p_ != end_) fail("trailing characters");
    return v;
  }

 private:
  const char* p_;
  const char* end_;
  [[noreturn]] void fail(const char* m) { throw std::runtime_error(std::string("json: ") + m); }
  void ws() { while (p_ < end_ && (*p_ == ' ' || *p_ == '\t' || *p_ == '\n' || *p_ == '\r')) ++p_; }
  bool lit(const char* s) {
    size_t n = strlen_(s);
    if ((size_t)(end_ - p_) >= n && std::equal(s, s + n, p_)) { p_ += n; return true; }
    return false;
  }
  static size_t strlen_(const char* s) { size_t n = 0; while (s[n]) ++n; return n; }
  static void utf8(std::string& o, uint32_t c) {
    if (c < 0x80) o += (char)c;
    else if (c < 0x800) { o += (char)(0xC0 | (c >> 6)); o += (char)(0x80 | (c & 0x3F)); }
    else if (c < 0x10000) { o += (char)(0xE0 | (c >> 12)); o += (char)(0x80 | ((c >> 6) & 0x3F)); o += (char)(0x80 | (c & 0x3F)); }
    else { o += (char)(0xF0 | (c >> 18)); o += (char)(0x80 | ((c >> 12) & 0x3F)); o += (char)(0x80 | ((c >> 6) & 0x3F)); o += (char)(0x80 | (c & 0x3F)); }
  }
  uint32_t hex4() {
    if (end_ - p_ < 4) fail("bad \\u escape");
    uint32_t v = 0;
    for (int i = 0; i < 4; ++i) {
      char c = *p_++;
      v <<= 4;
      if (c >= '0' && c <= '9') v |= c - '0';
      else if (c >= 'a' && c <= 'f') v |= c - 'a' + 10;
      else if (c >= 'A' && c <= 'F') v |= c - 'A' + 10;
      else fail("bad hex digit");
    }
    return v;
  }
  std::string string_() {
    if (p_ >= end_ || *p_ != '"') fail("expected string");
    ++p_;
    std::string o;
    while (true) {
      if (p_ >= end_) fail("unterminated string");
      char c = *p_++;
      if (c == '"') break;
      if (c == '\\') {
        if (p_ >= end_) fail("bad escape");
        char e = *p_++;
        switch (e) {
          case '"': o += '"'; break; case '\\': o += '\\'; break; case '/': o += '/'; break;
          case 'b': o += '\b'; break; case 'f': o += '\f'; break; case 'n': o += '\n'; break;
          case 'r': o += '\r'; break; case 't': o += '\t'; break;
          case 'u': {
            uint32_t c1 = hex4();
            if (c1 >= 0xD800 && c1 <= 0xDBFF && end_ - p_ >= 6 && p_[0] == '\\' && p_[1] == 'u') {
              p_ += 2;
              uint32_t c2 = hex4();
              if (c2 >= 0xDC00 && c2 <= 0xDFFF) c1 = 0x10000 + ((c1 - 0xD800) << 10) + (c2 - 0xDC00);
            }
            utf8(o, c1);
            break;
          }
          default: fail("unknown escape");
        }
      } else {
        o += c;
      }
    }
    return o;
  }
  Value value(int depth) {
    if (depth > 64) fail("nesting too deep");
    if (p_ >= end_) fail("unexpected end");
    char c = *p_;
    if (c == '{') {
      ++p_;
      Value v = Value::object();
      ws();
      if (p_ < end_ && *p_ == '}') { ++p_; return v; }
      while (true) {
        ws();
        std::string k = string_();
        ws();
        if (p_ >= end_ || *p_ != ':') fail("expected ':'");
        ++p_;
        ws();
        v.obj->emplace_back(std::move(k), value(depth + 1));
        ws();
        if (p_ < end_ && *p_ == ',') { ++p_; continue; }
        if (p_ < end_ && *p_ == '}') { ++p_; break; }
        fail("expected ',' or '}'");
      }
      return v;
    }
    if (c == '[') {
      ++p_;
      Value v = Value::array();
      ws();
      if (p_ < end_ && *p_ == ']') { ++p_; return v; }
      while (true) {
        ws();
        v.arr->push_back(value(depth + 1));
        ws();
        if (p_ < end_ && *p_ == ',') { ++p_; continue; }
        if (p_ < end_ && *p_ == ']') { ++p_; break; }
        fail("expected ',' or ']'");
      }
      return v;
    }
    if (c == '"') return Value(string_());
    if (lit("true")) return Value(true);
    if (lit("false")) return Value(false);
    if (lit("null")) return Value();
    // number
    const char* s = p_;
    bool isint = true;
    if (p_ < end_ && (*p_ == '-' || *p_ == '+')) ++p_;
    while (p_ < end_ && ((*p_ >= '0' && *p_ <= '9') || *p_ == '.' || *p_ == 'e' || *p_ == 'E' || *p_ == '-' || *p_ == '+')) {
      if (*p_ == '.' || *p_ == 'e' || *p_ == 'E') isint = false;
      ++p_;
    }
    if (p_ == s) fail("unexpected character");
    std::string t(s, p_);
    Value v;
    v.type = Value::Number;
    v.num = std::strtod(t.c_str(), nullptr);
    if (isint) { v.is_int = true; v.inum = std::strtoll(t.c_str(), nullptr, 10); }
    return v;
  }
};

inline Value parse(const std::string& s) { return Parser(s.data(), s.size()).parse(); }
inline Value parse(const char* p, size_t n) { return Parser(p, n).parse(); }

inline void escape_to(std::string& o, const std::string& s) {
  o += '"';
  for (unsigned char c : s) {
    switch (c) {
      case '"': o += "\\\""; break; case '\\': o += "\\\\"; break; case '\n': o += "\\n"; break;
      case '\r': o += "\\r"; break; case '\t': o += "\\t"; break; case '\b': o += "\\b"; break; case '\f': o += "\\f"; break;
      default:
        if (c < 0x20) { char b[8]; snprintf(b, sizeof b, "\\u%04x", c); o += b; }
        else o += (char)c;
    }
  }
  o += '"';
}

inline void dump_to(std::string& o, const Value& v) {
  switch (v.type) {
    case Value::Null: o += "null"; break;
    case Value::Bool: o += v.b ? "true" : "false"; break;
    case Value::Number: {
      char b[40];
      if (v.is_int) snprintf(b, sizeof b, "%lld", (long long)v.inum);
      else if (std::isfinite(v.num)) snprintf(b, sizeof b, "%.17g", v.num);
      else snprintf(b, sizeof b, "null");
      o += b;
      break;
    }
    case Value::String: escape_to(o, v.str); break;
    case Value::Arr: {
      o += '[';
      bool first = true;
      for (auto& e : *v.arr) { if (!first) o += ','; first = false; dump_to(o, e); }
      o += ']';
      break;
    }
    case Value::Obj: {
      o += '{';
      bool first = true;
      for (auto& kv : *v.obj) { if (!first) o += ','; first = false; escape_to(o, kv.first); o += ':'; dump_to(o, kv.second); }
      o += '}';
      break;
    }
  }
}
inline std::string dump(const Value& v) { std::string o; dump_to(o, v); return o; }

}  // namespace b2json
