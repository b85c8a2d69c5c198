// Minimal JSON reader for TF_CONFIG (pkg/trainer/replicas.go:45-53 is the writer).
// Objects keep insertion order; numbers are kept as doubles; no \u surrogate pairing
// beyond the BMP (TF_CONFIG holds DNS labels and ports).
#pragma once
#include <cstdlib>
#include <memory>
#include <string>
#include <utility>
#include <vector>

namespace tfg {
namespace json {

struct Value {
  enum Kind { Null, Bool, Num, Str, Arr, Obj } kind = Null;
  bool b = false;
  double num = 0;
  std::string str;
  std::vector<Value> arr;
  std::vector<std::pair<std::string, Value>> obj;
  const Value* get(const std::string& k) const {
    for (auto& kv : obj)
      if (kv.first == k) return &kv.second;
    return nullptr;
  }
};

class Parser {
 public:
  explicit Parser(const char* s) : p_(s) {}
  bool parse(Value* out, std::string* err) {
    if (!value(out, err)) return false;
    ws();
    if (*p_) { *err = "trailing characters after JSON value"; return false; }
    return true;
  }

 private:
  const char* p_;
  void ws() { while (*p_ == ' ' || *p_ == '\t' || *p_ == '\n' || *p_ == '\r') ++p_; }
  bool lit(const char* w) {
    const char* q = p_;
    while (*w) if (*q++ != *w++) return false;
    p_ = q;
    return true;
  }
  bool string(std::string* out, std::string* err) {
    if (*p_ != '"') { *err = "expected string"; return false; }
    ++p_;
    while (*p_ && *p_ != '"') {
      if (*p_ == '\\') {
        ++p_;
        switch (*p_) {
          case '"': out->push_back('"'); break;
          case '\\': out->push_back('\\'); break;
          case '/': out->push_back('/'); break;
          case 'b': out->push_back('\b'); break;
          case 'f': out->push_back('\f'); break;
          case 'n': out->push_back('\n'); break;
          case 'r': out->push_back('\r'); break;
          case 't': out->push_back('\t'); break;
          case 'u': {
            unsigned cp = 0;
            for (int i = 1; i <= 4; ++i) {
              char c = p_[i];
              cp <<= 4;
              if (c >= '0' && c <= '9') cp |= c - '0';
              else if (c >= 'a' && c <= 'f') cp |= c - 'a' + 10;
              else if (c >= 'A' && c <= 'F') cp |= c - 'A' + 10;
              else { *err = "bad \\u escape"; return false; }
            }
            p_ += 4;
            if (cp < 0x80) out->push_back((char)cp);
            else if (cp < 0x800) { out->push_back((char)(0xC0 | (cp >> 6))); out->push_back((char)(0x80 | (cp & 0x3F))); }
            else { out->push_back((char)(0xE0 | (cp >> 12))); out->push_back((char)(0x80 | ((cp >> 6) & 0x3F))); out->push_back((char)(0x80 | (cp & 0x3F))); }
            break;
          }
          default: *err = "bad escape"; return false;
        }
        ++p_;
      } else {
        out->push_back(*p_++);
      }
    }
    if (*p_ != '"') { *err = "unterminated string"; return false; }
    ++p_;
    return true;
  }
  bool value(Value* v, std::string* err) {
    ws();
    if (*p_ == '{') {
      v->kind = Value::Obj;
      ++p_; ws();
      if (*p_ == '}') { ++p_; return true; }
      for (;;) {
        ws();
        std::string k;
        if (!string(&k, err)) return false;
        ws();
        if (*p_ != ':') { *err = "expected ':'"; return false; }
        ++p_;
        Value child;
        if (!value(&child, err)) return false;
        v->obj.emplace_back(std::move(k), std::move(child));
        ws();
        if (*p_ == ',') { ++p_; continue; }
        if (*p_ == '}') { ++p_; return true; }
        *err = "expected ',' or '}'";
        return false;
      }
    }
    if (*p_ == '[') {
      v->kind = Value::Arr;
      ++p_; ws();
      if (*p_ == ']') { ++p_; return true; }
      for (;;) {
        Value child;
        if (!value(&child, err)) return false;
        v->arr.push_back(std::move(child));
        ws();
        if (*p_ == ',') { ++p_; continue; }
        if (*p_ == ']') { ++p_; return true; }
        *err = "expected ',' or ']'";
        return false;
      }
    }
    if (*p_ == '"') { v->kind = Value::Str; return string(&v->str, err); }
    if (lit("true")) { v->kind = Value::Bool; v->b = true; return true; }
    if (lit("false")) { v->kind = Value::Bool; v->b = false; return true; }
    if (lit("null")) { v->kind = Value::Null; return true; }
    char* end = nullptr;
    double d = strtod(p_, &end);
    if (end == p_) { *err = "unexpected character in JSON"; return false; }
    p_ = end;
    v->kind = Value::Num;
    v->num = d;
    return true;
  }
};

inline void escape(const std::string& s, std::string* out) {
  out->push_back('"');
  for (unsigned char c : s) {
    if (c == '"') *out += "\\\"";
    else if (c == '\\') *out += "\\\\";
    else if (c == '<') *out += "\\u003c";
    else if (c == '>') *out += "\\u003e";
    else if (c == '&') *out += "\\u0026";
    else if (c < 0x20) { char b[8]; snprintf(b, sizeof b, "\\u%04x", c); *out += b; }
    else out->push_back((char)c);
  }
  out->push_back('"');
}

}  // namespace json
}  // namespace tfg
