// json.h — the small JSON reader/writer llmd needs for Ollama's request and reply objects.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <vector>

namespace mini_json {

struct Value {
  enum Kind { Null, Bool, Num, Str, Arr, Obj } kind = Null;
  bool b = false;
  double num = 0;
  std::string str;
  std::vector<Value> arr;
  std::vector<std::pair<std::string, Value>> obj;

  const Value* get(const char* key) const {
    if (kind != Obj) return nullptr;
    for (auto& kv : obj)
      if (kv.first == key) return &kv.second;
    return nullptr;
  }
  std::string get_str(const char* key, const std::string& dflt = "") const {
    const Value* v = get(key);
    return v && v->kind == Str ? v->str : dflt;
  }
  double get_num(const char* key, double dflt) const {
    const Value* v = get(key);
    return v && v->kind == Num ? v->num : dflt;
  }
  bool get_bool(const char* key, bool dflt) const {
    const Value* v = get(key);
    return v && v->kind == Bool ? v->b : dflt;
  }
};

class Parser {
 public:
  explicit Parser(const std::string& s) : p_(s.data()), end_(s.data() + s.size()) {}
  bool parse(Value* out) {
    ws();
    if (!value(out, 0)) return false;
    ws();
    return p_ == end_;
  }

 private:
  const char *p_, *end_;
  void ws() {
    while (p_ < end_ && (*p_ == ' ' || *p_ == '\t' || *p_ == '\n' || *p_ == '\r')) ++p_;
  }
  bool lit(const char* s) {
    const size_t n = strlen(s);
    if ((size_t)(end_ - p_) < n || memcmp(p_, s, n)) return false;
    p_ += n;
    return true;
  }
  static void utf8(std::string& o, uint32_t c) {
    if (c < 0x80) o += (char)c;
    else if (c < 0x800) { o += (char)(0xC0 | (c >> 6)); o += (char)(0x80 | (c & 0x3F)); }
    else if (c < 0x10000) { o += (char)(0xE0 | (c >> 12)); o += (char)(0x80 | ((c >> 6) & 0x3F)); o += (char)(0x80 | (c & 0x3F)); }
    else { o += (char)(0xF0 | (c >> 18)); o += (char)(0x80 | ((c >> 12) & 0x3F)); o += (char)(0x80 | ((c >> 6) & 0x3F)); o += (char)(0x80 | (c & 0x3F)); }
  }
  bool hex4(uint32_t* v) {
    if (end_ - p_ < 4) return false;
    uint32_t r = 0;
    for (int i = 0; i < 4; ++i) {
      const char c = *p_++;
      r <<= 4;
      if (c >= '0' && c <= '9') r |= c - '0';
      else if (c >= 'a' && c <= 'f') r |= c - 'a' + 10;
      else if (c >= 'A' && c <= 'F') r |= c - 'A' + 10;
      else return false;
    }
    *v = r;
    return true;
  }
  bool string(std::string* out) {
    if (p_ >= end_ || *p_ != '"') return false;
    ++p_;
    while (p_ < end_ && *p_ != '"') {
      if (*p_ == '\\') {
        if (++p_ >= end_) return false;
        const char c = *p_++;
        switch (c) {
          case '"': *out += '"'; break;
          case '\\': *out += '\\'; break;
          case '/': *out += '/'; break;
          case 'b': *out += '\b'; break;
          case 'f': *out += '\f'; break;
          case 'n': *out += '\n'; break;
          case 'r': *out += '\r'; break;
          case 't': *out += '\t'; break;
          case 'u': {
            uint32_t cp;
            if (!hex4(&cp)) return false;
            if (cp >= 0xD800 && cp < 0xDC00 && end_ - p_ >= 6 && p_[0] == '\\' && p_[1] == 'u') {
              p_ += 2;
              uint32_t lo;
              if (!hex4(&lo)) return false;
              cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
            }
            utf8(*out, cp);
            break;
          }
          default: return false;
        }
      } else {
        *out += *p_++;
      }
    }
    if (p_ >= end_) return false;
    ++p_;
    return true;
  }
  bool value(Value* v, int depth) {
    if (depth > 32 || p_ >= end_) return false;
    if (*p_ == '{') {
      v->kind = Value::Obj;
      ++p_;
      ws();
      if (p_ < end_ && *p_ == '}') { ++p_; return true; }
      for (;;) {
        ws();
        std::string k;
        if (!string(&k)) return false;
        ws();
        if (p_ >= end_ || *p_++ != ':') return false;
        ws();
        Value c;
        if (!value(&c, depth + 1)) return false;
        v->obj.emplace_back(std::move(k), std::move(c));
        ws();
        if (p_ >= end_) return false;
        if (*p_ == ',') { ++p_; continue; }
        if (*p_ == '}') { ++p_; return true; }
        return false;
      }
    }
    if (*p_ == '[') {
      v->kind = Value::Arr;
      ++p_;
      ws();
      if (p_ < end_ && *p_ == ']') { ++p_; return true; }
      for (;;) {
        ws();
        Value c;
        if (!value(&c, depth + 1)) return false;
        v->arr.push_back(std::move(c));
        ws();
        if (p_ >= end_) return false;
        if (*p_ == ',') { ++p_; continue; }
        if (*p_ == ']') { ++p_; return true; }
        return false;
      }
    }
    if (*p_ == '"') { v->kind = Value::Str; return string(&v->str); }
    if (lit("true")) { v->kind = Value::Bool; v->b = true; return true; }
    if (lit("false")) { v->kind = Value::Bool; v->b = false; return true; }
    if (lit("null")) { v->kind = Value::Null; return true; }
    char* e = nullptr;
    std::string tmp(p_, (size_t)std::min<ptrdiff_t>(end_ - p_, 64));
    const double d = strtod(tmp.c_str(), &e);
    if (e == tmp.c_str()) return false;
    p_ += e - tmp.c_str();
    v->kind = Value::Num;
    v->num = d;
    return true;
  }
};

inline std::string quote(const std::string& s) {
  std::string o = "\"";
  for (unsigned char c : s) {
    switch (c) {
      case '"': o += "\\\""; break;
      case '\\': o += "\\\\"; break;
      case '\n': o += "\\n"; break;
      case '\r': o += "\\r"; break;
      case '\t': o += "\\t"; break;
      default:
        if (c < 0x20) { char b[8]; snprintf(b, sizeof b, "\\u%04x", c); o += b; }
        else o += (char)c;
    }
  }
  return o + "\"";
}

}  // namespace mini_json
