// Minimal JSON value + recursive-descent parser (config.json, safetensors headers, HTTP bodies)
// and a string escaper for responses.  Header-only, no dependencies.
#pragma once
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <map>
#include <memory>
#include <string>
#include <vector>

namespace kge {

struct Json {
  enum Type { Null, Bool, Num, Str, Arr, Obj } type = Null;
  bool b = false;
  double num = 0;
  std::string str;
  std::vector<Json> arr;
  std::vector<std::pair<std::string, Json>> obj;   // insertion order kept

  bool is_null() const { return type == Null; }
  bool is_num() const { return type == Num; }
  bool is_str() const { return type == Str; }
  bool is_arr() const { return type == Arr; }
  bool is_obj() const { return type == Obj; }
  const Json* get(const std::string& k) const {
    if (type != Obj) return nullptr;
    for (auto& kv : obj) if (kv.first == k) return &kv.second;
    return nullptr;
  }
  double num_or(const std::string& k, double d) const {
    const Json* j = get(k);
    return (j && j->type == Num) ? j->num : d;
  }
  // number -> integer without undefined behaviour on hostile input: NaN -> 0, out of range -> saturated
  static int64_t to_i64(double v) {
    if (!(v == v)) return 0;
    if (v >= 9.2e18) return INT64_MAX;
    if (v <= -9.2e18) return INT64_MIN;
    return (int64_t)v;
  }
  int64_t as_i64() const { return type == Num ? to_i64(num) : 0; }
  int64_t int_or(const std::string& k, int64_t d) const {
    const Json* j = get(k);
    return (j && j->type == Num) ? to_i64(j->num) : d;
  }
  bool bool_or(const std::string& k, bool d) const {
    const Json* j = get(k);
    if (!j) return d;
    if (j->type == Bool) return j->b;
    return d;
  }
  std::string str_or(const std::string& k, const std::string& d) const {
    const Json* j = get(k);
    return (j && j->type == Str) ? j->str : d;
  }
};

class JsonParser {
 public:
  JsonParser(const char* p, size_t n) : p_(p), e_(p + n) {}
  bool parse(Json* out, std::string* err) {
    ws();
    if (!value(out, 0)) { if (err) *err = err_.empty() ? "invalid JSON" : err_; return false; }
    ws();
    if (p_ != e_) { if (err) *err = "trailing characters after JSON value"; return false; }
    return true;
  }

 private:
  const char* p_;
  const char* e_;
  std::string err_;
  void ws() { while (p_ < e_ && (*p_ == ' ' || *p_ == '\n' || *p_ == '\t' || *p_ == '\r')) ++p_; }
  bool fail(const char* m) { if (err_.empty()) err_ = m; return false; }
  bool lit(const char* s) {
    size_t n = strlen(s);
    if ((size_t)(e_ - p_) < n || memcmp(p_, s, n) != 0) return false;
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
    if (e_ - p_ < 4) return false;
    uint32_t r = 0;
    for (int i = 0; i < 4; ++i) {
      char c = p_[i];
      r <<= 4;
      if (c >= '0' && c <= '9') r |= c - '0';
      else if (c >= 'a' && c <= 'f') r |= c - 'a' + 10;
      else if (c >= 'A' && c <= 'F') r |= c - 'A' + 10;
      else return false;
    }
    p_ += 4;
    *v = r;
    return true;
  }
  bool string(std::string* out) {
    if (p_ >= e_ || *p_ != '"') return fail("expected string");
    ++p_;
    while (p_ < e_ && *p_ != '"') {
      if (*p_ == '\\') {
        if (++p_ >= e_) return fail("bad escape");
        char c = *p_++;
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
            if (!hex4(&cp)) return fail("bad \\u escape");
            if (cp >= 0xD800 && cp < 0xDC00 && e_ - p_ >= 6 && p_[0] == '\\' && p_[1] == 'u') {
              p_ += 2;
              uint32_t lo;
              if (!hex4(&lo)) return fail("bad surrogate");
              cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
            }
            utf8(*out, cp);
            break;
          }
          default: return fail("bad escape");
        }
      } else {
        *out += *p_++;
      }
    }
    if (p_ >= e_) return fail("unterminated string");
    ++p_;
    return true;
  }
  bool value(Json* out, int depth) {
    if (depth > 64) return fail("nesting too deep");
    ws();
    if (p_ >= e_) return fail("unexpected end");
    char c = *p_;
    if (c == '{') {
      ++p_;
      out->type = Json::Obj;
      ws();
      if (p_ < e_ && *p_ == '}') { ++p_; return true; }
      while (true) {
        ws();
        std::string k;
        if (!string(&k)) return false;
        ws();
        if (p_ >= e_ || *p_ != ':') return fail("expected ':'");
        ++p_;
        Json v;
        if (!value(&v, depth + 1)) return false;
        out->obj.emplace_back(std::move(k), std::move(v));
        ws();
        if (p_ < e_ && *p_ == ',') { ++p_; continue; }
        if (p_ < e_ && *p_ == '}') { ++p_; return true; }
        return fail("expected ',' or '}'");
      }
    }
    if (c == '[') {
      ++p_;
      out->type = Json::Arr;
      ws();
      if (p_ < e_ && *p_ == ']') { ++p_; return true; }
      while (true) {
        Json v;
        if (!value(&v, depth + 1)) return false;
        out->arr.push_back(std::move(v));
        ws();
        if (p_ < e_ && *p_ == ',') { ++p_; continue; }
        if (p_ < e_ && *p_ == ']') { ++p_; return true; }
        return fail("expected ',' or ']'");
      }
    }
    if (c == '"') { out->type = Json::Str; return string(&out->str); }
    if (lit("true")) { out->type = Json::Bool; out->b = true; return true; }
    if (lit("false")) { out->type = Json::Bool; out->b = false; return true; }
    if (lit("null")) { out->type = Json::Null; return true; }
    if (c == '-' || (c >= '0' && c <= '9')) {
      const char* s = p_;
      while (p_ < e_ && (*p_ == '-' || *p_ == '+' || *p_ == '.' || *p_ == 'e' || *p_ == 'E' || (*p_ >= '0' && *p_ <= '9'))) ++p_;
      std::string t(s, p_ - s);
      char* endp = nullptr;
      out->num = strtod(t.c_str(), &endp);
      if (endp == t.c_str()) return fail("bad number");
      out->type = Json::Num;
      return true;
    }
    return fail("unexpected character");
  }
};

inline bool json_parse(const std::string& s, Json* out, std::string* err = nullptr) {
  JsonParser p(s.data(), s.size());
  return p.parse(out, err);
}

inline std::string json_escape(const std::string& s) {
  std::string o;
  o.reserve(s.size() + 2);
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
  return o;
}

}  // namespace kge
