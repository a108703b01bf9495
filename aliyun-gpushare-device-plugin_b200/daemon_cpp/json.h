// json.h -- a small JSON reader for the native daemon: PodList / Node documents from kubelet
// and the apiserver (the reference decodes them with encoding/json into client-go types,
// podmanager.go:187-245, client.go:119-134). DOM only; numbers keep their source text so a
// quantity written as a bare number reads the way its author wrote it.
#pragma once
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <utility>
#include <vector>

namespace gsd {

struct JVal {
  enum Type { Null, Bool, Num, Str, Arr, Obj } type = Null;
  bool b = false;
  std::string s;  // Str: decoded text; Num: the token as written
  std::vector<JVal> arr;
  std::vector<std::pair<std::string, JVal>> obj;

  const JVal *get(const char *key) const {
    if (type != Obj) return nullptr;
    for (const auto &kv : obj)
      if (kv.first == key) return &kv.second;
    return nullptr;
  }
  // path lookup through nested objects; nullptr when any level is absent or not an object
  const JVal *at(std::initializer_list<const char *> path) const {
    const JVal *v = this;
    for (const char *k : path) {
      v = v->get(k);
      if (!v) return nullptr;
    }
    return v;
  }
  bool is_str() const { return type == Str; }
  // text of a string or number ("" otherwise): what Python's str() gives for the JSON scalar
  std::string text() const { return type == Str || type == Num ? s : std::string(); }
};

class JsonParser {
 public:
  JsonParser(const char *p, size_t n) : p_(p), end_(p + n) {}
  bool parse(JVal &out) {
    skip();
    if (!value(out, 0)) return false;
    skip();
    return p_ == end_;
  }

 private:
  const char *p_, *end_;
  void skip() {
    while (p_ < end_ && (*p_ == ' ' || *p_ == '\t' || *p_ == '\n' || *p_ == '\r')) ++p_;
  }
  bool lit(const char *w) {
    const size_t n = strlen(w);
    if ((size_t)(end_ - p_) < n || memcmp(p_, w, n) != 0) return false;
    p_ += n;
    return true;
  }
  static void utf8(std::string &o, uint32_t cp) {
    if (cp < 0x80) {
      o.push_back((char)cp);
    } else if (cp < 0x800) {
      o.push_back((char)(0xc0 | cp >> 6));
      o.push_back((char)(0x80 | (cp & 0x3f)));
    } else if (cp < 0x10000) {
      o.push_back((char)(0xe0 | cp >> 12));
      o.push_back((char)(0x80 | ((cp >> 6) & 0x3f)));
      o.push_back((char)(0x80 | (cp & 0x3f)));
    } else {
      o.push_back((char)(0xf0 | cp >> 18));
      o.push_back((char)(0x80 | ((cp >> 12) & 0x3f)));
      o.push_back((char)(0x80 | ((cp >> 6) & 0x3f)));
      o.push_back((char)(0x80 | (cp & 0x3f)));
    }
  }
  bool hex4(uint32_t &v) {
    if (end_ - p_ < 4) return false;
    v = 0;
    for (int i = 0; i < 4; ++i) {
      const char c = *p_++;
      v <<= 4;
      if (c >= '0' && c <= '9') v |= (uint32_t)(c - '0');
      else if (c >= 'a' && c <= 'f') v |= (uint32_t)(c - 'a' + 10);
      else if (c >= 'A' && c <= 'F') v |= (uint32_t)(c - 'A' + 10);
      else return false;
    }
    return true;
  }
  bool string(std::string &o) {
    if (p_ >= end_ || *p_ != '"') return false;
    ++p_;
    while (p_ < end_) {
      const unsigned char c = (unsigned char)*p_++;
      if (c == '"') return true;
      if (c < 0x20) return false;
      if (c != '\\') {
        o.push_back((char)c);
        continue;
      }
      if (p_ >= end_) return false;
      const char e = *p_++;
      switch (e) {
        case '"': o.push_back('"'); break;
        case '\\': o.push_back('\\'); break;
        case '/': o.push_back('/'); break;
        case 'b': o.push_back('\b'); break;
        case 'f': o.push_back('\f'); break;
        case 'n': o.push_back('\n'); break;
        case 'r': o.push_back('\r'); break;
        case 't': o.push_back('\t'); break;
        case 'u': {
          uint32_t cp;
          if (!hex4(cp)) return false;
          if (cp >= 0xd800 && cp < 0xdc00) {  // surrogate pair
            uint32_t lo;
            if (end_ - p_ >= 6 && p_[0] == '\\' && p_[1] == 'u') {
              p_ += 2;
              if (!hex4(lo)) return false;
              if (lo >= 0xdc00 && lo < 0xe000) cp = 0x10000 + ((cp - 0xd800) << 10) + (lo - 0xdc00);
              else cp = 0xfffd;
            } else {
              cp = 0xfffd;
            }
          } else if (cp >= 0xdc00 && cp < 0xe000) {
            cp = 0xfffd;
          }
          utf8(o, cp);
          break;
        }
        default:
          return false;
      }
    }
    return false;
  }
  bool value(JVal &v, int depth) {
    if (depth > 64 || p_ >= end_) return false;
    const char c = *p_;
    if (c == '{') {
      ++p_;
      v.type = JVal::Obj;
      skip();
      if (p_ < end_ && *p_ == '}') {
        ++p_;
        return true;
      }
      for (;;) {
        skip();
        std::string k;
        if (!string(k)) return false;
        skip();
        if (p_ >= end_ || *p_++ != ':') return false;
        skip();
        v.obj.emplace_back(std::move(k), JVal());
        if (!value(v.obj.back().second, depth + 1)) return false;
        skip();
        if (p_ >= end_) return false;
        if (*p_ == ',') {
          ++p_;
          continue;
        }
        if (*p_ == '}') {
          ++p_;
          return true;
        }
        return false;
      }
    }
    if (c == '[') {
      ++p_;
      v.type = JVal::Arr;
      skip();
      if (p_ < end_ && *p_ == ']') {
        ++p_;
        return true;
      }
      for (;;) {
        skip();
        v.arr.emplace_back();
        if (!value(v.arr.back(), depth + 1)) return false;
        skip();
        if (p_ >= end_) return false;
        if (*p_ == ',') {
          ++p_;
          continue;
        }
        if (*p_ == ']') {
          ++p_;
          return true;
        }
        return false;
      }
    }
    if (c == '"') {
      v.type = JVal::Str;
      return string(v.s);
    }
    if (c == 't') {
      v.type = JVal::Bool;
      v.b = true;
      return lit("true");
    }
    if (c == 'f') {
      v.type = JVal::Bool;
      return lit("false");
    }
    if (c == 'n') return lit("null");
    const char *start = p_;
    if (*p_ == '-') ++p_;
    if (p_ >= end_ || *p_ < '0' || *p_ > '9') return false;
    while (p_ < end_ && ((*p_ >= '0' && *p_ <= '9') || *p_ == '.' || *p_ == 'e' || *p_ == 'E' || *p_ == '+' ||
                         *p_ == '-'))
      ++p_;
    v.type = JVal::Num;
    v.s.assign(start, (size_t)(p_ - start));
    return true;
  }
};

inline bool json_parse(const std::string &text, JVal &out) {
  JsonParser p(text.data(), text.size());
  return p.parse(out);
}

inline std::string json_quote(const std::string &s) {
  static const char hexd[] = "0123456789abcdef";
  std::string o = "\"";
  for (unsigned char c : s) {
    if (c == '"') o += "\\\"";
    else if (c == '\\') o += "\\\\";
    else if (c == '\n') o += "\\n";
    else if (c == '\r') o += "\\r";
    else if (c == '\t') o += "\\t";
    else if (c < 0x20) {
      o += "\\u00";
      o.push_back(hexd[c >> 4]);
      o.push_back(hexd[c & 15]);
    } else {
      o.push_back((char)c);
    }
  }
  o.push_back('"');
  return o;
}

}  // namespace gsd
