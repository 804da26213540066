// Minimal JSON reader for the two documents the host has to understand: nif_metadata.txt
// (src/neural_networks/NifMetaData.cpp:11-71 reads it with boost::property_tree) and the Keras
// `model_config` attribute (src/keras/Hdf5Model.cpp:8-60). Values keep their source text so that
// numbers can be read as the type the caller wants, like property_tree's get<T>().
#pragma once

#include <cstdlib>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace json {

struct Value {
  enum Kind { Null, Bool, Number, String, Array, Object } kind = Null;
  std::string text;                                            // Bool/Number/String payload
  std::vector<Value> items;                                    // Array
  std::vector<std::pair<std::string, Value>> members;          // Object, in file order

  bool has(const std::string& key) const {
    for (auto& m : members) if (m.first == key) return true;
    return false;
  }
  const Value& at(const std::string& key) const {
    for (auto& m : members) if (m.first == key) return m.second;
    throw std::runtime_error("No such node (" + key + ")");
  }
  double asDouble() const {
    if (kind != Number && kind != String) throw std::runtime_error("conversion of data to type \"double\" failed");
    char* end = nullptr;
    const double v = std::strtod(text.c_str(), &end);
    if (end == text.c_str()) throw std::runtime_error("conversion of data to type \"double\" failed");
    return v;
  }
  long long asInt() const { return (long long)asDouble(); }
  bool asBool() const {
    if (kind == Bool) return text == "true";
    if (kind == Number) return asDouble() != 0.0;
    throw std::runtime_error("conversion of data to type \"bool\" failed");
  }
  const std::string& asString() const { return text; }
};

class Parser {
public:
  explicit Parser(const std::string& s) : s_(s) {}
  Value parse() {
    Value v = value();
    ws();
    if (p_ != s_.size()) fail("trailing characters");
    return v;
  }

private:
  const std::string& s_;
  size_t p_ = 0;

  [[noreturn]] void fail(const std::string& what) const {
    throw std::runtime_error("JSON parse error at offset " + std::to_string(p_) + ": " + what);
  }
  void ws() { while (p_ < s_.size() && (s_[p_] == ' ' || s_[p_] == '\n' || s_[p_] == '\t' || s_[p_] == '\r')) ++p_; }
  bool lit(const char* w) {
    size_t n = 0; while (w[n]) ++n;
    if (s_.compare(p_, n, w) == 0) { p_ += n; return true; }
    return false;
  }
  static void utf8(std::string& out, unsigned cp) {
    if (cp < 0x80) out += (char)cp;
    else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
    else if (cp < 0x10000) { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
    else { out += (char)(0xF0 | (cp >> 18)); out += (char)(0x80 | ((cp >> 12) & 0x3F)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
  }
  unsigned hex4() {
    if (p_ + 4 > s_.size()) fail("short \\u escape");
    unsigned v = 0;
    for (int i = 0; i < 4; ++i) {
      const char c = s_[p_++];
      v <<= 4;
      if (c >= '0' && c <= '9') v |= (unsigned)(c - '0');
      else if (c >= 'a' && c <= 'f') v |= (unsigned)(c - 'a' + 10);
      else if (c >= 'A' && c <= 'F') v |= (unsigned)(c - 'A' + 10);
      else fail("bad \\u escape");
    }
    return v;
  }
  std::string str() {
    if (s_[p_] != '"') fail("expected string");
    ++p_;
    std::string out;
    while (true) {
      if (p_ >= s_.size()) fail("unterminated string");
      const char c = s_[p_++];
      if (c == '"') break;
      if (c != '\\') { out += c; continue; }
      if (p_ >= s_.size()) fail("unterminated escape");
      const char e = s_[p_++];
      switch (e) {
        case '"': out += '"'; break;
        case '\\': out += '\\'; break;
        case '/': out += '/'; break;
        case 'b': out += '\b'; break;
        case 'f': out += '\f'; break;
        case 'n': out += '\n'; break;
        case 'r': out += '\r'; break;
        case 't': out += '\t'; break;
        case 'u': {
          unsigned cp = hex4();
          if (cp >= 0xD800 && cp < 0xDC00 && s_.compare(p_, 2, "\\u") == 0) {
            p_ += 2;
            const unsigned lo = hex4();
            cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
          }
          utf8(out, cp);
          break;
        }
        default: fail("unknown escape");
      }
    }
    return out;
  }
  Value value() {
    ws();
    if (p_ >= s_.size()) fail("unexpected end");
    Value v;
    const char c = s_[p_];
    if (c == '{') {
      ++p_;
      v.kind = Value::Object;
      ws();
      if (p_ < s_.size() && s_[p_] == '}') { ++p_; return v; }
      while (true) {
        ws();
        std::string k = str();
        ws();
        if (p_ >= s_.size() || s_[p_] != ':') fail("expected ':'");
        ++p_;
        v.members.emplace_back(std::move(k), value());
        ws();
        if (p_ < s_.size() && s_[p_] == ',') { ++p_; continue; }
        if (p_ < s_.size() && s_[p_] == '}') { ++p_; break; }
        fail("expected ',' or '}'");
      }
    } else if (c == '[') {
      ++p_;
      v.kind = Value::Array;
      ws();
      if (p_ < s_.size() && s_[p_] == ']') { ++p_; return v; }
      while (true) {
        v.items.push_back(value());
        ws();
        if (p_ < s_.size() && s_[p_] == ',') { ++p_; continue; }
        if (p_ < s_.size() && s_[p_] == ']') { ++p_; break; }
        fail("expected ',' or ']'");
      }
    } else if (c == '"') {
      v.kind = Value::String;
      v.text = str();
    } else if (lit("true")) { v.kind = Value::Bool; v.text = "true"; }
    else if (lit("false")) { v.kind = Value::Bool; v.text = "false"; }
    else if (lit("null")) { v.kind = Value::Null; }
    else {
      const size_t b = p_;
      while (p_ < s_.size() && (std::string("+-0123456789.eE").find(s_[p_]) != std::string::npos)) ++p_;
      if (p_ == b) fail("unexpected character");
      v.kind = Value::Number;
      v.text = s_.substr(b, p_ - b);
    }
    return v;
  }
};

inline Value parse(const std::string& s) { return Parser(s).parse(); }

} // namespace json
