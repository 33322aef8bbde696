// veles_b200 native runtime — minimal JSON reader (the reference vendored rapidjson:
// libVeles/src/main_file_loader.cc:119-176).  Parses contents.json of an exported package.
#pragma once
#include <cstdlib>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace veles {

struct JsonValue {
  enum Type { Null, Bool, Number, String, Array, Object } type = Null;
  bool b = false;
  double num = 0.0;
  std::string str;
  std::vector<JsonValue> arr;
  std::vector<std::pair<std::string, JsonValue>> obj;  // keeps file order

  bool is_object() const { return type == Object; }
  bool is_array() const { return type == Array; }
  bool is_string() const { return type == String; }
  bool is_number() const { return type == Number; }
  bool is_bool() const { return type == Bool; }
  const JsonValue* find(const std::string& key) const {
    for (const auto& kv : obj)
      if (kv.first == key) return &kv.second;
    return nullptr;
  }
  const JsonValue& at(const std::string& key) const {
    const JsonValue* v = find(key);
    if (!v) throw std::runtime_error("json: missing key \"" + key + "\"");
    return *v;
  }
};

class JsonParser {
 public:
  explicit JsonParser(const std::string& text) : s_(text) {}
  JsonValue parse() {
    JsonValue v = value();
    ws();
    if (p_ != s_.size()) fail("trailing characters");
    return v;
  }

 private:
  const std::string& s_;
  size_t p_ = 0;
  [[noreturn]] void fail(const std::string& why) const {
    throw std::runtime_error("json: " + why + " at offset " + std::to_string(p_));
  }
  void ws() {
    while (p_ < s_.size() && (s_[p_] == ' ' || s_[p_] == '\n' || s_[p_] == '\t' || s_[p_] == '\r')) ++p_;
  }
  bool eat(char c) {
    ws();
    if (p_ < s_.size() && s_[p_] == c) {
      ++p_;
      return true;
    }
    return false;
  }
  JsonValue value() {
    ws();
    if (p_ >= s_.size()) fail("unexpected end");
    const char c = s_[p_];
    JsonValue v;
    if (c == '{') {
      ++p_;
      v.type = JsonValue::Object;
      if (eat('}')) return v;
      do {
        ws();
        std::string key = string();
        if (!eat(':')) fail("expected ':'");
        v.obj.emplace_back(key, value());
      } while (eat(','));
      if (!eat('}')) fail("expected '}'");
    } else if (c == '[') {
      ++p_;
      v.type = JsonValue::Array;
      if (eat(']')) return v;
      do {
        v.arr.push_back(value());
      } while (eat(','));
      if (!eat(']')) fail("expected ']'");
    } else if (c == '"') {
      v.type = JsonValue::String;
      v.str = string();
    } else if (s_.compare(p_, 4, "true") == 0) {
      v.type = JsonValue::Bool; v.b = true; p_ += 4;
    } else if (s_.compare(p_, 5, "false") == 0) {
      v.type = JsonValue::Bool; v.b = false; p_ += 5;
    } else if (s_.compare(p_, 4, "null") == 0) {
      p_ += 4;
    } else {
      char* end = nullptr;
      v.num = std::strtod(s_.c_str() + p_, &end);
      if (end == s_.c_str() + p_) fail("bad token");
      v.type = JsonValue::Number;
      p_ = end - s_.c_str();
    }
    return v;
  }
  std::string string() {
    if (p_ >= s_.size() || s_[p_] != '"') fail("expected string");
    ++p_;
    std::string out;
    while (p_ < s_.size() && s_[p_] != '"') {
      char c = s_[p_++];
      if (c == '\\' && p_ < s_.size()) {
        char e = s_[p_++];
        switch (e) {
          case 'n': out += '\n'; break;
          case 't': out += '\t'; break;
          case 'r': out += '\r'; break;
          case 'b': out += '\b'; break;
          case 'f': out += '\f'; break;
          case 'u': {
            if (p_ + 4 > s_.size()) fail("bad \\u escape");
            unsigned cp = std::strtoul(s_.substr(p_, 4).c_str(), nullptr, 16);
            p_ += 4;
            if (cp < 0x80) out += static_cast<char>(cp);
            else if (cp < 0x800) { out += static_cast<char>(0xC0 | (cp >> 6)); out += static_cast<char>(0x80 | (cp & 0x3F)); }
            else { out += static_cast<char>(0xE0 | (cp >> 12)); out += static_cast<char>(0x80 | ((cp >> 6) & 0x3F)); out += static_cast<char>(0x80 | (cp & 0x3F)); }
            break;
          }
          default: out += e;
        }
      } else {
        out += c;
      }
    }
    if (p_ >= s_.size()) fail("unterminated string");
    ++p_;
    return out;
  }
};

inline JsonValue parse_json(const std::string& text) { return JsonParser(text).parse(); }

}  // namespace veles
