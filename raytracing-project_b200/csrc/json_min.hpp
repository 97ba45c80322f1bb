// json_min.hpp — a small strict JSON (RFC 8259) reader for the scene format.
//
// The reference reads scenes with nlohmann/json (json_loader.cpp:21-22), a third-party
// header that is not part of the reference tree.  The scene format only needs objects,
// arrays, numbers, strings, booleans and null, so the host side carries its own reader
// instead of depending on that header.  Behaviour that the loader relies on:
//   * objects keep one value per key (a repeated key keeps the LAST value, as nlohmann does)
//   * numbers remember whether they were written as integers (get<int> on a float truncates)
//   * every malformed document raises json_min::parse_error (the loader maps it to the
//     reference's "JSON parse error: ..." message, json_loader.cpp:485-486)
#pragma once
#include <cmath>
#include <cstdlib>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace json_min {

struct parse_error : std::runtime_error { using std::runtime_error::runtime_error; };
struct type_error  : std::runtime_error { using std::runtime_error::runtime_error; };

class Value {
public:
    enum Kind { Null, Bool, Number, String, Array, Object };
    Kind kind{Null};
    bool b{false};
    double num{0.0};
    bool integral{false};
    std::string str;
    std::vector<Value> arr;
    std::vector<std::pair<std::string, Value>> obj;   // sorted by key, unique

    bool is_object() const { return kind == Object; }
    bool is_array()  const { return kind == Array; }
    bool is_number() const { return kind == Number; }
    bool is_string() const { return kind == String; }
    size_t size() const { return kind == Array ? arr.size() : kind == Object ? obj.size() : (kind == Null ? 0 : 1); }
    bool empty() const { return size() == 0; }

    bool contains(const char* key) const {
        if (kind != Object) return false;
        for (auto& kv : obj) if (kv.first == key) return true;
        return false;
    }
    const Value& at(const char* key) const {
        if (kind != Object) throw type_error(std::string("cannot use at() with a non-object (key '") + key + "')");
        for (auto& kv : obj) if (kv.first == key) return kv.second;
        throw type_error(std::string("key '") + key + "' not found");
    }
    const Value& at(size_t i) const {
        if (kind != Array) throw type_error("cannot use at(index) with a non-array");
        if (i >= arr.size()) throw type_error("array index out of range");
        return arr[i];
    }
    const Value& operator[](size_t i) const { return at(i); }

    double as_double() const {
        if (kind == Number) return num;
        if (kind == Bool) return b ? 1.0 : 0.0;      // nlohmann converts booleans to arithmetic types
        throw type_error("type must be number");
    }
    int as_int() const {
        if (kind == Number) return (int)num;          // truncation, like static_cast in nlohmann's get<int>
        if (kind == Bool) return b ? 1 : 0;
        throw type_error("type must be number");
    }
    const std::string& as_string() const {
        if (kind != String) throw type_error("type must be string");
        return str;
    }
};

class Parser {
public:
    explicit Parser(const std::string& text) : s_(text) {}
    Value parse_document() {
        skip_ws();
        Value v = parse_value(0);
        skip_ws();
        if (i_ != s_.size()) fail("unexpected trailing characters");
        return v;
    }
private:
    const std::string& s_;
    size_t i_{0};

    [[noreturn]] void fail(const std::string& what) const {
        throw parse_error("parse error at byte " + std::to_string(i_) + ": " + what);
    }
    void skip_ws() {
        while (i_ < s_.size() && (s_[i_] == ' ' || s_[i_] == '\t' || s_[i_] == '\n' || s_[i_] == '\r')) ++i_;
    }
    bool match(const char* lit) {
        size_t n = 0; while (lit[n]) ++n;
        if (s_.compare(i_, n, lit) == 0) { i_ += n; return true; }
        return false;
    }
    Value parse_value(int depth) {
        if (depth > 512) fail("nesting too deep");
        if (i_ >= s_.size()) fail("unexpected end of input");
        char c = s_[i_];
        Value v;
        if (c == '{') return parse_object(depth);
        if (c == '[') return parse_array(depth);
        if (c == '"') { v.kind = Value::String; v.str = parse_string(); return v; }
        if (c == 't') { if (!match("true")) fail("invalid literal"); v.kind = Value::Bool; v.b = true; return v; }
        if (c == 'f') { if (!match("false")) fail("invalid literal"); v.kind = Value::Bool; v.b = false; return v; }
        if (c == 'n') { if (!match("null")) fail("invalid literal"); v.kind = Value::Null; return v; }
        if (c == '-' || (c >= '0' && c <= '9')) return parse_number();
        fail(std::string("unexpected character '") + c + "'");
    }
    Value parse_number() {
        size_t start = i_;
        bool integral = true;
        if (s_[i_] == '-') ++i_;
        if (i_ >= s_.size()) fail("invalid number");
        if (s_[i_] == '0') { ++i_; }
        else if (s_[i_] >= '1' && s_[i_] <= '9') { while (i_ < s_.size() && isdigit((unsigned char)s_[i_])) ++i_; }
        else fail("invalid number");
        if (i_ < s_.size() && s_[i_] == '.') {
            integral = false; ++i_;
            if (i_ >= s_.size() || !isdigit((unsigned char)s_[i_])) fail("invalid number: digit expected after '.'");
            while (i_ < s_.size() && isdigit((unsigned char)s_[i_])) ++i_;
        }
        if (i_ < s_.size() && (s_[i_] == 'e' || s_[i_] == 'E')) {
            integral = false; ++i_;
            if (i_ < s_.size() && (s_[i_] == '+' || s_[i_] == '-')) ++i_;
            if (i_ >= s_.size() || !isdigit((unsigned char)s_[i_])) fail("invalid number: digit expected in exponent");
            while (i_ < s_.size() && isdigit((unsigned char)s_[i_])) ++i_;
        }
        Value v; v.kind = Value::Number; v.integral = integral;
        v.num = std::strtod(s_.substr(start, i_ - start).c_str(), nullptr);
        return v;
    }
    static void append_utf8(std::string& out, unsigned cp) {
        if (cp < 0x80) out += (char)cp;
        else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
        else if (cp < 0x10000) { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
        else { out += (char)(0xF0 | (cp >> 18)); out += (char)(0x80 | ((cp >> 12) & 0x3F)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
    }
    unsigned parse_hex4() {
        if (i_ + 4 > s_.size()) fail("invalid \\u escape");
        unsigned v = 0;
        for (int k = 0; k < 4; ++k) {
            char c = s_[i_++]; v <<= 4;
            if (c >= '0' && c <= '9') v |= (unsigned)(c - '0');
            else if (c >= 'a' && c <= 'f') v |= (unsigned)(c - 'a' + 10);
            else if (c >= 'A' && c <= 'F') v |= (unsigned)(c - 'A' + 10);
            else fail("invalid \\u escape");
        }
        return v;
    }
    std::string parse_string() {
        ++i_;  // opening quote
        std::string out;
        for (;;) {
            if (i_ >= s_.size()) fail("unterminated string");
            unsigned char c = (unsigned char)s_[i_++];
            if (c == '"') break;
            if (c < 0x20) fail("control character in string");
            if (c != '\\') { out += (char)c; continue; }
            if (i_ >= s_.size()) fail("unterminated escape");
            char e = s_[i_++];
            switch (e) {
                case '"': out += '"'; break;  case '\\': out += '\\'; break;  case '/': out += '/'; break;
                case 'b': out += '\b'; break; case 'f': out += '\f'; break;   case 'n': out += '\n'; break;
                case 'r': out += '\r'; break; case 't': out += '\t'; break;
                case 'u': {
                    unsigned cp = parse_hex4();
                    if (cp >= 0xD800 && cp <= 0xDBFF) {
                        if (!(i_ + 1 < s_.size() && s_[i_] == '\\' && s_[i_ + 1] == 'u')) fail("lone surrogate");
                        i_ += 2; unsigned lo = parse_hex4();
                        if (lo < 0xDC00 || lo > 0xDFFF) fail("invalid surrogate pair");
                        cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
                    } else if (cp >= 0xDC00 && cp <= 0xDFFF) fail("lone surrogate");
                    append_utf8(out, cp);
                } break;
                default: fail("invalid escape");
            }
        }
        return out;
    }
    Value parse_array(int depth) {
        Value v; v.kind = Value::Array; ++i_;
        skip_ws();
        if (i_ < s_.size() && s_[i_] == ']') { ++i_; return v; }
        for (;;) {
            skip_ws();
            v.arr.push_back(parse_value(depth + 1));
            skip_ws();
            if (i_ >= s_.size()) fail("unterminated array");
            if (s_[i_] == ',') { ++i_; continue; }
            if (s_[i_] == ']') { ++i_; break; }
            fail("expected ',' or ']'");
        }
        return v;
    }
    Value parse_object(int depth) {
        Value v; v.kind = Value::Object; ++i_;
        std::map<std::string, Value> m;
        skip_ws();
        if (i_ < s_.size() && s_[i_] == '}') { ++i_; return v; }
        for (;;) {
            skip_ws();
            if (i_ >= s_.size() || s_[i_] != '"') fail("expected string key");
            std::string key = parse_string();
            skip_ws();
            if (i_ >= s_.size() || s_[i_] != ':') fail("expected ':'");
            ++i_; skip_ws();
            m[key] = parse_value(depth + 1);
            skip_ws();
            if (i_ >= s_.size()) fail("unterminated object");
            if (s_[i_] == ',') { ++i_; continue; }
            if (s_[i_] == '}') { ++i_; break; }
            fail("expected ',' or '}'");
        }
        for (auto& kv : m) v.obj.emplace_back(kv.first, std::move(kv.second));
        return v;
    }
};

inline Value parse(const std::string& text) { return Parser(text).parse_document(); }

}  // namespace json_min
