// er_json.hpp — the small JSON subset the render-command protocol carries (headers, config, camera,
// texture metadata, materials, render info, device list; src/Managers.cpp:6-17,167-177,
// src/CommandManager.cpp:8-112,278-362). The reference uses Boost.JSON; the wire format is plain JSON, so any
// conforming reader/writer interoperates with ElevenBlender.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace erj {

struct Value {
    enum Kind { Null, Bool, Number, String, Array, Object } kind = Null;
    bool b = false;
    double num = 0.0;
    bool is_int = false;
    long long inum = 0;
    std::string str;
    std::vector<Value> arr;
    std::vector<std::pair<std::string, Value>> obj;  // insertion order kept (serialisation is stable)

    static Value make_object() { Value v; v.kind = Object; return v; }
    static Value make_array() { Value v; v.kind = Array; return v; }
    static Value of(const std::string& s) { Value v; v.kind = String; v.str = s; return v; }
    static Value of(const char* s) { return of(std::string(s)); }
    static Value of(bool x) { Value v; v.kind = Bool; v.b = x; return v; }
    static Value of(double x) { Value v; v.kind = Number; v.num = x; return v; }
    static Value of(long long x) { Value v; v.kind = Number; v.is_int = true; v.inum = x; v.num = (double)x; return v; }
    static Value of(unsigned long long x) { return of((long long)x); }
    static Value of(int x) { return of((long long)x); }
    static Value of(unsigned x) { return of((long long)x); }

    const Value* find(const std::string& k) const {
        if (kind != Object) return nullptr;
        for (const auto& kv : obj)
            if (kv.first == k) return &kv.second;
        return nullptr;
    }
    bool contains(const std::string& k) const { return find(k) != nullptr; }
    const Value& at(const std::string& k) const {
        const Value* v = find(k);
        if (!v) throw std::runtime_error("json: missing key '" + k + "'");
        return *v;
    }
    void set(const std::string& k, Value v) {
        for (auto& kv : obj)
            if (kv.first == k) { kv.second = std::move(v); return; }
        obj.emplace_back(k, std::move(v));
    }
    // Boost.JSON's as_double / as_int64 / as_bool / as_string throw on a kind mismatch; the reference's
    // handlers catch that and log "Invalid ... format" (src/CommandManager.cpp:121-128)
    double as_double() const {
        if (kind != Number || is_int) throw std::runtime_error("json: not a double");
        return num;
    }
    long long as_int64() const {
        if (kind != Number || !is_int) throw std::runtime_error("json: not an int64");
        return inum;
    }
    double as_number() const {  // lenient: ints are accepted where a float is meant
        if (kind != Number) throw std::runtime_error("json: not a number");
        return num;
    }
    bool as_bool() const {
        if (kind != Bool) throw std::runtime_error("json: not a bool");
        return b;
    }
    const std::string& as_string() const {
        if (kind != String) throw std::runtime_error("json: not a string");
        return str;
    }
};

class Parser {
   public:
    explicit Parser(const std::string& s) : s_(s) {}
    Value parse() {
        Value v = value();
        ws();
        if (i_ != s_.size()) throw std::runtime_error("json: trailing characters");
        return v;
    }

   private:
    const std::string& s_;
    size_t i_ = 0;
    void ws() { while (i_ < s_.size() && (s_[i_] == ' ' || s_[i_] == '\t' || s_[i_] == '\n' || s_[i_] == '\r')) i_++; }
    char peek() { return i_ < s_.size() ? s_[i_] : '\0'; }
    void expect(char c) {
        if (peek() != c) throw std::runtime_error(std::string("json: expected '") + c + "'");
        i_++;
    }
    Value value() {
        ws();
        const char c = peek();
        if (c == '{') return object();
        if (c == '[') return array();
        if (c == '"') return Value::of(string());
        if (s_.compare(i_, 4, "true") == 0) { i_ += 4; return Value::of(true); }
        if (s_.compare(i_, 5, "false") == 0) { i_ += 5; return Value::of(false); }
        if (s_.compare(i_, 4, "null") == 0) { i_ += 4; return Value(); }
        return number();
    }
    Value object() {
        Value v = Value::make_object();
        expect('{');
        ws();
        if (peek() == '}') { i_++; return v; }
        for (;;) {
            ws();
            std::string k = string();
            ws();
            expect(':');
            v.obj.emplace_back(std::move(k), value());
            ws();
            if (peek() == ',') { i_++; continue; }
            expect('}');
            return v;
        }
    }
    Value array() {
        Value v = Value::make_array();
        expect('[');
        ws();
        if (peek() == ']') { i_++; return v; }
        for (;;) {
            v.arr.push_back(value());
            ws();
            if (peek() == ',') { i_++; continue; }
            expect(']');
            return v;
        }
    }
    std::string string() {
        expect('"');
        std::string out;
        while (i_ < s_.size() && s_[i_] != '"') {
            char c = s_[i_++];
            if (c == '\\' && i_ < s_.size()) {
                c = s_[i_++];
                switch (c) {
                    case 'n': out += '\n'; break;
                    case 't': out += '\t'; break;
                    case 'r': out += '\r'; break;
                    case 'b': out += '\b'; break;
                    case 'f': out += '\f'; break;
                    case 'u': {
                        unsigned cp = (unsigned)strtoul(s_.substr(i_, 4).c_str(), nullptr, 16);
                        i_ += 4;
                        if (cp < 0x80) out += (char)cp;
                        else if (cp < 0x800) { out += (char)(0xc0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3f)); }
                        else { out += (char)(0xe0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3f)); out += (char)(0x80 | (cp & 0x3f)); }
                        break;
                    }
                    default: out += c;
                }
            } else {
                out += c;
            }
        }
        expect('"');
        return out;
    }
    Value number() {
        const size_t b = i_;
        bool integral = true;
        if (peek() == '-' || peek() == '+') i_++;
        while (i_ < s_.size()) {
            const char c = s_[i_];
            if (c >= '0' && c <= '9') { i_++; continue; }
            if (c == '.' || c == 'e' || c == 'E' || c == '-' || c == '+') { integral = false; i_++; continue; }
            break;
        }
        if (b == i_) throw std::runtime_error("json: bad value");
        const std::string t = s_.substr(b, i_ - b);
        if (integral) return Value::of((long long)strtoll(t.c_str(), nullptr, 10));
        return Value::of(strtod(t.c_str(), nullptr));
    }
};

inline Value parse(const std::string& s) { return Parser(s).parse(); }

inline void write(const Value& v, std::string& out) {
    switch (v.kind) {
        case Value::Null: out += "null"; break;
        case Value::Bool: out += v.b ? "true" : "false"; break;
        case Value::Number: {
            char buf[64];
            if (v.is_int) snprintf(buf, sizeof buf, "%lld", v.inum);
            else snprintf(buf, sizeof buf, "%.17g", v.num);
            out += buf;
            break;
        }
        case Value::String: {
            out += '"';
            for (char c : v.str) {
                if (c == '"' || c == '\\') { out += '\\'; out += c; }
                else if (c == '\n') out += "\\n";
                else if (c == '\t') out += "\\t";
                else if (c == '\r') out += "\\r";
                else if ((unsigned char)c < 0x20) { char b[8]; snprintf(b, sizeof b, "\\u%04x", c); out += b; }
                else out += c;
            }
            out += '"';
            break;
        }
        case Value::Array: {
            out += '[';
            for (size_t i = 0; i < v.arr.size(); i++) { if (i) out += ','; write(v.arr[i], out); }
            out += ']';
            break;
        }
        case Value::Object: {
            out += '{';
            for (size_t i = 0; i < v.obj.size(); i++) {
                if (i) out += ',';
                write(Value::of(v.obj[i].first), out);
                out += ':';
                write(v.obj[i].second, out);
            }
            out += '}';
            break;
        }
    }
}
inline std::string serialize(const Value& v) { std::string s; write(v, s); return s; }

}  // namespace erj
