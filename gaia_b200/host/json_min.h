// Minimal JSON reader/writer for the host adapter (Models.json / Parameters.json / recovery states / statistics).
// Stands in for the nlohmann::json + MF::BaseJsonConfig pair the reference uses (M/Json/json.hpp,
// 3P/MeshFrame2/MeshFrame/Utility/Parser.h:104-120: a missing key keeps the C++ default).
#pragma once
#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace jsonmin {

struct Value {
    enum Type { Null, Bool, Number, String, Array, Object } type = Null;
    bool b = false;
    double num = 0.0;
    std::string str;
    std::vector<Value> arr;
    std::vector<std::pair<std::string, Value>> obj;

    bool is_null() const { return type == Null; }
    bool is_object() const { return type == Object; }
    bool is_array() const { return type == Array; }
    size_t size() const { return type == Array ? arr.size() : obj.size(); }
    const Value* find(const std::string& k) const {
        if (type != Object) return nullptr;
        for (auto& kv : obj) if (kv.first == k) return &kv.second;
        return nullptr;
    }
    const Value& operator[](size_t i) const { return arr.at(i); }
    double as_double() const { return type == Bool ? (b ? 1.0 : 0.0) : num; }
    bool as_bool() const { return type == Bool ? b : num != 0.0; }
    // helpers mirroring EXTRACT_FROM_JSON: assign when present, keep the default otherwise
    template <class T> bool get(const std::string& k, T& out) const {
        const Value* v = find(k);
        if (!v || v->is_null()) return false;
        out = (T)v->as_double();
        return true;
    }
    bool get(const std::string& k, bool& out) const { const Value* v = find(k); if (!v || v->is_null()) return false; out = v->as_bool(); return true; }
    bool get(const std::string& k, std::string& out) const { const Value* v = find(k); if (!v || v->type != String) return false; out = v->str; return true; }
    bool get_vec3(const std::string& k, float out[3]) const {
        const Value* v = find(k);
        if (!v || v->type != Array || v->arr.size() < 3) return false;
        for (int i = 0; i < 3; i++) out[i] = (float)v->arr[i].as_double();
        return true;
    }
    bool get_ints(const std::string& k, std::vector<int>& out) const {
        const Value* v = find(k);
        if (!v || v->type != Array) return false;
        out.clear();
        for (auto& e : v->arr) out.push_back((int)e.as_double());
        return true;
    }
};

class Parser {
public:
    explicit Parser(const std::string& text) : s_(text) {}
    Value parse() { Value v = value(); ws(); if (p_ != s_.size()) fail("trailing characters"); return v; }
private:
    const std::string& s_;
    size_t p_ = 0;
    [[noreturn]] void fail(const char* what) { throw std::runtime_error(std::string("JSON: ") + what + " at offset " + std::to_string(p_)); }
    void ws() { while (p_ < s_.size() && std::isspace((unsigned char)s_[p_])) p_++; }
    bool eat(char c) { ws(); if (p_ < s_.size() && s_[p_] == c) { p_++; return true; } return false; }
    Value value() {
        ws();
        if (p_ >= s_.size()) fail("unexpected end");
        const char c = s_[p_];
        Value v;
        if (c == '{') {
            p_++; v.type = Value::Object;
            if (eat('}')) return v;
            do { ws(); Value k = string_(); if (!eat(':')) fail("expected ':'"); v.obj.emplace_back(k.str, value()); } while (eat(','));
            if (!eat('}')) fail("expected '}'");
        } else if (c == '[') {
            p_++; v.type = Value::Array;
            if (eat(']')) return v;
            do { v.arr.push_back(value()); } while (eat(','));
            if (!eat(']')) fail("expected ']'");
        } else if (c == '"') {
            v = string_();
        } else if (!s_.compare(p_, 4, "true")) { v.type = Value::Bool; v.b = true; p_ += 4; }
        else if (!s_.compare(p_, 5, "false")) { v.type = Value::Bool; v.b = false; p_ += 5; }
        else if (!s_.compare(p_, 4, "null")) { p_ += 4; }
        else if (!s_.compare(p_, 3, "NaN")) { v.type = Value::Number; v.num = std::strtod("nan", nullptr); p_ += 3; }
        else {
            char* end = nullptr;
            v.type = Value::Number;
            v.num = std::strtod(s_.c_str() + p_, &end);
            if (end == s_.c_str() + p_) fail("bad token");
            p_ = (size_t)(end - s_.c_str());
        }
        return v;
    }
    Value string_() {
        if (p_ >= s_.size() || s_[p_] != '"') fail("expected string");
        p_++;
        Value v; v.type = Value::String;
        while (p_ < s_.size() && s_[p_] != '"') {
            char c = s_[p_++];
            if (c == '\\' && p_ < s_.size()) {
                const char e = s_[p_++];
                switch (e) {
                    case 'n': c = '\n'; break; case 't': c = '\t'; break; case 'r': c = '\r'; break;
                    case 'b': c = '\b'; break; case 'f': c = '\f'; break;
                    case 'u': { c = (char)std::strtol(s_.substr(p_, 4).c_str(), nullptr, 16); p_ += 4; } break;
                    default: c = e;   // \" \\ \/
                }
            }
            v.str.push_back(c);
        }
        if (p_ >= s_.size()) fail("unterminated string");
        p_++;
        return v;
    }
};

inline Value parse_file(const std::string& path) {
    std::ifstream f(path);
    if (!f) throw std::runtime_error("cannot open " + path);
    std::stringstream ss;
    ss << f.rdbuf();
    const std::string text = ss.str();
    return Parser(text).parse();
}

}  // namespace jsonmin
