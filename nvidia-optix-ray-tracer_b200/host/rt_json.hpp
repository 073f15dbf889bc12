// rt_json.hpp — a small recursive-descent JSON reader (RFC 8259 subset: no \u surrogate pairs
// beyond the BMP) for config.json and .vtk.series files.  The reference uses the vendored
// nlohmann/json (include/JSON/json.hpp) for parsing only; nothing else of it is needed here.
#pragma once
#include <cmath>
#include <cstdlib>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace rtjson {

struct Value;
using Array = std::vector<Value>;
using Object = std::vector<std::pair<std::string, Value>>;  // keeps file order

struct Value {
    enum Kind { Null, Bool, Number, String, Arr, Obj } kind = Null;
    bool b = false;
    double num = 0.0;
    std::string str;
    std::shared_ptr<Array> arr;
    std::shared_ptr<Object> obj;

    bool is_object() const { return kind == Obj; }
    bool is_array() const { return kind == Arr; }
    bool contains(const std::string &k) const {
        if (kind != Obj) return false;
        for (auto &kv : *obj)
            if (kv.first == k) return true;
        return false;
    }
    // nlohmann's .at(): a missing key is an error (the reference aborts on it)
    const Value &at(const std::string &k) const {
        if (kind != Obj) throw std::runtime_error("JSON: '" + k + "' requested from a non-object");
        for (auto &kv : *obj)
            if (kv.first == k) return kv.second;
        throw std::runtime_error("JSON: missing key '" + k + "'");
    }
    const Value &at(size_t i) const {
        if (kind != Arr || i >= arr->size()) throw std::runtime_error("JSON: array index out of range");
        return (*arr)[i];
    }
    size_t size() const { return kind == Arr ? arr->size() : kind == Obj ? obj->size() : 0; }
    double as_number() const {
        if (kind != Number) throw std::runtime_error("JSON: number expected");
        return num;
    }
    bool as_bool() const {
        if (kind != Bool) throw std::runtime_error("JSON: boolean expected");
        return b;
    }
    const std::string &as_string() const {
        if (kind != String) throw std::runtime_error("JSON: string expected");
        return str;
    }
    std::vector<float> as_floats() const {
        if (kind != Arr) throw std::runtime_error("JSON: array expected");
        std::vector<float> r;
        for (auto &v : *arr) r.push_back((float)v.as_number());
        return r;
    }
};

class Parser {
public:
    explicit Parser(const std::string &text) : s_(text) {}
    Value parse() {
        Value v = value();
        ws();
        if (p_ != s_.size()) fail("trailing characters");
        return v;
    }

private:
    const std::string &s_;
    size_t p_ = 0;

    [[noreturn]] void fail(const std::string &what) const {
        size_t line = 1;
        for (size_t i = 0; i < p_ && i < s_.size(); i++)
            if (s_[i] == '\n') line++;
        throw std::runtime_error("JSON parse error at line " + std::to_string(line) + ": " + what);
    }
    void ws() {
        while (p_ < s_.size() && (s_[p_] == ' ' || s_[p_] == '\t' || s_[p_] == '\n' || s_[p_] == '\r')) p_++;
    }
    bool lit(const char *w) {
        size_t n = 0;
        while (w[n]) n++;
        if (s_.compare(p_, n, w) == 0) {
            p_ += n;
            return true;
        }
        return false;
    }
    Value value() {
        ws();
        if (p_ >= s_.size()) fail("unexpected end of input");
        Value v;
        const char c = s_[p_];
        if (c == '{') {
            v.kind = Value::Obj;
            v.obj = std::make_shared<Object>();
            p_++;
            ws();
            if (p_ < s_.size() && s_[p_] == '}') {
                p_++;
                return v;
            }
            for (;;) {
                ws();
                if (p_ >= s_.size() || s_[p_] != '"') fail("object key expected");
                std::string k = string();
                ws();
                if (p_ >= s_.size() || s_[p_] != ':') fail("':' expected");
                p_++;
                v.obj->emplace_back(std::move(k), value());
                ws();
                if (p_ < s_.size() && s_[p_] == ',') {
                    p_++;
                    continue;
                }
                if (p_ < s_.size() && s_[p_] == '}') {
                    p_++;
                    return v;
                }
                fail("',' or '}' expected");
            }
        }
        if (c == '[') {
            v.kind = Value::Arr;
            v.arr = std::make_shared<Array>();
            p_++;
            ws();
            if (p_ < s_.size() && s_[p_] == ']') {
                p_++;
                return v;
            }
            for (;;) {
                v.arr->push_back(value());
                ws();
                if (p_ < s_.size() && s_[p_] == ',') {
                    p_++;
                    continue;
                }
                if (p_ < s_.size() && s_[p_] == ']') {
                    p_++;
                    return v;
                }
                fail("',' or ']' expected");
            }
        }
        if (c == '"') {
            v.kind = Value::String;
            v.str = string();
            return v;
        }
        if (lit("true")) {
            v.kind = Value::Bool;
            v.b = true;
            return v;
        }
        if (lit("false")) {
            v.kind = Value::Bool;
            v.b = false;
            return v;
        }
        if (lit("null")) return v;
        if (c == '-' || (c >= '0' && c <= '9')) {
            const char *b = s_.c_str() + p_;
            char *e = nullptr;
            v.num = std::strtod(b, &e);
            if (e == b) fail("bad number");
            p_ += (size_t)(e - b);
            v.kind = Value::Number;
            return v;
        }
        fail(std::string("unexpected character '") + c + "'");
    }
    std::string string() {
        std::string r;
        p_++;  // opening quote
        while (p_ < s_.size() && s_[p_] != '"') {
            char c = s_[p_++];
            if (c != '\\') {
                r.push_back(c);
                continue;
            }
            if (p_ >= s_.size()) fail("bad escape");
            c = s_[p_++];
            switch (c) {
                case '"': r.push_back('"'); break;
                case '\\': r.push_back('\\'); break;
                case '/': r.push_back('/'); break;
                case 'b': r.push_back('\b'); break;
                case 'f': r.push_back('\f'); break;
                case 'n': r.push_back('\n'); break;
                case 'r': r.push_back('\r'); break;
                case 't': r.push_back('\t'); break;
                case 'u': {
                    if (p_ + 4 > s_.size()) fail("bad \\u escape");
                    unsigned cp = (unsigned)std::strtoul(s_.substr(p_, 4).c_str(), nullptr, 16);
                    p_ += 4;
                    if (cp < 0x80) r.push_back((char)cp);
                    else if (cp < 0x800) {
                        r.push_back((char)(0xC0 | (cp >> 6)));
                        r.push_back((char)(0x80 | (cp & 0x3F)));
                    } else {
                        r.push_back((char)(0xE0 | (cp >> 12)));
                        r.push_back((char)(0x80 | ((cp >> 6) & 0x3F)));
                        r.push_back((char)(0x80 | (cp & 0x3F)));
                    }
                    break;
                }
                default: fail("bad escape");
            }
        }
        if (p_ >= s_.size()) fail("unterminated string");
        p_++;  // closing quote
        return r;
    }
};

inline Value parse(const std::string &text) { return Parser(text).parse(); }

}  // namespace rtjson
