// Minimal JSON reader for WOMD-shaped scenario files (host side, load-once).
// Numbers are parsed with strtod (double) and narrowed on read, which is what the
// reference gets from nlohmann::json `.get<float>()` (src/json_serialization.hpp:12-16).
#pragma once
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace gd {

struct JsonValue {
    enum Kind : uint8_t { Null, Bool, Int, Float, String, Array, Object };
    Kind kind = Null;
    bool b = false;
    int64_t i = 0;      // valid when kind == Int
    double d = 0.0;     // valid when kind == Int or Float
    std::string s;
    std::vector<JsonValue> arr;
    std::vector<std::pair<std::string, JsonValue>> obj;

    bool isNumber() const { return kind == Int || kind == Float; }
    size_t size() const { return kind == Array ? arr.size() : (kind == Object ? obj.size() : 0); }

    const JsonValue *find(const char *key) const {
        if (kind != Object) return nullptr;
        for (const auto &kv : obj)
            if (kv.first == key) return &kv.second;
        return nullptr;
    }
    bool contains(const char *key) const { return find(key) != nullptr; }
    const JsonValue &at(const char *key) const {
        const JsonValue *v = find(key);
        if (!v) throw std::runtime_error(std::string("json: missing key '") + key + "'");
        return *v;
    }
    const JsonValue &operator[](size_t idx) const {
        if (kind != Array || idx >= arr.size()) throw std::runtime_error("json: index out of range");
        return arr[idx];
    }
    float asFloat() const {
        if (!isNumber()) throw std::runtime_error("json: number expected");
        return kind == Int ? (float)i : (float)d;
    }
    int64_t asInt() const {
        if (kind == Int) return i;
        if (kind == Float) return (int64_t)d;
        if (kind == Bool) return b ? 1 : 0;
        throw std::runtime_error("json: integer expected");
    }
    bool asBool() const {
        if (kind != Bool) throw std::runtime_error("json: boolean expected");
        return b;
    }
    const std::string &asString() const {
        if (kind != String) throw std::runtime_error("json: string expected");
        return s;
    }
};

class JsonParser {
public:
    JsonParser(const char *begin, const char *end) : p_(begin), end_(end) {}

    JsonValue parse() {
        JsonValue v = value();
        ws();
        if (p_ != end_) fail("trailing characters");
        return v;
    }

private:
    const char *p_, *end_;

    [[noreturn]] void fail(const char *msg) const {
        throw std::runtime_error(std::string("json: ") + msg);
    }
    void ws() {
        while (p_ < end_ && (*p_ == ' ' || *p_ == '\n' || *p_ == '\t' || *p_ == '\r')) ++p_;
    }
    bool lit(const char *s) {
        size_t n = strlen(s);
        if ((size_t)(end_ - p_) >= n && memcmp(p_, s, n) == 0) { p_ += n; return true; }
        return false;
    }

    JsonValue value() {
        ws();
        if (p_ >= end_) fail("unexpected end");
        JsonValue v;
        char c = *p_;
        if (c == '{') {
            ++p_;
            v.kind = JsonValue::Object;
            ws();
            if (p_ < end_ && *p_ == '}') { ++p_; return v; }
            for (;;) {
                ws();
                if (p_ >= end_ || *p_ != '"') fail("object key expected");
                std::string k = str();
                ws();
                if (p_ >= end_ || *p_ != ':') fail("':' expected");
                ++p_;
                v.obj.emplace_back(std::move(k), value());
                ws();
                if (p_ < end_ && *p_ == ',') { ++p_; continue; }
                if (p_ < end_ && *p_ == '}') { ++p_; break; }
                fail("',' or '}' expected");
            }
        } else if (c == '[') {
            ++p_;
            v.kind = JsonValue::Array;
            ws();
            if (p_ < end_ && *p_ == ']') { ++p_; return v; }
            for (;;) {
                v.arr.emplace_back(value());
                ws();
                if (p_ < end_ && *p_ == ',') { ++p_; continue; }
                if (p_ < end_ && *p_ == ']') { ++p_; break; }
                fail("',' or ']' expected");
            }
        } else if (c == '"') {
            v.kind = JsonValue::String;
            v.s = str();
        } else if (lit("true")) {
            v.kind = JsonValue::Bool; v.b = true;
        } else if (lit("false")) {
            v.kind = JsonValue::Bool; v.b = false;
        } else if (lit("null")) {
            v.kind = JsonValue::Null;
        } else if (lit("NaN")) {       // python json.dump emits these; be lenient
            v.kind = JsonValue::Float; v.d = std::strtod("nan", nullptr);
        } else {
            number(v);
        }
        return v;
    }

    void number(JsonValue &v) {
        const char *s = p_;
        bool is_float = false;
        if (p_ < end_ && (*p_ == '-' || *p_ == '+')) ++p_;
        if (lit("Infinity")) {
            v.kind = JsonValue::Float;
            v.d = (*s == '-') ? -std::strtod("inf", nullptr) : std::strtod("inf", nullptr);
            return;
        }
        while (p_ < end_) {
            char c = *p_;
            if (c >= '0' && c <= '9') { ++p_; }
            else if (c == '.' || c == 'e' || c == 'E' || c == '+' || c == '-') { is_float = true; ++p_; }
            else break;
        }
        if (p_ == s) fail("value expected");
        std::string tok(s, p_);
        if (!is_float) {
            errno = 0;
            char *e = nullptr;
            long long iv = std::strtoll(tok.c_str(), &e, 10);
            if (errno == 0 && e && *e == 0) {
                v.kind = JsonValue::Int; v.i = iv; v.d = (double)iv;
                return;
            }
        }
        char *e = nullptr;
        v.d = std::strtod(tok.c_str(), &e);
        if (!e || *e != 0) fail("bad number");
        v.kind = JsonValue::Float;
    }

    std::string str() {
        ++p_;  // opening quote
        std::string out;
        while (p_ < end_ && *p_ != '"') {
            char c = *p_++;
            if (c != '\\') { out.push_back(c); continue; }
            if (p_ >= end_) fail("bad escape");
            char e = *p_++;
            switch (e) {
            case 'n': out.push_back('\n'); break;
            case 't': out.push_back('\t'); break;
            case 'r': out.push_back('\r'); break;
            case 'b': out.push_back('\b'); break;
            case 'f': out.push_back('\f'); break;
            case 'u': {
                if (end_ - p_ < 4) fail("bad \\u escape");
                unsigned cp = (unsigned)std::strtoul(std::string(p_, p_ + 4).c_str(), nullptr, 16);
                p_ += 4;
                if (cp < 0x80) out.push_back((char)cp);
                else if (cp < 0x800) { out.push_back((char)(0xC0 | (cp >> 6))); out.push_back((char)(0x80 | (cp & 0x3F))); }
                else { out.push_back((char)(0xE0 | (cp >> 12))); out.push_back((char)(0x80 | ((cp >> 6) & 0x3F))); out.push_back((char)(0x80 | (cp & 0x3F))); }
            } break;
            default: out.push_back(e); break;   // \" \\ \/
            }
        }
        if (p_ >= end_) fail("unterminated string");
        ++p_;  // closing quote
        return out;
    }
};

} // namespace gd
