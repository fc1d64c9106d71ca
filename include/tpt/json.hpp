// json.hpp — a small JSON DOM reader, enough for ASCII .gltf files (RFC 8259 values, UTF-8 passed through,
// \uXXXX escapes decoded to UTF-8).  The reference parses glTF with the tinygltf + nlohmann bundle it vendors
// (include/tinygltf/); this repository carries its own reader instead of a third-party copy.
#pragma once
#include <cstdint>
#include <cstdlib>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace tpt {
namespace json {

struct Value;
using Array = std::vector<Value>;
using Object = std::map<std::string, Value>;

struct Value {
    enum Kind { Null, Bool, Number, String, ArrayK, ObjectK } kind = Null;
    bool b = false;
    double num = 0.0;
    std::string str;
    std::shared_ptr<Array> arr;
    std::shared_ptr<Object> obj;

    bool is_object() const { return kind == ObjectK; }
    bool is_array() const { return kind == ArrayK; }
    bool is_number() const { return kind == Number; }
    bool is_string() const { return kind == String; }
    bool has(const std::string& key) const { return kind == ObjectK && obj->count(key) != 0; }
    // member access that never throws: a missing member is a Null value
    const Value& operator[](const std::string& key) const {
        static const Value none;
        if (kind != ObjectK) return none;
        auto it = obj->find(key);
        return it == obj->end() ? none : it->second;
    }
    const Value& operator[](size_t i) const {
        static const Value none;
        return (kind == ArrayK && i < arr->size()) ? (*arr)[i] : none;
    }
    size_t size() const { return kind == ArrayK ? arr->size() : (kind == ObjectK ? obj->size() : 0); }
    double number(double dflt) const { return kind == Number ? num : dflt; }
    int integer(int dflt) const { return kind == Number ? (int)num : dflt; }
    const std::string& string(const std::string& dflt) const { return kind == String ? str : dflt; }
};

class Parser {
public:
    explicit Parser(const std::string& text) : s_(text) {}
    Value parse() {
        Value v = value();
        ws();
        if (p_ != s_.size()) fail("trailing characters");
        return v;
    }

private:
    const std::string& s_;
    size_t p_ = 0;

    [[noreturn]] void fail(const char* what) const { throw std::runtime_error(std::string("JSON: ") + what + " at byte " + std::to_string(p_)); }
    void ws() {
        while (p_ < s_.size() && (s_[p_] == ' ' || s_[p_] == '\n' || s_[p_] == '\r' || s_[p_] == '\t')) ++p_;
    }
    bool eat(char c) {
        ws();
        if (p_ < s_.size() && s_[p_] == c) {
            ++p_;
            return true;
        }
        return false;
    }
    void expect(const char* lit) {
        for (const char* c = lit; *c; ++c, ++p_)
            if (p_ >= s_.size() || s_[p_] != *c) fail("bad literal");
    }
    static void utf8(std::string& out, uint32_t cp) {
        if (cp < 0x80) out += (char)cp;
        else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
        else if (cp < 0x10000) { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
        else { out += (char)(0xF0 | (cp >> 18)); out += (char)(0x80 | ((cp >> 12) & 0x3F)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
    }
    uint32_t hex4() {
        if (p_ + 4 > s_.size()) fail("short \\u escape");
        uint32_t v = 0;
        for (int i = 0; i < 4; ++i) {
            const char c = s_[p_++];
            v <<= 4;
            if (c >= '0' && c <= '9') v |= (uint32_t)(c - '0');
            else if (c >= 'a' && c <= 'f') v |= (uint32_t)(c - 'a' + 10);
            else if (c >= 'A' && c <= 'F') v |= (uint32_t)(c - 'A' + 10);
            else fail("bad \\u escape");
        }
        return v;
    }
    std::string string_body() {
        std::string out;
        // long base64 payloads: copy runs between escapes in one go
        while (true) {
            const size_t start = p_;
            while (p_ < s_.size() && s_[p_] != '"' && s_[p_] != '\\') ++p_;
            out.append(s_, start, p_ - start);
            if (p_ >= s_.size()) fail("unterminated string");
            if (s_[p_] == '"') {
                ++p_;
                return out;
            }
            ++p_;   // backslash
            if (p_ >= s_.size()) fail("unterminated escape");
            const char c = s_[p_++];
            switch (c) {
                case '"': out += '"'; break;
                case '\\': out += '\\'; break;
                case '/': out += '/'; break;
                case 'b': out += '\b'; break;
                case 'f': out += '\f'; break;
                case 'n': out += '\n'; break;
                case 'r': out += '\r'; break;
                case 't': out += '\t'; break;
                case 'u': {
                    uint32_t cp = hex4();
                    if (cp >= 0xD800 && cp <= 0xDBFF && p_ + 1 < s_.size() && s_[p_] == '\\' && s_[p_ + 1] == 'u') {
                        p_ += 2;
                        const uint32_t lo = hex4();
                        cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
                    }
                    utf8(out, cp);
                    break;
                }
                default: fail("bad escape");
            }
        }
    }
    Value value() {
        ws();
        if (p_ >= s_.size()) fail("unexpected end");
        Value v;
        const char c = s_[p_];
        if (c == '{') {
            ++p_;
            v.kind = Value::ObjectK;
            v.obj = std::make_shared<Object>();
            if (eat('}')) return v;
            do {
                ws();
                if (p_ >= s_.size() || s_[p_] != '"') fail("object key expected");
                ++p_;
                std::string key = string_body();
                if (!eat(':')) fail("':' expected");
                (*v.obj)[key] = value();
            } while (eat(','));
            if (!eat('}')) fail("'}' expected");
        } else if (c == '[') {
            ++p_;
            v.kind = Value::ArrayK;
            v.arr = std::make_shared<Array>();
            if (eat(']')) return v;
            do v.arr->push_back(value());
            while (eat(','));
            if (!eat(']')) fail("']' expected");
        } else if (c == '"') {
            ++p_;
            v.kind = Value::String;
            v.str = string_body();
        } else if (c == 't') {
            expect("true");
            v.kind = Value::Bool;
            v.b = true;
        } else if (c == 'f') {
            expect("false");
            v.kind = Value::Bool;
        } else if (c == 'n') {
            expect("null");
        } else {
            const char* begin = s_.c_str() + p_;
            char* end = nullptr;
            v.num = std::strtod(begin, &end);
            if (end == begin) fail("value expected");
            p_ += (size_t)(end - begin);
            v.kind = Value::Number;
        }
        return v;
    }
};

inline Value parse(const std::string& text) { return Parser(text).parse(); }

}   // namespace json
}   // namespace tpt
