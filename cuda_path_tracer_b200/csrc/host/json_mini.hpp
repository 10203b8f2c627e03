// json_mini.hpp -- the JSON subset scene.json needs (objects, arrays, numbers, strings, bools).
//
// The reference parses scenes with nlohmann::json 3.11.3 (src/scene.cpp:15,91), whose default
// object type is std::map: iterating "Materials" visits keys in SORTED order, and that order
// is what assigns material ids (src/scene.cpp:302-327).  This DOM keeps objects in a std::map
// for the same reason.  Numbers go through strtod like nlohmann's lexer, then narrow to float
// at the point of use.
#pragma once
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace ptb {

struct JsonValue {
    enum Kind { Null, Bool, Number, String, Array, Object } kind = Null;
    bool b = false;
    double num = 0.0;
    std::string str;
    std::vector<JsonValue> arr;
    std::map<std::string, JsonValue> obj;

    bool contains(const std::string& k) const { return kind == Object && obj.count(k) > 0; }
    const JsonValue& at(const std::string& k) const {
        auto it = obj.find(k);
        if (kind != Object || it == obj.end()) throw std::runtime_error("json: missing key '" + k + "'");
        return it->second;
    }
    const JsonValue& at(size_t i) const {
        if (kind != Array || i >= arr.size()) throw std::runtime_error("json: array index out of range");
        return arr[i];
    }
    double number() const {
        if (kind != Number) throw std::runtime_error("json: value is not a number");
        return num;
    }
    const std::string& string() const {
        if (kind != String) throw std::runtime_error("json: value is not a string");
        return str;
    }
    float value_f(const std::string& k, float dflt) const { return contains(k) ? (float)at(k).number() : dflt; }
    std::string value_s(const std::string& k, const std::string& dflt) const {
        return contains(k) ? at(k).string() : dflt;
    }
};

class JsonParser {
public:
    explicit JsonParser(const std::string& text) : s_(text), p_(0) {}
    JsonValue parse() {
        JsonValue v = value();
        ws();
        if (p_ != s_.size()) fail("trailing characters");
        return v;
    }

private:
    const std::string& s_;
    size_t p_;

    [[noreturn]] void fail(const char* what) const {
        throw std::runtime_error(std::string("json: ") + what + " at byte " + std::to_string(p_));
    }
    void ws() {
        while (p_ < s_.size() && (s_[p_] == ' ' || s_[p_] == '\t' || s_[p_] == '\n' || s_[p_] == '\r')) p_++;
    }
    JsonValue value() {
        ws();
        if (p_ >= s_.size()) fail("unexpected end");
        char c = s_[p_];
        JsonValue v;
        if (c == '{') {
            v.kind = JsonValue::Object;
            p_++;
            ws();
            if (p_ < s_.size() && s_[p_] == '}') { p_++; return v; }
            for (;;) {
                ws();
                if (p_ >= s_.size() || s_[p_] != '"') fail("expected object key");
                std::string k = str();
                ws();
                if (p_ >= s_.size() || s_[p_] != ':') fail("expected ':'");
                p_++;
                v.obj[k] = value();   // duplicate keys: last wins, as in nlohmann
                ws();
                if (p_ < s_.size() && s_[p_] == ',') { p_++; continue; }
                if (p_ < s_.size() && s_[p_] == '}') { p_++; break; }
                fail("expected ',' or '}'");
            }
        } else if (c == '[') {
            v.kind = JsonValue::Array;
            p_++;
            ws();
            if (p_ < s_.size() && s_[p_] == ']') { p_++; return v; }
            for (;;) {
                v.arr.push_back(value());
                ws();
                if (p_ < s_.size() && s_[p_] == ',') { p_++; continue; }
                if (p_ < s_.size() && s_[p_] == ']') { p_++; break; }
                fail("expected ',' or ']'");
            }
        } else if (c == '"') {
            v.kind = JsonValue::String;
            v.str = str();
        } else if (!s_.compare(p_, 4, "true")) {
            v.kind = JsonValue::Bool; v.b = true; p_ += 4;
        } else if (!s_.compare(p_, 5, "false")) {
            v.kind = JsonValue::Bool; v.b = false; p_ += 5;
        } else if (!s_.compare(p_, 4, "null")) {
            p_ += 4;
        } else if (c == '-' || (c >= '0' && c <= '9')) {
            const char* b = s_.c_str() + p_;
            char* e = nullptr;
            v.kind = JsonValue::Number;
            v.num = strtod(b, &e);
            if (e == b) fail("bad number");
            p_ += (size_t)(e - b);
        } else {
            fail("unexpected character");
        }
        return v;
    }
    std::string str() {
        std::string out;
        p_++;  // opening quote
        while (p_ < s_.size() && s_[p_] != '"') {
            char c = s_[p_++];
            if (c == '\\') {
                if (p_ >= s_.size()) fail("bad escape");
                char e = s_[p_++];
                switch (e) {
                    case 'n': out += '\n'; break;
                    case 't': out += '\t'; break;
                    case 'r': out += '\r'; break;
                    case 'b': out += '\b'; break;
                    case 'f': out += '\f'; break;
                    case 'u': {
                        if (p_ + 4 > s_.size()) fail("bad \\u escape");
                        unsigned cp = (unsigned)strtoul(s_.substr(p_, 4).c_str(), nullptr, 16);
                        p_ += 4;
                        if (cp < 0x80) out += (char)cp;
                        else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
                        else { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
                        break;
                    }
                    default: out += e; break;  // \" \\ \/
                }
            } else {
                out += c;
            }
        }
        if (p_ >= s_.size()) fail("unterminated string");
        p_++;  // closing quote
        return out;
    }
};

}  // namespace ptb
