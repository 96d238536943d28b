// json_min.hpp -- a small JSON reader for the scene files (host side, header only).
//
// The reference parses its scenes with nlohmann::json (path_tracer/src/scene.cpp:33-37).  What
// the loader relies on, and what this reader therefore reproduces:
//   * objects iterate in lexicographic (byte-wise) key order -- nlohmann's default object type is
//     a std::map -- which is what makes material ids follow the alphabetical order of their names
//     (SURVEY.md F12);
//   * numbers: a literal without '.', 'e' or 'E' is an integer (int64), anything else a double
//     parsed by strtod; conversion to float is a static_cast at the point of use;
//   * `value(key, default)` semantics: a missing key yields the default.
// Not supported (not needed by the format): \u escapes beyond Latin-1, comments, NaN/Infinity.
#pragma once
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace b200host {

class Json {
public:
    enum Kind { Null, Bool, Int, Real, String, Array, Object };

    Kind kind = Null;
    bool b = false;
    int64_t i = 0;
    double d = 0.0;
    std::string s;
    std::vector<Json> arr;
    std::map<std::string, Json> obj;

    static Json parse(const std::string& text)
    {
        Parser p{text.c_str(), text.c_str() + text.size()};
        Json v = p.value();
        p.skip_ws();
        if (p.cur != p.end) p.fail("trailing characters");
        return v;
    }

    bool is_number() const { return kind == Int || kind == Real; }
    bool contains(const std::string& k) const { return kind == Object && obj.count(k) != 0; }

    const Json& at(const std::string& k) const
    {
        if (kind != Object) throw std::runtime_error("json: not an object (looking up \"" + k + "\")");
        auto it = obj.find(k);
        if (it == obj.end()) throw std::runtime_error("json: missing key \"" + k + "\"");
        return it->second;
    }
    const Json& at(size_t idx) const
    {
        if (kind != Array || idx >= arr.size()) throw std::runtime_error("json: array index out of range");
        return arr[idx];
    }

    double as_double() const
    {
        if (kind == Int) return (double)i;
        if (kind == Real) return d;
        throw std::runtime_error("json: not a number");
    }
    float as_float() const { return kind == Int ? (float)i : (float)as_double(); }
    int as_int() const
    {
        if (kind == Int) return (int)i;
        if (kind == Real) return (int)d;
        throw std::runtime_error("json: not a number");
    }
    const std::string& as_string() const
    {
        if (kind != String) throw std::runtime_error("json: not a string");
        return s;
    }

    float value(const std::string& k, float dflt) const { return contains(k) ? at(k).as_float() : dflt; }
    std::string value(const std::string& k, const std::string& dflt) const
    {
        return contains(k) ? at(k).as_string() : dflt;
    }

private:
    struct Parser {
        const char* cur;
        const char* end;

        [[noreturn]] void fail(const char* what) const { throw std::runtime_error(std::string("json: ") + what); }
        void skip_ws()
        {
            while (cur != end && (*cur == ' ' || *cur == '\t' || *cur == '\n' || *cur == '\r')) ++cur;
        }
        bool eat(const char* lit)
        {
            const size_t n = strlen(lit);
            if ((size_t)(end - cur) >= n && memcmp(cur, lit, n) == 0) { cur += n; return true; }
            return false;
        }
        Json value()
        {
            skip_ws();
            if (cur == end) fail("unexpected end of input");
            Json v;
            switch (*cur) {
            case '{': {
                ++cur;
                v.kind = Object;
                skip_ws();
                if (cur != end && *cur == '}') { ++cur; return v; }
                while (true) {
                    skip_ws();
                    if (cur == end || *cur != '"') fail("expected a key");
                    std::string key = string();
                    skip_ws();
                    if (cur == end || *cur != ':') fail("expected ':'");
                    ++cur;
                    v.obj[key] = value();          // a repeated key keeps the last value, as nlohmann does
                    skip_ws();
                    if (cur != end && *cur == ',') { ++cur; continue; }
                    if (cur != end && *cur == '}') { ++cur; return v; }
                    fail("expected ',' or '}'");
                }
            }
            case '[': {
                ++cur;
                v.kind = Array;
                skip_ws();
                if (cur != end && *cur == ']') { ++cur; return v; }
                while (true) {
                    v.arr.push_back(value());
                    skip_ws();
                    if (cur != end && *cur == ',') { ++cur; continue; }
                    if (cur != end && *cur == ']') { ++cur; return v; }
                    fail("expected ',' or ']'");
                }
            }
            case '"':
                v.kind = String;
                v.s = string();
                return v;
            case 't':
                if (!eat("true")) fail("bad literal");
                v.kind = Bool; v.b = true;
                return v;
            case 'f':
                if (!eat("false")) fail("bad literal");
                v.kind = Bool; v.b = false;
                return v;
            case 'n':
                if (!eat("null")) fail("bad literal");
                return v;
            default:
                return number();
            }
        }
        std::string string()
        {
            ++cur;                                   // opening quote
            std::string out;
            while (true) {
                if (cur == end) fail("unterminated string");
                const char c = *cur++;
                if (c == '"') return out;
                if (c != '\\') { out += c; continue; }
                if (cur == end) fail("unterminated escape");
                const char e = *cur++;
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
                    if (end - cur < 4) fail("bad \\u escape");
                    unsigned cp = 0;
                    for (int k = 0; k < 4; ++k) {
                        const char h = *cur++;
                        cp <<= 4;
                        if (h >= '0' && h <= '9') cp |= (unsigned)(h - '0');
                        else if (h >= 'a' && h <= 'f') cp |= (unsigned)(h - 'a' + 10);
                        else if (h >= 'A' && h <= 'F') cp |= (unsigned)(h - 'A' + 10);
                        else fail("bad \\u escape");
                    }
                    if (cp < 0x80) out += (char)cp;                       // UTF-8 encode (BMP only)
                    else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
                    else { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
                    break;
                }
                default: fail("bad escape");
                }
            }
        }
        Json number()
        {
            const char* start = cur;
            bool real = false;
            if (cur != end && *cur == '-') ++cur;
            if (cur == end || *cur < '0' || *cur > '9') fail("bad number");
            while (cur != end && ((*cur >= '0' && *cur <= '9') || *cur == '.' || *cur == 'e' || *cur == 'E' || *cur == '+' || *cur == '-')) {
                if (*cur == '.' || *cur == 'e' || *cur == 'E') real = true;
                ++cur;
            }
            const std::string lit(start, cur);
            Json v;
            if (real) { v.kind = Real; v.d = strtod(lit.c_str(), nullptr); }
            else { v.kind = Int; v.i = strtoll(lit.c_str(), nullptr, 10); }
            return v;
        }
    };
};

}  // namespace b200host
