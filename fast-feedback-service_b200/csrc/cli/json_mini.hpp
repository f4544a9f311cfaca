// json_mini.hpp -- the small JSON subset the spotfinder executable needs: parsing the
// /dev/shm stream header (start_1, spotfinder/shmread.cc:13-58) and the --detector argument
// (spotfinder/kernels/masking.cuh:42-61), and writing one result object per line
// (spotfinder.cc:997-1008).  nlohmann_json is not available offline.
#pragma once
#include <charconv>
#include <cmath>
#include <cstdint>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace jsonmini {

struct Value {
    enum Kind { Null, Bool, Number, String, Array, Object } kind = Null;
    bool b = false;
    double num = 0;
    std::string str;
    std::vector<Value> arr;
    std::map<std::string, Value> obj;

    bool contains(const std::string& k) const { return kind == Object && obj.count(k); }
    const Value& at(const std::string& k) const {
        auto it = obj.find(k);
        if (kind != Object || it == obj.end()) throw std::runtime_error("JSON key not found: " + k);
        return it->second;
    }
    double number() const {
        if (kind != Number) throw std::runtime_error("JSON value is not a number");
        return num;
    }
};

class Parser {
    const char* p;
    const char* end;
    void ws() {
        while (p < end && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) ++p;
    }
    [[noreturn]] void fail(const char* what) { throw std::runtime_error(std::string("JSON parse error: ") + what); }
    std::string string() {
        if (*p != '"') fail("expected string");
        ++p;
        std::string s;
        while (p < end && *p != '"') {
            if (*p == '\\') {
                ++p;
                if (p >= end) fail("bad escape");
                switch (*p) {
                case 'n': s += '\n'; break;
                case 't': s += '\t'; break;
                case 'r': s += '\r'; break;
                case 'b': s += '\b'; break;
                case 'f': s += '\f'; break;
                case 'u': {  // keep BMP code points as UTF-8
                    if (end - p < 5) fail("bad \\u escape");
                    unsigned cp = (unsigned)std::stoul(std::string(p + 1, p + 5), nullptr, 16);
                    p += 4;
                    if (cp < 0x80) s += (char)cp;
                    else if (cp < 0x800) { s += (char)(0xC0 | (cp >> 6)); s += (char)(0x80 | (cp & 0x3F)); }
                    else { s += (char)(0xE0 | (cp >> 12)); s += (char)(0x80 | ((cp >> 6) & 0x3F)); s += (char)(0x80 | (cp & 0x3F)); }
                    break;
                }
                default: s += *p;
                }
                ++p;
            } else {
                s += *p++;
            }
        }
        if (p >= end) fail("unterminated string");
        ++p;
        return s;
    }
    Value value() {
        ws();
        if (p >= end) fail("unexpected end");
        Value v;
        if (*p == '{') {
            v.kind = Value::Object;
            ++p;
            ws();
            if (*p == '}') { ++p; return v; }
            for (;;) {
                ws();
                std::string k = string();
                ws();
                if (*p != ':') fail("expected ':'");
                ++p;
                v.obj[k] = value();
                ws();
                if (*p == ',') { ++p; continue; }
                if (*p == '}') { ++p; return v; }
                fail("expected ',' or '}'");
            }
        } else if (*p == '[') {
            v.kind = Value::Array;
            ++p;
            ws();
            if (*p == ']') { ++p; return v; }
            for (;;) {
                v.arr.push_back(value());
                ws();
                if (*p == ',') { ++p; continue; }
                if (*p == ']') { ++p; return v; }
                fail("expected ',' or ']'");
            }
        } else if (*p == '"') {
            v.kind = Value::String;
            v.str = string();
        } else if (end - p >= 4 && std::string(p, p + 4) == "true") {
            v.kind = Value::Bool; v.b = true; p += 4;
        } else if (end - p >= 5 && std::string(p, p + 5) == "false") {
            v.kind = Value::Bool; p += 5;
        } else if (end - p >= 4 && std::string(p, p + 4) == "null") {
            p += 4;
        } else {
            const char* q = p;
            while (q < end && (std::isdigit((unsigned char)*q) || *q == '-' || *q == '+' || *q == '.' || *q == 'e' || *q == 'E')) ++q;
            if (q == p) fail("unexpected character");
            v.kind = Value::Number;
            v.num = std::stod(std::string(p, q));
            p = q;
        }
        return v;
    }

  public:
    static Value parse(const std::string& text) {
        Parser ps;
        ps.p = text.data();
        ps.end = text.data() + text.size();
        Value v = ps.value();
        ps.ws();
        if (ps.p != ps.end) ps.fail("trailing characters");
        return v;
    }
};

inline std::string escape(const std::string& s) {
    std::string o = "\"";
    for (char ch : s) {
        switch (ch) {
        case '"': o += "\\\""; break;
        case '\\': o += "\\\\"; break;
        case '\n': o += "\\n"; break;
        case '\t': o += "\\t"; break;
        case '\r': o += "\\r"; break;
        default:
            if ((unsigned char)ch < 0x20) { char b[8]; snprintf(b, sizeof b, "\\u%04x", ch); o += b; }
            else o += ch;
        }
    }
    return o + "\"";
}

// nlohmann dumps a float as the shortest round-trip representation of its DOUBLE value,
// always with a ".0" when the text would otherwise look like an integer.
inline std::string number(double d) {
    if (!std::isfinite(d)) return "null";
    char buf[40];
    auto r = std::to_chars(buf, buf + sizeof buf, d);
    std::string s(buf, r.ptr);
    if (s.find_first_of(".eE") == std::string::npos) s += ".0";
    return s;
}

}  // namespace jsonmini
