// Minimal self-contained JSON reader/writer for model specifications (the reference uses nlohmann::json for the same job:
// /root/reference/interface/CPP/teqp_impl_factory.cpp:104-130 reads j["kind"], j["model"]).  Host only.
#pragma once
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace tq {

struct JsonError : std::runtime_error { using std::runtime_error::runtime_error; };

class Json {
public:
    enum Kind { Null, Bool, Number, String, Array, Object };
    Kind kind = Null;
    bool b = false;
    double num = 0;
    bool is_int = false;
    std::string str;
    std::vector<Json> arr;
    std::vector<std::pair<std::string, Json>> obj;  // insertion order kept

    bool is_null() const { return kind == Null; }
    bool is_bool() const { return kind == Bool; }
    bool is_number() const { return kind == Number; }
    bool is_string() const { return kind == String; }
    bool is_array() const { return kind == Array; }
    bool is_object() const { return kind == Object; }
    size_t size() const { return kind == Array ? arr.size() : kind == Object ? obj.size() : 0; }

    bool contains(const std::string& k) const {
        if (kind != Object) return false;
        for (auto& kv : obj) if (kv.first == k) return true;
        return false;
    }
    const Json& at(const std::string& k) const {
        if (kind != Object) throw JsonError("JSON value is not an object (looking up key '" + k + "')");
        for (auto& kv : obj) if (kv.first == k) return kv.second;
        throw JsonError("key '" + k + "' not found");
    }
    const Json& at(size_t i) const {
        if (kind != Array) throw JsonError("JSON value is not an array");
        if (i >= arr.size()) throw JsonError("array index out of range");
        return arr[i];
    }
    const Json& operator[](const std::string& k) const { return at(k); }
    const Json& operator[](size_t i) const { return at(i); }

    double as_double() const {
        if (kind == Number) return num;
        throw JsonError("JSON value is not a number");
    }
    long long as_int() const { return static_cast<long long>(as_double()); }
    bool as_bool() const { if (kind == Bool) return b; throw JsonError("JSON value is not a boolean"); }
    const std::string& as_string() const { if (kind == String) return str; throw JsonError("JSON value is not a string"); }
    std::vector<double> as_doubles() const {
        if (kind != Array) throw JsonError("JSON value is not an array of numbers");
        std::vector<double> v; v.reserve(arr.size());
        for (auto& e : arr) v.push_back(e.as_double());
        return v;
    }
    std::vector<std::string> as_strings() const {
        if (kind != Array) throw JsonError("JSON value is not an array of strings");
        std::vector<std::string> v; v.reserve(arr.size());
        for (auto& e : arr) v.push_back(e.as_string());
        return v;
    }
    double value(const std::string& k, double dflt) const { return contains(k) ? at(k).as_double() : dflt; }
    std::string value(const std::string& k, const std::string& dflt) const { return contains(k) ? at(k).as_string() : dflt; }
    bool value_bool(const std::string& k, bool dflt) const { return contains(k) ? at(k).as_bool() : dflt; }

    static Json parse(const std::string& text) {
        Parser p{text.data(), text.data() + text.size(), text.data()};
        Json j = p.value();
        p.ws();
        if (p.cur != p.end) p.fail("trailing characters after JSON document");
        return j;
    }
    static Json parse_file(const std::string& path) {
        std::ifstream f(path, std::ios::binary);
        if (!f) throw JsonError("cannot open JSON file: " + path);
        std::stringstream ss; ss << f.rdbuf();
        return parse(ss.str());
    }

    std::string dump() const { std::string s; dump_to(s); return s; }

private:
    struct Parser {
        const char* begin; const char* end; const char* cur;
        [[noreturn]] void fail(const std::string& m) const {
            throw JsonError("JSON parse error at offset " + std::to_string(cur - begin) + ": " + m);
        }
        void ws() { while (cur != end && (*cur == ' ' || *cur == '\t' || *cur == '\n' || *cur == '\r')) ++cur; }
        bool lit(const char* s) {
            const char* p = cur;
            while (*s) { if (p == end || *p != *s) return false; ++p; ++s; }
            cur = p; return true;
        }
        Json value() {
            ws();
            if (cur == end) fail("unexpected end of input");
            Json j;
            char c = *cur;
            if (c == '{') {
                ++cur; j.kind = Object; ws();
                if (cur != end && *cur == '}') { ++cur; return j; }
                for (;;) {
                    ws();
                    if (cur == end || *cur != '"') fail("expected string key");
                    std::string k = string();
                    ws();
                    if (cur == end || *cur != ':') fail("expected ':'");
                    ++cur;
                    j.obj.emplace_back(std::move(k), value());
                    ws();
                    if (cur == end) fail("unterminated object");
                    if (*cur == ',') { ++cur; continue; }
                    if (*cur == '}') { ++cur; break; }
                    fail("expected ',' or '}'");
                }
                return j;
            }
            if (c == '[') {
                ++cur; j.kind = Array; ws();
                if (cur != end && *cur == ']') { ++cur; return j; }
                for (;;) {
                    j.arr.push_back(value());
                    ws();
                    if (cur == end) fail("unterminated array");
                    if (*cur == ',') { ++cur; continue; }
                    if (*cur == ']') { ++cur; break; }
                    fail("expected ',' or ']'");
                }
                return j;
            }
            if (c == '"') { j.kind = String; j.str = string(); return j; }
            if (lit("true")) { j.kind = Bool; j.b = true; return j; }
            if (lit("false")) { j.kind = Bool; j.b = false; return j; }
            if (lit("null")) { return j; }
            if (lit("NaN")) { j.kind = Number; j.num = std::nan(""); return j; }
            if (c == '-' || (c >= '0' && c <= '9')) {
                const char* s = cur;
                bool isint = true;
                if (*cur == '-') ++cur;
                while (cur != end && ((*cur >= '0' && *cur <= '9') || *cur == '.' || *cur == 'e' || *cur == 'E' || *cur == '+' || *cur == '-')) {
                    if (*cur == '.' || *cur == 'e' || *cur == 'E') isint = false;
                    ++cur;
                }
                std::string t(s, cur);
                char* ep = nullptr;
                j.kind = Number; j.num = std::strtod(t.c_str(), &ep); j.is_int = isint;
                if (ep == t.c_str()) fail("bad number");
                return j;
            }
            fail(std::string("unexpected character '") + c + "'");
        }
        static void utf8(std::string& out, unsigned cp) {
            if (cp < 0x80) out += char(cp);
            else if (cp < 0x800) { out += char(0xC0 | (cp >> 6)); out += char(0x80 | (cp & 0x3F)); }
            else if (cp < 0x10000) { out += char(0xE0 | (cp >> 12)); out += char(0x80 | ((cp >> 6) & 0x3F)); out += char(0x80 | (cp & 0x3F)); }
            else { out += char(0xF0 | (cp >> 18)); out += char(0x80 | ((cp >> 12) & 0x3F)); out += char(0x80 | ((cp >> 6) & 0x3F)); out += char(0x80 | (cp & 0x3F)); }
        }
        unsigned hex4() {
            unsigned v = 0;
            for (int i = 0; i < 4; ++i) {
                if (cur == end) fail("bad \\u escape");
                char c = *cur++;
                v <<= 4;
                if (c >= '0' && c <= '9') v |= c - '0';
                else if (c >= 'a' && c <= 'f') v |= c - 'a' + 10;
                else if (c >= 'A' && c <= 'F') v |= c - 'A' + 10;
                else fail("bad \\u escape");
            }
            return v;
        }
        std::string string() {
            ++cur;  // opening quote
            std::string out;
            for (;;) {
                if (cur == end) fail("unterminated string");
                char c = *cur++;
                if (c == '"') break;
                if (c != '\\') { out += c; continue; }
                if (cur == end) fail("unterminated escape");
                char e = *cur++;
                switch (e) {
                    case '"': out += '"'; break; case '\\': out += '\\'; break; case '/': out += '/'; break;
                    case 'b': out += '\b'; break; case 'f': out += '\f'; break; case 'n': out += '\n'; break;
                    case 'r': out += '\r'; break; case 't': out += '\t'; break;
                    case 'u': {
                        unsigned cp = hex4();
                        if (cp >= 0xD800 && cp < 0xDC00 && cur + 1 < end && cur[0] == '\\' && cur[1] == 'u') {
                            cur += 2; unsigned lo = hex4();
                            cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
                        }
                        utf8(out, cp); break;
                    }
                    default: fail("bad escape");
                }
            }
            return out;
        }
    };

    void dump_to(std::string& s) const {
        switch (kind) {
            case Null: s += "null"; break;
            case Bool: s += b ? "true" : "false"; break;
            case Number: { char buf[40]; if (is_int && std::fabs(num) < 9e15) std::snprintf(buf, sizeof buf, "%lld", (long long)num); else std::snprintf(buf, sizeof buf, "%.17g", num); s += buf; break; }
            case String: s += '"'; for (char c : str) { if (c == '"' || c == '\\') { s += '\\'; s += c; } else if (c == '\n') s += "\\n"; else s += c; } s += '"'; break;
            case Array: s += '['; for (size_t i = 0; i < arr.size(); ++i) { if (i) s += ','; arr[i].dump_to(s); } s += ']'; break;
            case Object: s += '{'; for (size_t i = 0; i < obj.size(); ++i) { if (i) s += ','; s += '"'; s += obj[i].first; s += "\":"; obj[i].second.dump_to(s); } s += '}'; break;
        }
    }
};

}  // namespace tq
