// Minimal JSON value + recursive-descent parser + string escaper for the native API server.
// (The reference vendors nlohmann/json, src/json.hpp, 24.7k lines; the server needs only request
// parsing and a handful of fixed response shapes, so this is a ~150-line reader and an escaper.)
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <utility>
#include <vector>

namespace dljson {

struct Value {
    enum Type { NUL, BOOL, NUM, STR, ARR, OBJ } type = NUL;
    bool b = false;
    double num = 0.0;
    long long inum = 0;
    bool is_int = false;
    std::string str;
    std::vector<Value> arr;
    std::vector<std::pair<std::string, Value>> obj;

    const Value* get(const char* key) const {
        if (type != OBJ) return nullptr;
        for (const auto& kv : obj)
            if (kv.first == key) return &kv.second;
        return nullptr;
    }
    bool is_null() const { return type == NUL; }
    bool truthy() const {
        switch (type) {
            case NUL: return false;
            case BOOL: return b;
            case NUM: return num != 0.0;
            case STR: return !str.empty();
            case ARR: return !arr.empty();
            case OBJ: return !obj.empty();
        }
        return false;
    }
};

class Parser {
public:
    Parser(const char* p, size_t n) : p_(p), end_(p + n) {}
    // returns false and sets err on malformed input (trailing non-space characters included)
    bool parse(Value& out, std::string& err) {
        if (!value(out, 0)) { err = err_.empty() ? "invalid JSON" : err_; return false; }
        ws();
        if (p_ != end_) { err = "invalid JSON: trailing characters"; return false; }
        return true;
    }

private:
    const char* p_;
    const char* end_;
    std::string err_;

    void ws() { while (p_ < end_ && (*p_ == ' ' || *p_ == '\t' || *p_ == '\n' || *p_ == '\r')) p_++; }
    bool fail(const char* m) { if (err_.empty()) err_ = m; return false; }
    bool lit(const char* s) {
        size_t n = strlen(s);
        if ((size_t)(end_ - p_) < n || memcmp(p_, s, n) != 0) return fail("invalid JSON literal");
        p_ += n;
        return true;
    }
    static void put_utf8(std::string& o, uint32_t c) {
        if (c < 0x80) o.push_back((char)c);
        else if (c < 0x800) { o.push_back((char)(0xc0 | (c >> 6))); o.push_back((char)(0x80 | (c & 0x3f))); }
        else if (c < 0x10000) {
            o.push_back((char)(0xe0 | (c >> 12))); o.push_back((char)(0x80 | ((c >> 6) & 0x3f)));
            o.push_back((char)(0x80 | (c & 0x3f)));
        } else {
            o.push_back((char)(0xf0 | (c >> 18))); o.push_back((char)(0x80 | ((c >> 12) & 0x3f)));
            o.push_back((char)(0x80 | ((c >> 6) & 0x3f))); o.push_back((char)(0x80 | (c & 0x3f)));
        }
    }
    bool hex4(uint32_t& v) {
        if (end_ - p_ < 4) return fail("invalid \\u escape");
        v = 0;
        for (int i = 0; i < 4; i++) {
            char c = *p_++;
            v <<= 4;
            if (c >= '0' && c <= '9') v |= c - '0';
            else if (c >= 'a' && c <= 'f') v |= c - 'a' + 10;
            else if (c >= 'A' && c <= 'F') v |= c - 'A' + 10;
            else return fail("invalid \\u escape");
        }
        return true;
    }
    bool string(std::string& o) {
        if (p_ >= end_ || *p_ != '"') return fail("expected string");
        p_++;
        while (p_ < end_) {
            unsigned char c = (unsigned char)*p_++;
            if (c == '"') return true;
            if (c == '\\') {
                if (p_ >= end_) break;
                char e = *p_++;
                switch (e) {
                    case '"': o.push_back('"'); break;
                    case '\\': o.push_back('\\'); break;
                    case '/': o.push_back('/'); break;
                    case 'b': o.push_back('\b'); break;
                    case 'f': o.push_back('\f'); break;
                    case 'n': o.push_back('\n'); break;
                    case 'r': o.push_back('\r'); break;
                    case 't': o.push_back('\t'); break;
                    case 'u': {
                        uint32_t v;
                        if (!hex4(v)) return false;
                        if (v >= 0xd800 && v <= 0xdbff && end_ - p_ >= 6 && p_[0] == '\\' && p_[1] == 'u') {
                            const char* save = p_;
                            p_ += 2;
                            uint32_t lo;
                            if (!hex4(lo)) return false;
                            if (lo >= 0xdc00 && lo <= 0xdfff) v = 0x10000 + ((v - 0xd800) << 10) + (lo - 0xdc00);
                            else { p_ = save; v = 0xfffd; }
                        } else if (v >= 0xd800 && v <= 0xdfff) {
                            v = 0xfffd;                     // lone surrogate
                        }
                        put_utf8(o, v);
                        break;
                    }
                    default: return fail("invalid escape");
                }
            } else if (c < 0x20) {
                return fail("control character in string");
            } else {
                o.push_back((char)c);
            }
        }
        return fail("unterminated string");
    }
    bool number(Value& v) {
        const char* s = p_;
        if (p_ < end_ && *p_ == '-') p_++;
        if (p_ >= end_ || *p_ < '0' || *p_ > '9') return fail("invalid number");
        while (p_ < end_ && *p_ >= '0' && *p_ <= '9') p_++;
        bool integral = true;
        if (p_ < end_ && *p_ == '.') { integral = false; p_++; while (p_ < end_ && *p_ >= '0' && *p_ <= '9') p_++; }
        if (p_ < end_ && (*p_ == 'e' || *p_ == 'E')) {
            integral = false;
            p_++;
            if (p_ < end_ && (*p_ == '+' || *p_ == '-')) p_++;
            while (p_ < end_ && *p_ >= '0' && *p_ <= '9') p_++;
        }
        std::string t(s, p_ - s);
        v.type = Value::NUM;
        v.num = strtod(t.c_str(), nullptr);
        v.is_int = integral && t.size() <= 19;
        v.inum = v.is_int ? strtoll(t.c_str(), nullptr, 10) : (long long)v.num;
        return true;
    }
    bool value(Value& v, int depth) {
        if (depth > 64) return fail("JSON nested too deeply");
        ws();
        if (p_ >= end_) return fail("unexpected end of JSON");
        char c = *p_;
        if (c == '{') {
            p_++;
            v.type = Value::OBJ;
            ws();
            if (p_ < end_ && *p_ == '}') { p_++; return true; }
            while (true) {
                ws();
                std::string k;
                if (!string(k)) return false;
                ws();
                if (p_ >= end_ || *p_ != ':') return fail("expected ':'");
                p_++;
                Value child;
                if (!value(child, depth + 1)) return false;
                v.obj.emplace_back(std::move(k), std::move(child));
                ws();
                if (p_ < end_ && *p_ == ',') { p_++; continue; }
                if (p_ < end_ && *p_ == '}') { p_++; return true; }
                return fail("expected ',' or '}'");
            }
        }
        if (c == '[') {
            p_++;
            v.type = Value::ARR;
            ws();
            if (p_ < end_ && *p_ == ']') { p_++; return true; }
            while (true) {
                Value child;
                if (!value(child, depth + 1)) return false;
                v.arr.push_back(std::move(child));
                ws();
                if (p_ < end_ && *p_ == ',') { p_++; continue; }
                if (p_ < end_ && *p_ == ']') { p_++; return true; }
                return fail("expected ',' or ']'");
            }
        }
        if (c == '"') { v.type = Value::STR; return string(v.str); }
        if (c == 't') { v.type = Value::BOOL; v.b = true; return lit("true"); }
        if (c == 'f') { v.type = Value::BOOL; v.b = false; return lit("false"); }
        if (c == 'n') { v.type = Value::NUL; return lit("null"); }
        return number(v);
    }
};

// appends `s` (UTF-8) as a quoted JSON string; invalid UTF-8 bytes become U+FFFD so the output is
// always valid JSON (the reference's json.dump throws on them)
inline void append_quoted(std::string& o, const char* s, size_t n) {
    o.push_back('"');
    size_t i = 0;
    while (i < n) {
        unsigned char c = (unsigned char)s[i];
        if (c == '"') { o += "\\\""; i++; }
        else if (c == '\\') { o += "\\\\"; i++; }
        else if (c == '\n') { o += "\\n"; i++; }
        else if (c == '\r') { o += "\\r"; i++; }
        else if (c == '\t') { o += "\\t"; i++; }
        else if (c < 0x20) { char b[8]; snprintf(b, sizeof b, "\\u%04x", c); o += b; i++; }
        else if (c < 0x80) { o.push_back((char)c); i++; }
        else {
            int need = (c >= 0xc2 && c <= 0xdf) ? 2 : (c >= 0xe0 && c <= 0xef) ? 3 : (c >= 0xf0 && c <= 0xf4) ? 4 : 0;
            bool ok = need != 0 && i + need <= n;
            for (int k = 1; ok && k < need; k++) ok = ((unsigned char)s[i + k] & 0xc0) == 0x80;
            if (ok) { o.append(s + i, need); i += need; }
            else { o += "\xef\xbf\xbd"; i++; }
        }
    }
    o.push_back('"');
}
inline void append_quoted(std::string& o, const std::string& s) { append_quoted(o, s.data(), s.size()); }

// serialises a Value (numbers: integers exactly, the rest with 17 significant digits; non-finite -> null)
inline void dump(const Value& v, std::string& o) {
    switch (v.type) {
        case Value::NUL: o += "null"; break;
        case Value::BOOL: o += v.b ? "true" : "false"; break;
        case Value::NUM: {
            char b[40];
            if (v.is_int) snprintf(b, sizeof b, "%lld", v.inum);
            else if (!std::isfinite(v.num)) snprintf(b, sizeof b, "null");
            else snprintf(b, sizeof b, "%.17g", v.num);
            o += b;
            break;
        }
        case Value::STR: append_quoted(o, v.str); break;
        case Value::ARR:
            o.push_back('[');
            for (size_t i = 0; i < v.arr.size(); i++) { if (i) o.push_back(','); dump(v.arr[i], o); }
            o.push_back(']');
            break;
        case Value::OBJ:
            o.push_back('{');
            for (size_t i = 0; i < v.obj.size(); i++) {
                if (i) o.push_back(',');
                append_quoted(o, v.obj[i].first);
                o.push_back(':');
                dump(v.obj[i].second, o);
            }
            o.push_back('}');
            break;
    }
}

}  // namespace dljson
