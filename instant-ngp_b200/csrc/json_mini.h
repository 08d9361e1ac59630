// Minimal JSON reader for the tcnn-style network configs (configs/<mode>/*.json in the reference): objects, arrays,
// strings, numbers, booleans, null, with // and /* */ comments allowed (the reference parses with
// json::parse(..., allow_exceptions=true, ignore_comments=true), testbed.cu:268) and "parent"-style merge_patch
// (testbed.cu:95-106) done by the caller via merge_patch().  Binary values (msgpack bin, used by snapshots: msgpack_mini.h)
// and an integer/float distinction for numbers are carried so that a config read from a snapshot is written back unchanged.
#pragma once
#include <map>
#include <memory>
#include <string>
#include <vector>
#include <stdexcept>
#include <cstdlib>
#include <cstdint>
#include <cctype>

namespace ngp {

struct Json {
    enum Type { Null, Bool, Number, String, Array, Object, Binary } type = Null;
    bool b = false; double num = 0; bool is_float = false; std::string str;   // str also holds the bytes of a Binary
    Json() = default;
    static Json number(double v, bool as_float = true) { Json j; j.type = Number; j.num = v; j.is_float = as_float; return j; }
    static Json integer(int64_t v) { return number((double)v, false); }
    static Json string(const std::string& v) { Json j; j.type = String; j.str = v; return j; }
    static Json boolean(bool v) { Json j; j.type = Bool; j.b = v; return j; }
    static Json binary(const void* p, size_t n) { Json j; j.type = Binary; j.str.assign((const char*)p, n); return j; }
    static Json array_of(const float* v, size_t n) { Json j; j.type = Array; for (size_t i = 0; i < n; ++i) j.arr.push_back(number(v[i])); return j; }
    std::vector<Json> arr; std::map<std::string, Json> obj;

    bool is_object() const { return type == Object; }
    bool contains(const std::string& k) const { return type == Object && obj.count(k); }
    const Json& at(const std::string& k) const {
        auto it = obj.find(k);
        if (type != Object || it == obj.end()) throw std::runtime_error("json: missing key '" + k + "'");
        return it->second;
    }
    Json& operator[](const std::string& k) { if (type != Object) { *this = Json(); type = Object; } return obj[k]; }
    double value(const std::string& k, double def) const { return contains(k) && at(k).type == Number ? at(k).num : def; }
    std::string value(const std::string& k, const char* def) const { return contains(k) && at(k).type == String ? at(k).str : std::string(def); }
    bool value(const std::string& k, bool def) const { return contains(k) && at(k).type == Bool ? at(k).b : def; }

    // RFC 7386 merge patch: objects merge recursively, null deletes, everything else replaces
    void merge_patch(const Json& patch) {
        if (patch.type != Object) { *this = patch; return; }
        if (type != Object) { *this = Json(); type = Object; }
        for (auto& kv : patch.obj) {
            if (kv.second.type == Null) obj.erase(kv.first);
            else obj[kv.first].merge_patch(kv.second);
        }
    }

    static Json parse(const std::string& s) { size_t i = 0; Json j = parse_value(s, i); skip_ws(s, i); if (i != s.size()) fail(s, i, "trailing characters"); return j; }

private:
    [[noreturn]] static void fail(const std::string&, size_t i, const char* what) { throw std::runtime_error(std::string("json parse error at offset ") + std::to_string(i) + ": " + what); }
    static void skip_ws(const std::string& s, size_t& i) {
        for (;;) {
            while (i < s.size() && std::isspace((unsigned char)s[i])) ++i;
            if (i + 1 < s.size() && s[i] == '/' && s[i + 1] == '/') { while (i < s.size() && s[i] != '\n') ++i; continue; }
            if (i + 1 < s.size() && s[i] == '/' && s[i + 1] == '*') { i += 2; while (i + 1 < s.size() && !(s[i] == '*' && s[i + 1] == '/')) ++i; i += 2; continue; }
            break;
        }
    }
    static std::string parse_string(const std::string& s, size_t& i) {
        if (s[i] != '"') fail(s, i, "expected string");
        ++i; std::string out;
        while (i < s.size() && s[i] != '"') {
            if (s[i] == '\\' && i + 1 < s.size()) {
                ++i;
                switch (s[i]) { case 'n': out += '\n'; break; case 't': out += '\t'; break; case 'r': out += '\r'; break;
                    case 'b': out += '\b'; break; case 'f': out += '\f'; break; default: out += s[i]; }
                ++i;
            } else out += s[i++];
        }
        if (i >= s.size()) fail(s, i, "unterminated string");
        ++i; return out;
    }
    static Json parse_value(const std::string& s, size_t& i) {
        skip_ws(s, i);
        if (i >= s.size()) fail(s, i, "unexpected end");
        Json j;
        char c = s[i];
        if (c == '{') {
            j.type = Object; ++i; skip_ws(s, i);
            if (i < s.size() && s[i] == '}') { ++i; return j; }
            for (;;) {
                skip_ws(s, i);
                std::string k = parse_string(s, i);
                skip_ws(s, i);
                if (i >= s.size() || s[i] != ':') fail(s, i, "expected ':'");
                ++i;
                j.obj[k] = parse_value(s, i);
                skip_ws(s, i);
                if (i < s.size() && s[i] == ',') { ++i; continue; }
                if (i < s.size() && s[i] == '}') { ++i; break; }
                fail(s, i, "expected ',' or '}'");
            }
        } else if (c == '[') {
            j.type = Array; ++i; skip_ws(s, i);
            if (i < s.size() && s[i] == ']') { ++i; return j; }
            for (;;) {
                j.arr.push_back(parse_value(s, i));
                skip_ws(s, i);
                if (i < s.size() && s[i] == ',') { ++i; continue; }
                if (i < s.size() && s[i] == ']') { ++i; break; }
                fail(s, i, "expected ',' or ']'");
            }
        } else if (c == '"') { j.type = String; j.str = parse_string(s, i);
        } else if (s.compare(i, 4, "true") == 0) { j.type = Bool; j.b = true; i += 4;
        } else if (s.compare(i, 5, "false") == 0) { j.type = Bool; j.b = false; i += 5;
        } else if (s.compare(i, 4, "null") == 0) { j.type = Null; i += 4;
        } else {
            const char* start = s.c_str() + i; char* end = nullptr;
            j.num = std::strtod(start, &end);
            if (end == start) fail(s, i, "unexpected token");
            j.type = Number; j.is_float = std::string(start, (size_t)(end - start)).find_first_of(".eEnN") != std::string::npos; i += (size_t)(end - start);
        }
        return j;
    }
};

inline std::string to_lower(std::string s) { for (auto& c : s) c = (char)std::tolower((unsigned char)c); return s; }

}  // namespace ngp
