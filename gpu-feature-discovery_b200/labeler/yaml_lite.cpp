#include "yaml_lite.hpp"

#include <cctype>
#include <cerrno>
#include <cmath>
#include <cstdlib>
#include <cstring>

namespace yamllite {

const Value *Value::get(const std::string &key) const {
    for (const auto &kv : obj)
        if (kv.first == key) return &kv.second;
    return nullptr;
}

const Value *Value::get_fold(const std::string &key) const {
    if (const Value *v = get(key)) return v;
    for (const auto &kv : obj) {
        if (kv.first.size() != key.size()) continue;
        bool eq = true;
        for (size_t c = 0; c < key.size() && eq; c++)
            eq = std::tolower((unsigned char)kv.first[c]) == std::tolower((unsigned char)key[c]);
        if (eq) return &kv.second;
    }
    return nullptr;
}

namespace {

[[noreturn]] void fail(const std::string &m) { throw ParseError("yaml: " + m); }

bool all_digits(const std::string &s, size_t from) {
    if (from >= s.size()) return false;
    for (size_t i = from; i < s.size(); i++)
        if (!(std::isdigit((unsigned char)s[i]) || s[i] == '_')) return false;
    return true;
}

// yaml.v2 resolve(): plain scalar -> typed value
Value resolve_plain(const std::string &s) {
    Value v;
    if (s.empty() || s == "~" || s == "null" || s == "Null" || s == "NULL") return v;
    static const char *T[] = {"y", "Y", "yes", "Yes", "YES", "true", "True", "TRUE", "on", "On", "ON"};
    static const char *F[] = {"n", "N", "no", "No", "NO", "false", "False", "FALSE", "off", "Off", "OFF"};
    for (const char *t : T) if (s == t) { v.kind = Value::Bool; v.b = true; return v; }
    for (const char *f : F) if (s == f) { v.kind = Value::Bool; v.b = false; return v; }
    size_t st = (s[0] == '-' || s[0] == '+') ? 1 : 0;
    std::string plain;
    for (char c : s) if (c != '_') plain += c;
    if (all_digits(s, st)) {
        errno = 0;
        char *end = nullptr;
        long long x = strtoll(plain.c_str(), &end, 10);
        if (errno == 0 && end && *end == 0) { v.kind = Value::Int; v.i = x; v.s = s; return v; }
    }
    if (s.size() > st + 2 && s[st] == '0' && (s[st + 1] == 'x' || s[st + 1] == 'X')) {
        errno = 0;
        char *end = nullptr;
        long long x = strtoll(plain.c_str(), &end, 16);
        if (errno == 0 && end && *end == 0) { v.kind = Value::Int; v.i = x; v.s = s; return v; }
    }
    // float: needs a digit and only float characters
    bool digit = false, ok = true;
    for (size_t i = st; i < s.size() && ok; i++) {
        char c = s[i];
        if (std::isdigit((unsigned char)c)) digit = true;
        else if (!(c == '.' || c == 'e' || c == 'E' || c == '-' || c == '+' || c == '_')) ok = false;
    }
    if (ok && digit) {
        char *end = nullptr;
        double d = strtod(plain.c_str(), &end);
        if (end && *end == 0) { v.kind = Value::Float; v.f = d; v.s = s; return v; }
    }
    if (s == ".inf" || s == ".Inf" || s == ".INF" || s == "+.inf") { v.kind = Value::Float; v.f = INFINITY; v.s = s; return v; }
    if (s == "-.inf" || s == "-.Inf" || s == "-.INF") { v.kind = Value::Float; v.f = -INFINITY; v.s = s; return v; }
    if (s == ".nan" || s == ".NaN" || s == ".NAN") { v.kind = Value::Float; v.f = NAN; v.s = s; return v; }
    v.kind = Value::String;
    v.s = s;
    return v;
}

void append_utf8(std::string &o, uint32_t cp) {
    if (cp < 0x80) o += (char)cp;
    else if (cp < 0x800) { o += (char)(0xC0 | (cp >> 6)); o += (char)(0x80 | (cp & 0x3F)); }
    else if (cp < 0x10000) { o += (char)(0xE0 | (cp >> 12)); o += (char)(0x80 | ((cp >> 6) & 0x3F)); o += (char)(0x80 | (cp & 0x3F)); }
    else { o += (char)(0xF0 | (cp >> 18)); o += (char)(0x80 | ((cp >> 12) & 0x3F)); o += (char)(0x80 | ((cp >> 6) & 0x3F)); o += (char)(0x80 | (cp & 0x3F)); }
}

// ---- flow / JSON parser over a character range ----
struct Flow {
    const std::string &t;
    size_t p;
    explicit Flow(const std::string &text, size_t pos = 0) : t(text), p(pos) {}

    void ws() {
        while (p < t.size()) {
            char c = t[p];
            if (c == ' ' || c == '\t' || c == '\n' || c == '\r') p++;
            else if (c == '#' && (p == 0 || std::isspace((unsigned char)t[p - 1]))) { while (p < t.size() && t[p] != '\n') p++; }
            else break;
        }
    }

    std::string dquoted() {
        std::string o;
        p++;  // opening quote
        while (true) {
            if (p >= t.size()) fail("unterminated double-quoted scalar");
            char c = t[p++];
            if (c == '"') break;
            if (c != '\\') { o += c; continue; }
            if (p >= t.size()) fail("bad escape");
            char e = t[p++];
            switch (e) {
            case 'n': o += '\n'; break;
            case 't': o += '\t'; break;
            case 'r': o += '\r'; break;
            case 'b': o += '\b'; break;
            case 'f': o += '\f'; break;
            case '0': o += '\0'; break;
            case '/': o += '/'; break;
            case '\\': o += '\\'; break;
            case '"': o += '"'; break;
            case ' ': o += ' '; break;
            case 'x': case 'u': case 'U': {
                int n = e == 'x' ? 2 : e == 'u' ? 4 : 8;
                if (p + n > t.size()) fail("bad escape");
                uint32_t cp = (uint32_t)strtoul(t.substr(p, n).c_str(), nullptr, 16);
                p += n;
                if (e == 'u' && cp >= 0xD800 && cp <= 0xDBFF && p + 6 <= t.size() && t[p] == '\\' && t[p + 1] == 'u') {
                    uint32_t lo = (uint32_t)strtoul(t.substr(p + 2, 4).c_str(), nullptr, 16);
                    if (lo >= 0xDC00 && lo <= 0xDFFF) { cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00); p += 6; }
                }
                append_utf8(o, cp);
                break;
            }
            default: fail(std::string("unknown escape \\") + e);
            }
        }
        return o;
    }

    std::string squoted() {
        std::string o;
        p++;
        while (true) {
            if (p >= t.size()) fail("unterminated single-quoted scalar");
            char c = t[p++];
            if (c == '\'') {
                if (p < t.size() && t[p] == '\'') { o += '\''; p++; continue; }
                break;
            }
            o += c;
        }
        return o;
    }

    Value value() {
        ws();
        if (p >= t.size()) fail("unexpected end of input");
        char c = t[p];
        Value v;
        if (c == '{') {
            p++;
            v.kind = Value::Object;
            ws();
            if (p < t.size() && t[p] == '}') { p++; return v; }
            while (true) {
                ws();
                std::string key;
                if (p < t.size() && t[p] == '"') key = dquoted();
                else if (p < t.size() && t[p] == '\'') key = squoted();
                else {
                    size_t s = p;
                    while (p < t.size() && t[p] != ':' && t[p] != ',' && t[p] != '}' && t[p] != '\n') p++;
                    key = t.substr(s, p - s);
                    while (!key.empty() && std::isspace((unsigned char)key.back())) key.pop_back();
                }
                ws();
                if (p >= t.size() || t[p] != ':') fail("expected ':' in flow mapping");
                p++;
                Value val = value();
                bool dup = false;
                for (auto &kv : v.obj) if (kv.first == key) { kv.second = val; dup = true; }
                if (!dup) v.obj.emplace_back(key, val);
                ws();
                if (p < t.size() && t[p] == ',') { p++; ws(); if (p < t.size() && t[p] == '}') { p++; break; } continue; }
                if (p < t.size() && t[p] == '}') { p++; break; }
                fail("expected ',' or '}' in flow mapping");
            }
            return v;
        }
        if (c == '[') {
            p++;
            v.kind = Value::Array;
            ws();
            if (p < t.size() && t[p] == ']') { p++; return v; }
            while (true) {
                v.arr.push_back(value());
                ws();
                if (p < t.size() && t[p] == ',') { p++; ws(); if (p < t.size() && t[p] == ']') { p++; break; } continue; }
                if (p < t.size() && t[p] == ']') { p++; break; }
                fail("expected ',' or ']' in flow sequence");
            }
            return v;
        }
        if (c == '"') { v.kind = Value::String; v.s = dquoted(); v.quoted = true; return v; }
        if (c == '\'') { v.kind = Value::String; v.s = squoted(); v.quoted = true; return v; }
        size_t s = p;
        while (p < t.size() && t[p] != ',' && t[p] != '}' && t[p] != ']' && t[p] != '\n') {
            if (t[p] == ':' && (p + 1 >= t.size() || t[p + 1] == ' ')) break;
            if (t[p] == '#' && p > s && std::isspace((unsigned char)t[p - 1])) break;
            p++;
        }
        std::string plain = t.substr(s, p - s);
        while (!plain.empty() && std::isspace((unsigned char)plain.back())) plain.pop_back();
        return resolve_plain(plain);
    }
};

// ---- block parser ----
struct Line {
    int indent;
    std::string text;  // without indentation, comments and trailing blanks
};

// strip a trailing comment that is outside quotes
std::string strip_comment(const std::string &s) {
    bool in_s = false, in_d = false;
    for (size_t i = 0; i < s.size(); i++) {
        char c = s[i];
        if (in_d) { if (c == '\\') i++; else if (c == '"') in_d = false; continue; }
        if (in_s) { if (c == '\'') in_s = false; continue; }
        if (c == '"') in_d = true;
        else if (c == '\'') in_s = true;
        else if (c == '#' && (i == 0 || s[i - 1] == ' ' || s[i - 1] == '\t')) return s.substr(0, i);
    }
    return s;
}

struct Block {
    std::vector<Line> lines;
    size_t pos = 0;

    Value scalar_or_flow(const std::string &text) {
        if (text.empty()) return Value();
        char c = text[0];
        if (c == '&' || c == '*' || c == '!' || c == '|' || c == '>') fail("unsupported YAML feature: " + text);
        Flow f(text);
        if (c == '[' || c == '{' || c == '"' || c == '\'') {
            Value v = f.value();
            f.ws();
            if (f.p != text.size()) fail("trailing characters after value: " + text);
            return v;
        }
        return resolve_plain(text);
    }

    // find "key:" split honoring quotes; returns false if the line is not a mapping entry
    static bool split_key(const std::string &t, std::string *key, std::string *rest) {
        size_t i = 0;
        if (!t.empty() && (t[0] == '"' || t[0] == '\'')) {
            Flow f(t);
            *key = t[0] == '"' ? f.dquoted() : f.squoted();
            i = f.p;
            while (i < t.size() && t[i] == ' ') i++;
            if (i >= t.size() || t[i] != ':') return false;
        } else {
            if (!t.empty() && (t[0] == '[' || t[0] == '{')) return false;
            for (; i < t.size(); i++)
                if (t[i] == ':' && (i + 1 == t.size() || t[i + 1] == ' ' || t[i + 1] == '\t')) break;
            if (i >= t.size()) return false;
            *key = t.substr(0, i);
            while (!key->empty() && std::isspace((unsigned char)key->back())) key->pop_back();
        }
        *rest = t.substr(i + 1);
        size_t b = 0;
        while (b < rest->size() && (rest->at(b) == ' ' || rest->at(b) == '\t')) b++;
        *rest = rest->substr(b);
        return true;
    }

    Value parse_node(int indent) {
        if (pos >= lines.size()) return Value();
        const Line &l = lines[pos];
        if (l.text == "-" || l.text.compare(0, 2, "- ") == 0) return parse_seq(l.indent);
        std::string k, r;
        if (split_key(l.text, &k, &r)) return parse_map(l.indent);
        (void)indent;
        Value v = scalar_or_flow(l.text);
        pos++;
        return v;
    }

    Value parse_seq(int indent) {
        Value v;
        v.kind = Value::Array;
        while (pos < lines.size() && lines[pos].indent == indent &&
               (lines[pos].text == "-" || lines[pos].text.compare(0, 2, "- ") == 0)) {
            std::string rest = lines[pos].text.size() > 1 ? lines[pos].text.substr(2) : "";
            size_t lead = 0;
            while (lead < rest.size() && rest[lead] == ' ') lead++;
            rest = rest.substr(lead);
            if (rest.empty()) {
                pos++;
                if (pos < lines.size() && lines[pos].indent > indent) v.arr.push_back(parse_node(lines[pos].indent));
                else v.arr.push_back(Value());
                continue;
            }
            std::string k, r;
            if (rest == "-" || rest.compare(0, 2, "- ") == 0 || split_key(rest, &k, &r)) {
                // nested collection starting on the dash line: re-home the remainder as a virtual line
                lines[pos].indent = indent + 2 + (int)lead;
                lines[pos].text = rest;
                v.arr.push_back(parse_node(lines[pos].indent));
                continue;
            }
            v.arr.push_back(scalar_or_flow(rest));
            pos++;
        }
        return v;
    }

    Value parse_map(int indent) {
        Value v;
        v.kind = Value::Object;
        while (pos < lines.size() && lines[pos].indent == indent) {
            std::string key, rest;
            if (lines[pos].text == "-" || lines[pos].text.compare(0, 2, "- ") == 0) break;
            if (!split_key(lines[pos].text, &key, &rest)) fail("expected 'key: value' at: " + lines[pos].text);
            pos++;
            Value val;
            if (!rest.empty()) {
                val = scalar_or_flow(rest);
            } else if (pos < lines.size() &&
                       (lines[pos].indent > indent ||
                        (lines[pos].indent == indent && (lines[pos].text == "-" || lines[pos].text.compare(0, 2, "- ") == 0)))) {
                val = parse_node(lines[pos].indent);
            }
            bool dup = false;
            for (auto &kv : v.obj) if (kv.first == key) { kv.second = val; dup = true; }
            if (!dup) v.obj.emplace_back(key, val);
        }
        if (pos < lines.size() && lines[pos].indent > indent) fail("bad indentation at: " + lines[pos].text);
        return v;
    }
};

}  // namespace

Value parse(const std::string &text) {
    // whole-document flow (JSON) form
    size_t first = 0;
    while (first < text.size() && std::isspace((unsigned char)text[first])) first++;
    if (first < text.size() && (text[first] == '{' || text[first] == '[')) {
        Flow f(text, first);
        Value v = f.value();
        f.ws();
        if (f.p != text.size()) fail("trailing characters after document");
        return v;
    }
    Block b;
    size_t i = 0;
    while (i <= text.size()) {
        size_t e = text.find('\n', i);
        if (e == std::string::npos) e = text.size();
        std::string raw = text.substr(i, e - i);
        i = e + 1;
        if (!raw.empty() && raw.back() == '\r') raw.pop_back();
        std::string s = strip_comment(raw);
        while (!s.empty() && (s.back() == ' ' || s.back() == '\t')) s.pop_back();
        size_t ind = 0;
        while (ind < s.size() && s[ind] == ' ') ind++;
        if (ind < s.size() && s[ind] == '\t') fail("tabs are not allowed for indentation");
        std::string body = s.substr(ind);
        if (body.empty()) { if (e == text.size()) break; continue; }
        if (body == "---") { if (!b.lines.empty()) fail("multiple documents are not supported"); continue; }
        if (body == "...") break;
        b.lines.push_back({(int)ind, body});
        if (e == text.size()) break;
    }
    if (b.lines.empty()) return Value();
    Value v = b.parse_node(b.lines[0].indent);
    if (b.pos != b.lines.size()) fail("could not parse line: " + b.lines[b.pos].text);
    return v;
}

}  // namespace yamllite
