#include "yaml_lite.hpp"

#include <cctype>
#include <cerrno>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>

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

// strconv.ParseInt(s, 0, 64) / ParseUint: optional sign (ParseInt only), then 0x / 0b / 0o prefixes, a leading 0 = octal,
// else decimal; at least one digit; every digit valid for the base; no overflow.  (Underscores are gone by now: yaml.v2
// strips them before it calls ParseInt, resolve.go:147.)
bool go_parse_int(const std::string &s, bool allow_sign, bool *negative, unsigned long long *magnitude) {
    size_t i = 0;
    *negative = false;
    if (allow_sign && i < s.size() && (s[i] == '+' || s[i] == '-')) { *negative = s[i] == '-'; i++; }
    int base = 10;
    if (i + 1 < s.size() && s[i] == '0') {
        char p = s[i + 1];
        if (p == 'x' || p == 'X') { base = 16; i += 2; }
        else if (p == 'b' || p == 'B') { base = 2; i += 2; }
        else if (p == 'o' || p == 'O') { base = 8; i += 2; }
        else { base = 8; i += 1; }
    }
    if (i >= s.size()) return false;
    unsigned long long acc = 0;
    for (; i < s.size(); i++) {
        char c = s[i];
        int d = (c >= '0' && c <= '9') ? c - '0' : (c >= 'a' && c <= 'f') ? c - 'a' + 10 : (c >= 'A' && c <= 'F') ? c - 'A' + 10 : 99;
        if (d >= base) return false;
        if (acc > (~0ull - (unsigned)d) / (unsigned)base) return false;  // overflow of uint64
        acc = acc * (unsigned)base + (unsigned)d;
    }
    *magnitude = acc;
    return true;
}

// yamlStyleFloat (resolve.go:84): ^[-+]?(\.[0-9]+|[0-9]+(\.[0-9]*)?)([eE][-+]?[0-9]+)?$
bool yaml_style_float(const std::string &s) {
    size_t i = 0, n = s.size();
    if (i < n && (s[i] == '+' || s[i] == '-')) i++;
    size_t d0 = i;
    while (i < n && std::isdigit((unsigned char)s[i])) i++;
    if (i > d0) {
        if (i < n && s[i] == '.') { i++; while (i < n && std::isdigit((unsigned char)s[i])) i++; }
    } else {
        if (i >= n || s[i] != '.') return false;
        i++;
        size_t f0 = i;
        while (i < n && std::isdigit((unsigned char)s[i])) i++;
        if (i == f0) return false;
    }
    if (i < n && (s[i] == 'e' || s[i] == 'E')) {
        i++;
        if (i < n && (s[i] == '+' || s[i] == '-')) i++;
        size_t e0 = i;
        while (i < n && std::isdigit((unsigned char)s[i])) i++;
        if (i == e0) return false;
    }
    return i == n;
}

// yaml.v2 parseTimestamp (resolve.go:224-258): an unquoted scalar that starts "dddd-" and parses under one of four Go layouts
// ("2006-1-2T15:4:5.999999999Z07:00", the same with 't', "2006-1-2 15:4:5.999999999", "2006-1-2") is a time value; on its way
// through sigs.k8s.io/yaml it becomes the JSON string time.Time marshals to (RFC 3339, nanoseconds trimmed).  Returns false if
// `s` is not such a timestamp.
bool go_timestamp(const std::string &s, std::string *rfc3339) {
    size_t i = 0;
    while (i < s.size() && std::isdigit((unsigned char)s[i])) i++;
    if (i != 4 || i == s.size() || s[i] != '-') return false;
    auto num = [&](size_t &p, int min_digits, int max_digits, int *out) {
        int n = 0, v = 0;
        while (p < s.size() && n < max_digits && std::isdigit((unsigned char)s[p])) { v = v * 10 + (s[p] - '0'); p++; n++; }
        *out = v;
        return n >= min_digits;
    };
    size_t p = 0;
    int year, month, day, hour = 0, minute = 0, second = 0;
    std::string frac, zone = "Z";
    if (!num(p, 4, 4, &year) || p >= s.size() || s[p++] != '-') return false;
    if (!num(p, 1, 2, &month) || p >= s.size() || s[p++] != '-') return false;
    if (!num(p, 1, 2, &day)) return false;
    if (p < s.size()) {
        const char sep = s[p++];
        if (sep != 'T' && sep != 't' && sep != ' ') return false;
        if (!num(p, 1, 2, &hour) || p >= s.size() || s[p++] != ':') return false;
        if (!num(p, 1, 2, &minute) || p >= s.size() || s[p++] != ':') return false;
        if (!num(p, 1, 2, &second)) return false;
        if (p < s.size() && s[p] == '.' && p + 1 < s.size() && std::isdigit((unsigned char)s[p + 1])) {
            p++;
            while (p < s.size() && std::isdigit((unsigned char)s[p])) frac += s[p++];
            if (frac.size() > 9) frac.resize(9);  // time.Parse keeps nanoseconds
            while (!frac.empty() && frac.back() == '0') frac.pop_back();
        }
        if (sep == ' ') {
            if (p != s.size()) return false;  // the space form carries no zone
        } else {
            if (p >= s.size()) return false;  // the T forms need one
            if (s[p] == 'Z') {
                p++;
            } else if (s[p] == '+' || s[p] == '-') {
                const char sign = s[p++];
                int zh, zm;
                size_t q = p;
                if (!num(q, 2, 2, &zh) || q >= s.size() || s[q++] != ':' || !num(q, 2, 2, &zm)) return false;
                if (zh > 24 || zm > 60) return false;
                p = q;
                if (zh != 0 || zm != 0) {
                    char buf[32];
                    snprintf(buf, sizeof buf, "%c%02d:%02d", sign, zh, zm);
                    zone = buf;
                }
            } else {
                return false;
            }
            if (p != s.size()) return false;
        }
    }
    static const int kDays[] = {31, 28, 31, 30, 31, 30, 31, 31, 30, 31, 30, 31};
    const bool leap = (year % 4 == 0 && year % 100 != 0) || year % 400 == 0;
    if (month < 1 || month > 12 || day < 1 || day > kDays[month - 1] + ((month == 2 && leap) ? 1 : 0)) return false;
    if (hour > 23 || minute > 59 || second > 59) return false;
    char buf[96];
    snprintf(buf, sizeof buf, "%04d-%02d-%02dT%02d:%02d:%02d", year, month, day, hour, minute, second);
    *rfc3339 = std::string(buf) + (frac.empty() ? "" : "." + frac) + zone;
    return true;
}

// yaml.v2 resolve() (resolve.go:86-207) for a plain scalar with no tag: the fixed map of null / bool / inf / nan words, then by
// first character: a digit or a sign -> int (Go base-0 syntax, so 017 is 15 and 0b11 is 3), unsigned, float by yamlStyleFloat;
// a dot -> float; anything else a string.  Timestamps (2001-12-14 ...) come out as the RFC 3339 string the reference's
// YAML -> JSON step turns them into.
Value resolve_plain(const std::string &s) {
    Value v;
    if (s.empty() || s == "~" || s == "null" || s == "Null" || s == "NULL") return v;
    static const char *T[] = {"y", "Y", "yes", "Yes", "YES", "true", "True", "TRUE", "on", "On", "ON"};
    static const char *F[] = {"n", "N", "no", "No", "NO", "false", "False", "FALSE", "off", "Off", "OFF"};
    for (const char *t : T) if (s == t) { v.kind = Value::Bool; v.b = true; return v; }
    for (const char *f : F) if (s == f) { v.kind = Value::Bool; v.b = false; return v; }
    if (s == ".inf" || s == ".Inf" || s == ".INF" || s == "+.inf" || s == "+.Inf" || s == "+.INF") { v.kind = Value::Float; v.f = INFINITY; v.s = s; return v; }
    if (s == "-.inf" || s == "-.Inf" || s == "-.INF") { v.kind = Value::Float; v.f = -INFINITY; v.s = s; return v; }
    if (s == ".nan" || s == ".NaN" || s == ".NAN") { v.kind = Value::Float; v.f = NAN; v.s = s; return v; }
    auto as_string = [&] { v.kind = Value::String; v.s = s; return v; };
    auto as_float = [&](const std::string &text) {
        char *end = nullptr;
        errno = 0;
        double d = strtod(text.c_str(), &end);
        if (!end || *end != 0) return false;
        v.kind = Value::Float;
        v.f = d;
        v.s = s;
        return true;
    };
    const char c0 = s[0];
    if (c0 == '.') {  // hint '.': strconv.ParseFloat(in) on the text as it stands
        if (yaml_style_float(s) && as_float(s)) return v;
        return as_string();
    }
    if (!(std::isdigit((unsigned char)c0) || c0 == '+' || c0 == '-')) return as_string();
    std::string stamp;
    if (go_timestamp(s, &stamp)) {
        v.kind = Value::String;
        v.s = stamp;
        v.quoted = true;  // a JSON string from here on: no further coercion
        return v;
    }
    std::string plain;
    for (char c : s) if (c != '_') plain += c;
    bool neg = false;
    unsigned long long mag = 0;
    if (go_parse_int(plain, true, &neg, &mag) && (neg ? mag <= 9223372036854775808ull : mag <= 9223372036854775807ull)) {
        v.kind = Value::Int;
        v.i = neg ? (int64_t)(0ull - mag) : (int64_t)mag;
        v.s = s;
        return v;
    }
    if (go_parse_int(plain, false, &neg, &mag)) {  // ParseUint: above int64 -- kept as a float here (no field of the config takes one)
        v.kind = Value::Float;
        v.f = (double)mag;
        v.s = s;
        return v;
    }
    if (yaml_style_float(plain) && as_float(plain)) return v;
    return as_string();
}

// A mapping key ends up as a JSON object key: yaml.v2 resolves an unquoted key like any plain scalar and sigs.k8s.io/yaml turns
// the typed key back into a string (yaml.go:176-206): 010 becomes "8", yes becomes "true"; a null key is refused there.
// (Float keys keep their spelling here; the reference would print them with strconv 'g'.)
std::string key_string(const std::string &raw, bool quoted) {
    if (quoted) return raw;
    Value v = resolve_plain(raw);
    switch (v.kind) {
    case Value::Int: return std::to_string(v.i);
    case Value::Bool: return v.b ? "true" : "false";
    case Value::Null: fail("unsupported map key of type null: " + raw);
    default: return raw;
    }
}

void append_utf8(std::string &o, uint32_t cp) {
    if (cp < 0x80) o += (char)cp;
    else if (cp < 0x800) { o += (char)(0xC0 | (cp >> 6)); o += (char)(0x80 | (cp & 0x3F)); }
    else if (cp < 0x10000) { o += (char)(0xE0 | (cp >> 12)); o += (char)(0x80 | ((cp >> 6) & 0x3F)); o += (char)(0x80 | (cp & 0x3F)); }
    else { o += (char)(0xF0 | (cp >> 18)); o += (char)(0x80 | ((cp >> 12) & 0x3F)); o += (char)(0x80 | ((cp >> 6) & 0x3F)); o += (char)(0x80 | (cp & 0x3F)); }
}

using Anchors = std::map<std::string, Value>;

// What a tag does to a SCALAR (yaml.v2 resolve(): an explicit core tag re-types the text; any other tag leaves a string).
// `text` is the scalar's content (quotes already removed); collections keep their shape whatever the tag says.
Value apply_scalar_tag(const std::string &tag, const std::string &text, Value v) {
    if (tag.empty() || v.kind == Value::Array || v.kind == Value::Object) return v;
    if (tag == "!!int" || tag == "!!float" || tag == "!!bool" || tag == "!!null") {
        Value r = resolve_plain(text);
        if (tag == "!!float" && r.kind == Value::Int) { r.kind = Value::Float; r.f = (double)r.i; }  // yaml.v2 resolve.go: int under !!float
        return r;
    }
    if (tag == "!!binary") fail("unsupported YAML feature: !!binary");
    Value out;  // !!str, !local, !<verbatim>: a string
    out.kind = Value::String;
    out.s = text;
    out.quoted = true;
    return out;
}

// ---- flow / JSON parser over a character range ----
struct Flow {
    const std::string &t;
    size_t p;
    Anchors *anchors;
    explicit Flow(const std::string &text, size_t pos = 0, Anchors *a = nullptr) : t(text), p(pos), anchors(a) {}

    std::string token() {  // an anchor / alias / tag name: up to a blank or a flow indicator
        size_t s = p;
        while (p < t.size() && !std::isspace((unsigned char)t[p]) && t[p] != ',' && t[p] != ']' && t[p] != '}' && t[p] != '[' && t[p] != '{') p++;
        return t.substr(s, p - s);
    }

    void ws() {
        while (p < t.size()) {
            char c = t[p];
            if (c == ' ' || c == '\t' || c == '\n' || c == '\r') p++;
            else if (c == '#' && (p == 0 || std::isspace((unsigned char)t[p - 1]))) { while (p < t.size() && t[p] != '\n') p++; }
            else break;
        }
    }

    // A line break inside a quoted scalar (YAML line folding): white space around it goes, one break reads as a space,
    // k + 1 breaks in a row (k empty lines) read as k newlines.  p stands just after the first '\n'.
    void fold_break(std::string &o) {
        while (!o.empty() && (o.back() == ' ' || o.back() == '\t')) o.pop_back();
        size_t breaks = 1;
        while (p < t.size() && (t[p] == ' ' || t[p] == '\t' || t[p] == '\n' || t[p] == '\r')) {
            if (t[p] == '\n') breaks++;
            p++;
        }
        if (breaks == 1) o += ' ';
        else o.append(breaks - 1, '\n');
    }

    std::string dquoted() {
        std::string o;
        p++;  // opening quote
        while (true) {
            if (p >= t.size()) fail("unterminated double-quoted scalar");
            char c = t[p++];
            if (c == '"') break;
            if (c == '\r' && p < t.size() && t[p] == '\n') continue;
            if (c == '\n') { fold_break(o); continue; }
            if (c != '\\') { o += c; continue; }
            if (p >= t.size()) fail("bad escape");
            char e = t[p++];
            if (e == '\r' && p < t.size() && t[p] == '\n') e = t[p++];
            if (e == '\n') {  // escaped line break: the lines join with nothing in between, leading blanks of the next line go
                while (p < t.size() && (t[p] == ' ' || t[p] == '\t')) p++;
                continue;
            }
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
            if (c == '\r' && p < t.size() && t[p] == '\n') continue;
            if (c == '\n') { fold_break(o); continue; }
            o += c;
        }
        return o;
    }

    Value value() {
        ws();
        if (p >= t.size()) fail("unexpected end of input");
        if (t[p] == '*') {  // alias
            p++;
            std::string name = token();
            if (!anchors || !anchors->count(name)) fail("unknown anchor '" + name + "' referenced");
            return (*anchors)[name];
        }
        if (t[p] == '&' || t[p] == '!') {  // node properties, then the node
            std::string anchor, tag;
            while (p < t.size() && (t[p] == '&' || t[p] == '!')) {
                bool is_anchor = t[p] == '&';
                std::string tok = token();
                if (is_anchor) anchor = tok.substr(1); else tag = tok;
                ws();
            }
            Value v;
            if (p < t.size() && t[p] != ',' && t[p] != ']' && t[p] != '}') {
                size_t s = p;
                v = value();
                std::string text = v.quoted || v.kind == Value::String ? v.s : t.substr(s, p - s);
                while (!text.empty() && std::isspace((unsigned char)text.back())) text.pop_back();
                v = apply_scalar_tag(tag, text, v);
            } else {
                v = apply_scalar_tag(tag, "", v);
            }
            if (!anchor.empty() && anchors) (*anchors)[anchor] = v;
            return v;
        }
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
                bool key_quoted = true;
                if (p < t.size() && t[p] == '"') key = dquoted();
                else if (p < t.size() && t[p] == '\'') key = squoted();
                else {
                    size_t s = p;
                    while (p < t.size() && t[p] != ':' && t[p] != ',' && t[p] != '}' && t[p] != '\n') p++;
                    key = t.substr(s, p - s);
                    while (!key.empty() && std::isspace((unsigned char)key.back())) key.pop_back();
                    key_quoted = false;
                }
                ws();
                if (p >= t.size() || t[p] != ':') fail("expected ':' in flow mapping");
                p++;
                Value val = value();
                if (key == "<<" && !key_quoted) {  // merge key, applied where it stands (yaml.v2 decode.go:660-664)
                    std::vector<const Value *> srcs;
                    if (val.kind == Value::Object) srcs.push_back(&val);
                    else if (val.kind == Value::Array) for (size_t i = val.arr.size(); i-- > 0;) srcs.push_back(&val.arr[i]);
                    else fail("map merge requires map or sequence of maps as the value");
                    for (const Value *src : srcs) {
                        if (src->kind != Value::Object) fail("map merge requires map or sequence of maps as the value");
                        for (const auto &m : src->obj) {
                            bool dup = false;
                            for (auto &kv : v.obj) if (kv.first == m.first) { kv.second = m.second; dup = true; }
                            if (!dup) v.obj.emplace_back(m.first, m.second);
                        }
                    }
                } else {
                    const std::string ks = key_string(key, key_quoted);
                    bool dup = false;
                    for (auto &kv : v.obj) if (kv.first == ks) { kv.second = val; dup = true; }
                    if (!dup) v.obj.emplace_back(ks, val);
                }
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
        // plain scalar in flow context; it may continue over line breaks (folded: one break = a space, k empty lines = k newlines)
        std::string plain;
        bool multi = false;
        while (true) {
            size_t s = p;
            while (p < t.size() && t[p] != ',' && t[p] != '}' && t[p] != ']' && t[p] != '\n') {
                if (t[p] == ':' && (p + 1 >= t.size() || t[p + 1] == ' ' || t[p + 1] == '\n')) break;
                if (t[p] == '#' && p > s && std::isspace((unsigned char)t[p - 1])) break;
                p++;
            }
            std::string seg = t.substr(s, p - s);
            while (!seg.empty() && std::isspace((unsigned char)seg.back())) seg.pop_back();
            plain += seg;
            if (p >= t.size() || t[p] != '\n') break;
            size_t q = p, breaks = 0;
            while (q < t.size() && (t[q] == ' ' || t[q] == '\t' || t[q] == '\n' || t[q] == '\r')) { if (t[q] == '\n') breaks++; q++; }
            if (q >= t.size() || t[q] == ',' || t[q] == ']' || t[q] == '}' || t[q] == '#' || plain.empty()) break;
            if (breaks == 1) plain += ' '; else plain.append(breaks - 1, '\n');
            p = q;
            multi = true;
        }
        if (multi) { Value sv; sv.kind = Value::String; sv.s = plain; return sv; }
        return resolve_plain(plain);
    }
};


// ---- block parser ----
struct Line {
    int indent;
    std::string text;  // without indentation, comments and trailing blanks
    size_t raw;        // index of this line in Block::raw (block scalars read the untouched text)
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

void ltrim(std::string &t) {
    size_t b = 0;
    while (b < t.size() && (t[b] == ' ' || t[b] == '\t')) b++;
    t = t.substr(b);
}

struct Props {
    std::string anchor, tag;
};

// "&a !!str rest" -> anchor "a", tag "!!str", t = "rest"
void take_props(std::string &t, Props *p) {
    while (!t.empty() && (t[0] == '&' || t[0] == '!')) {
        size_t e = t.find_first_of(" \t");
        std::string tok = t.substr(0, e);
        if (t[0] == '&') {
            if (tok.size() < 2) fail("anchor without a name");
            p->anchor = tok.substr(1);
        } else {
            p->tag = tok;
        }
        t = e == std::string::npos ? "" : t.substr(e);
        ltrim(t);
    }
}

struct Block {
    std::vector<Line> lines;
    std::vector<std::string> raw;  // every physical line of the first document, untouched
    size_t pos = 0;
    Anchors anchors;

    bool is_dash(const std::string &t) const { return t == "-" || t.compare(0, 2, "- ") == 0; }

    // Is `text` (a value that starts with a quote or a flow bracket) still open at its end -- inside quotes, or with brackets unbalanced?
    static bool still_open(const std::string &text) {
        int depth = 0;
        char quote = 0;
        for (size_t i = 0; i < text.size(); i++) {
            char c = text[i];
            if (quote == '"') { if (c == '\\') i++; else if (c == '"') quote = 0; continue; }
            if (quote == '\'') { if (c == '\'') { if (i + 1 < text.size() && text[i + 1] == '\'') i++; else quote = 0; } continue; }
            if (c == '"' || c == '\'') quote = c;
            else if (c == '[' || c == '{') depth++;
            else if (c == ']' || c == '}') depth--;
            else if (c == '#' && (i == 0 || text[i - 1] == ' ' || text[i - 1] == '\t' || text[i - 1] == '\n')) {
                while (i < text.size() && text[i] != '\n') i++;  // a comment inside a flow collection
            }
        }
        return quote != 0 || depth > 0;
    }

    // A quoted scalar or a flow collection may run over several physical lines: append the following RAW lines (comment
    // stripping is not safe inside quotes) until it closes, and step pos past them.  `head` = index of the line it started on.
    std::string gather(std::string text, size_t head) {
        if (text.empty() || !(text[0] == '"' || text[0] == '\'' || text[0] == '[' || text[0] == '{') || !still_open(text)) return text;
        // restart from the raw line: the stripped text may have lost a '#...' that was inside quotes further on
        size_t r = lines[head].raw + 1;
        while (r < raw.size() && still_open(text)) {
            text += "\n" + raw[r];
            r++;
        }
        while (pos < lines.size() && lines[pos].raw < r) pos++;
        return text;
    }

    Value scalar_or_flow(const std::string &text, const std::string &tag) {
        if (text.empty()) return apply_scalar_tag(tag, "", Value());
        char c = text[0];
        if (c == '*') {
            std::string name = text.substr(1);
            if (!anchors.count(name)) fail("unknown anchor '" + name + "' referenced");
            return anchors[name];
        }
        if (c == '|' || c == '>') fail("block scalar not allowed here: " + text);
        Flow f(text, 0, &anchors);
        if (c == '[' || c == '{' || c == '"' || c == '\'') {
            Value v = f.value();
            f.ws();
            if (f.p != text.size()) fail("trailing characters after value: " + text);
            return apply_scalar_tag(tag, v.s, v);
        }
        return apply_scalar_tag(tag, text, resolve_plain(text));
    }

    // find "key:" split honoring quotes; returns false if the line is not a mapping entry
    static bool split_key(const std::string &t, std::string *key, std::string *rest, bool *quoted = nullptr) {
        size_t i = 0;
        if (quoted) *quoted = false;
        if (!t.empty() && (t[0] == '"' || t[0] == '\'')) {
            Flow f(t);
            try {
                *key = t[0] == '"' ? f.dquoted() : f.squoted();
            } catch (const ParseError &) {
                return false;  // a quoted SCALAR that continues on the next line, not a key
            }
            if (quoted) *quoted = true;
            i = f.p;
            while (i < t.size() && t[i] == ' ') i++;
            if (i >= t.size() || t[i] != ':') return false;
        } else {
            if (!t.empty() && (t[0] == '[' || t[0] == '{' || t[0] == '&' || t[0] == '*' || t[0] == '!' || t[0] == '|' || t[0] == '>')) return false;
            for (; i < t.size(); i++)
                if (t[i] == ':' && (i + 1 == t.size() || t[i + 1] == ' ' || t[i + 1] == '\t')) break;
            if (i >= t.size()) return false;
            *key = t.substr(0, i);
            while (!key->empty() && std::isspace((unsigned char)key->back())) key->pop_back();
        }
        *rest = t.substr(i + 1);
        ltrim(*rest);
        return true;
    }

    // Literal ('|') / folded ('>') block scalar whose header sits on line `head` (a Line index); the content is every following
    // physical line that is blank or indented deeper than `parent_indent`.  Advances pos past the content.
    Value block_scalar(const std::string &header, size_t head, int parent_indent) {
        const bool folded = header[0] == '>';
        int explicit_indent = 0;
        char chomp = 0;  // 0 clip, '-' strip, '+' keep
        for (size_t i = 1; i < header.size(); i++) {
            char c = header[i];
            if (c == '-' || c == '+') chomp = c;
            else if (c >= '1' && c <= '9') explicit_indent = c - '0';
            else fail("bad block scalar header: " + header);
        }
        size_t r = lines[head].raw + 1;
        std::vector<std::string> body;
        int content_indent = explicit_indent ? (parent_indent < 0 ? 0 : parent_indent) + explicit_indent : -1;
        for (; r < raw.size(); r++) {
            const std::string &ln = raw[r];
            size_t ind = 0;
            while (ind < ln.size() && ln[ind] == ' ') ind++;
            bool blank = ind == ln.size();
            if (!blank && (int)ind <= parent_indent) break;
            if (content_indent < 0 && !blank) content_indent = (int)ind;
            if (!blank && content_indent >= 0 && (int)ind < content_indent) break;
            body.push_back(blank ? (content_indent >= 0 && (int)ln.size() > content_indent ? ln.substr((size_t)content_indent) : "")
                                 : ln.substr((size_t)content_indent));
        }
        while (pos < lines.size() && lines[pos].raw < r) pos++;
        // trailing blank lines belong to the chomping decision
        size_t n = body.size();
        while (n > 0 && body[n - 1].empty()) n--;
        std::string out;
        if (!folded) {
            for (size_t i = 0; i < n; i++) { out += body[i]; if (i + 1 < n) out += '\n'; }
        } else {
            // folding: a single line break between two non-indented text lines becomes a space; blank lines give newlines;
            // "more indented" lines (and the breaks around them) are kept literally
            bool prev_more = false, prev_text = false;
            size_t pending_blank = 0;
            for (size_t i = 0; i < n; i++) {
                const std::string &ln = body[i];
                if (ln.empty()) { pending_blank++; continue; }
                bool more = ln[0] == ' ' || ln[0] == '\t';
                if (prev_text) {
                    if (pending_blank) out.append(pending_blank + ((prev_more || more) ? 1 : 0), '\n');
                    else out += (prev_more || more) ? '\n' : ' ';
                } else {
                    out.append(pending_blank, '\n');
                }
                pending_blank = 0;
                out += ln;
                prev_more = more;
                prev_text = true;
            }
        }
        if (n > 0 || chomp == '+') {
            if (chomp == 0 && n > 0) out += '\n';
            else if (chomp == '+') out.append((n > 0 ? 1 : 0) + (body.size() - n), '\n');
        }
        Value v;
        v.kind = Value::String;
        v.s = out;
        v.quoted = true;
        return v;
    }

    // the value that follows "key:" or "- " on the same line (`rest`), or the nested block when rest carries nothing but
    // properties; `indent` is the indentation of the line the key / dash sits on, `head` its index in `lines` (pos is already past it)
    Value value_after(std::string rest, int indent, size_t head, bool allow_same_indent_seq) {
        Props props;
        take_props(rest, &props);
        Value v;
        if (!rest.empty() && (rest[0] == '|' || rest[0] == '>')) {
            v = apply_scalar_tag(props.tag, "", Value());
            Value bs = block_scalar(rest, head, indent);
            v = props.tag.empty() ? bs : apply_scalar_tag(props.tag, bs.s, bs);
        } else if (!rest.empty()) {
            const char c = rest[0];
            if (c == '"' || c == '\'' || c == '[' || c == '{' || c == '*') {
                v = scalar_or_flow(gather(rest, head), props.tag);
            } else {
                // a plain scalar may continue on following, deeper-indented lines: one break folds to a space, k empty lines
                // in between to k newlines; a multi-line plain scalar is always a string
                std::string text = rest;
                bool multi = false;
                size_t last_raw = lines[head].raw;
                while (pos < lines.size() && lines[pos].indent > indent) {
                    const std::string &t = lines[pos].text;
                    std::string k, r;
                    if (is_dash(t) ? false : split_key(t, &k, &r)) fail("mapping values are not allowed in this context: " + t);
                    size_t blanks = lines[pos].raw - last_raw - 1;
                    if (blanks) text.append(blanks, '\n'); else text += ' ';
                    text += t;
                    last_raw = lines[pos].raw;
                    multi = true;
                    pos++;
                }
                if (multi) {
                    v.kind = Value::String;
                    v.s = text;
                    v = props.tag.empty() ? v : apply_scalar_tag(props.tag, text, v);
                } else {
                    v = scalar_or_flow(text, props.tag);
                }
            }
        } else if (pos < lines.size() &&
                   (lines[pos].indent > indent || (allow_same_indent_seq && lines[pos].indent == indent && is_dash(lines[pos].text)))) {
            v = parse_node(lines[pos].indent);
        } else {
            v = apply_scalar_tag(props.tag, "", Value());
        }
        if (!props.anchor.empty()) anchors[props.anchor] = v;
        return v;
    }

    Value parse_node(int indent) {
        if (pos >= lines.size()) return Value();
        const Line &l = lines[pos];
        if (is_dash(l.text)) return parse_seq(l.indent);
        std::string k, r;
        if (split_key(l.text, &k, &r)) return parse_map(l.indent);
        // a scalar / flow collection / alias / block scalar standing alone on its line (document root or a sequence item's body)
        size_t head = pos;
        std::string text = l.text;
        pos++;
        return value_after(text, indent - 1, head, false);
    }

    Value parse_seq(int indent) {
        Value v;
        v.kind = Value::Array;
        while (pos < lines.size() && lines[pos].indent == indent && is_dash(lines[pos].text)) {
            std::string rest = lines[pos].text.size() > 1 ? lines[pos].text.substr(2) : "";
            size_t lead = 0;
            while (lead < rest.size() && rest[lead] == ' ') lead++;
            rest = rest.substr(lead);
            std::string k, r;
            if (!rest.empty() && (is_dash(rest) || split_key(rest, &k, &r))) {
                // nested collection starting on the dash line: re-home the remainder as a virtual line
                lines[pos].indent = indent + 2 + (int)lead;
                lines[pos].text = rest;
                v.arr.push_back(parse_node(lines[pos].indent));
                continue;
            }
            size_t head = pos;
            pos++;
            v.arr.push_back(value_after(rest, indent, head, false));
        }
        return v;
    }

    static void map_put(Value &m, const std::string &key, const Value &val) {
        for (auto &kv : m.obj)
            if (kv.first == key) { kv.second = val; return; }
        m.obj.emplace_back(key, val);
    }
    // "<<": yaml.v2 applies a merge WHERE IT STANDS in the mapping (decode.go:660-664): it overwrites keys set before it and is
    // overwritten by keys after it; in a sequence of maps earlier entries win (decode.go:795-807).
    static void map_merge(Value &m, const Value &src) {
        if (src.kind == Value::Object) {
            for (const auto &kv : src.obj) map_put(m, kv.first, kv.second);
        } else if (src.kind == Value::Array) {
            for (size_t i = src.arr.size(); i-- > 0;) {
                if (src.arr[i].kind != Value::Object) fail("map merge requires map or sequence of maps as the value");
                for (const auto &kv : src.arr[i].obj) map_put(m, kv.first, kv.second);
            }
        } else {
            fail("map merge requires map or sequence of maps as the value");
        }
    }

    Value parse_map(int indent) {
        Value v;
        v.kind = Value::Object;
        while (pos < lines.size() && lines[pos].indent == indent) {
            std::string key, rest;
            bool quoted = false;
            if (is_dash(lines[pos].text)) break;
            if (!split_key(lines[pos].text, &key, &rest, &quoted)) fail("expected 'key: value' at: " + lines[pos].text);
            size_t head = pos;
            pos++;
            Value val = value_after(rest, indent, head, true);
            if (key == "<<" && !quoted) map_merge(v, val);
            else map_put(v, key_string(key, quoted), val);
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
        Anchors anchors;
        Flow f(text, first, &anchors);
        Value v = f.value();
        f.ws();
        if (f.p != text.size()) fail("trailing characters after document");
        return v;
    }
    // Only the FIRST document is read, as yaml.v2's Unmarshal does; directives (%YAML, %TAG) in front of it are skipped.
    Block b;
    size_t i = 0;
    bool started = false;
    while (i <= text.size()) {
        size_t e = text.find('\n', i);
        if (e == std::string::npos) e = text.size();
        std::string rawline = text.substr(i, e - i);
        i = e + 1;
        if (!rawline.empty() && rawline.back() == '\r') rawline.pop_back();
        if (rawline == "..." || rawline.compare(0, 4, "... ") == 0) { if (started) break; if (e == text.size()) break; continue; }
        if (rawline == "---" || rawline.compare(0, 4, "--- ") == 0) {
            if (started) break;  // a second document: not ours
            started = true;
            rawline = rawline.size() > 4 ? "    " + rawline.substr(4) : "";  // "--- value" keeps its value, at a column of its own
        }
        if (!started && !rawline.empty() && rawline[0] == '%') { if (e == text.size()) break; continue; }
        b.raw.push_back(rawline);
        std::string s = strip_comment(rawline);
        while (!s.empty() && (s.back() == ' ' || s.back() == '\t')) s.pop_back();
        size_t ind = 0;
        while (ind < s.size() && s[ind] == ' ') ind++;
        if (ind < s.size() && s[ind] == '\t') fail("tabs are not allowed for indentation");
        std::string body = s.substr(ind);
        if (!body.empty()) { b.lines.push_back({(int)ind, body, b.raw.size() - 1}); started = true; }
        if (e == text.size()) break;
    }
    if (b.lines.empty()) return Value();
    Value v = b.parse_node(b.lines[0].indent);
    if (b.pos != b.lines.size()) fail("could not parse line: " + b.lines[b.pos].text);
    return v;
}

}  // namespace yamllite
