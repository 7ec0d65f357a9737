#include "go_compat.hpp"

#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <ctime>

namespace gocompat {

namespace {

// utf8.DecodeRune: returns rune and width; invalid encodings give (0xFFFD, 1)
uint32_t decode_rune(const unsigned char *p, size_t n, size_t *w) {
    unsigned char c = p[0];
    *w = 1;
    if (c < 0x80) return c;
    int need;
    uint32_t r, lo;
    if ((c & 0xE0) == 0xC0) { need = 2; r = c & 0x1F; lo = 0x80; }
    else if ((c & 0xF0) == 0xE0) { need = 3; r = c & 0x0F; lo = 0x800; }
    else if ((c & 0xF8) == 0xF0) { need = 4; r = c & 0x07; lo = 0x10000; }
    else return 0xFFFD;
    if ((size_t)need > n) return 0xFFFD;
    for (int k = 1; k < need; k++) {
        if ((p[k] & 0xC0) != 0x80) return 0xFFFD;
        r = (r << 6) | (p[k] & 0x3F);
    }
    if (r < lo || r > 0x10FFFF || (r >= 0xD800 && r <= 0xDFFF)) return 0xFFFD;
    *w = (size_t)need;
    return r;
}

// unicode.IsSpace
bool is_space(uint32_t r) {
    switch (r) {
    case '\t': case '\n': case '\v': case '\f': case '\r': case ' ': case 0x85: case 0xA0:
    case 0x1680: case 0x2028: case 0x2029: case 0x202F: case 0x205F: case 0x3000:
        return true;
    default:
        return r >= 0x2000 && r <= 0x200A;
    }
}

}  // namespace

std::string trim_space(const std::string &s) {
    const unsigned char *p = reinterpret_cast<const unsigned char *>(s.data());
    size_t start = 0, end = s.size();
    while (start < end) {
        size_t w;
        uint32_t r = decode_rune(p + start, end - start, &w);
        if (!is_space(r)) break;
        start += w;
    }
    while (end > start) {
        // utf8.DecodeLastRune: back up over at most three continuation bytes
        size_t lim = end >= 4 ? end - 4 : 0;
        if (lim < start) lim = start;
        size_t k = end - 1;
        while (k > lim && (p[k] & 0xC0) == 0x80) k--;
        size_t w;
        uint32_t r = decode_rune(p + k, end - k, &w);
        if (k + w != end) { r = 0xFFFD; k = end - 1; }
        if (!is_space(r)) break;
        end = k;
    }
    return s.substr(start, end - start);
}

bool parse_bool(const std::string &s, bool *out) {
    static const char *t[] = {"1", "t", "T", "TRUE", "true", "True"};
    static const char *f[] = {"0", "f", "F", "FALSE", "false", "False"};
    for (const char *x : t) if (s == x) { *out = true; return true; }
    for (const char *x : f) if (s == x) { *out = false; return true; }
    return false;
}

bool parse_duration(const std::string &in, int64_t *ns_out) {
    // time.ParseDuration: [-+]?([0-9]*(\.[0-9]*)?[a-z]+)+  ("0" is valid on its own)
    std::string s = in;
    const std::string orig = s;
    uint64_t d = 0;
    bool neg = false;
    if (!s.empty() && (s[0] == '-' || s[0] == '+')) { neg = s[0] == '-'; s.erase(0, 1); }
    if (s == "0") { *ns_out = 0; return true; }
    if (s.empty()) return false;
    while (!s.empty()) {
        uint64_t v = 0, f = 0;
        double scale = 1;
        if (!(s[0] == '.' || (s[0] >= '0' && s[0] <= '9'))) return false;
        size_t pl = s.size();
        // leading integer
        size_t i = 0;
        while (i < s.size() && s[i] >= '0' && s[i] <= '9') {
            if (v > ((uint64_t)1 << 63) / 10) return false;
            v = v * 10 + (uint64_t)(s[i] - '0');
            if (v > ((uint64_t)1 << 63)) return false;
            i++;
        }
        s.erase(0, i);
        bool pre = pl != s.size();
        bool post = false;
        if (!s.empty() && s[0] == '.') {
            s.erase(0, 1);
            pl = s.size();
            i = 0;
            bool overflow = false;
            while (i < s.size() && s[i] >= '0' && s[i] <= '9') {
                if (!overflow) {
                    if (f > (((uint64_t)1 << 63) - 1) / 10) overflow = true;
                    else {
                        uint64_t y = f * 10 + (uint64_t)(s[i] - '0');
                        if (y > ((uint64_t)1 << 63)) overflow = true;
                        else { f = y; scale *= 10; }
                    }
                }
                i++;
            }
            s.erase(0, i);
            post = pl != s.size();
        }
        if (!pre && !post) return false;
        i = 0;
        while (i < s.size() && s[i] != '.' && !(s[i] >= '0' && s[i] <= '9')) i++;
        if (i == 0) return false;  // missing unit
        std::string u = s.substr(0, i);
        s.erase(0, i);
        uint64_t unit;
        if (u == "ns") unit = 1;
        else if (u == "us" || u == "\xC2\xB5s" || u == "\xCE\xBCs") unit = 1000;
        else if (u == "ms") unit = 1000000;
        else if (u == "s") unit = 1000000000ull;
        else if (u == "m") unit = 60ull * 1000000000ull;
        else if (u == "h") unit = 3600ull * 1000000000ull;
        else return false;
        if (v > ((uint64_t)1 << 63) / unit) return false;
        v *= unit;
        if (f > 0) {
            v += (uint64_t)((double)f * ((double)unit / scale));
            if (v > ((uint64_t)1 << 63)) return false;
        }
        d += v;
        if (d > ((uint64_t)1 << 63)) return false;
    }
    if (neg) { *ns_out = -(int64_t)d; return true; }
    if (d > ((uint64_t)1 << 63) - 1) return false;
    *ns_out = (int64_t)d;
    return true;
}

std::string duration_string(int64_t d) {
    // time.Duration.String
    if (d == 0) return "0s";
    char buf[40];
    int w = (int)sizeof buf;
    uint64_t u = d < 0 ? (uint64_t)(-(d + 1)) + 1 : (uint64_t)d;
    bool neg = d < 0;
    auto fmt_frac = [&](uint64_t v, int prec) -> uint64_t {
        bool print = false;
        for (int i = 0; i < prec; i++) {
            int digit = (int)(v % 10);
            print = print || digit != 0;
            if (print) buf[--w] = (char)('0' + digit);
            v /= 10;
        }
        if (print) buf[--w] = '.';
        return v;
    };
    auto fmt_int = [&](uint64_t v) {
        if (v == 0) buf[--w] = '0';
        else while (v > 0) { buf[--w] = (char)('0' + v % 10); v /= 10; }
    };
    if (u < 1000000000ull) {
        int prec;
        buf[--w] = 's';
        if (u < 1000) { prec = 0; buf[--w] = 'n'; }
        else if (u < 1000000) { prec = 3; buf[--w] = (char)0xB5; buf[--w] = (char)0xC2; }  // "µs"
        else { prec = 6; buf[--w] = 'm'; }
        u = fmt_frac(u, prec);
        fmt_int(u);
    } else {
        buf[--w] = 's';
        u = fmt_frac(u, 9);
        fmt_int(u % 60);
        u /= 60;
        if (u > 0) {
            buf[--w] = 'm';
            fmt_int(u % 60);
            u /= 60;
            if (u > 0) { buf[--w] = 'h'; fmt_int(u); }
        }
    }
    if (neg) buf[--w] = '-';
    return std::string(buf + w, sizeof buf - (size_t)w);
}

std::string spaces_to_dashes(const std::string &s) {
    std::string o = s;
    for (char &c : o) if (c == ' ') c = '-';
    return o;
}

void logf(const char *fmt, ...) {
    char msg[8192];
    va_list ap;
    va_start(ap, fmt);
    int n = vsnprintf(msg, sizeof msg, fmt, ap);
    va_end(ap);
    if (n < 0) return;
    if ((size_t)n >= sizeof msg) n = (int)sizeof msg - 1;
    time_t t = time(nullptr);
    struct tm tm;
    localtime_r(&t, &tm);
    char ts[32];
    strftime(ts, sizeof ts, "%Y/%m/%d %H:%M:%S", &tm);
    bool nl = n > 0 && msg[n - 1] == '\n';
    fprintf(stderr, "%s %s%s", ts, msg, nl ? "" : "\n");
    fflush(stderr);
}

}  // namespace gocompat
