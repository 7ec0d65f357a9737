// Test helper (not product): parse stdin with the labeler's YAML reader and print the value tree as JSON, so that
// tests/test_yaml_reader.py can compare it with what PyYAML reads from the same text.  Exit 1 + message on a parse error.
#include <cmath>
#include <cstdio>
#include <iostream>
#include <iterator>
#include <string>

#include "../gpu-feature-discovery_b200/labeler/yaml_lite.hpp"

static void put_string(const std::string &s) {
    std::fputc('"', stdout);
    for (unsigned char c : s) {
        if (c == '"' || c == '\\') std::printf("\\%c", c);
        else if (c < 0x20) std::printf("\\u%04x", c);
        else std::fputc(c, stdout);
    }
    std::fputc('"', stdout);
}

static void put(const yamllite::Value &v) {
    using V = yamllite::Value;
    switch (v.kind) {
    case V::Null: std::fputs("null", stdout); break;
    case V::Bool: std::fputs(v.b ? "true" : "false", stdout); break;
    case V::Int: std::printf("%lld", (long long)v.i); break;
    case V::Float:
        if (std::isnan(v.f)) std::fputs("\"__nan__\"", stdout);
        else if (std::isinf(v.f)) std::fputs(v.f > 0 ? "\"__inf__\"" : "\"__-inf__\"", stdout);
        else std::printf("{\"__float__\": %.17g}", v.f);
        break;
    case V::String: put_string(v.s); break;
    case V::Array:
        std::fputc('[', stdout);
        for (size_t i = 0; i < v.arr.size(); i++) { if (i) std::fputc(',', stdout); put(v.arr[i]); }
        std::fputc(']', stdout);
        break;
    case V::Object:
        std::fputc('{', stdout);
        for (size_t i = 0; i < v.obj.size(); i++) { if (i) std::fputc(',', stdout); put_string(v.obj[i].first); std::fputc(':', stdout); put(v.obj[i].second); }
        std::fputc('}', stdout);
        break;
    }
}

int main() {
    std::string text((std::istreambuf_iterator<char>(std::cin)), std::istreambuf_iterator<char>());
    try {
        put(yamllite::parse(text));
        std::fputc('\n', stdout);
    } catch (const std::exception &e) {
        std::fprintf(stderr, "%s\n", e.what());
        return 1;
    }
    return 0;
}
