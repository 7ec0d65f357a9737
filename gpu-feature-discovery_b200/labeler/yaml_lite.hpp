// A small YAML-subset / JSON reader producing a JSON-like value tree.
//
// The reference reads its config file with sigs.k8s.io/yaml (YAML -> JSON -> encoding/json,
// V/spec/config.go:74-99).  This reader covers what such config files use: block mappings and sequences,
// flow sequences/mappings (so any JSON document parses), single/double-quoted and plain scalars, comments,
// YAML 1.1 scalar resolution as yaml.v2 does it (null/bool/int/float/string), anchors / aliases / merge keys ("<<", applied
// where they stand, as yaml.v2 does), core and local tags on scalars, literal and folded block scalars with chomping and
// indentation indicators, directives, and "first document only" for multi-document files (yaml.v2's Unmarshal).
// Plain and quoted scalars and flow collections may continue on following lines (scalars with YAML's line folding).
// Still refused with a parse error: complex keys ("? "), properties in front of an implicit key ("&a key: v"), !!binary,
// tab indentation.
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace yamllite {

struct Value {
    enum Kind { Null, Bool, Int, Float, String, Array, Object } kind = Null;
    bool b = false;
    int64_t i = 0;
    double f = 0;
    std::string s;       // String payload; for Int/Float the source text
    bool quoted = false; // String came from a quoted scalar
    std::vector<Value> arr;
    std::vector<std::pair<std::string, Value>> obj;

    const Value *get(const std::string &key) const;     // exact key
    const Value *get_fold(const std::string &key) const; // exact, else ASCII case-insensitive (encoding/json)
};

struct ParseError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

Value parse(const std::string &text);

}  // namespace yamllite
