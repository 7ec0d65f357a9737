#include "config.hpp"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <stdexcept>

#include "go_compat.hpp"
#include "yaml_lite.hpp"

namespace gfdcfg {

using yamllite::Value;

#ifndef GFD_VERSION
#define GFD_VERSION "v0.7.0-b200"
#endif

// ---------------------------------------------------------------------------------------------------
// Flag table (cmd/gpu-feature-discovery/main.go:33-82) + the new opt-in flags
// ---------------------------------------------------------------------------------------------------
enum class Kind { String, Bool, Duration, Int };

struct FlagDef {
    const char *name;
    std::vector<const char *> aliases;
    Kind kind;
    const char *def;  // textual default
    std::vector<const char *> env;
    const char *usage;
    bool is_new;  // not in the reference
};

static const std::vector<FlagDef> &flag_defs() {
    static const std::vector<FlagDef> defs = {
        {"mig-strategy", {}, Kind::String, "none", {"GFD_MIG_STRATEGY", "MIG_STRATEGY"},
         "the desired strategy for exposing MIG devices on GPUs that support it:\n\t\t[none | single | mixed]", false},
        {"fail-on-init-error", {}, Kind::Bool, "true", {"GFD_FAIL_ON_INIT_ERROR", "FAIL_ON_INIT_ERROR"},
         "fail the plugin if an error is encountered during initialization, otherwise block indefinitely", false},
        {"oneshot", {}, Kind::Bool, "false", {"GFD_ONESHOT"}, "Label once and exit", false},
        {"no-timestamp", {}, Kind::Bool, "false", {"GFD_NO_TIMESTAMP"}, "Do not add the timestamp to the labels", false},
        {"sleep-interval", {}, Kind::Duration, "1m0s", {"GFD_SLEEP_INTERVAL"}, "Time to sleep between labeling", false},
        {"output-file", {"output", "o"}, Kind::String, "/etc/kubernetes/node-feature-discovery/features.d/gfd", {"GFD_OUTPUT_FILE"}, "", false},
        {"machine-type-file", {}, Kind::String, "/sys/class/dmi/id/product_name", {"GFD_MACHINE_TYPE_FILE"},
         "a path to a file that contains the DMI (SMBIOS) information for the node", false},
        {"config-file", {}, Kind::String, "", {"GFD_CONFIG_FILE", "CONFIG_FILE"},
         "the path to a config file as an alternative to command line options or environment variables", false},
        // ---- new, opt-in; absent from the reference ----
        {"probe-hbm", {}, Kind::Bool, "false", {"GFD_PROBE_HBM"},
         "[b200, opt-in] measure HBM read / copy / write bandwidth per GPU with sm_100a streaming kernels (data verified) and add nvidia.com/gpu.probe.hbm.* labels", true},
        {"probe-nvlink", {}, Kind::Bool, "false", {"GFD_PROBE_NVLINK"},
         "[b200, opt-in] measure peer-to-peer NVLink bandwidth per GPU pair and add nvidia.com/gpu.probe.nvlink.* labels", true},
        {"probe-size-mib", {}, Kind::Int, "4096", {"GFD_PROBE_SIZE_MIB"}, "[b200, opt-in] probe buffer size in MiB", true},
        {"prefetch-threads", {}, Kind::Int, "0", {"GFD_PREFETCH_THREADS"},
         "[b200] host threads used to attach and query GPUs concurrently (0 = one per GPU, 1 = serial); never changes the labels", true},
        {"keep-nvml-attached", {}, Kind::Bool, "false", {"GFD_KEEP_NVML_ATTACHED"},
         "[b200, opt-in] keep NVML initialised between labeling passes instead of re-attaching every pass", true},
        {"current-tables", {}, Kind::Bool, "false", {"GFD_CURRENT_TABLES"},
         "[b200, opt-in] use up-to-date tables where the reference's are stale: gpu.family=blackwell for compute 10.x/12.x, MIG profile ids of CUDA 12.9 (e.g. 1g.45gb) resolved instead of failing; CHANGES labels", true},
        {"timing", {}, Kind::Bool, "false", {"GFD_TIMING"}, "[b200, opt-in] log the wall-clock time of every labeling pass", true},
        {"passes", {}, Kind::Int, "0", {"GFD_PASSES"},
         "[b200, opt-in] label this many times (sleeping --sleep-interval in between), keep the file and exit; 0 = run until signalled", true},
        {"help", {"h"}, Kind::Bool, "false", {}, "show help", false},
        {"version", {"v"}, Kind::Bool, "false", {}, "print the version", false},
    };
    return defs;
}

struct FlagState {
    std::string value;  // textual
    bool set_cli = false;
    bool set_env = false;
};

struct ParsedFlags {
    std::map<std::string, FlagState> st;  // canonical name -> state
    std::vector<std::string> args;        // positional

    const FlagState &at(const std::string &n) const { return st.at(n); }
    bool is_set(const std::string &n) const { const auto &s = st.at(n); return s.set_cli || s.set_env; }  // cli.Context.IsSet (V/cli/context.go:53-74)
    std::string str(const std::string &n) const { return st.at(n).value; }
    bool boolean(const std::string &n) const { bool b = false; gocompat::parse_bool(st.at(n).value, &b); return b; }
    int64_t duration(const std::string &n) const { int64_t d = 0; gocompat::parse_duration(st.at(n).value, &d); return d; }
};

static std::string quote(const std::string &s) {  // Go %q for the plain strings that reach error messages
    std::string o = "\"";
    for (char c : s) {
        if (c == '"' || c == '\\') { o += '\\'; o += c; }
        else if (c == '\n') o += "\\n";
        else if (c == '\t') o += "\\t";
        else o += c;
    }
    return o + "\"";
}

std::shared_ptr<ParsedFlags> parse_cli(int argc, char **argv, CliResult *res) {
    auto pf = std::make_shared<ParsedFlags>();
    std::map<std::string, const FlagDef *> by_name;  // every spelling -> def
    for (const FlagDef &d : flag_defs()) {
        by_name[d.name] = &d;
        for (const char *a : d.aliases) by_name[a] = &d;
        pf->st[d.name].value = d.def;
    }
    // ---- environment first (urfave applies it while building the flag set: V/cli/flag_*.go Apply) ----
    for (const FlagDef &d : flag_defs()) {
        for (const char *e : d.env) {
            const char *val = getenv(e);
            if (!val) continue;  // flagFromEnvOrFile: the FIRST variable that is set wins, even when empty
            std::string v = val;
            FlagState &s = pf->st[d.name];
            if (d.kind == Kind::String) {
                s.value = v;
                s.set_env = true;
            } else if (!v.empty()) {
                bool okv = true;
                if (d.kind == Kind::Bool) {
                    bool b;
                    okv = gocompat::parse_bool(v, &b);
                    if (!okv) {
                        res->action = CliAction::Error;
                        res->error = "could not parse " + quote(v) + " as bool value from environment variable " + quote(e) +
                                     " for flag " + d.name + ": strconv.ParseBool: parsing " + quote(v) + ": invalid syntax";
                        return pf;
                    }
                    v = b ? "true" : "false";
                } else if (d.kind == Kind::Duration) {
                    int64_t ns;
                    okv = gocompat::parse_duration(v, &ns);
                    if (!okv) {
                        res->action = CliAction::Error;
                        res->error = "could not parse " + quote(v) + " as duration value from environment variable " + quote(e) +
                                     " for flag " + d.name + ": time: invalid duration " + quote(v);
                        return pf;
                    }
                } else {
                    char *end = nullptr;
                    strtoll(v.c_str(), &end, 0);
                    if (!end || *end) {
                        res->action = CliAction::Error;
                        res->error = "could not parse " + quote(v) + " as int value from environment variable " + quote(e) + " for flag " + d.name;
                        return pf;
                    }
                }
                s.value = v;
                s.set_env = true;
            }
            break;
        }
    }
    // ---- command line: Go's flag.FlagSet.Parse ----
    std::map<std::string, std::string> seen_spelling;  // canonical -> spelling used (normalizeFlags, V/cli/flag.go:164-195)
    int i = 1;
    std::string perr;
    while (i < argc) {
        std::string s = argv[i];
        if (s.size() < 2 || s[0] != '-') break;
        size_t minuses = 1;
        if (s[1] == '-') {
            minuses = 2;
            if (s.size() == 2) { i++; break; }  // "--" terminates the flags
        }
        std::string name = s.substr(minuses);
        if (name.empty() || name[0] == '-' || name[0] == '=') { perr = "bad flag syntax: " + s; break; }
        i++;
        bool has_value = false;
        std::string value;
        for (size_t k = 1; k < name.size(); k++)
            if (name[k] == '=') { value = name.substr(k + 1); has_value = true; name = name.substr(0, k); break; }
        auto it = by_name.find(name);
        if (it == by_name.end()) { perr = "flag provided but not defined: -" + name; break; }
        const FlagDef &d = *it->second;
        FlagState &st = pf->st[d.name];
        if (d.kind == Kind::Bool) {
            if (has_value) {
                bool b;
                if (!gocompat::parse_bool(value, &b)) { perr = "invalid boolean value " + quote(value) + " for -" + name + ": parse error"; break; }
                st.value = b ? "true" : "false";
            } else {
                st.value = "true";
            }
        } else {
            if (!has_value && i < argc) { has_value = true; value = argv[i++]; }
            if (!has_value) { perr = "flag needs an argument: -" + name; break; }
            if (d.kind == Kind::Duration) {
                int64_t ns;
                if (!gocompat::parse_duration(value, &ns)) { perr = "invalid value " + quote(value) + " for flag -" + name + ": parse error"; break; }
            } else if (d.kind == Kind::Int) {
                char *end = nullptr;
                strtoll(value.c_str(), &end, 0);
                if (value.empty() || !end || *end) { perr = "invalid value " + quote(value) + " for flag -" + name + ": parse error"; break; }
            }
            st.value = value;
        }
        st.set_cli = true;
        auto sp = seen_spelling.find(d.name);
        if (sp != seen_spelling.end() && sp->second != name && res->error.empty()) {
            res->action = CliAction::Error;
            res->error = "Cannot use two forms of the same flag: " + name + " " + sp->second;
        }
        seen_spelling[d.name] = name;
    }
    for (; i < argc; i++) pf->args.push_back(argv[i]);
    if (res->action == CliAction::Error) { res->usage_error = res->error; return pf; }
    if (!perr.empty()) {
        res->action = CliAction::Error;
        res->error = perr;
        res->usage_error = "Incorrect Usage. " + perr;
        return pf;
    }
    if (pf->boolean("help")) { res->action = CliAction::Help; return pf; }
    if (pf->boolean("version")) { res->action = CliAction::Version; return pf; }
    if (!pf->args.empty() && (pf->args[0] == "help" || pf->args[0] == "h")) { res->action = CliAction::Help; return pf; }
    return pf;
}

std::string version_text() { return std::string("GPU Feature Discovery version ") + GFD_VERSION + "\n"; }

std::string help_text() {
    std::ostringstream o;
    o << "NAME:\n   GPU Feature Discovery - generate labels for NVIDIA devices\n\n"
      << "USAGE:\n   gpu-feature-discovery [global options] command [command options] [arguments...]\n\n"
      << "VERSION:\n   " << GFD_VERSION << "\n\n"
      << "COMMANDS:\n   help, h  Shows a list of commands or help for one command\n\n"
      << "GLOBAL OPTIONS:\n";
    for (const FlagDef &d : flag_defs()) {
        std::string names = std::string("--") + d.name;
        for (const char *a : d.aliases) names += std::string(", ") + (strlen(a) == 1 ? "-" : "--") + a;
        if (d.kind != Kind::Bool) names += " value";
        o << "   " << names << "  " << d.usage;
        if (d.kind != Kind::Bool && d.def[0]) o << " (default: " << (d.kind == Kind::String ? quote(d.def) : std::string(d.def)) << ")";
        else if (d.kind == Kind::Bool && strcmp(d.name, "help") && strcmp(d.name, "version")) o << " (default: " << d.def << ")";
        if (!d.env.empty()) {
            o << " [";
            for (size_t k = 0; k < d.env.size(); k++) o << (k ? ", " : "") << "$" << d.env[k];
            o << "]";
        }
        o << "\n";
    }
    o << "\n";
    return o.str();
}

// ---------------------------------------------------------------------------------------------------
// Config file -> Config (encoding/json semantics over the YAML->JSON tree)
// ---------------------------------------------------------------------------------------------------
namespace {

[[noreturn]] void jerr(const std::string &m) { throw std::runtime_error(m); }

const char *kind_name(const Value &v) {
    switch (v.kind) {
    case Value::Null: return "null";
    case Value::Bool: return "bool";
    case Value::Int: case Value::Float: return "number";
    case Value::String: return "string";
    case Value::Array: return "array";
    default: return "object";
    }
}

std::string format_float_g32(double f) {  // strconv.FormatFloat(f, 'g', -1, 32) for the YAML->string coercion
    char buf[64];
    for (int prec = 1; prec <= 9; prec++) {
        snprintf(buf, sizeof buf, "%.*g", prec, f);
        if ((float)strtod(buf, nullptr) == (float)f) break;
    }
    return buf;
}

// target: plain Go string (sigs.k8s.io/yaml coerces YAML scalars to strings for string-kind targets)
bool as_plain_string(const Value &v, const char *field, std::string *out) {
    switch (v.kind) {
    case Value::Null: return false;
    case Value::String: *out = v.s; return true;
    case Value::Bool: *out = v.b ? "true" : "false"; return true;
    case Value::Int: *out = std::to_string(v.i); return true;
    case Value::Float: *out = format_float_g32(v.f); return true;
    default: jerr(std::string("unmarshal error: json: cannot unmarshal ") + kind_name(v) + " into Go struct field " + field + " of type string");
    }
}

// target with its own UnmarshalJSON taking a JSON string (no YAML coercion)
bool as_json_string(const Value &v, const char *field, std::string *out) {
    if (v.kind == Value::Null) return false;
    if (v.kind != Value::String) jerr(std::string("unmarshal error: json: cannot unmarshal ") + kind_name(v) + " into Go value of type string (" + field + ")");
    *out = v.s;
    return true;
}

bool as_bool(const Value &v, const char *field, bool *out) {
    if (v.kind == Value::Null) return false;
    if (v.kind != Value::Bool) jerr(std::string("unmarshal error: json: cannot unmarshal ") + kind_name(v) + " into Go struct field " + field + " of type bool");
    *out = v.b;
    return true;
}

bool as_int(const Value &v, const char *field, int64_t *out) {
    if (v.kind == Value::Null) return false;
    if (v.kind == Value::Int) { *out = v.i; return true; }
    if (v.kind == Value::Float && std::floor(v.f) == v.f && std::fabs(v.f) < 9e15 && v.s.find_first_of(".eE") == std::string::npos) { *out = (int64_t)v.f; return true; }
    jerr(std::string("unmarshal error: json: cannot unmarshal ") + kind_name(v) + " into Go value of type int (" + field + ")");
}

const Value &want_object(const Value &v, const char *field) {
    if (v.kind != Value::Object) jerr(std::string("unmarshal error: json: cannot unmarshal ") + kind_name(v) + " into Go struct field " + field);
    return v;
}

// k8s.io/apimachinery validation.NameIsDNSSubdomain(name, false)
bool is_dns_subdomain(const std::string &s) {
    if (s.empty() || s.size() > 253) return false;
    size_t i = 0;
    while (true) {
        size_t e = s.find('.', i);
        std::string label = s.substr(i, e == std::string::npos ? std::string::npos : e - i);
        if (label.empty()) return false;
        auto alnum = [](char c) { return (c >= 'a' && c <= 'z') || (c >= '0' && c <= '9'); };
        if (!alnum(label.front()) || !alnum(label.back())) return false;
        for (char c : label) if (!alnum(c) && c != '-') return false;
        if (e == std::string::npos) break;
        i = e + 1;
    }
    return true;
}

// NewResourceName (V/spec/resources.go:48-64)
std::string new_resource_name(std::string n) {
    if (n.compare(0, 11, "nvidia.com/") != 0) n = "nvidia.com/" + n;
    if (n.size() > 63) jerr("fully-qualified resource name must be 63 characters or less: " + n);
    size_t slash = n.find('/');
    std::string name = n.substr(slash + 1);
    if (!is_dns_subdomain(name)) jerr("incorrect format for resource name '" + n + "'");
    return n;
}

bool is_uint_str(const std::string &s) {
    if (s.empty()) return false;
    size_t i = (s[0] == '+') ? 1 : 0;
    if (i >= s.size()) return false;
    for (; i < s.size(); i++) if (s[i] < '0' || s[i] > '9') return false;
    return s.size() <= 19;
}

bool is_uuid(const std::string &s) {  // github.com/google/uuid Parse: the 36-char canonical form (also urn:/braces/32-hex)
    std::string u = s;
    if (u.size() == 45 && u.compare(0, 9, "urn:uuid:") == 0) u = u.substr(9);
    if (u.size() == 38 && u.front() == '{' && u.back() == '}') u = u.substr(1, 36);
    auto hex = [](char c) { return std::isxdigit((unsigned char)c) != 0; };
    if (u.size() == 32) { for (char c : u) if (!hex(c)) return false; return true; }
    if (u.size() != 36) return false;
    for (size_t i = 0; i < 36; i++) {
        if (i == 8 || i == 13 || i == 18 || i == 23) { if (u[i] != '-') return false; }
        else if (!hex(u[i])) return false;
    }
    return true;
}

bool is_gpu_uuid(const std::string &s) { return s.compare(0, 4, "GPU-") == 0 && is_uuid(s.substr(4)); }

bool is_mig_uuid(const std::string &s) {  // V/spec/replicas.go:93-118
    if (s.compare(0, 4, "MIG-") != 0) return false;
    std::string suffix = s.substr(4);
    if (is_uuid(suffix)) return true;
    size_t a = suffix.find('/');
    if (a == std::string::npos) return false;
    size_t b = suffix.find('/', a + 1);
    if (b == std::string::npos) return false;
    return is_gpu_uuid(suffix.substr(0, a)) && is_uint_str(suffix.substr(a + 1, b - a - 1)) && is_uint_str(suffix.substr(b + 1));
}

bool is_mig_index(const std::string &s) {
    size_t c = s.find(':');
    return c != std::string::npos && is_uint_str(s.substr(0, c)) && is_uint_str(s.substr(c + 1));
}

// ReplicatedDevices.UnmarshalJSON (V/spec/replicas.go:226-281)
ReplicatedDevices parse_devices(const Value &v) {
    ReplicatedDevices d;
    if (v.kind == Value::String) {
        if (v.s != "all") jerr("devices set as '" + v.s + "' but the only valid string input is 'all'");
        d.all = true;
        return d;
    }
    if (v.kind == Value::Int || (v.kind == Value::Float && std::floor(v.f) == v.f)) {
        int64_t c = v.kind == Value::Int ? v.i : (int64_t)v.f;
        if (c <= 0) jerr("devices set as '" + std::to_string(c) + "' but a count of devices must be > 0");
        d.count = (int)c;
        return d;
    }
    if (v.kind == Value::Array) {
        for (const Value &e : v.arr) {
            if (e.kind == Value::Int && e.i >= 0) { d.list.push_back(std::to_string(e.i)); continue; }
            if (e.kind == Value::String && (is_uint_str(e.s) || is_mig_index(e.s) || is_gpu_uuid(e.s) || is_mig_uuid(e.s))) { d.list.push_back(e.s); continue; }
            jerr("unsupported type for device in devices list: " + (e.kind == Value::String ? e.s : std::string()) + ", string");
        }
        d.list_set = true;
        return d;
    }
    jerr("unrecognized type for devices spec");
}

// ReplicatedResource.UnmarshalJSON (V/spec/replicas.go:168-224)
ReplicatedResource parse_replicated_resource(const Value &v) {
    if (v.kind != Value::Object && v.kind != Value::Null) jerr(std::string("json: cannot unmarshal ") + kind_name(v) + " into Go value of type map[string]json.RawMessage");
    ReplicatedResource r;
    const Value *name = v.get("name");
    if (!name) jerr("no resource name specified");
    std::string raw;
    if (as_json_string(*name, "name", &raw)) r.name = new_resource_name(raw);
    const Value *devices = v.get("devices");
    if (!devices) r.devices.all = true;
    else r.devices = parse_devices(*devices);
    const Value *replicas = v.get("replicas");
    if (!replicas) jerr("no replicas specified");
    as_int(*replicas, "replicas", &r.replicas);
    if (r.replicas < 2) jerr("number of replicas must be >= 2");
    if (const Value *rename = v.get("rename")) {
        std::string rr;
        if (as_json_string(*rename, "rename", &rr)) r.rename = new_resource_name(rr);
    }
    return r;
}

// TimeSlicing.UnmarshalJSON (V/spec/replicas.go:121-166)
TimeSlicing parse_time_slicing(const Value &v) {
    if (v.kind != Value::Object && v.kind != Value::Null) jerr(std::string("json: cannot unmarshal ") + kind_name(v) + " into Go value of type map[string]json.RawMessage");
    TimeSlicing t;
    if (const Value *x = v.get("renameByDefault")) as_bool(*x, "renameByDefault", &t.rename_by_default);
    if (const Value *x = v.get("failRequestsGreaterThanOne")) as_bool(*x, "failRequestsGreaterThanOne", &t.fail_requests_greater_than_one);
    const Value *res = v.get("resources");
    if (!res) jerr("no resources specified");
    if (res->kind != Value::Array && res->kind != Value::Null) jerr(std::string("json: cannot unmarshal ") + kind_name(*res) + " into Go value of type []v1.ReplicatedResource");
    for (const Value &e : res->arr) t.resources.push_back(parse_replicated_resource(e));
    if (t.resources.empty()) jerr("no resources specified");
    for (auto &r : t.resources)
        if (t.rename_by_default && r.rename.empty()) r.rename = r.name + ".shared";
    return t;
}

// Resource.UnmarshalJSON (V/spec/resources.go:120-150)
Resource parse_resource(const Value &v) {
    if (v.kind != Value::Object && v.kind != Value::Null) jerr(std::string("json: cannot unmarshal ") + kind_name(v) + " into Go value of type map[string]json.RawMessage");
    const Value *pattern = v.get("pattern"), *name = v.get("name");
    if (!pattern) jerr("resources must have a 'pattern' field set");
    if (!name) jerr("resources must have a 'name' field set");
    Resource r;
    as_json_string(*pattern, "pattern", &r.pattern);
    std::string raw;
    if (as_json_string(*name, "name", &raw)) r.name = new_resource_name(raw);
    return r;
}

void parse_resource_list(const Value &v, const char *field, std::vector<Resource> *out, bool *present) {
    if (v.kind == Value::Null) return;
    if (v.kind != Value::Array) jerr(std::string("unmarshal error: json: cannot unmarshal ") + kind_name(v) + " into Go struct field " + field + " of type []v1.Resource");
    *present = true;
    for (const Value &e : v.arr) out->push_back(parse_resource(e));
}

}  // namespace

Config parse_config_text(const std::string &text) {
    // parseConfigFrom (V/spec/config.go:74-99)
    Value root;
    try {
        root = yamllite::parse(text);
    } catch (const yamllite::ParseError &e) {
        jerr(std::string("unmarshal error: error converting YAML to JSON: ") + e.what());
    }
    Config c;
    c.version.clear();
    if (root.kind == Value::Null) { c.version = "v1"; return c; }
    if (root.kind != Value::Object) jerr(std::string("unmarshal error: json: cannot unmarshal ") + kind_name(root) + " into Go value of type v1.Config");
    try {
        if (const Value *v = root.get_fold("version")) as_plain_string(*v, "Config.version", &c.version);
        if (const Value *fv = root.get_fold("flags")) {
            if (fv->kind != Value::Null) {
                const Value &f = want_object(*fv, "Config.flags");
                std::string s;
                bool b;
                if (const Value *v = f.get_fold("migStrategy")) if (as_plain_string(*v, "Config.flags.migStrategy", &s)) c.mig_strategy = s;
                if (const Value *v = f.get_fold("failOnInitError")) if (as_bool(*v, "Config.flags.failOnInitError", &b)) c.fail_on_init_error = b;
                if (const Value *v = f.get_fold("nvidiaDriverRoot")) if (as_plain_string(*v, "Config.flags.nvidiaDriverRoot", &s)) c.nvidia_driver_root = s;
                if (const Value *v = f.get_fold("gdsEnabled")) if (as_bool(*v, "Config.flags.gdsEnabled", &b)) c.gds_enabled = b;
                if (const Value *v = f.get_fold("mofedEnabled")) if (as_bool(*v, "Config.flags.mofedEnabled", &b)) c.mofed_enabled = b;
                if (const Value *pv = f.get_fold("plugin")) {
                    if (pv->kind != Value::Null) {
                        const Value &p = want_object(*pv, "Config.flags.plugin");
                        if (const Value *v = p.get_fold("passDeviceSpecs")) as_bool(*v, "Config.flags.plugin.passDeviceSpecs", &b);
                        if (const Value *v = p.get_fold("deviceListStrategy")) as_plain_string(*v, "Config.flags.plugin.deviceListStrategy", &s);
                        if (const Value *v = p.get_fold("deviceIDStrategy")) as_plain_string(*v, "Config.flags.plugin.deviceIDStrategy", &s);
                    }
                }
                if (const Value *gv = f.get_fold("gfd")) {
                    if (gv->kind != Value::Null) {
                        const Value &g = want_object(*gv, "Config.flags.gfd");
                        if (const Value *v = g.get_fold("oneshot")) if (as_bool(*v, "Config.flags.gfd.oneshot", &b)) c.oneshot = b;
                        if (const Value *v = g.get_fold("noTimestamp")) if (as_bool(*v, "Config.flags.gfd.noTimestamp", &b)) c.no_timestamp = b;
                        if (const Value *v = g.get_fold("sleepInterval")) {
                            // Duration.UnmarshalJSON (V/spec/duration.go:33-51)
                            if (v->kind == Value::Int) c.sleep_interval_ns = v->i;
                            else if (v->kind == Value::Float) c.sleep_interval_ns = (int64_t)v->f;
                            else if (v->kind == Value::String) {
                                int64_t ns;
                                if (!gocompat::parse_duration(v->s, &ns)) jerr("unmarshal error: time: invalid duration " + quote(v->s));
                                c.sleep_interval_ns = ns;
                            } else if (v->kind != Value::Null) jerr("unmarshal error: invalid duration");
                            else jerr("unmarshal error: invalid duration");
                        }
                        if (const Value *v = g.get_fold("outputFile")) if (as_plain_string(*v, "Config.flags.gfd.outputFile", &s)) c.output_file = s;
                        if (const Value *v = g.get_fold("machineTypeFile")) if (as_plain_string(*v, "Config.flags.gfd.machineTypeFile", &s)) c.machine_type_file = s;
                    }
                }
            }
        }
        if (const Value *rv = root.get_fold("resources")) {
            if (rv->kind != Value::Null) {
                const Value &r = want_object(*rv, "Config.resources");
                if (const Value *v = r.get_fold("gpus")) parse_resource_list(*v, "Config.resources.gpus", &c.gpus, &c.gpus_present);
                if (const Value *v = r.get_fold("mig")) parse_resource_list(*v, "Config.resources.mig", &c.migs, &c.migs_present);
            }
        }
        if (const Value *sv = root.get_fold("sharing")) {
            if (sv->kind != Value::Null) {
                const Value &s = want_object(*sv, "Config.sharing");
                if (const Value *v = s.get_fold("timeSlicing")) c.time_slicing = parse_time_slicing(*v);
            }
        }
    } catch (const std::runtime_error &e) {
        std::string m = e.what();
        if (m.compare(0, 15, "unmarshal error") != 0) m = "unmarshal error: " + m;
        jerr(m);
    }
    if (c.version.empty()) c.version = "v1";
    if (c.version != "v1") jerr("unknown version: " + c.version);
    return c;
}

Config load_config(const ParsedFlags &f) {
    Config c;
    std::string path = f.str("config-file");
    if (!path.empty()) {
        // spec.NewConfig -> parseConfig (V/spec/config.go:43-72)
        std::ifstream in(path, std::ios::binary);
        if (!in) throw std::runtime_error("unable to load config: unable to finalize config: unable to parse config file: error opening config file: open " + path + ": " + strerror(errno));
        std::stringstream ss;
        ss << in.rdbuf();
        try {
            c = parse_config_text(ss.str());
        } catch (const std::runtime_error &e) {
            throw std::runtime_error(std::string("unable to load config: unable to finalize config: unable to parse config file: error parsing config file: ") + e.what());
        }
    }
    // Flags.UpdateFromCLIFlags (V/spec/flags.go:75-121): a flag overrides the file iff it was set (CLI or env) or the file left it unset
    if (f.is_set("mig-strategy") || !c.mig_strategy) c.mig_strategy = f.str("mig-strategy");
    if (f.is_set("fail-on-init-error") || !c.fail_on_init_error) c.fail_on_init_error = f.boolean("fail-on-init-error");
    if (f.is_set("oneshot") || !c.oneshot) c.oneshot = f.boolean("oneshot");
    if (f.is_set("output-file") || !c.output_file) c.output_file = f.str("output-file");
    if (f.is_set("sleep-interval") || !c.sleep_interval_ns) c.sleep_interval_ns = f.duration("sleep-interval");
    if (f.is_set("no-timestamp") || !c.no_timestamp) c.no_timestamp = f.boolean("no-timestamp");
    if (f.is_set("machine-type-file") || !c.machine_type_file) c.machine_type_file = f.str("machine-type-file");

    // disableResourceRenamingInConfig (main.go:236-270)
    if (!c.gpus.empty() || !c.migs.empty())
        gocompat::logf("Customizing the 'resources' field is not yet supported in the config. Ignoring...");
    c.gpus.clear();
    c.migs.clear();
    c.gpus_present = c.migs_present = false;
    bool non_default_rename = false, sets_devices = false;
    for (auto &r : c.time_slicing.resources) {
        if (!c.time_slicing.rename_by_default && !r.rename.empty()) { non_default_rename = true; r.rename.clear(); }
        if (c.time_slicing.rename_by_default && r.rename != r.name + ".shared") { non_default_rename = true; r.rename = r.name + ".shared"; }
        if (!r.devices.all) { sets_devices = true; r.devices = ReplicatedDevices(); r.devices.all = true; }
    }
    if (non_default_rename) gocompat::logf("Setting the 'rename' field in sharing.timeSlicing.resources is not yet supported in the config. Ignoring...");
    if (sets_devices) gocompat::logf("Customizing the 'devices' field in sharing.timeSlicing.resources is not yet supported in the config. Ignoring...");

    // new, opt-in knobs
    c.probe_hbm = f.boolean("probe-hbm");
    c.probe_nvlink = f.boolean("probe-nvlink");
    c.probe_bytes = strtoll(f.str("probe-size-mib").c_str(), nullptr, 0) << 20;
    if (c.probe_bytes <= 0) c.probe_bytes = 4ll << 30;
    c.prefetch_threads = (int)strtoll(f.str("prefetch-threads").c_str(), nullptr, 0);
    c.keep_nvml_attached = f.boolean("keep-nvml-attached");
    c.timing = f.boolean("timing");
    c.current_tables = f.boolean("current-tables");
    c.passes = (int)strtoll(f.str("passes").c_str(), nullptr, 0);
    if (const char *root = getenv("GFD_B200_PCI_DEVICES_ROOT")) c.pci_devices_root = root;
    return c;
}

// ---------------------------------------------------------------------------------------------------
// json.MarshalIndent(config, "", "  ")
// ---------------------------------------------------------------------------------------------------
static std::string jstr(const std::string &s) {  // encoding/json string encoding (HTML-safe)
    std::string o = "\"";
    const unsigned char *p = reinterpret_cast<const unsigned char *>(s.data());
    for (size_t i = 0; i < s.size(); i++) {
        unsigned char c = p[i];
        char buf[8];
        if (c == '"' || c == '\\') { o += '\\'; o += (char)c; }
        else if (c == '\n') o += "\\n";
        else if (c == '\r') o += "\\r";
        else if (c == '\t') o += "\\t";
        else if (c < 0x20 || c == '<' || c == '>' || c == '&') { snprintf(buf, sizeof buf, "\\u%04x", c); o += buf; }
        else if (c == 0xE2 && i + 2 < s.size() && p[i + 1] == 0x80 && (p[i + 2] == 0xA8 || p[i + 2] == 0xA9)) { o += p[i + 2] == 0xA8 ? "\\u2028" : "\\u2029"; i += 2; }
        else o += (char)c;
    }
    return o + "\"";
}

std::string Config::to_json() const {
    std::ostringstream o;
    auto optb = [](const std::optional<bool> &b) { return !b ? std::string("null") : std::string(*b ? "true" : "false"); };
    auto opts = [](const std::optional<std::string> &s) { return !s ? std::string("null") : jstr(*s); };
    o << "{\n";
    o << "  \"version\": " << jstr(version) << ",\n";
    o << "  \"flags\": {\n";
    o << "    \"migStrategy\": " << opts(mig_strategy) << ",\n";
    o << "    \"failOnInitError\": " << optb(fail_on_init_error) << ",\n";
    if (nvidia_driver_root) o << "    \"nvidiaDriverRoot\": " << jstr(*nvidia_driver_root) << ",\n";
    o << "    \"gdsEnabled\": " << optb(gds_enabled) << ",\n";
    o << "    \"mofedEnabled\": " << optb(mofed_enabled) << ",\n";
    o << "    \"gfd\": {\n";
    o << "      \"oneshot\": " << optb(oneshot) << ",\n";
    o << "      \"noTimestamp\": " << optb(no_timestamp) << ",\n";
    o << "      \"sleepInterval\": " << (sleep_interval_ns ? jstr(gocompat::duration_string(*sleep_interval_ns)) : std::string("null")) << ",\n";
    o << "      \"outputFile\": " << opts(output_file) << ",\n";
    o << "      \"machineTypeFile\": " << opts(machine_type_file) << "\n";
    o << "    }\n";
    o << "  },\n";
    o << "  \"resources\": {\n";
    o << "    \"gpus\": null\n";  // disableResourceRenamingInConfig nils both lists; `mig` is omitempty
    o << "  },\n";
    o << "  \"sharing\": {\n";
    const TimeSlicing &t = time_slicing;
    std::vector<std::string> fields;
    if (t.rename_by_default) fields.push_back("      \"renameByDefault\": true");
    if (t.fail_requests_greater_than_one) fields.push_back("      \"failRequestsGreaterThanOne\": true");
    if (!t.resources.empty()) {
        std::ostringstream r;
        r << "      \"resources\": [\n";
        for (size_t i = 0; i < t.resources.size(); i++) {
            const auto &x = t.resources[i];
            r << "        {\n";
            r << "          \"name\": " << jstr(x.name) << ",\n";
            if (!x.rename.empty()) r << "          \"rename\": " << jstr(x.rename) << ",\n";
            r << "          \"devices\": \"all\",\n";
            r << "          \"replicas\": " << x.replicas << "\n";
            r << "        }" << (i + 1 < t.resources.size() ? "," : "") << "\n";
        }
        r << "      ]";
        fields.push_back(r.str());
    }
    if (fields.empty()) {
        o << "    \"timeSlicing\": {}\n";
    } else {
        o << "    \"timeSlicing\": {\n";
        for (size_t i = 0; i < fields.size(); i++) o << fields[i] << (i + 1 < fields.size() ? "," : "") << "\n";
        o << "    }\n";
    }
    o << "  }\n";
    o << "}";
    return o.str();
}

}  // namespace gfdcfg
