// Configuration of the labeler: CLI flags (+ environment) merged over an optional YAML/JSON config file.
//
// Replaces cmd/gpu-feature-discovery/main.go:22-107,236-270 (flag table, loadConfig,
// disableResourceRenamingInConfig) and the parts of the vendored k8s-device-plugin config package GFD uses:
// V/spec/config.go:33-99, flags.go:29-121, replicas.go:29-281, resources.go:48-64,120-165, duration.go.
// Flag grammar and env handling follow urfave/cli v2.8.1 (V/cli/flag.go:400-415, flag_bool.go:44-67,
// flag_string.go, flag_duration.go, context.go:53-74) on top of Go's `flag` package.
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <optional>
#include <string>
#include <vector>

namespace gfdcfg {

struct ReplicatedDevices {  // V/spec/replicas.go:43-49
    bool all = false;
    int count = 0;
    std::vector<std::string> list;
    bool list_set = false;
};

struct ReplicatedResource {  // V/spec/replicas.go:36-41
    std::string name;
    std::string rename;
    ReplicatedDevices devices;
    int64_t replicas = 0;
};

struct TimeSlicing {  // V/spec/replicas.go:29-33
    bool rename_by_default = false;
    bool fail_requests_greater_than_one = false;
    std::vector<ReplicatedResource> resources;
};

struct Resource {  // V/spec/resources.go:35-38
    std::string pattern, name;
};

struct Config {  // V/spec/config.go:33-38, flags.go:43-72
    std::string version = "v1";
    std::optional<std::string> mig_strategy;
    std::optional<bool> fail_on_init_error;
    std::optional<std::string> nvidia_driver_root;
    std::optional<bool> gds_enabled, mofed_enabled;
    // flags.gfd.*
    std::optional<bool> oneshot, no_timestamp;
    std::optional<int64_t> sleep_interval_ns;
    std::optional<std::string> output_file, machine_type_file;
    // resources / sharing
    std::vector<Resource> gpus, migs;
    bool gpus_present = false;  // distinguishes `null` from `[]` in the JSON dump
    bool migs_present = false;
    TimeSlicing time_slicing;

    // ---- NEW, opt-in (no reference counterpart; never part of the dumped reference config) ----
    bool probe_hbm = false;          // --probe-hbm: emit nvidia.com/gpu.probe.hbm.* labels
    bool probe_nvlink = false;       // --probe-nvlink: emit nvidia.com/gpu.probe.nvlink.* labels
    int64_t probe_bytes = 4ll << 30; // --probe-size-mib
    int prefetch_threads = 0;        // --prefetch-threads: per-GPU attach/query threads (0 = one per GPU, 1 = serial)
    bool keep_nvml_attached = false; // --keep-nvml-attached
    bool current_tables = false;     // --current-tables: blackwell family + CUDA 12.9 MIG profile ids (changes labels)
    bool timing = false;             // --timing: log per-phase wall-clock of every pass
    int passes = 0;                  // --passes N: exit (keeping the file) after N passes; 0 = daemon
    std::string pci_devices_root = "/sys/bus/pci/devices";  // GFD_B200_PCI_DEVICES_ROOT (test hook)

    std::string to_json() const;  // json.MarshalIndent(config, "", "  ") as main.go:127 prints it
};

enum class CliAction { Run, Help, Version, Error };

struct CliResult {
    CliAction action = CliAction::Run;
    std::string error;        // for Error: the message main() prints after "Error: "
    std::string usage_error;  // "Incorrect Usage. ..." preamble printed before the help text, if any
};

// Parse argv + environment into `flags` (values and which of them were explicitly set); config-file path too.
struct ParsedFlags;
std::shared_ptr<ParsedFlags> parse_cli(int argc, char **argv, CliResult *res);

// loadConfig (main.go:96-107): config file (if any) first, then CLI/env overlay, then
// disableResourceRenamingInConfig (main.go:236-270).  Throws std::runtime_error with the reference's message chain.
Config load_config(const ParsedFlags &flags);

std::string help_text();
std::string version_text();

// Exposed for tests (--dump-config-file): parse a YAML/JSON config document.
Config parse_config_text(const std::string &text);

}  // namespace gfdcfg
