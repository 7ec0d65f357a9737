#include "labeler.hpp"

#include <dirent.h>
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <cerrno>
#include <climits>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <system_error>
#include <thread>

#include "go_compat.hpp"

using gocompat::logf;

namespace gfdlm {

namespace {

const char *kFullGpuResource = "nvidia.com/gpu";  // internal/lm/resource.go:27

[[noreturn]] void fail(const std::string &what, int status) {
    throw PassError(what + ": " + gfd_strerror(status));
}
[[noreturn]] void fail(const std::string &what) { throw PassError(what); }
// fmt.Errorf("<prefix>: %v", err) around a callee, so that the logged error chain reads like the reference's
template <class F>
void wrapped(const char *prefix, F &&fn) {
    try {
        fn();
    } catch (const PassError &e) {
        throw PassError(std::string(prefix) + ": " + e.what());
    }
}

// os.ReadFile: whole file or an error (directories and unreadable files are errors; empty files are not)
bool read_file(const std::string &path, std::string *out, std::string *err) {
    int fd = open(path.c_str(), O_RDONLY | O_CLOEXEC);
    if (fd < 0) { *err = "open " + path + ": " + strerror(errno); return false; }
    out->clear();
    char buf[4096];
    while (true) {
        ssize_t n = read(fd, buf, sizeof buf);
        if (n < 0) {
            if (errno == EINTR) continue;
            *err = "read " + path + ": " + strerror(errno);
            close(fd);
            return false;
        }
        if (n == 0) break;
        out->append(buf, (size_t)n);
    }
    close(fd);
    return true;
}

// ---- thin throwing wrappers over the C ABI (resource.Device, internal/resource/types.go:32-42) ----
std::string dev_name(gfd_dev *d, const char *ctx) {
    char buf[256];
    int rc = gfd_dev_name(d, buf, sizeof buf);
    if (rc) fail(ctx, rc);
    return buf;
}
uint64_t dev_memory_mb(gfd_dev *d, const char *ctx) {
    uint64_t mb = 0;
    int rc = gfd_dev_total_memory_mb(d, &mb);
    if (rc) fail(ctx, rc);
    return mb;
}
bool dev_mig_enabled(gfd_dev *d) {
    int v = 0;
    int rc = gfd_dev_is_mig_enabled(d, &v);
    if (rc) fail("error getting MIG mode", rc);
    return v != 0;
}
std::vector<gfd_dev *> dev_migs(gfd_dev *d) {
    int n = 0;
    int rc = gfd_dev_mig_count(d, &n);
    if (rc) fail("error getting MIG devices", rc);
    std::vector<gfd_dev *> out;
    for (int i = 0; i < n; i++) {
        gfd_dev *m = nullptr;
        if ((rc = gfd_dev_mig_get(d, i, &m))) fail("error getting MIG devices", rc);
        out.push_back(m);
    }
    return out;
}

// ---- resourceLabeler (internal/lm/resource.go:113-226) ----
struct ResourceLabeler {
    std::string resource_name;
    const gfdcfg::Config *config;  // nullptr == sharing disabled (resource.go:30-33,194-196)

    std::string key(const std::string &suffix) const { return resource_name + "." + suffix; }

    const gfdcfg::ReplicatedResource *replication_info() const {  // resource.go:215-226
        if (!config) return nullptr;
        for (const auto &r : config->time_slicing.resources)
            if (r.name == resource_name) return &r;
        return nullptr;
    }
    bool is_shared() const { auto *r = replication_info(); return r && r->replicas > 1; }
    bool is_renamed() const { auto *r = replication_info(); return r && !r->rename.empty(); }

    void product(Labels &out, std::initializer_list<std::string> parts) const {  // resource.go:151-168
        std::vector<std::string> stripped;
        for (const std::string &p : parts)
            if (!p.empty()) stripped.push_back(gocompat::spaces_to_dashes(p));
        if (stripped.empty()) return;
        if (is_shared() && !is_renamed()) stripped.push_back("SHARED");
        std::string v;
        for (size_t i = 0; i < stripped.size(); i++) v += (i ? "-" : "") + stripped[i];
        out[key("product")] = v;
    }
    void base(Labels &out, int count, std::initializer_list<std::string> parts) const {  // resource.go:142-191
        product(out, parts);
        out[key("count")] = std::to_string(count);
        long long replicas = 1;
        if (!config) replicas = 0;
        else if (auto *r = replication_info(); r && r->replicas > 1) replicas = r->replicas;
        out[key("replicas")] = std::to_string(replicas);
    }
};

// getArchFamily (resource.go:261-284): the reference's table ends at hopper; anything newer is "undefined"
const char *arch_family(int major, int minor, bool current_tables) {
    if (current_tables && (major == 10 || major == 12)) return "blackwell";  // NEW, opt-in (--current-tables)
    if (current_tables && major == 8 && minor == 9) return "ada-lovelace";  // same name newer GFD releases use
    switch (major) {
    case 1: return "tesla";
    case 2: return "fermi";
    case 3: return "kepler";
    case 5: return "maxwell";
    case 6: return "pascal";
    case 7: return minor < 5 ? "volta" : "turing";
    case 8: return "ampere";
    case 9: return "hopper";
    }
    return "undefined";
}

// NewGPUResourceLabeler (resource.go:36-73)
void gpu_resource_labels(Labels &out, const gfdcfg::Config *config, gfd_dev *device, int count, bool current_tables) {
    if (count == 0) return;
    std::string model = dev_name(device, "failed to get device model");
    uint64_t total_mb = dev_memory_mb(device, "failed to get memory info for device");
    ResourceLabeler rl{kFullGpuResource, config};
    int major = 0, minor = 0;
    int rc = gfd_dev_cuda_compute_capability(device, &major, &minor);
    if (rc) fail("failed to create architecture labels: failed to determine CUDA compute capability", rc);
    rl.base(out, count, {model});
    if (total_mb != 0) out[rl.key("memory")] = std::to_string(total_mb);
    if (major != 0) {  // newArchitectureLabels (resource.go:239-258)
        out[rl.key("family")] = arch_family(major, minor, current_tables);
        out[rl.key("compute.major")] = std::to_string(major);
        out[rl.key("compute.minor")] = std::to_string(minor);
    }
}

// NewMIGResourceLabeler (resource.go:76-111)
void mig_resource_labels(Labels &out, const std::string &resource_name, const gfdcfg::Config *config, gfd_dev *mig, int count) {
    if (count == 0) return;
    gfd_dev *parent = nullptr;
    int rc = gfd_dev_parent(mig, &parent);
    if (rc) fail("failed to get parent of MIG device", rc);
    std::string model = dev_name(parent, "failed to get device model");
    std::string profile = dev_name(mig, "failed to get MIG profile name");
    ResourceLabeler rl{resource_name, config};
    gfd_mig_attributes a;
    if ((rc = gfd_dev_attributes(mig, &a))) fail("faled to get MIG attribute labels: unable to get attributes of MIG device", rc);
    rl.base(out, count, {model, "MIG", profile});
    out[rl.key("memory")] = std::to_string(a.memory_mb);  // nvml-mig-device.go:40-50
    out[rl.key("multiprocessors")] = std::to_string(a.multiprocessors);
    out[rl.key("slices.gi")] = std::to_string(a.slices_gi);
    out[rl.key("slices.ci")] = std::to_string(a.slices_ci);
    out[rl.key("engines.copy")] = std::to_string(a.engines_copy);
    out[rl.key("engines.decoder")] = std::to_string(a.engines_decoder);
    out[rl.key("engines.encoder")] = std::to_string(a.engines_encoder);
    out[rl.key("engines.jpeg")] = std::to_string(a.engines_jpeg);
    out[rl.key("engines.ofa")] = std::to_string(a.engines_ofa);
}

// mig.DeviceInfo (internal/mig/mig.go:24-124).  The reference builds a fresh one in three places; the
// answers are identical because the shim caches per bracket, so one instance per pass is enough.
struct DeviceInfo {
    std::vector<gfd_dev *> enabled, disabled;
    bool any = false;
    explicit DeviceInfo(Manager &mgr) {
        for (gfd_dev *d : mgr.devices()) {
            any = true;
            (dev_mig_enabled(d) ? enabled : disabled).push_back(d);
        }
    }
    bool any_enabled_is_empty() const {  // mig.go:85-106
        if (enabled.empty()) return true;
        for (gfd_dev *d : enabled)
            if (dev_migs(d).empty()) return true;
        return false;
    }
    std::vector<gfd_dev *> all_migs() const {  // mig.go:109-124
        std::vector<gfd_dev *> out;
        for (gfd_dev *d : enabled)
            for (gfd_dev *m : dev_migs(d)) out.push_back(m);
        return out;
    }
};

// newGPULabelers (mig-strategy.go:113-179).  Go ranges over two maps here; on a heterogeneous node the
// winner of nvidia.com/gpu.* is therefore random in the reference.  We iterate model names in sorted order.
void gpu_labels(Labels &out, const DeviceInfo &di, const gfdcfg::Config &config) {
    if (!di.any) fail("no GPU devices detected");
    std::map<std::string, int> counts;
    std::map<std::string, gfd_dev *> mig_enabled, full;
    for (gfd_dev *d : di.enabled) { std::string n = dev_name(d, "error getting device name"); mig_enabled[n] = d; counts[n]++; }
    for (gfd_dev *d : di.disabled) { std::string n = dev_name(d, "error getting device name"); full[n] = d; counts[n]++; }
    if (counts.size() > 1) {
        std::string names;
        for (auto &kv : counts) names += (names.empty() ? "" : " ") + kv.first;
        logf("WARNING: Multiple device types detected: [%s]", names.c_str());
    }
    wrapped("failed to construct labeler", [&] {  // mig-strategy.go:162,173
        for (auto &kv : mig_enabled) gpu_resource_labels(out, nullptr, kv.second, counts[kv.first], config.current_tables);  // WithoutSharing
        for (auto &kv : full) gpu_resource_labels(out, &config, kv.second, counts[kv.first], config.current_tables);
    });
}

// newInvalidMigStrategyLabeler (mig-strategy.go:243-262)
void invalid_single_labels(Labels &out, gfd_dev *device, const char *reason) {
    logf("WARNING: Invalid configuration detected for mig-strategy=single: %s", reason);
    std::string model = dev_name(device, "failed to get device model");
    ResourceLabeler rl{kFullGpuResource, nullptr};
    rl.product(out, {model, "MIG", "INVALID"});
    out[rl.key("count")] = "0";
    out[rl.key("replicas")] = "0";
    out[rl.key("memory")] = "0";
}

struct MigResource {  // migResource (mig-strategy.go:37-43)
    std::string resource_name;
    gfd_dev *device;
    int count;
};

std::map<std::string, MigResource> group_migs(const DeviceInfo &di, bool mixed) {
    std::map<std::string, MigResource> resources;
    std::vector<gfd_dev *> migs;
    wrapped("unable to retrieve list of MIG devices", [&] { migs = di.all_migs(); });  // mig-strategy.go:210-213,270-273
    for (gfd_dev *m : migs) {
        std::string n = dev_name(m, "unable to get MIG device name: failed to get MIG profile");
        auto it = resources.find(n);
        if (it == resources.end())
            it = resources.emplace(n, MigResource{mixed ? "nvidia.com/mig-" + n : std::string(kFullGpuResource), m, 0}).first;
        it->second.count++;
    }
    return resources;
}

// newMIGDeviceLabelers (mig-strategy.go:297-309)
void mig_device_labels(Labels &out, const std::map<std::string, MigResource> &resources, const gfdcfg::Config &config) {
    wrapped("failed to construct labeler", [&] {
        for (auto &kv : resources) mig_resource_labels(out, kv.second.resource_name, &config, kv.second.device, kv.second.count);
    });
}

// newMigStrategySingleLabeler (mig-strategy.go:181-241)
void single_labels(Labels &out, const DeviceInfo &di, const gfdcfg::Config &config) {
    if (di.enabled.empty()) return;
    bool has_empty = false;
    wrapped("failed to check for empty MIG-enabled devices", [&] { has_empty = di.any_enabled_is_empty(); });
    if (has_empty) return invalid_single_labels(out, di.enabled[0], "at least one MIG device is enabled but empty");
    if (!di.disabled.empty()) return invalid_single_labels(out, di.enabled[0], "devices with MIG enabled and disable detected");
    auto resources = group_migs(di, false);
    if (resources.size() != 1) return invalid_single_labels(out, di.enabled[0], "more than one MIG device type present on node");
    mig_device_labels(out, resources, config);
}

// newMigStrategyMixedLabeler (mig-strategy.go:264-295)
void mixed_labels(Labels &out, const DeviceInfo &di, const gfdcfg::Config &config) {
    mig_device_labels(out, group_migs(di, true), config);
}

// NewResourceLabeler (mig-strategy.go:45-77) + newMigLabeler (:84-110) + migStrategyLabeler (strategy.go:20-28)
void resource_labels(Labels &out, Manager &mgr, const gfdcfg::Config &config) {
    bool none = true;
    wrapped("error getting devices", [&] { none = mgr.devices().empty(); });
    if (none) return;
    std::unique_ptr<DeviceInfo> di;
    wrapped("failed to construct GPU labeler", [&] {
        wrapped("error getting map of devices", [&] { di.reset(new DeviceInfo(mgr)); });  // mig.go:41-64 via mig-strategy.go:116-119
        gpu_labels(out, *di, config);
    });
    const std::string &s = *config.mig_strategy;
    if (s == "none") return;
    Labels mig;
    wrapped("failed to construct MIG resource labeler", [&] {
        if (s == "single") wrapped("failed to create labeler for mig-strategy=single", [&] { single_labels(mig, *di, config); });
        else if (s == "mixed") wrapped("failed to create labeler for mig-strategy=mixed", [&] { mixed_labels(mig, *di, config); });
        else fail("unknown strategy: " + s);
    });
    out["nvidia.com/mig.strategy"] = s;
    for (auto &kv : mig) out[kv.first] = kv.second;
}

// newMachineTypeLabeler / getMachineType (internal/lm/machine-type.go:30-51)
void machine_type_label(Labels &out, const std::string &path) {
    std::string mt = "unknown";
    if (!path.empty()) {
        std::string data, err;
        if (read_file(path, &data, &err)) mt = gocompat::trim_space(data);
        else logf("WARNING: error getting machine type from %s: could not open machine type file: %s", path.c_str(), err.c_str());
    }
    out["nvidia.com/gpu.machine"] = gocompat::spaces_to_dashes(mt);
}

// newVersionLabeler (internal/lm/nvml.go:75-106)
void version_labels(Labels &out, Manager &mgr) {
    std::string dv = mgr.driver_version();
    std::vector<std::string> parts;
    size_t i = 0;
    while (true) {
        size_t e = dv.find('.', i);
        parts.push_back(dv.substr(i, e == std::string::npos ? std::string::npos : e - i));
        if (e == std::string::npos) break;
        i = e + 1;
    }
    if (parts.size() > 3 || parts.size() < 2)
        fail("failed to construct version labeler: error getting driver version: Version \"" + dv + "\" does not match format \"X.Y[.Z]\"");
    unsigned major = 0, minor = 0;
    mgr.cuda_driver_version(&major, &minor);
    out["nvidia.com/cuda.driver.major"] = parts[0];
    out["nvidia.com/cuda.driver.minor"] = parts[1];
    out["nvidia.com/cuda.driver.rev"] = parts.size() > 2 ? parts[2] : "";
    out["nvidia.com/cuda.runtime.major"] = std::to_string(major);
    out["nvidia.com/cuda.runtime.minor"] = std::to_string(minor);
}

// newMigCapabilityLabeler (internal/lm/nvml.go:110-137): true iff any device is MIG capable; stops at the first
void mig_capability_label(Labels &out, Manager &mgr) {
    bool capable = false;
    for (gfd_dev *d : mgr.devices()) {
        int v = 0;
        int rc = gfd_dev_is_mig_capable(d, &v);
        if (rc) fail("error creating mig capability labeler: error getting mig capability: error getting MIG mode", rc);
        capable = v != 0;
        if (capable) break;
    }
    out["nvidia.com/mig.capable"] = capable ? "true" : "false";
}

}  // namespace

// ---- Manager -----------------------------------------------------------------------------------------
Manager::Manager(int backend, unsigned flags, bool fallback_to_null) : fallback_(fallback_to_null) {
    int rc = gfd_mgr_new(backend, flags, &primary_);
    if (rc) fail("failed to create resource manager", rc);
    if ((rc = gfd_mgr_new(GFD_BACKEND_NULL, 0, &null_))) fail("failed to create null resource manager", rc);
}

Manager::~Manager() {
    gfd_mgr_free(primary_);
    gfd_mgr_free(null_);
}

void Manager::init() {
    int rc = gfd_init(active());
    if (rc == 0) return;
    if (!fallback_) fail("failed to initialize NVML", rc);
    // withFallBack.Init (fallback.go:37-44): log, become the null manager for good
    logf("Warning: failed to initialize resource manager: %s", gfd_strerror(rc));
    fell_back_ = true;
    gfd_init(null_);
}

void Manager::shutdown() { gfd_shutdown(active()); }

std::vector<gfd_dev *> Manager::devices() {
    int n = 0;
    int rc = gfd_device_count(active(), &n);
    if (rc) fail("error getting devices", rc);
    std::vector<gfd_dev *> out;
    for (int i = 0; i < n; i++) {
        gfd_dev *d = nullptr;
        if ((rc = gfd_device_get(active(), i, &d))) fail("error getting devices", rc);
        out.push_back(d);
    }
    return out;
}

std::string Manager::driver_version() {
    char buf[128];
    int rc = gfd_driver_version(active(), buf, sizeof buf);
    if (rc) fail("failed to construct version labeler: error getting driver version", rc);
    return buf;
}

void Manager::cuda_driver_version(unsigned *major, unsigned *minor) {
    int rc = gfd_cuda_driver_version(active(), major, minor);
    if (rc) fail("failed to construct version labeler: error getting cuda driver version", rc);
}

void Manager::prefetch(int threads) { gfd_prefetch(active(), threads); }

std::unique_ptr<Manager> new_manager(const gfdcfg::Config &config) {
    // resource.getManager (factory.go:42-74)
    char reason[1024];
    bool has_nvml = gfd_has_nvml(reason, sizeof reason) != 0;
    logf("Detected %sNVML platform: %s", has_nvml ? "" : "non-", reason);
    const char *tegra_root = getenv("GFD_B200_TEGRA_ROOT");  // test hook: prefix for the two Tegra probe paths
    bool is_tegra = gfd_is_tegra_system(tegra_root ? tegra_root : "", reason, sizeof reason) != 0;
    logf("Detected %sTegra platform: %s", is_tegra ? "" : "non-", reason);
    if (has_nvml && is_tegra) {
        logf("WARNING: Disabling Tegra-based resources on NVML system");
        is_tegra = false;
    }
    int backend;
    if (has_nvml) { logf("Using NVML manager"); backend = GFD_BACKEND_NVML; }
    else if (is_tegra) { logf("Using CUDA manager"); backend = GFD_BACKEND_CUDA; }
    else { logf("WARNING: No valid resources detected; using empty manager."); backend = GFD_BACKEND_NULL; }
    unsigned flags = (config.keep_nvml_attached ? GFD_MGR_KEEP_ATTACHED : 0) | (config.current_tables ? GFD_MGR_CURRENT_TABLES : 0);
    return std::unique_ptr<Manager>(new Manager(backend, flags, !*config.fail_on_init_error));  // WithConfig (factory.go:33-39)
}

// ---- NewNVMLLabeler (internal/lm/nvml.go:29-72) ----------------------------------------------------
Labels nvml_labels(Manager &mgr, const gfdcfg::Config &config, PhaseTimes *times) {
    auto now = [] {
        timespec ts;
        clock_gettime(CLOCK_MONOTONIC, &ts);
        return ts.tv_sec * 1e3 + ts.tv_nsec / 1e6;
    };
    PhaseTimes local;
    PhaseTimes &t = times ? *times : local;
    t = PhaseTimes();
    double t0 = now();
    try {
        mgr.init();
    } catch (const PassError &e) {
        throw PassError(std::string("error creating NVML labeler: ") + e.what());
    }
    t.init_ms = now() - t0;
    struct Guard {  // defer manager.Shutdown()
        Manager &m;
        PhaseTimes &t;
        double (*clock)();
        ~Guard() {
            double a = clock();
            m.shutdown();
            t.shutdown_ms = clock() - a;
        }
    } guard{mgr, t, +now};
    Labels out;
    try {
        double t1 = now();
        mgr.prefetch(config.prefetch_threads);  // 0 = one host thread per GPU (no-op on 0/1-GPU nodes), 1 = serial
        bool none = mgr.devices().empty();
        t.attach_ms = now() - t1;
        if (none) return out;
        struct Tail {  // everything after the walk: label building from (mostly cached) answers
            PhaseTimes &t;
            double start;
            double (*clock)();
            ~Tail() { t.labels_ms = clock() - start; }
        } tail{t, now(), +now};
        machine_type_label(out, *config.machine_type_file);
        version_labels(out, mgr);
        mig_capability_label(out, mgr);
        try {
            resource_labels(out, mgr, config);
        } catch (const PassError &e) {
            throw PassError(std::string("error creating resource labeler: ") + e.what());
        }
    } catch (const PassError &e) {
        throw PassError(std::string("error creating NVML labeler: ") + e.what());
    }
    return out;
}

// ---- vGPU (internal/vgpu/pciutil.go:70-112, vgpu.go:66-90, internal/lm/vgpu.go:37-55) -------------------
Labels vgpu_labels(const gfdcfg::Config &config) {
    Labels out;
    const std::string &root = config.pci_devices_root;
    std::vector<std::string> names;
    DIR *dir = opendir(root.c_str());
    if (!dir) fail("unable to get vGPU devices: error getting NVIDIA specific PCI devices: unable to read PCI bus devices: open " + root + ": " + strerror(errno));
    while (struct dirent *e = readdir(dir)) {
        if (!strcmp(e->d_name, ".") || !strcmp(e->d_name, "..")) continue;
        names.push_back(e->d_name);
    }
    closedir(dir);
    std::sort(names.begin(), names.end());  // ioutil.ReadDir sorts by filename
    struct Vgpu { std::string address, version, branch; };
    std::vector<Vgpu> vgpus;
    struct Pci { std::string address, config; size_t cap; };
    std::vector<Pci> nvidia;
    const std::string pre = "unable to get vGPU devices: error getting NVIDIA specific PCI devices: ";
    for (const std::string &address : names) {
        std::string path = root + "/" + address, vendor, klass, cfg, err;
        if (!read_file(path + "/vendor", &vendor, &err)) fail(pre + "unable to read PCI device vendor id for " + address + ": " + err);
        if (gocompat::trim_space(vendor) != "0x10de") continue;
        if (!read_file(path + "/class", &klass, &err)) fail(pre + "unable to read PCI device class for " + address + ": " + err);
        if (!read_file(path + "/config", &cfg, &err)) fail(pre + "unable to read PCI configuration space for " + address + ": " + err);
        if (klass.size() < 4) fail(pre + "PCI device class for " + address + " is shorter than 4 bytes");  // string(class)[0:4] panics
        // capacity os.ReadFile would have allocated: max(512, st_size + 1), at least what was read
        struct stat st;
        size_t cap = 512;
        if (stat((path + "/config").c_str(), &st) == 0 && (size_t)st.st_size + 1 > cap) cap = (size_t)st.st_size + 1;
        if (cfg.size() > cap) cap = cfg.size();
        nvidia.push_back({address, cfg, cap});
    }
    for (const Pci &p : nvidia) {  // Lib.Devices (vgpu.go:66-90)
        gfd_vgpu_info info;
        int rc = gfd_vgpu_parse(reinterpret_cast<const uint8_t *>(p.config.data()), p.config.size(), p.cap, &info);
        if (rc == GFD_ERR_PCI_CONFIG_SHORT || (rc && !info.is_vgpu))
            fail("unable to get vGPU devices: unable to read vendor specific capability for " + p.address, rc);
        if (rc) fail("error getting vGPU device info for device " + p.address, rc);
        if (info.is_vgpu)
            vgpus.push_back({p.address, std::string(info.host_driver_version, (size_t)info.host_driver_version_len),
                             std::string(info.host_driver_branch, (size_t)info.host_driver_branch_len)});
    }
    if (!vgpus.empty()) out["nvidia.com/vgpu.present"] = "true";
    for (const Vgpu &v : vgpus) {  // the last vGPU device wins (internal/lm/vgpu.go:46-53)
        out["nvidia.com/vgpu.host-driver-version"] = v.version;
        out["nvidia.com/vgpu.host-driver-branch"] = v.branch;
    }
    return out;
}

// ---- NEW opt-in probe labels ---------------------------------------------------------------------------
Labels probe_labels(const gfdcfg::Config &config) {
    Labels out;
    if (!config.probe_hbm && !config.probe_nvlink) return out;
    int n = 0;
    int rc = probe_device_count(&n);
    if (rc) fail("probe requested but no CUDA device is usable (the probes have no CPU fallback)", rc);
    if (config.probe_hbm) {
        // A probe is a guest on the GPU: when the requested buffer does not fit beside what already lives in HBM (a MIG
        // slice, a running job) it halves the size down to 256 MiB -- still >> the 126 MB L2 -- before giving up.
        const int kOutOfMemory = GFD_ERR_CUDA + 2;  // cudaErrorMemoryAllocation
        auto run = [&](int d, size_t bytes, int mode, gfd_probe_result *r, size_t *used) {
            int e;
            while ((e = probe_hbm(d, bytes, 3, 20, mode, r)) == kOutOfMemory && bytes > ((size_t)256 << 20)) bytes /= 2;
            *used = bytes;
            return e;
        };
        // The GPUs are independent (replicas): one host thread per GPU runs its three probes; results and the
        // first error (lowest GPU index) are folded afterwards on this thread.
        struct PerGpu { int rc = 0; const char *what = ""; double read = 0, copy = 0, write = 0; size_t used = 0; };
        std::vector<PerGpu> res((size_t)n);
        std::vector<std::thread> pool;
        auto probe_gpu = [&](int d) {
                PerGpu &g = res[(size_t)d];
                gfd_probe_result r;
                size_t used = 0;
                if ((g.rc = run(d, (size_t)config.probe_bytes, GFD_PROBE_READ, &r, &g.used))) { g.what = "HBM read probe failed on GPU "; return; }
                g.read = r.gbps;
                if ((g.rc = run(d, (size_t)config.probe_bytes / 2, GFD_PROBE_COPY, &r, &used))) { g.what = "HBM copy probe failed on GPU "; return; }
                g.copy = r.gbps;
                if ((g.rc = run(d, (size_t)config.probe_bytes, GFD_PROBE_WRITE, &r, &used))) { g.what = "HBM write probe failed on GPU "; return; }
                g.write = r.gbps;
            };
        for (int d = 0; d < n; d++) {
            try {
                pool.emplace_back(probe_gpu, d);
            } catch (const std::system_error &) {
                probe_gpu(d);  // no thread to be had: this GPU is probed here, after the others were started
            }
        }
        for (auto &t : pool) t.join();
        double rmin = 1e30, cmin = 1e30, wmin = 1e30;
        size_t size_min = (size_t)config.probe_bytes;
        for (int d = 0; d < n; d++) {
            const PerGpu &g = res[(size_t)d];
            if (g.rc) fail(g.what + std::to_string(d), g.rc);
            rmin = std::min(rmin, g.read);
            cmin = std::min(cmin, g.copy);
            wmin = std::min(wmin, g.write);
            size_min = std::min(size_min, g.used);
        }
        out["nvidia.com/gpu.probe.hbm.read-gbps"] = std::to_string((long long)rmin);
        out["nvidia.com/gpu.probe.hbm.copy-gbps"] = std::to_string((long long)cmin);
        out["nvidia.com/gpu.probe.hbm.write-gbps"] = std::to_string((long long)wmin);
        out["nvidia.com/gpu.probe.hbm.size-mib"] = std::to_string((long long)(size_min >> 20));
    }
    if (config.probe_nvlink && n >= 2) {
        // all directed pairs, concurrently: n-1 rounds of n/2 disjoint pairs, both directions at once
        size_t bytes = (size_t)std::min<int64_t>(config.probe_bytes, 1ll << 30);
        // Two figures per node, each the MINIMUM over all directed pairs under that schedule, each transfer verified against the
        // pattern's checksum: what a kernel's peer LOADS get (p2p-read) and what its peer STORES get (p2p-write).  With both
        // directions of a link busy they differ: a load puts a request packet on the opposite direction for every response
        // it receives (nvltx bytes on the pulling GPU, profiles/r2_ncu_p2p_summary.md), a posted store does not.
        std::vector<double> m((size_t)n * n, 0.0);
        double ring = 0.0, ring_w = 0.0;
        auto min_pair = [&](int *pairs) {
            double pmin = 1e30;
            *pairs = 0;
            for (int a = 0; a < n; a++)
                for (int b = 0; b < n; b++)
                    if (a != b && m[(size_t)a * n + b] > 0) { pmin = std::min(pmin, m[(size_t)a * n + b]); (*pairs)++; }
            return pmin;
        };
        rc = probe_p2p_matrix(n, bytes, 2, 10, GFD_PROBE_READ, m.data(), &ring);
        if (rc == GFD_ERR_PEER_ACCESS) {
            out["nvidia.com/gpu.probe.nvlink.pairs"] = "0";
        } else {
            if (rc) fail("NVLink probe failed", rc);
            int pairs = 0;
            double rmin = min_pair(&pairs);
            out["nvidia.com/gpu.probe.nvlink.pairs"] = std::to_string(pairs);
            if (pairs) out["nvidia.com/gpu.probe.nvlink.p2p-read-gbps"] = std::to_string((long long)rmin);
            out["nvidia.com/gpu.probe.nvlink.ring-gbps"] = std::to_string((long long)ring);
            std::fill(m.begin(), m.end(), 0.0);
            rc = probe_p2p_matrix(n, bytes, 2, 10, GFD_PROBE_PUSH, m.data(), &ring_w);
            if (rc) fail("NVLink write probe failed", rc);
            int wpairs = 0;
            double wmin = min_pair(&wpairs);
            if (wpairs) out["nvidia.com/gpu.probe.nvlink.p2p-write-gbps"] = std::to_string((long long)wmin);
            // and what the copy engine gets over the same links under the same schedule (cudaMemcpy between peers: no kernel
            // of ours) -- the figure a framework's peer copies will see, and the yardstick for the two kernel figures above
            std::fill(m.begin(), m.end(), 0.0);
            double ring_d = 0.0;
            rc = probe_p2p_matrix(n, bytes, 2, 10, GFD_PROBE_DMA, m.data(), &ring_d);
            if (rc) fail("NVLink copy-engine probe failed", rc);
            int dpairs = 0;
            double dmin = min_pair(&dpairs);
            if (dpairs) out["nvidia.com/gpu.probe.nvlink.p2p-dma-gbps"] = std::to_string((long long)dmin);
        }
    }
    return out;
}

// ---- output file (internal/lm/labels.go:37-114, cmd/gpu-feature-discovery/main.go:212-232) ------------
namespace {
std::string abs_path(const std::string &p) {  // filepath.Abs: Join(cwd, p) + Clean
    std::string full = p;
    if (p.empty() || p[0] != '/') {
        char cwd[PATH_MAX];
        if (!getcwd(cwd, sizeof cwd)) fail(std::string("failed to retrieve absolute path of output file: getwd: ") + strerror(errno));
        full = std::string(cwd) + "/" + p;
    }
    std::vector<std::string> parts;  // filepath.Clean
    size_t i = 0;
    while (i <= full.size()) {
        size_t e = full.find('/', i);
        if (e == std::string::npos) e = full.size();
        std::string seg = full.substr(i, e - i);
        if (seg == "..") { if (!parts.empty()) parts.pop_back(); }
        else if (!seg.empty() && seg != ".") parts.push_back(seg);
        i = e + 1;
    }
    std::string out;
    for (auto &s : parts) out += "/" + s;
    return out.empty() ? "/" : out;
}
std::string dir_of(const std::string &abs) {
    size_t s = abs.rfind('/');
    return s == 0 ? "/" : abs.substr(0, s);
}
int mkdir_p(const std::string &path) {
    std::string cur;
    size_t i = 0;
    while (i < path.size()) {
        size_t e = path.find('/', i + 1);
        if (e == std::string::npos) e = path.size();
        cur = path.substr(0, e);
        if (mkdir(cur.c_str(), 0777) != 0 && errno != EEXIST) return -1;
        i = e;
    }
    struct stat st;
    if (stat(path.c_str(), &st) != 0 || !S_ISDIR(st.st_mode)) { errno = ENOTDIR; return -1; }
    return 0;
}
void remove_all(const std::string &path) {
    DIR *d = opendir(path.c_str());
    if (d) {
        while (struct dirent *e = readdir(d)) {
            if (!strcmp(e->d_name, ".") || !strcmp(e->d_name, "..")) continue;
            std::string p = path + "/" + e->d_name;
            struct stat st;
            if (lstat(p.c_str(), &st) == 0 && S_ISDIR(st.st_mode)) remove_all(p);
            else unlink(p.c_str());
        }
        closedir(d);
        rmdir(path.c_str());
    } else {
        unlink(path.c_str());
    }
}
}  // namespace

void write_labels(const Labels &labels, const std::string &path) {
    std::string body;
    for (const auto &kv : labels) body += kv.first + "=" + kv.second + "\n";  // Labels.WriteTo, keys sorted
    if (path.empty()) {
        if (fwrite(body.data(), 1, body.size(), stdout) != body.size() || fflush(stdout) != 0) fail(std::string("error writing labels to stdout: ") + strerror(errno));
        return;
    }
    const std::string pre = "error atomically writing file '" + path + "': ";
    std::string abs = abs_path(path), tmp_dir = dir_of(abs) + "/gfd-tmp";
    if (mkdir_p(tmp_dir) != 0) fail(pre + "failed to create temporary directory: mkdir " + tmp_dir + ": " + strerror(errno));
    std::string tmpl = tmp_dir + "/gfd-XXXXXX";
    std::vector<char> name(tmpl.begin(), tmpl.end());
    name.push_back(0);
    int fd = mkstemp(name.data());  // ioutil.TempFile(tmpDir, "gfd-"): created 0600
    if (fd < 0) {
        std::string e = strerror(errno);
        remove_all(tmp_dir);
        fail(pre + "fail to create temporary output file: " + e);
    }
    auto bail = [&](const std::string &msg) {
        close(fd);
        unlink(name.data());
        remove_all(tmp_dir);
        fail(pre + msg);
    };
    size_t off = 0;
    while (off < body.size()) {
        ssize_t n = write(fd, body.data() + off, body.size() - off);
        if (n < 0) {
            if (errno == EINTR) continue;
            bail(std::string("error writing temporary file '") + name.data() + "': " + strerror(errno));
        }
        off += (size_t)n;
    }
    if (rename(name.data(), path.c_str()) != 0) bail("error moving temporary file to '" + path + "': " + strerror(errno));
    if (chmod(path.c_str(), 0644) != 0) {
        std::string e = strerror(errno);
        close(fd);
        remove_all(tmp_dir);
        fail(pre + "error setting permissions on '" + path + "': " + e);
    }
    close(fd);
}

void remove_output_file(const std::string &path) {
    std::string abs = abs_path(path), tmp_dir = dir_of(abs) + "/gfd-tmp";
    remove_all(tmp_dir);
    if (unlink(abs.c_str()) != 0) fail("failed to remove output file: remove " + abs + ": " + strerror(errno));
}

}  // namespace gfdlm
