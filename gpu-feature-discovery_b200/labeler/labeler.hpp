// Label builders of the B200-native labeler, driving the C ABI of libgfd_b200.so.
//
// Same observable contract as the reference's internal/lm + internal/mig packages; citations are on the
// definitions in labeler.cpp.  Internals differ on purpose: labels live in one sorted map (deterministic
// output instead of Go's random map order), devices are walked once, and all NVML answers come from the
// shim's per-bracket cache.
#pragma once
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "config.hpp"
#include "gfd_b200.h"

namespace gfdlm {

using Labels = std::map<std::string, std::string>;  // lm.Labels (internal/lm/labels.go:29), sorted

struct PassError : std::runtime_error {  // a Go `error` returned from run(): "Error: ..." + exit 1
    using std::runtime_error::runtime_error;
};

// resource.Manager with the init-error fallback (internal/resource/factory.go:27-74, fallback.go:23-64)
class Manager {
  public:
    Manager(int backend, unsigned flags, bool fallback_to_null);
    ~Manager();
    Manager(const Manager &) = delete;
    Manager &operator=(const Manager &) = delete;

    void init();      // never throws when fallback_to_null: swaps to the null manager instead
    void shutdown();  // errors are the caller's to ignore (`defer manager.Shutdown()`)
    gfd_mgr *active() const { return fell_back_ ? null_ : primary_; }
    std::vector<gfd_dev *> devices();
    std::string driver_version();
    void cuda_driver_version(unsigned *major, unsigned *minor);
    void prefetch(int threads);

  private:
    gfd_mgr *primary_ = nullptr, *null_ = nullptr;
    bool fallback_ = false, fell_back_ = false;
};

// resource.NewManager: platform detection + backend pick, with the reference's log lines
std::unique_ptr<Manager> new_manager(const gfdcfg::Config &config);

// lm.NewNVMLLabeler ... (everything between Init and the deferred Shutdown); may throw PassError.
struct PhaseTimes {  // wall-clock of the NVML bracket's phases, ms (--timing)
    double init_ms = 0;      // dlopen + nvmlInit_v2
    double attach_ms = 0;    // device walk (+ per-GPU prefetch threads): where the driver attaches each GPU
    double labels_ms = 0;    // label building from the cached answers
    double shutdown_ms = 0;  // nvmlShutdown + dlclose
};
Labels nvml_labels(Manager &mgr, const gfdcfg::Config &config, PhaseTimes *times = nullptr);
// lm.NewVGPULabeler(...).Labels()
Labels vgpu_labels(const gfdcfg::Config &config);
// Probe entry points of libgfd_b200.so, bound lazily (probe_loader.cpp); throw PassError if the library is missing.
int probe_device_count(int *n);
int probe_hbm(int dev, size_t bytes, int warmup, int iters, int mode, gfd_probe_result *out);
int probe_p2p(int src, int dst, size_t bytes, int warmup, int iters, int mode, gfd_probe_result *out);
int probe_p2p_matrix(int n, size_t bytes, int warmup, int iters, int mode, double *matrix, double *ring);
// NEW opt-in probe labels (empty unless --probe-hbm / --probe-nvlink)
Labels probe_labels(const gfdcfg::Config &config);

// Labels.WriteToFile (internal/lm/labels.go:37-114): "" = stdout, else atomic replace via <dir>/gfd-tmp
void write_labels(const Labels &labels, const std::string &path);
// removeOutputFile (cmd/gpu-feature-discovery/main.go:212-232)
void remove_output_file(const std::string &path);

}  // namespace gfdlm
