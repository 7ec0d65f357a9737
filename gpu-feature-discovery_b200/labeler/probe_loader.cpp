// Lazy binding of the opt-in GPU probes.
//
// The labeler binary carries the host-only part of the C ABI (NVML / libcuda managers, MIG arithmetic, vGPU parser:
// csrc/gfd_nvml.cpp, gfd_mig_profile.cpp, gfd_cuda_mgr.cpp) linked in statically, so the default labeling pass pays
// neither the dynamic-loader cost of a 2 MB CUDA library nor the CUDA runtime's start-up (measured: 3.3 -> 1.5 ms of
// process start-up).  The sm_100a kernels live in libgfd_b200.so next to the executable; it is dlopen()ed only when a
// probe flag asks for GPU work.  No fallback: if it cannot be loaded the pass fails with the loader's message.
#include <dlfcn.h>
#include <unistd.h>

#include <climits>
#include <string>

#include "labeler.hpp"

namespace gfdlm {

namespace {
struct ProbeApi {
    int (*device_count)(int *) = nullptr;
    int (*probe_hbm)(int, size_t, int, int, int, gfd_probe_result *) = nullptr;
    int (*probe_p2p)(int, int, size_t, int, int, int, gfd_probe_result *) = nullptr;
    int (*probe_p2p_matrix)(int, size_t, int, int, int, double *, double *) = nullptr;
    bool loaded = false;
};
ProbeApi g_api;

std::string exe_dir() {
    char buf[PATH_MAX];
    ssize_t n = readlink("/proc/self/exe", buf, sizeof buf - 1);
    if (n <= 0) return ".";
    buf[n] = 0;
    std::string p(buf);
    size_t s = p.rfind('/');
    return s == std::string::npos ? "." : p.substr(0, s);
}

void load() {
    if (g_api.loaded) return;
    std::string path = exe_dir() + "/libgfd_b200.so";
    void *lib = dlopen(path.c_str(), RTLD_NOW | RTLD_LOCAL);
    if (!lib) throw PassError(std::string("probe requested but the probe library cannot be loaded: ") + dlerror());
    *(void **)&g_api.device_count = dlsym(lib, "gfd_cuda_device_count");
    *(void **)&g_api.probe_hbm = dlsym(lib, "gfd_probe_hbm");
    *(void **)&g_api.probe_p2p = dlsym(lib, "gfd_probe_p2p");
    *(void **)&g_api.probe_p2p_matrix = dlsym(lib, "gfd_probe_p2p_matrix");
    if (!g_api.device_count || !g_api.probe_hbm || !g_api.probe_p2p || !g_api.probe_p2p_matrix)
        throw PassError("probe requested but " + path + " does not export the probe entry points of include/gfd_b200.h");
    g_api.loaded = true;
}
}  // namespace

int probe_device_count(int *n) { load(); return g_api.device_count(n); }
int probe_hbm(int dev, size_t bytes, int warmup, int iters, int mode, gfd_probe_result *out) {
    load();
    return g_api.probe_hbm(dev, bytes, warmup, iters, mode, out);
}
int probe_p2p(int src, int dst, size_t bytes, int warmup, int iters, int mode, gfd_probe_result *out) {
    load();
    return g_api.probe_p2p(src, dst, bytes, warmup, iters, mode, out);
}

int probe_p2p_matrix(int n, size_t bytes, int warmup, int iters, int mode, double *matrix, double *ring) {
    load();
    return g_api.probe_p2p_matrix(n, bytes, warmup, iters, mode, matrix, ring);
}

}  // namespace gfdlm
