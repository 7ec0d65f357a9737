// Internal declarations shared by the translation units of libgfd_b200.so.
#pragma once
#include <atomic>
#include <cstdint>
#include <map>
#include <memory>
#include <string>
#include <utility>
#include <vector>

#define NVML_NO_UNVERSIONED_FUNC_DEFS
#include <nvml.h>

#include "gfd_b200.h"

#define GFD_API extern "C" __attribute__((visibility("default")))

namespace gfd {

// Function table resolved from libnvidia-ml.so.1 with dlsym -- the ~24 symbols the reference reaches
// (SURVEY.md 8(b)).  A missing symbol stays null and its call site reports NVML_ERROR_FUNCTION_NOT_FOUND.
struct NvmlApi {
    void *lib = nullptr;
    nvmlReturn_t (*Init_v2)() = nullptr;
    nvmlReturn_t (*Init_v1)() = nullptr;  // "nvmlInit": go-nvml falls back to it when _v2 is absent (init.go:127-131)
    nvmlReturn_t (*Shutdown)() = nullptr;
    const char *(*ErrorString)(nvmlReturn_t) = nullptr;
    nvmlReturn_t (*SystemGetDriverVersion)(char *, unsigned) = nullptr;
    nvmlReturn_t (*SystemGetCudaDriverVersion)(int *) = nullptr;
    nvmlReturn_t (*DeviceGetCount)(unsigned *) = nullptr;           // _v2 preferred, v1 fallback
    nvmlReturn_t (*DeviceGetHandleByIndex)(unsigned, nvmlDevice_t *) = nullptr;  // _v2 preferred
    nvmlReturn_t (*DeviceGetName)(nvmlDevice_t, char *, unsigned) = nullptr;
    nvmlReturn_t (*DeviceGetMemoryInfo)(nvmlDevice_t, nvmlMemory_t *) = nullptr;  // v1 struct, as the reference
    nvmlReturn_t (*DeviceGetCudaComputeCapability)(nvmlDevice_t, int *, int *) = nullptr;
    nvmlReturn_t (*DeviceGetMigMode)(nvmlDevice_t, unsigned *, unsigned *) = nullptr;
    nvmlReturn_t (*DeviceGetMaxMigDeviceCount)(nvmlDevice_t, unsigned *) = nullptr;
    nvmlReturn_t (*DeviceGetMigDeviceHandleByIndex)(nvmlDevice_t, unsigned, nvmlDevice_t *) = nullptr;
    nvmlReturn_t (*DeviceIsMigDeviceHandle)(nvmlDevice_t, unsigned *) = nullptr;
    nvmlReturn_t (*DeviceGetDeviceHandleFromMigDeviceHandle)(nvmlDevice_t, nvmlDevice_t *) = nullptr;
    nvmlReturn_t (*DeviceGetAttributes)(nvmlDevice_t, nvmlDeviceAttributes_t *) = nullptr;  // _v2 preferred
    nvmlReturn_t (*DeviceGetGpuInstanceId)(nvmlDevice_t, unsigned *) = nullptr;
    nvmlReturn_t (*DeviceGetComputeInstanceId)(nvmlDevice_t, unsigned *) = nullptr;
    nvmlReturn_t (*DeviceGetGpuInstanceById)(nvmlDevice_t, unsigned, nvmlGpuInstance_t *) = nullptr;
    nvmlReturn_t (*GpuInstanceGetInfo)(nvmlGpuInstance_t, nvmlGpuInstanceInfo_t *) = nullptr;
    nvmlReturn_t (*GpuInstanceGetComputeInstanceById)(nvmlGpuInstance_t, unsigned, nvmlComputeInstance_t *) = nullptr;
    nvmlReturn_t (*ComputeInstanceGetInfo)(nvmlComputeInstance_t, nvmlComputeInstanceInfo_t *) = nullptr;  // _v2 preferred
    nvmlReturn_t (*DeviceGetGpuInstanceProfileInfo)(nvmlDevice_t, unsigned, nvmlGpuInstanceProfileInfo_t *) = nullptr;
    nvmlReturn_t (*GpuInstanceGetComputeInstanceProfileInfo)(nvmlGpuInstance_t, unsigned, unsigned,
                                                             nvmlComputeInstanceProfileInfo_t *) = nullptr;
    nvmlReturn_t (*DeviceGetUUID)(nvmlDevice_t, char *, unsigned) = nullptr;  // probes only
};

// One cached answer: the NVML (or shim) status it produced and, when 0, the value.
template <class T>
struct Cached {
    bool done = false;
    int rc = 0;
    T v{};
};

struct GiEntry {  // one GPU instance of a parent device
    Cached<nvmlGpuInstance_t> handle;
    Cached<unsigned> profile_id;                              // nvmlGpuInstanceGetInfo().profileId
    std::map<unsigned, Cached<unsigned>> ci_profile_info_id;  // CI profile idx -> .id
};

}  // namespace gfd

struct gfd_dev {
    gfd_mgr *mgr = nullptr;
    bool is_mig = false;
    nvmlDevice_t h = nullptr;
    gfd_dev *parent = nullptr;  // for MIG devices: the device they were enumerated from

    gfd::Cached<std::string> name;          // full GPU: model; MIG: dotted profile string
    gfd::Cached<uint64_t> mem_total_bytes;  // nvmlDeviceGetMemoryInfo(v1).total
    gfd::Cached<std::pair<int, int>> cc;
    struct MigMode { bool supported = false; unsigned current = 0, pending = 0; };
    gfd::Cached<MigMode> mig_mode;
    gfd::Cached<std::vector<std::unique_ptr<gfd_dev>>> migs;
    gfd::Cached<std::string> uuid;
    // parent side of profile resolution
    std::map<unsigned, gfd::Cached<unsigned>> gi_profile_info_id;  // GI profile idx -> .id
    std::map<unsigned, gfd::GiEntry> gis;                           // GI id -> entry
    // MIG side
    gfd::Cached<nvmlDeviceAttributes_t> attrs;
    gfd::Cached<nvmlDevice_t> parent_handle;  // nvmlDeviceGetDeviceHandleFromMigDeviceHandle
    gfd::Cached<std::unique_ptr<gfd_dev>> parent_dev;
};

struct gfd_mgr {
    int backend = GFD_BACKEND_NVML;
    unsigned flags = 0;
    bool initialized = false;
    bool attached = false;  // KEEP_ATTACHED: an extra nvmlInit reference is held
    gfd::NvmlApi nv;
    gfd::Cached<std::vector<std::unique_ptr<gfd_dev>>> devices;
    gfd::Cached<std::string> driver_version;  // nvmlSystemGetDriverVersion, once per bracket
    gfd::Cached<int> cuda_version;            // nvmlSystemGetCudaDriverVersion (NVML dlopens libcuda for it: ~0.4 ms)
    // Device objects of the previous bracket.  A handle used after gfd_shutdown is a caller bug, but it must get
    // GFD_ERR_NOT_INITIALIZED (what NVML answers for a stale handle), not a use-after-free: the objects stay
    // allocated until the next gfd_init or gfd_mgr_free.
    std::vector<std::unique_ptr<gfd_dev>> retired;
    std::atomic<uint64_t> nvml_calls{0}, dl_probes{0};
    // CUDA (Tegra) backend state lives in gfd_cuda_mgr.cpp
    void *cuda_state = nullptr;
};

namespace gfd {
int copy_out(const std::string &s, char *buf, size_t len);
// implemented in gfd_mig_profile.cpp
int mig_profile_string(int gi_profile_id, int ci_profile_id, uint64_t mig_mb, uint64_t dev_bytes, std::string *out,
                       bool current_tables = false);
// implemented in gfd_cuda_mgr.cpp
int cuda_mgr_init(gfd_mgr *m);
int cuda_mgr_shutdown(gfd_mgr *m);
int cuda_mgr_driver_version(gfd_mgr *m, unsigned *major, unsigned *minor);
int cuda_mgr_enumerate(gfd_mgr *m);
int cuda_dev_name(gfd_dev *d);
int cuda_dev_mem(gfd_dev *d);
int cuda_dev_cc(gfd_dev *d);
void cuda_mgr_free(gfd_mgr *m);
const char *nvml_error_string_if_loaded(int rc);
}  // namespace gfd
