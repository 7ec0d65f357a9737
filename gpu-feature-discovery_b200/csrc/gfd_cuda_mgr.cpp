// CUDA-driver-API resource.Manager (Tegra / no-NVML systems) for libgfd_b200.so.
//
// Replaces internal/resource/cuda-lib.go:25-88, cuda-device.go:25-98 and the cgo binding
// internal/cuda/{api.go:23-118, cuda.go:103-109}: the same 7 libcuda.so.1 entry points, resolved with
// dlsym at Init (the reference leaves them unresolved at link time and lets ld.so bind them lazily).
// Reachable only when NVML is absent and the host is a Tegra (factory.go:59-70); never on a B200 node.
#include <dlfcn.h>

#include <cstring>

#include "gfd_internal.h"

namespace {
struct CudaState {
    void *lib = nullptr;
    int (*cuInit)(unsigned) = nullptr;
    int (*cuDriverGetVersion)(int *) = nullptr;
    int (*cuDeviceGet)(int *, int) = nullptr;
    int (*cuDeviceGetAttribute)(int *, int, int) = nullptr;
    int (*cuDeviceGetCount)(int *) = nullptr;
    // NB: the reference binds the UNVERSIONED symbol (internal/cuda/cuda.go:108), i.e. the legacy entry point
    // whatever width it writes, into a zeroed uint64.  Bound the same way here so the value is the same.
    int (*cuDeviceTotalMem)(void *, int) = nullptr;
    int (*cuDeviceGetName)(char *, int, int) = nullptr;
};
CudaState *st(gfd_mgr *m) { return static_cast<CudaState *>(m->cuda_state); }
int cu(int r) { return r == 0 ? 0 : GFD_ERR_CUDA + r; }
const int kErrorUnknown = 999;
}  // namespace

namespace gfd {

int cuda_mgr_init(gfd_mgr *m) {
    // cuda.Init (internal/cuda/api.go:32-44)
    if (!m->cuda_state) m->cuda_state = new CudaState();
    CudaState *s = st(m);
    s->lib = dlopen("libcuda.so.1", RTLD_LAZY | RTLD_GLOBAL);
    if (!s->lib) return cu(kErrorUnknown);
    *(void **)&s->cuInit = dlsym(s->lib, "cuInit");
    if (!s->cuInit) return cu(kErrorUnknown);
    *(void **)&s->cuDriverGetVersion = dlsym(s->lib, "cuDriverGetVersion");
    *(void **)&s->cuDeviceGet = dlsym(s->lib, "cuDeviceGet");
    *(void **)&s->cuDeviceGetAttribute = dlsym(s->lib, "cuDeviceGetAttribute");
    *(void **)&s->cuDeviceGetCount = dlsym(s->lib, "cuDeviceGetCount");
    *(void **)&s->cuDeviceTotalMem = dlsym(s->lib, "cuDeviceTotalMem");
    *(void **)&s->cuDeviceGetName = dlsym(s->lib, "cuDeviceGetName");
    return cu(s->cuInit(0));
}

int cuda_mgr_shutdown(gfd_mgr *m) {
    // cuda.Shutdown (api.go:47-55)
    CudaState *s = st(m);
    if (!s || !s->lib) return 0;
    int r = dlclose(s->lib);
    s->lib = nullptr;
    return r == 0 ? 0 : cu(kErrorUnknown);
}

void cuda_mgr_free(gfd_mgr *m) {
    delete st(m);
    m->cuda_state = nullptr;
}

int cuda_mgr_driver_version(gfd_mgr *m, unsigned *major, unsigned *minor) {
    // cuda-lib.go:53-63 -- note `% 100`, unlike the NVML manager's `% 1000`
    CudaState *s = st(m);
    if (!s->cuDriverGetVersion) return cu(kErrorUnknown);
    int v = 0;
    int r = s->cuDriverGetVersion(&v);
    if (r) return cu(r);
    uint64_t uv = (uint64_t)(int64_t)v;  // uint(version) on a 64-bit Go
    *major = (unsigned)(uv / 1000);
    *minor = (unsigned)(uv % 100 / 10);
    return 0;
}

int cuda_mgr_enumerate(gfd_mgr *m) {
    // cudaLib.GetDevices (cuda-lib.go:35-50)
    if (m->devices.done) return m->devices.rc;
    m->devices.done = true;
    CudaState *s = st(m);
    if (!s->cuDeviceGetCount || !s->cuDeviceGet) return m->devices.rc = cu(kErrorUnknown);
    int count = 0;
    int r = s->cuDeviceGetCount(&count);
    if (r) return m->devices.rc = cu(r);
    for (int i = 0; i < count; i++) {
        int dev = 0;
        r = s->cuDeviceGet(&dev, i);
        if (r) { m->devices.v.clear(); return m->devices.rc = cu(r); }
        std::unique_ptr<gfd_dev> d(new gfd_dev());
        d->mgr = m;
        d->h = (nvmlDevice_t)(intptr_t)(dev + 1);  // CUdevice ordinal, biased so the handle is never null
        m->devices.v.push_back(std::move(d));
    }
    return m->devices.rc = 0;
}

static int ordinal(gfd_dev *d) { return (int)(intptr_t)d->h - 1; }

int cuda_dev_name(gfd_dev *d) {
    // cuda.DeviceGetName: 96-byte buffer (api.go:94-101)
    if (!d->name.done) {
        d->name.done = true;
        CudaState *s = st(d->mgr);
        char buf[96];
        memset(buf, 0, sizeof buf);
        int r = s->cuDeviceGetName ? s->cuDeviceGetName(buf, (int)sizeof buf, ordinal(d)) : kErrorUnknown;
        d->name.rc = cu(r);
        size_t n = 0;
        while (n < sizeof buf && buf[n]) n++;
        d->name.v.assign(buf, n);
    }
    return d->name.rc;
}

int cuda_dev_mem(gfd_dev *d) {
    if (!d->mem_total_bytes.done) {
        d->mem_total_bytes.done = true;
        CudaState *s = st(d->mgr);
        uint64_t bytes = 0;
        int r = s->cuDeviceTotalMem ? s->cuDeviceTotalMem(&bytes, ordinal(d)) : kErrorUnknown;
        d->mem_total_bytes.rc = cu(r);
        d->mem_total_bytes.v = bytes;
    }
    return d->mem_total_bytes.rc;
}

int cuda_dev_cc(gfd_dev *d) {
    if (!d->cc.done) {
        d->cc.done = true;
        CudaState *s = st(d->mgr);
        int major = 0, minor = 0;
        int r = s->cuDeviceGetAttribute ? s->cuDeviceGetAttribute(&major, 75, ordinal(d)) : kErrorUnknown;
        if (r == 0) r = s->cuDeviceGetAttribute(&minor, 76, ordinal(d));
        d->cc.rc = cu(r);
        d->cc.v = {major, minor};
    }
    return d->cc.rc;
}

}  // namespace gfd
