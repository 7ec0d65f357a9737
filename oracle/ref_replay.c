/*
 * ref_replay -- CPU BASELINE + second oracle.  TEST/BENCH INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * The reference (Go + cgo) cannot be built on this image (no Go toolchain), so the "reference GFD timed on the
 * same host" that BASELINE.json asks for is replaced by this: a plain-C, CALL-SEQUENCE-FAITHFUL restatement of
 * one reference labeling pass.  It issues, in the same order, every NVML call and every dlopen/dlsym/dlclose
 * cycle the reference issues:
 *     Init = dlopen + 21 versioned-symbol dlsym probes + nvmlInit_v2     V/go-nvml/nvml/init.go:33-44,127-218
 *     4 full device walks per pass (5 with a MIG strategy)                internal/lm/nvml.go:35,113; mig-strategy.go:46; internal/mig/mig.go:46
 *     one dlopen/dlsym/dlclose per IsMigCapable / IsMigEnabled            V/nvlib/nvlib/device/device.go:326-338
 *     GetProfile per MIG-device wrapper (cache per wrapper only)          V/nvlib/nvlib/device/mig_device.go:52-148
 *     Shutdown = nvmlShutdown + dlclose                                    init.go:60-72
 * and builds the same labels (internal/lm/ *.go) and writes them the same way (internal/lm/labels.go:68-114).
 * It has none of Go's runtime cost (cgo transitions, GC, scheduler), so its time is a LOWER bound on the
 * reference's.  tests/test_ref_replay.py checks (a) its NVML call trace equals the Python oracle's call for
 * call and (b) its label file equals the oracle's, on every fixture.
 *
 *   ref_replay [--mig-strategy S] [--no-timestamp] [--machine-type-file F] [-o OUT] [--passes N] [--pci-root DIR]
 * prints one JSON line: {"passes":N,"ms":[...],"init_ms":[...],"shutdown_ms":[...],"nvml_calls":C,"dl_cycles":D,"ok":true}
 *
 * Map-order caveat: the reference ranges over Go maps; like the product and the oracle this uses sorted order.
 * Not restated here: config file / time-slicing (replicas are always 1 / 0) and the vGPU sysfs scan beyond
 * vendor-id filtering (non-NVIDIA functions are skipped exactly as the reference does; an NVIDIA function's
 * capability walk follows internal/vgpu/pciutil.go:115-149, vgpu.go:92-153).
 */
#define _GNU_SOURCE
#include <dirent.h>
#include <dlfcn.h>
#include <errno.h>
#include <fcntl.h>
#include <limits.h>
#include <math.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#define NVML_NO_UNVERSIONED_FUNC_DEFS
#include <nvml.h>

/* ---- counters ---- */
static unsigned long g_calls, g_dl;
static double g_init_ms;

static double now_ms(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e3 + ts.tv_nsec / 1e6;
}

/* ---- NVML binding (lazy dlsym per call site, as cgo's lazy binding does) ---- */
static void *g_lib;
#define NV(ret, name, ...) ((g_calls++), ((ret(*)(__VA_ARGS__))must_sym(name)))
static void *must_sym(const char *name) {
    void *p = dlsym(g_lib, name);
    if (!p) { fprintf(stderr, "ref_replay: missing symbol %s\n", name); exit(3); }
    return p;
}

static int fail(const char *what, int rc) {
    fprintf(stderr, "Error: %s: %d\n", what, rc);
    return 1;
}

static const char *kVersioned[] = {  /* updateVersionedSymbols (init.go:127-218): 21 lookups */
    "nvmlInit_v2", "nvmlDeviceGetPciInfo_v2", "nvmlDeviceGetPciInfo_v3", "nvmlDeviceGetCount_v2", "nvmlDeviceGetHandleByIndex_v2",
    "nvmlDeviceGetHandleByPciBusId_v2", "nvmlDeviceGetNvLinkRemotePciInfo_v2", "nvmlDeviceGetGridLicensableFeatures_v2",
    "nvmlDeviceGetGridLicensableFeatures_v3", "nvmlDeviceGetGridLicensableFeatures_v4", "nvmlEventSetWait_v2",
    "nvmlDeviceGetAttributes_v2", "nvmlComputeInstanceGetInfo_v2", "nvmlDeviceGetComputeRunningProcesses_v2",
    "nvmlDeviceGetComputeRunningProcesses_v3", "nvmlDeviceGetGraphicsRunningProcesses_v2", "nvmlDeviceGetGraphicsRunningProcesses_v3",
    "nvmlDeviceGetMPSComputeRunningProcesses_v2", "nvmlDeviceGetMPSComputeRunningProcesses_v3",
    "nvmlDeviceGetGpuInstancePossiblePlacements_v2", "nvmlVgpuInstanceGetLicenseInfo_v2"};

static int ref_init(void) {
    double t0 = now_ms();
    g_lib = dlopen("libnvidia-ml.so.1", RTLD_LAZY | RTLD_GLOBAL);
    if (!g_lib) return 12;
    for (size_t i = 0; i < sizeof kVersioned / sizeof *kVersioned; i++) { dlerror(); (void)dlsym(g_lib, kVersioned[i]); (void)dlerror(); }
    int rc = NV(nvmlReturn_t, "nvmlInit_v2", void)();
    g_init_ms = now_ms() - t0;
    return rc;
}

static double g_shutdown_ms;
static void ref_shutdown(void) {
    double t0 = now_ms();
    if (NV(nvmlReturn_t, "nvmlShutdown", void)() == NVML_SUCCESS) dlclose(g_lib);
    g_shutdown_ms = now_ms() - t0;
}

/* nvmlLookupSymbol (device.go:326-338): a full dlopen/dlsym/dlclose cycle per call */
static int lookup_symbol(const char *sym) {
    g_dl++;
    void *lib = dlopen("libnvidia-ml.so.1", RTLD_LAZY | RTLD_GLOBAL);
    if (!lib) return 0;
    dlerror();
    (void)dlsym(lib, sym);
    int ok = dlerror() == NULL;
    dlclose(lib);
    return ok;
}

/* ---- labels: ordered map emulation ---- */
typedef struct { char k[128]; char v[160]; } kv_t;
typedef struct { kv_t *e; int n, cap; } labels_t;
static void put(labels_t *l, const char *k, const char *v) {
    for (int i = 0; i < l->n; i++)
        if (!strcmp(l->e[i].k, k)) { snprintf(l->e[i].v, sizeof l->e[i].v, "%s", v); return; }
    if (l->n == l->cap) { l->cap = l->cap ? l->cap * 2 : 64; l->e = realloc(l->e, (size_t)l->cap * sizeof(kv_t)); }
    snprintf(l->e[l->n].k, sizeof l->e[l->n].k, "%s", k);
    snprintf(l->e[l->n].v, sizeof l->e[l->n].v, "%s", v);
    l->n++;
}
static void putf(labels_t *l, const char *res, const char *suffix, const char *fmt, ...) __attribute__((format(printf, 4, 5)));
static void putf(labels_t *l, const char *res, const char *suffix, const char *fmt, ...) {
    char k[160], v[160];
    snprintf(k, sizeof k, "%s.%s", res, suffix);
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(v, sizeof v, fmt, ap);
    va_end(ap);
    put(l, k, v);
}
static void merge(labels_t *dst, const labels_t *src) {
    for (int i = 0; i < src->n; i++) put(dst, src->e[i].k, src->e[i].v);
}
static int kvcmp(const void *a, const void *b) { return strcmp(((const kv_t *)a)->k, ((const kv_t *)b)->k); }

static void dashes(char *s) { for (; *s; s++) if (*s == ' ') *s = '-'; }

/* ---- device walk (devicelib.VisitDevices, device.go:206-236) ---- */
#define MAXDEV 64
#define MAXMIG 16
static int get_devices(nvmlDevice_t *out, int *n) {
    unsigned count = 0;
    int rc = NV(nvmlReturn_t, "nvmlDeviceGetCount_v2", unsigned *)(&count);
    if (rc) return rc;
    {   /* bench-only: bench.py --gpus N labels the first N indices in BOTH arms (same variable as the product's shim) */
        const char *cap = getenv("GFD_B200_BENCH_MAX_DEVICES");
        long c = cap && *cap ? strtol(cap, NULL, 10) : 0;
        if (c > 0 && (unsigned long)c < count) count = (unsigned)c;
    }
    *n = 0;
    for (unsigned i = 0; i < count; i++) {
        nvmlDevice_t h;
        rc = NV(nvmlReturn_t, "nvmlDeviceGetHandleByIndex_v2", unsigned, nvmlDevice_t *)(i, &h);
        if (rc) return rc;
        char name[96] = {0};
        rc = NV(nvmlReturn_t, "nvmlDeviceGetName", nvmlDevice_t, char *, unsigned)(h, name, 96);  /* isSkipped */
        if (rc) return rc;
        if (!strcmp(name, "NVIDIA DGX Display")) continue;
        if (*n < MAXDEV) out[(*n)++] = h;
    }
    return 0;
}

static int mig_mode(nvmlDevice_t h, int *capable, int *enabled) {  /* device.go:55-88 */
    *capable = *enabled = 0;
    if (!lookup_symbol("nvmlDeviceGetMigMode")) return 0;
    unsigned cur = 0, pend = 0;
    int rc = NV(nvmlReturn_t, "nvmlDeviceGetMigMode", nvmlDevice_t, unsigned *, unsigned *)(h, &cur, &pend);
    if (rc == NVML_ERROR_NOT_SUPPORTED) return 0;
    if (rc) return rc;
    *capable = 1;
    *enabled = cur == NVML_DEVICE_MIG_ENABLE;
    return 0;
}

static int get_mig_devices(nvmlDevice_t h, nvmlDevice_t *out, int *n) {  /* device.go:91-118 + mig_device.go:40-49 */
    unsigned count = 0;
    int rc = NV(nvmlReturn_t, "nvmlDeviceGetMaxMigDeviceCount", nvmlDevice_t, unsigned *)(h, &count);
    if (rc) return rc;
    *n = 0;
    for (unsigned i = 0; i < count; i++) {
        nvmlDevice_t m;
        rc = NV(nvmlReturn_t, "nvmlDeviceGetMigDeviceHandleByIndex", nvmlDevice_t, unsigned, nvmlDevice_t *)(h, i, &m);
        if (rc == NVML_ERROR_NOT_FOUND || rc == NVML_ERROR_INVALID_ARGUMENT) continue;
        if (rc) return rc;
        unsigned is_mig = 0;
        rc = NV(nvmlReturn_t, "nvmlDeviceIsMigDeviceHandle", nvmlDevice_t, unsigned *)(m, &is_mig);
        if (rc) return rc;
        if (!is_mig) return -1;
        if (*n < MAXMIG) out[(*n)++] = m;
    }
    return 0;
}

static double go_round(double x) {  /* math.Round */
    double t = trunc(x);
    if (fabs(x - t) >= 0.5) return t + copysign(1.0, x);
    return t;
}

static unsigned long long mig_memory_gb(unsigned long long total, unsigned long long mig_mb) {  /* mig_profile.go:324-332 */
    volatile double frac = ((double)mig_mb * (1024.0 * 1024.0)) / (double)total;
    frac = ceil(frac * 8.0) / 8.0;
    volatile double total_gb = (double)((total + (1ull << 30) - 1) / (1ull << 30));
    volatile double p = frac * total_gb;
    return (unsigned long long)go_round(p);
}

static int slices_of(int id, int is_gi) {
    switch (id) {
    case 0: return 1; case 1: return 2; case 2: return 3; case 3: return 4; case 4: return 7; case 5: return 8; case 6: return 6;
    case 7: return is_gi ? 1 : 0;
    default: return 0;
    }
}

/* migdevice.GetProfile (mig_device.go:52-148) -> dotted name; 0 ok, <0 shim-level failure, >0 NVML rc */
static int mig_profile_name(nvmlDevice_t m, char *out, size_t n) {
    nvmlDevice_t parent;
    int rc;
    if ((rc = NV(nvmlReturn_t, "nvmlDeviceGetDeviceHandleFromMigDeviceHandle", nvmlDevice_t, nvmlDevice_t *)(m, &parent))) return rc;
    nvmlMemory_t mem;
    if ((rc = NV(nvmlReturn_t, "nvmlDeviceGetMemoryInfo", nvmlDevice_t, nvmlMemory_t *)(parent, &mem))) return rc;
    nvmlDeviceAttributes_t attr;
    if ((rc = NV(nvmlReturn_t, "nvmlDeviceGetAttributes_v2", nvmlDevice_t, nvmlDeviceAttributes_t *)(m, &attr))) return rc;
    unsigned gi_id, ci_id;
    if ((rc = NV(nvmlReturn_t, "nvmlDeviceGetGpuInstanceId", nvmlDevice_t, unsigned *)(m, &gi_id))) return rc;
    if ((rc = NV(nvmlReturn_t, "nvmlDeviceGetComputeInstanceId", nvmlDevice_t, unsigned *)(m, &ci_id))) return rc;
    nvmlGpuInstance_t gi;
    if ((rc = NV(nvmlReturn_t, "nvmlDeviceGetGpuInstanceById", nvmlDevice_t, unsigned, nvmlGpuInstance_t *)(parent, gi_id, &gi))) return rc;
    nvmlComputeInstance_t ci;
    if ((rc = NV(nvmlReturn_t, "nvmlGpuInstanceGetComputeInstanceById", nvmlGpuInstance_t, unsigned, nvmlComputeInstance_t *)(gi, ci_id, &ci))) return rc;
    nvmlGpuInstanceInfo_t gi_info;
    if ((rc = NV(nvmlReturn_t, "nvmlGpuInstanceGetInfo", nvmlGpuInstance_t, nvmlGpuInstanceInfo_t *)(gi, &gi_info))) return rc;
    nvmlComputeInstanceInfo_t ci_info;
    if ((rc = NV(nvmlReturn_t, "nvmlComputeInstanceGetInfo_v2", nvmlComputeInstance_t, nvmlComputeInstanceInfo_t *)(ci, &ci_info))) return rc;
    for (unsigned i = 0; i < 8; i++) {  /* GPU_INSTANCE_PROFILE_COUNT of the reference's header */
        nvmlGpuInstanceProfileInfo_t gp;
        rc = NV(nvmlReturn_t, "nvmlDeviceGetGpuInstanceProfileInfo", nvmlDevice_t, unsigned, nvmlGpuInstanceProfileInfo_t *)(parent, i, &gp);
        if (rc == NVML_ERROR_NOT_SUPPORTED || rc == NVML_ERROR_INVALID_ARGUMENT) continue;
        if (rc) return rc;
        if (gp.id != gi_info.profileId) continue;
        for (unsigned j = 0; j < 7; j++) {  /* COMPUTE_INSTANCE_PROFILE_COUNT */
            nvmlComputeInstanceProfileInfo_t cp;
            rc = NV(nvmlReturn_t, "nvmlGpuInstanceGetComputeInstanceProfileInfo", nvmlGpuInstance_t, unsigned, unsigned,
                    nvmlComputeInstanceProfileInfo_t *)(gi, j, 0, &cp);
            if (rc == NVML_ERROR_NOT_SUPPORTED || rc == NVML_ERROR_INVALID_ARGUMENT) continue;
            if (rc) return rc;
            if (cp.id != ci_info.profileId) continue;
            int g = slices_of((int)i, 1), c = slices_of((int)j, 0);
            if (!g || !c) return -2;
            unsigned long long gb = mig_memory_gb(mem.total, attr.memorySizeMB);
            const char *me = i == 7 ? ".me" : "";  /* "+me" with '+' -> '.' (nvml-mig-device.go:103) */
            if (c == g) snprintf(out, n, "%dg.%llugb%s", g, gb, me);
            else snprintf(out, n, "%dc.%dg.%llugb%s", c, g, gb, me);
            return 0;
        }
    }
    return -3; /* no matching profile IDs found */
}

static const char *arch_family(int major, int minor) {  /* resource.go:261-284 */
    switch (major) {
    case 1: return "tesla"; case 2: return "fermi"; case 3: return "kepler"; case 5: return "maxwell"; case 6: return "pascal";
    case 7: return minor < 5 ? "volta" : "turing";
    case 8: return "ampere"; case 9: return "hopper";
    }
    return "undefined";
}

/* NewGPUResourceLabeler (resource.go:36-73); sharing: replicas 0 when `shared_disabled`, else 1 (no config file here) */
static int gpu_resource_labels(labels_t *l, nvmlDevice_t h, int count, int sharing_disabled) {
    if (!count) return 0;
    char model[96] = {0};
    int rc;
    if ((rc = NV(nvmlReturn_t, "nvmlDeviceGetName", nvmlDevice_t, char *, unsigned)(h, model, 96))) return rc;
    nvmlMemory_t mem;
    if ((rc = NV(nvmlReturn_t, "nvmlDeviceGetMemoryInfo", nvmlDevice_t, nvmlMemory_t *)(h, &mem))) return rc;
    int major = 0, minor = 0;
    if ((rc = NV(nvmlReturn_t, "nvmlDeviceGetCudaComputeCapability", nvmlDevice_t, int *, int *)(h, &major, &minor))) return rc;
    dashes(model);
    if (model[0]) putf(l, "nvidia.com/gpu", "product", "%s", model);
    putf(l, "nvidia.com/gpu", "count", "%d", count);
    putf(l, "nvidia.com/gpu", "replicas", "%d", sharing_disabled ? 0 : 1);
    unsigned long long mb = mem.total / (1024ull * 1024ull);
    if (mb) putf(l, "nvidia.com/gpu", "memory", "%llu", mb);
    if (major) {
        putf(l, "nvidia.com/gpu", "family", "%s", arch_family(major, minor));
        putf(l, "nvidia.com/gpu", "compute.major", "%d", major);
        putf(l, "nvidia.com/gpu", "compute.minor", "%d", minor);
    }
    return 0;
}

typedef struct { char name[96]; nvmlDevice_t dev; int count; } group_t;
static group_t *group_find(group_t *g, int *n, const char *name) {
    for (int i = 0; i < *n; i++) if (!strcmp(g[i].name, name)) return &g[i];
    snprintf(g[*n].name, sizeof g[*n].name, "%s", name);
    g[*n].count = 0;
    g[*n].dev = NULL;
    return &g[(*n)++];
}
static int groupcmp(const void *a, const void *b) { return strcmp(((const group_t *)a)->name, ((const group_t *)b)->name); }

/* mig.DeviceInfo.GetDevicesMap (mig.go:41-64): a fresh walk + IsMigEnabled per device */
static int devices_by_mig(nvmlDevice_t *en, int *nen, nvmlDevice_t *dis, int *ndis) {
    nvmlDevice_t devs[MAXDEV];
    int n, rc;
    if ((rc = get_devices(devs, &n))) return rc;
    *nen = *ndis = 0;
    for (int i = 0; i < n; i++) {
        int cap, e;
        if ((rc = mig_mode(devs[i], &cap, &e))) return rc;
        if (e) en[(*nen)++] = devs[i]; else dis[(*ndis)++] = devs[i];
    }
    return 0;
}

static int invalid_labels(labels_t *l, nvmlDevice_t dev) {  /* mig-strategy.go:243-262 */
    char model[96] = {0};
    int rc;
    if ((rc = NV(nvmlReturn_t, "nvmlDeviceGetName", nvmlDevice_t, char *, unsigned)(dev, model, 96))) return rc;
    dashes(model);
    if (model[0]) putf(l, "nvidia.com/gpu", "product", "%s-MIG-INVALID", model);
    else putf(l, "nvidia.com/gpu", "product", "MIG-INVALID");
    putf(l, "nvidia.com/gpu", "count", "0");
    putf(l, "nvidia.com/gpu", "replicas", "0");
    putf(l, "nvidia.com/gpu", "memory", "0");
    return 0;
}

/* NewMIGResourceLabeler (resource.go:76-111); profile name comes from the wrapper's cache */
static int mig_resource_labels(labels_t *l, const char *res, nvmlDevice_t mig, const char *profile, int count) {
    nvmlDevice_t parent;
    int rc;
    if ((rc = NV(nvmlReturn_t, "nvmlDeviceGetDeviceHandleFromMigDeviceHandle", nvmlDevice_t, nvmlDevice_t *)(mig, &parent))) return rc;
    char model[96] = {0};
    if ((rc = NV(nvmlReturn_t, "nvmlDeviceGetName", nvmlDevice_t, char *, unsigned)(parent, model, 96))) return rc;
    nvmlDeviceAttributes_t a;
    if ((rc = NV(nvmlReturn_t, "nvmlDeviceGetAttributes_v2", nvmlDevice_t, nvmlDeviceAttributes_t *)(mig, &a))) return rc;
    dashes(model);
    char prof[96];
    snprintf(prof, sizeof prof, "%s", profile);
    dashes(prof);
    putf(l, res, "product", "%s%sMIG-%s", model, model[0] ? "-" : "", prof);
    putf(l, res, "count", "%d", count);
    putf(l, res, "replicas", "1");
    putf(l, res, "memory", "%llu", a.memorySizeMB);
    putf(l, res, "multiprocessors", "%u", a.multiprocessorCount);
    putf(l, res, "slices.gi", "%u", a.gpuInstanceSliceCount);
    putf(l, res, "slices.ci", "%u", a.computeInstanceSliceCount);
    putf(l, res, "engines.copy", "%u", a.sharedCopyEngineCount);
    putf(l, res, "engines.decoder", "%u", a.sharedDecoderCount);
    putf(l, res, "engines.encoder", "%u", a.sharedEncoderCount);
    putf(l, res, "engines.jpeg", "%u", a.sharedJpegCount);
    putf(l, res, "engines.ofa", "%u", a.sharedOfaCount);
    return 0;
}

/* GetAllMigDevices + group by profile name (mig-strategy.go:210-232 / 275-292) */
static int group_all_migs(nvmlDevice_t *en, int nen, group_t *groups, int *ng) {
    static nvmlDevice_t all[MAXDEV * MAXMIG];
    int total = 0;
    for (int i = 0; i < nen; i++) {  /* first every GPU's MIG walk ... */
        nvmlDevice_t migs[MAXMIG];
        int nm, rc;
        if ((rc = get_mig_devices(en[i], migs, &nm))) return rc;
        for (int k = 0; k < nm; k++) all[total++] = migs[k];
    }
    *ng = 0;
    for (int k = 0; k < total; k++) {  /* ... then GetName (= GetProfile) per wrapper, in collection order */
        char name[96];
        int rc = mig_profile_name(all[k], name, sizeof name);
        if (rc) return rc;
        group_t *g = group_find(groups, ng, name);
        if (!g->dev) g->dev = all[k];
        g->count++;
    }
    return 0;
}

static int go_is_space(unsigned c) { return c == ' ' || (c >= 9 && c <= 13); }  /* ASCII subset: DMI strings are ASCII */

static int nvml_labels(labels_t *out, const char *strategy, const char *machine_file) {
    int rc = ref_init();
    if (rc) return fail("failed to initialize NVML", rc);
    int ret = 1;
    nvmlDevice_t devs[MAXDEV];
    int n;
    if ((rc = get_devices(devs, &n))) { fail("error getting devices", rc); goto done; }  /* walk #1 */
    if (n == 0) { ret = 0; goto done; }
    {   /* machine type (machine-type.go:30-51) */
        char mt[256] = "unknown";
        if (machine_file[0]) {
            FILE *f = fopen(machine_file, "rb");
            if (f) {
                char buf[256];
                size_t got = fread(buf, 1, sizeof buf - 1, f);
                int err = ferror(f);
                fclose(f);
                if (!err) {
                    buf[got] = 0;
                    char *b = buf, *e = buf + got;
                    while (b < e && go_is_space((unsigned char)*b)) b++;
                    while (e > b && go_is_space((unsigned char)e[-1])) e--;
                    *e = 0;
                    snprintf(mt, sizeof mt, "%s", b);
                }
            }
        }
        dashes(mt);
        put(out, "nvidia.com/gpu.machine", mt);
    }
    {   /* version labeler (nvml.go:75-106) */
        char dv[80] = {0};
        if ((rc = NV(nvmlReturn_t, "nvmlSystemGetDriverVersion", char *, unsigned)(dv, 80))) { fail("error getting driver version", rc); goto done; }
        char *parts[5];
        int np = 0;
        char tmp[80];
        snprintf(tmp, sizeof tmp, "%s", dv);
        char *p = tmp;
        parts[np++] = p;
        for (; *p && np < 5; p++) if (*p == '.') { *p = 0; parts[np++] = p + 1; }
        if (np > 3 || np < 2) { fail("driver version does not match X.Y[.Z]", 0); goto done; }
        int cv = 0;
        if ((rc = NV(nvmlReturn_t, "nvmlSystemGetCudaDriverVersion", int *)(&cv))) { fail("error getting cuda driver version", rc); goto done; }
        put(out, "nvidia.com/cuda.driver.major", parts[0]);
        put(out, "nvidia.com/cuda.driver.minor", parts[1]);
        put(out, "nvidia.com/cuda.driver.rev", np > 2 ? parts[2] : "");
        char b[32];
        snprintf(b, sizeof b, "%d", cv / 1000); put(out, "nvidia.com/cuda.runtime.major", b);
        snprintf(b, sizeof b, "%d", cv % 1000 / 10); put(out, "nvidia.com/cuda.runtime.minor", b);
    }
    {   /* mig capability (nvml.go:110-137): walk #2 */
        nvmlDevice_t d2[MAXDEV];
        int n2, capable = 0;
        if ((rc = get_devices(d2, &n2))) { fail("error getting devices", rc); goto done; }
        for (int i = 0; i < n2 && !capable; i++) {
            int e;
            if ((rc = mig_mode(d2[i], &capable, &e))) { fail("error getting mig capability", rc); goto done; }
        }
        put(out, "nvidia.com/mig.capable", capable ? "true" : "false");
    }
    {   /* NewResourceLabeler (mig-strategy.go:45-77): walk #3, then newGPULabelers: walk #4 */
        nvmlDevice_t d3[MAXDEV];
        int n3;
        if ((rc = get_devices(d3, &n3))) { fail("error getting devices", rc); goto done; }
        if (n3 == 0) { ret = 0; goto done; }
        nvmlDevice_t en[MAXDEV], dis[MAXDEV];
        int nen, ndis;
        if ((rc = devices_by_mig(en, &nen, dis, &ndis))) { fail("error getting map of devices", rc); goto done; }
        static group_t counts[MAXDEV], gen[MAXDEV], gfull[MAXDEV];
        int nc = 0, ngen = 0, ngfull = 0;
        for (int i = 0; i < nen; i++) {
            char name[96] = {0};
            if ((rc = NV(nvmlReturn_t, "nvmlDeviceGetName", nvmlDevice_t, char *, unsigned)(en[i], name, 96))) { fail("error getting device name", rc); goto done; }
            group_find(gen, &ngen, name)->dev = en[i];
            group_find(counts, &nc, name)->count++;
        }
        for (int i = 0; i < ndis; i++) {
            char name[96] = {0};
            if ((rc = NV(nvmlReturn_t, "nvmlDeviceGetName", nvmlDevice_t, char *, unsigned)(dis[i], name, 96))) { fail("error getting device name", rc); goto done; }
            group_find(gfull, &ngfull, name)->dev = dis[i];
            group_find(counts, &nc, name)->count++;
        }
        qsort(gen, (size_t)ngen, sizeof *gen, groupcmp);
        qsort(gfull, (size_t)ngfull, sizeof *gfull, groupcmp);
        for (int i = 0; i < ngen; i++)
            if ((rc = gpu_resource_labels(out, gen[i].dev, group_find(counts, &nc, gen[i].name)->count, 1))) { fail("failed to construct labeler", rc); goto done; }
        for (int i = 0; i < ngfull; i++)
            if ((rc = gpu_resource_labels(out, gfull[i].dev, group_find(counts, &nc, gfull[i].name)->count, 0))) { fail("failed to construct labeler", rc); goto done; }

        if (strcmp(strategy, "none")) {
            int single = !strcmp(strategy, "single"), mixed = !strcmp(strategy, "mixed");
            if (!single && !mixed) { fail("unknown strategy", 0); goto done; }
            labels_t mig = {0};
            /* a fresh DeviceInfo: walk #5 */
            if ((rc = devices_by_mig(en, &nen, dis, &ndis))) { fail("unable to retrieve list of MIG-enabled devices", rc); goto done; }
            static group_t groups[MAXDEV * MAXMIG];
            int ng = 0;
            if (single) {
                if (nen > 0) {
                    int has_empty = 0;  /* AnyMigEnabledDeviceIsEmpty (mig.go:85-106) */
                    for (int i = 0; i < nen && !has_empty; i++) {
                        nvmlDevice_t migs[MAXMIG];
                        int nm;
                        if ((rc = get_mig_devices(en[i], migs, &nm))) { fail("failed to check for empty MIG-enabled devices", rc); goto done; }
                        if (nm == 0) has_empty = 1;
                    }
                    if (has_empty || ndis != 0) {
                        if ((rc = invalid_labels(&mig, en[0]))) { fail("failed to get device model", rc); goto done; }
                    } else {
                        if ((rc = group_all_migs(en, nen, groups, &ng))) { fail("unable to get MIG device name", rc); goto done; }
                        if (ng != 1) {
                            if ((rc = invalid_labels(&mig, en[0]))) { fail("failed to get device model", rc); goto done; }
                        } else if ((rc = mig_resource_labels(&mig, "nvidia.com/gpu", groups[0].dev, groups[0].name, groups[0].count))) {
                            fail("failed to construct labeler", rc); goto done;
                        }
                    }
                }
            } else {
                if ((rc = group_all_migs(en, nen, groups, &ng))) { fail("unable to get MIG device name", rc); goto done; }
                qsort(groups, (size_t)ng, sizeof *groups, groupcmp);
                for (int i = 0; i < ng; i++) {
                    char res[160];
                    snprintf(res, sizeof res, "nvidia.com/mig-%.140s", groups[i].name);
                    if ((rc = mig_resource_labels(&mig, res, groups[i].dev, groups[i].name, groups[i].count))) { fail("failed to construct labeler", rc); goto done; }
                }
            }
            put(out, "nvidia.com/mig.strategy", strategy);
            merge(out, &mig);
            free(mig.e);
        }
    }
    ret = 0;
done:
    ref_shutdown();  /* defer manager.Shutdown() */
    return ret;
}

/* ---- vGPU (pciutil.go:70-149, vgpu.go:66-153) ---- */
static int read_all(const char *path, unsigned char *buf, size_t cap, size_t *len) {
    int fd = open(path, O_RDONLY);
    if (fd < 0) return -1;
    *len = 0;
    while (*len < cap) {
        ssize_t n = read(fd, buf + *len, cap - *len);
        if (n < 0) { if (errno == EINTR) continue; close(fd); return -1; }
        if (n == 0) break;
        *len += (size_t)n;
    }
    close(fd);
    return 0;
}
static int namecmp(const void *a, const void *b) { return strcmp(*(char *const *)a, *(char *const *)b); }

static int vgpu_labels(labels_t *out, const char *root) {
    DIR *d = opendir(root);
    if (!d) return fail("unable to read PCI bus devices", errno);
    char *names[4096];
    int n = 0;
    struct dirent *e;
    while ((e = readdir(d)) && n < 4096) if (strcmp(e->d_name, ".") && strcmp(e->d_name, "..")) names[n++] = strdup(e->d_name);
    closedir(d);
    qsort(names, (size_t)n, sizeof *names, namecmp);
    int present = 0, ret = 0;
    char ver[64] = "", br[64] = "";
    for (int i = 0; i < n && !ret; i++) {
        char path[PATH_MAX];
        unsigned char buf[8192];
        size_t len;
        snprintf(path, sizeof path, "%s/%s/vendor", root, names[i]);
        if (read_all(path, buf, sizeof buf - 1, &len)) { ret = fail("unable to read PCI device vendor id", errno); break; }
        buf[len] = 0;
        char *v = (char *)buf;
        while (*v && go_is_space((unsigned char)*v)) v++;
        size_t vl = strlen(v);
        while (vl && go_is_space((unsigned char)v[vl - 1])) v[--vl] = 0;
        if (strcmp(v, "0x10de")) continue;
        snprintf(path, sizeof path, "%s/%s/class", root, names[i]);
        if (read_all(path, buf, sizeof buf, &len)) { ret = fail("unable to read PCI device class", errno); break; }
        if (len < 4) { ret = fail("class shorter than 4 bytes (Go would panic)", 0); break; }
        snprintf(path, sizeof path, "%s/%s/config", root, names[i]);
        unsigned char cfg[8192] = {0};
        if (read_all(path, cfg, 4096, &len)) { ret = fail("unable to read PCI configuration space", errno); break; }
        struct stat st;
        size_t cap = 512;
        if (stat(path, &st) == 0 && (size_t)st.st_size + 1 > cap) cap = (size_t)st.st_size + 1;
        if (len > cap) cap = len;
        if (len < 256) { ret = fail("entire PCI configuration is not read; run privileged", 0); break; }
        if (!(cfg[0x06] & 0x10)) continue;
        unsigned char visited[256] = {0};
        size_t pos = cfg[0x34], cpos = 0, clen = 0;
        int found = 0;
        while (pos) {
            if (pos + 2 >= len) { ret = fail("index out of range (Go would panic)", 0); break; }
            unsigned id = cfg[pos], next = cfg[pos + 1], length = cfg[pos + 2];
            if (visited[pos]) break;
            if (id == 0xff) break;
            if (id == 0x09) {
                if (pos + length > cap) { ret = fail("slice bounds out of range (Go would panic)", 0); break; }
                cpos = pos; clen = length; found = 1;
                break;
            }
            visited[pos]++;
            pos = next;
        }
        if (ret || !found) continue;
        unsigned char c[256];
        for (size_t k = 0; k < clen; k++) c[k] = cpos + k < len ? cfg[cpos + k] : 0;
        if (clen < 5 || c[3] != 0x56 || c[4] != 0x46) continue;
        present = 1;
        if (clen <= 5) { ret = fail("index out of range (Go would panic)", 0); break; }
        size_t p = 5;
        unsigned rec = c[5];
        while (rec != 0 && p < clen) {
            if (p + 1 >= clen) { ret = fail("index out of range (Go would panic)", 0); break; }
            unsigned rl = c[p + 1];
            if (!rl) { ret = fail("zero-length record (the reference would spin forever)", 0); break; }
            p += rl;
            if (p >= clen) { ret = fail("index out of range (Go would panic)", 0); break; }
            rec = c[p];
        }
        if (ret) break;
        if (rec == 0 && p + 22 <= clen) {
            for (int f = 0; f < 2; f++) {
                char tmp[24];
                size_t tn = 0;
                for (int k = 0; k < 10; k++) {  /* Go string(uint8): UTF-8 of the code point */
                    unsigned ch = c[p + 2 + (size_t)f * 10 + (size_t)k];
                    if (ch < 0x80) tmp[tn++] = (char)ch;
                    else { tmp[tn++] = (char)(0xC0 | (ch >> 6)); tmp[tn++] = (char)(0x80 | (ch & 0x3F)); }
                }
                size_t b = 0, en = tn;
                while (b < en && !tmp[b]) b++;
                while (en > b && !tmp[en - 1]) en--;
                char *dst = f ? br : ver;
                memcpy(dst, tmp + b, en - b);
                dst[en - b] = 0;
            }
        } else { ret = fail("cannot find driver version record", 0); }
    }
    for (int i = 0; i < n; i++) free(names[i]);
    if (ret) return ret;
    if (present) {
        put(out, "nvidia.com/vgpu.present", "true");
        put(out, "nvidia.com/vgpu.host-driver-version", ver);
        put(out, "nvidia.com/vgpu.host-driver-branch", br);
    }
    return 0;
}

/* ---- writeFileAtomically (labels.go:68-114) ---- */
static int write_file(const labels_t *l, const char *path) {
    qsort(l->e, (size_t)l->n, sizeof(kv_t), kvcmp);
    size_t cap = 1, off = 0;
    for (int i = 0; i < l->n; i++) cap += strlen(l->e[i].k) + strlen(l->e[i].v) + 2;
    char *body = malloc(cap);
    for (int i = 0; i < l->n; i++) off += (size_t)sprintf(body + off, "%s=%s\n", l->e[i].k, l->e[i].v);
    if (!path[0]) { fwrite(body, 1, off, stdout); free(body); return 0; }
    char abs[2 * PATH_MAX + 2];
    if (path[0] == '/') snprintf(abs, sizeof abs, "%s", path);
    else { char cwd[PATH_MAX]; if (!getcwd(cwd, sizeof cwd)) return 1; snprintf(abs, sizeof abs, "%s/%s", cwd, path); }
    char *slash = strrchr(abs, '/');
    char dir[sizeof abs + 16], tmp[sizeof abs + 32];
    snprintf(dir, sizeof dir, "%.*s/gfd-tmp", (int)(slash - abs), abs);
    if (mkdir(dir, 0777) != 0 && errno != EEXIST) { free(body); return fail("failed to create temporary directory", errno); }
    snprintf(tmp, sizeof tmp, "%s/gfd-XXXXXX", dir);
    int fd = mkstemp(tmp);
    if (fd < 0) { free(body); return fail("fail to create temporary output file", errno); }
    size_t w = 0;
    while (w < off) { ssize_t k = write(fd, body + w, off - w); if (k < 0) { if (errno == EINTR) continue; break; } w += (size_t)k; }
    close(fd);
    free(body);
    if (w != off) return fail("error writing temporary file", errno);
    if (rename(tmp, path)) return fail("error moving temporary file", errno);
    if (chmod(path, 0644)) return fail("error setting permissions", errno);
    return 0;
}

int main(int argc, char **argv) {
    const char *strategy = "none", *machine = "/sys/class/dmi/id/product_name", *out = "", *pci = "/sys/bus/pci/devices";
    int no_ts = 0, passes = 1;
    for (int i = 1; i < argc; i++) {
        if (!strcmp(argv[i], "--mig-strategy") && i + 1 < argc) strategy = argv[++i];
        else if (!strcmp(argv[i], "--machine-type-file") && i + 1 < argc) machine = argv[++i];
        else if (!strcmp(argv[i], "-o") && i + 1 < argc) out = argv[++i];
        else if (!strcmp(argv[i], "--pci-root") && i + 1 < argc) pci = argv[++i];
        else if (!strcmp(argv[i], "--passes") && i + 1 < argc) passes = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--no-timestamp")) no_ts = 1;
        else { fprintf(stderr, "ref_replay: unknown argument %s\n", argv[i]); return 2; }
    }
    double *ms = calloc((size_t)passes, sizeof(double)), *init = calloc((size_t)passes, sizeof(double)), *shut = calloc((size_t)passes, sizeof(double));
    long long ts = (long long)time(NULL);  /* NewTimestampLabeler: once per run() */
    unsigned long calls_per_pass = 0, dl_per_pass = 0;
    for (int p = 0; p < passes; p++) {
        g_calls = g_dl = 0;
        double t0 = now_ms();
        labels_t labels = {0}, nv = {0}, vg = {0};
        if (!no_ts) { char b[32]; snprintf(b, sizeof b, "%lld", ts); put(&labels, "nvidia.com/gfd.timestamp", b); }
        if (nvml_labels(&nv, strategy, machine)) return 1;
        if (vgpu_labels(&vg, pci)) return 1;
        merge(&labels, &nv);
        merge(&labels, &vg);
        if (labels.n <= 1) fprintf(stderr, "Warning: no labels generated from any source\n");
        if (write_file(&labels, out)) return 1;
        ms[p] = now_ms() - t0;
        init[p] = g_init_ms;
        shut[p] = g_shutdown_ms;
        calls_per_pass = g_calls;
        dl_per_pass = g_dl;
        free(labels.e); free(nv.e); free(vg.e);
    }
    FILE *js = out[0] ? stdout : stderr;
    fprintf(js, "{\"passes\":%d,\"ms\":[", passes);
    for (int p = 0; p < passes; p++) fprintf(js, "%s%.4f", p ? "," : "", ms[p]);
    fprintf(js, "],\"init_ms\":[");
    for (int p = 0; p < passes; p++) fprintf(js, "%s%.4f", p ? "," : "", init[p]);
    fprintf(js, "],\"shutdown_ms\":[");
    for (int p = 0; p < passes; p++) fprintf(js, "%s%.4f", p ? "," : "", shut[p]);
    fprintf(js, "],\"nvml_calls\":%lu,\"dl_cycles\":%lu,\"ok\":true}\n", calls_per_pass, dl_per_pass);
    return 0;
}
