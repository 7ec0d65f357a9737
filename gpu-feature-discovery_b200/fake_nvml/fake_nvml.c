/*
 * fake libnvidia-ml.so.1 -- TEST DOUBLE for hosts without a GPU/driver.
 *
 * Exports exactly the NVML entry points the reference GFD reaches (SURVEY.md 8(b);
 * vendor/github.com/NVIDIA/go-nvml/pkg/nvml/nvml.go lines 31..2364) and answers them from a
 * `.nvmlfix` fixture named by $GFD_FAKE_NVML (grammar: oracle/nvml_model.py docstring; the Python
 * interpreter of the same grammar, RawNvml, must agree with this file call for call).
 *
 *   GFD_FAKE_NVML=<fixture>          node description (re-read at every nvmlInit_v2)
 *   GFD_FAKE_NVML_TRACE=<file>       append one line per NVML call (symbol name)
 *   GFD_FAKE_NVML_LATENCY_US=<n>     sleep n us in every per-device call (models the ioctl round trip)
 *   GFD_FAKE_NVML_INIT_LATENCY_US=<n> sleep n us in nvmlInit_v2 (models driver attach)
 *
 * Build variants: -DFAKE_NO_MIG_SYMBOLS drops nvmlDeviceGetMigMode (pre-R450 driver; `hide_symbol`).
 * Not linked into the product: the shim dlopen()s "libnvidia-ml.so.1" and finds this one only when a
 * test puts its directory on LD_LIBRARY_PATH.
 */
#define _GNU_SOURCE
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#define NVML_NO_UNVERSIONED_FUNC_DEFS
#include <nvml.h>

#define MAX_GPUS 64
#define MAX_MIGS 16
#define MAX_GI_IDX 32
#define MAX_CI_IDX 16
#define MAX_RC 16

enum { K_GPU = 1, K_MIG = 2, K_GI = 3, K_CI = 4 };

typedef struct { char sym[64]; int code; } rc_t;
typedef struct { rc_t v[MAX_RC]; int n; } rcset_t;

struct gpu_s;
struct mig_s;
typedef struct { int kind; struct mig_s *m; } inst_t; /* GPU-instance / compute-instance pseudo handle */
typedef struct mig_s {
    int kind; /* K_MIG -- must be first: handles are pointers to this tag */
    struct gpu_s *gpu;
    int slot, gi_id, ci_id, gi_profile_id, ci_profile_id;
    unsigned long long attrs[9];
    rcset_t rc;
    inst_t gi, ci; /* GI / CI pseudo-handles for this MIG device */
} mig_t;

typedef struct gpu_s {
    int kind; /* K_GPU */
    int index;
    char name[128];
    unsigned long long memory_total;
    int cc_major, cc_minor;
    int has_mig_mode, mig_cur, mig_pend;
    int max_mig;
    struct { int present, id, slices; unsigned long long mb; } gi_prof[MAX_GI_IDX];
    struct { int present, id, slices; } ci_prof[MAX_GI_IDX][MAX_CI_IDX];
    mig_t migs[MAX_MIGS];
    int nmigs;
    rcset_t rc;
} gpu_t;

static struct {
    char driver_version[96];
    int cuda_driver_version;
    gpu_t gpus[MAX_GPUS];
    int ngpus;
    rcset_t rc;
} N;

static int g_init = 0;
static pthread_mutex_t g_mu = PTHREAD_MUTEX_INITIALIZER;

static void rc_add(rcset_t *s, const char *sym, int code) {
    if (s->n < MAX_RC) { snprintf(s->v[s->n].sym, sizeof s->v[s->n].sym, "%s", sym); s->v[s->n].code = code; s->n++; }
}
static int rc_find(const rcset_t *s, const char *sym, int *code) {
    for (int i = 0; i < s->n; i++) if (!strcmp(s->v[i].sym, sym)) { *code = s->v[i].code; return 1; }
    return 0;
}

static void load_fixture(void) {
    memset(&N, 0, sizeof N);
    snprintf(N.driver_version, sizeof N.driver_version, "575.57.08");
    N.cuda_driver_version = 12090;
    const char *path = getenv("GFD_FAKE_NVML");
    if (!path) return;
    FILE *f = fopen(path, "r");
    if (!f) { fprintf(stderr, "fake_nvml: cannot open fixture %s\n", path); return; }
    char line[512];
    gpu_t *g = NULL; mig_t *m = NULL;
    while (fgets(line, sizeof line, f)) {
        char *hash = strchr(line, '#'); if (hash) *hash = 0;
        char *p = line; while (*p == ' ' || *p == '\t') p++;
        size_t L = strlen(p); while (L && (p[L-1] == '\n' || p[L-1] == '\r' || p[L-1] == ' ' || p[L-1] == '\t')) p[--L] = 0;
        if (!L) continue;
        char k[64]; int off = 0;
        if (sscanf(p, "%63s%n", k, &off) != 1) continue;
        char *rest = p + off; while (*rest == ' ' || *rest == '\t') rest++;
        if (!strcmp(k, "driver_version")) snprintf(N.driver_version, sizeof N.driver_version, "%s", rest);
        else if (!strcmp(k, "cuda_driver_version")) N.cuda_driver_version = atoi(rest);
        else if (!strcmp(k, "hide_symbol")) { /* honoured by the *_nomig build only */ }
        else if (!strcmp(k, "rc")) {
            char sym[64]; int code;
            if (sscanf(rest, "%63s %d", sym, &code) == 2) rc_add(m ? &m->rc : g ? &g->rc : &N.rc, sym, code);
        } else if (!strcmp(k, "gpu")) {
            if (N.ngpus >= MAX_GPUS) break;
            g = &N.gpus[N.ngpus]; g->kind = K_GPU; g->index = N.ngpus++; m = NULL;
        } else if (!g) { fprintf(stderr, "fake_nvml: directive outside gpu block: %s\n", p); }
        else if (!strcmp(k, "name")) snprintf(g->name, sizeof g->name, "%s", rest);
        else if (!strcmp(k, "memory_total")) g->memory_total = strtoull(rest, NULL, 10);
        else if (!strcmp(k, "cc")) sscanf(rest, "%d %d", &g->cc_major, &g->cc_minor);
        else if (!strcmp(k, "mig_mode")) { g->has_mig_mode = 1; sscanf(rest, "%d %d", &g->mig_cur, &g->mig_pend); }
        else if (!strcmp(k, "max_mig_count")) g->max_mig = atoi(rest);
        else if (!strcmp(k, "gi_profile")) {
            int idx, id, sl; unsigned long long mb;
            if (sscanf(rest, "%d id %d slices %d memory_mb %llu", &idx, &id, &sl, &mb) == 4 && idx >= 0 && idx < MAX_GI_IDX) {
                g->gi_prof[idx].present = 1; g->gi_prof[idx].id = id; g->gi_prof[idx].slices = sl; g->gi_prof[idx].mb = mb;
            }
        } else if (!strcmp(k, "ci_profile")) {
            int gp, idx, id, sl;
            if (sscanf(rest, "%d %d id %d slices %d", &gp, &idx, &id, &sl) == 4 && gp >= 0 && gp < MAX_GI_IDX && idx >= 0 && idx < MAX_CI_IDX) {
                g->ci_prof[gp][idx].present = 1; g->ci_prof[gp][idx].id = id; g->ci_prof[gp][idx].slices = sl;
            }
        } else if (!strcmp(k, "mig")) {
            if (g->nmigs >= MAX_MIGS) continue;
            m = &g->migs[g->nmigs++]; m->kind = K_MIG; m->gpu = g; m->slot = atoi(rest);
            m->gi.kind = K_GI; m->gi.m = m; m->ci.kind = K_CI; m->ci.m = m;
        } else if (!m) { fprintf(stderr, "fake_nvml: unknown directive: %s\n", p); }
        else if (!strcmp(k, "gi_id")) m->gi_id = atoi(rest);
        else if (!strcmp(k, "ci_id")) m->ci_id = atoi(rest);
        else if (!strcmp(k, "gi_profile_id")) m->gi_profile_id = atoi(rest);
        else if (!strcmp(k, "ci_profile_id")) m->ci_profile_id = atoi(rest);
        else if (!strcmp(k, "attrs")) sscanf(rest, "%llu %llu %llu %llu %llu %llu %llu %llu %llu", &m->attrs[0], &m->attrs[1],
                                           &m->attrs[2], &m->attrs[3], &m->attrs[4], &m->attrs[5], &m->attrs[6], &m->attrs[7], &m->attrs[8]);
        else fprintf(stderr, "fake_nvml: unknown directive: %s\n", p);
    }
    fclose(f);
}

static void trace(const char *sym) {
    const char *t = getenv("GFD_FAKE_NVML_TRACE");
    if (!t) return;
    pthread_mutex_lock(&g_mu);
    FILE *f = fopen(t, "a");
    if (f) { fprintf(f, "%s\n", sym); fclose(f); }
    pthread_mutex_unlock(&g_mu);
}

static void latency(void) {
    static int us = -1;
    if (us < 0) { const char *e = getenv("GFD_FAKE_NVML_LATENCY_US"); us = e ? atoi(e) : 0; }
    if (us > 0) usleep((useconds_t)us);
}

/* rc lookup: MIG scope -> GPU scope -> node scope -> SUCCESS (== RawNvml._rc) */
static nvmlReturn_t rc_for(const char *sym, gpu_t *g, mig_t *m) {
    trace(sym);
    if (strcmp(sym, "nvmlInit_v2") && !g_init) return NVML_ERROR_UNINITIALIZED;
    if (g || m) latency();
    int code;
    if (m && rc_find(&m->rc, sym, &code)) return (nvmlReturn_t)code;
    if (g && rc_find(&g->rc, sym, &code)) return (nvmlReturn_t)code;
    if (rc_find(&N.rc, sym, &code)) return (nvmlReturn_t)code;
    return NVML_SUCCESS;
}

static int kind_of(void *h) { return h ? *(int *)h : 0; }
static gpu_t *gpu_of(void *h) {
    switch (kind_of(h)) {
    case K_GPU: return (gpu_t *)h;
    case K_MIG: return ((mig_t *)h)->gpu;
    default: return NULL;
    }
}
static mig_t *mig_of(void *h) { return kind_of(h) == K_MIG ? (mig_t *)h : NULL; }

#define EXPORT __attribute__((visibility("default")))

EXPORT nvmlReturn_t nvmlInit_v2(void) {
    pthread_mutex_lock(&g_mu);
    int first = (g_init == 0);
    pthread_mutex_unlock(&g_mu);
    if (first) load_fixture();
    nvmlReturn_t rc = rc_for("nvmlInit_v2", NULL, NULL);
    const char *e = getenv("GFD_FAKE_NVML_INIT_LATENCY_US");
    if (e && first) usleep((useconds_t)atoi(e));
    if (rc == NVML_SUCCESS) { pthread_mutex_lock(&g_mu); g_init++; pthread_mutex_unlock(&g_mu); }
    return rc;
}

EXPORT nvmlReturn_t nvmlShutdown(void) {
    nvmlReturn_t rc = rc_for("nvmlShutdown", NULL, NULL);
    /* an injected error code still tears the library down, so one test's fixture cannot leak into the next */
    if (rc != NVML_ERROR_UNINITIALIZED) { pthread_mutex_lock(&g_mu); if (g_init > 0) g_init--; pthread_mutex_unlock(&g_mu); }
    return rc;
}

EXPORT const char *nvmlErrorString(nvmlReturn_t r) {
    trace("nvmlErrorString");
    switch ((int)r) {
    case 0: return "SUCCESS"; case 1: return "ERROR_UNINITIALIZED"; case 2: return "ERROR_INVALID_ARGUMENT";
    case 3: return "ERROR_NOT_SUPPORTED"; case 4: return "ERROR_NO_PERMISSION"; case 5: return "ERROR_ALREADY_INITIALIZED";
    case 6: return "ERROR_NOT_FOUND"; case 7: return "ERROR_INSUFFICIENT_SIZE"; case 8: return "ERROR_INSUFFICIENT_POWER";
    case 9: return "ERROR_DRIVER_NOT_LOADED"; case 10: return "ERROR_TIMEOUT"; case 11: return "ERROR_IRQ_ISSUE";
    case 12: return "ERROR_LIBRARY_NOT_FOUND"; case 13: return "ERROR_FUNCTION_NOT_FOUND"; case 14: return "ERROR_CORRUPTED_INFOROM";
    case 15: return "ERROR_GPU_IS_LOST"; case 16: return "ERROR_RESET_REQUIRED"; case 17: return "ERROR_OPERATING_SYSTEM";
    case 18: return "ERROR_LIB_RM_VERSION_MISMATCH"; case 19: return "ERROR_IN_USE"; case 20: return "ERROR_MEMORY";
    case 21: return "ERROR_NO_DATA"; case 22: return "ERROR_VGPU_ECC_NOT_SUPPORTED"; case 23: return "ERROR_INSUFFICIENT_RESOURCES";
    case 999: return "ERROR_UNKNOWN";
    default: return "Unknown return value";
    }
}

EXPORT nvmlReturn_t nvmlSystemGetDriverVersion(char *version, unsigned int length) {
    nvmlReturn_t rc = rc_for("nvmlSystemGetDriverVersion", NULL, NULL);
    if (version && length) snprintf(version, length, "%s", N.driver_version);
    return rc;
}

EXPORT nvmlReturn_t nvmlSystemGetCudaDriverVersion(int *v) {
    nvmlReturn_t rc = rc_for("nvmlSystemGetCudaDriverVersion", NULL, NULL);
    if (v) *v = N.cuda_driver_version;
    return rc;
}

EXPORT nvmlReturn_t nvmlDeviceGetCount_v2(unsigned int *count) {
    nvmlReturn_t rc = rc_for("nvmlDeviceGetCount_v2", NULL, NULL);
    if (count) *count = (unsigned)N.ngpus;
    return rc;
}

EXPORT nvmlReturn_t nvmlDeviceGetHandleByIndex_v2(unsigned int index, nvmlDevice_t *device) {
    if ((int)index >= N.ngpus) { trace("nvmlDeviceGetHandleByIndex_v2"); return NVML_ERROR_INVALID_ARGUMENT; }
    gpu_t *g = &N.gpus[index];
    nvmlReturn_t rc = rc_for("nvmlDeviceGetHandleByIndex_v2", g, NULL);
    if (device) *device = (nvmlDevice_t)g;
    return rc;
}

EXPORT nvmlReturn_t nvmlDeviceGetName(nvmlDevice_t d, char *name, unsigned int length) {
    gpu_t *g = gpu_of(d); mig_t *m = mig_of(d);
    if (!g) { trace("nvmlDeviceGetName"); return NVML_ERROR_INVALID_ARGUMENT; }
    nvmlReturn_t rc = rc_for("nvmlDeviceGetName", g, m);
    if (name && length) snprintf(name, length, m ? "%s MIG" : "%s", g->name);
    return rc;
}

/* For pynvml-driven checkers only (tests/test_gpu_labels.py and tools/mig_hw_stage.py can then be dry-run against a simulated node);
 * the reference's call sequence contains neither symbol, so the traces of the parity tests are unaffected. */
EXPORT nvmlReturn_t nvmlInitWithFlags(unsigned int flags) { (void)flags; return nvmlInit_v2(); }
EXPORT nvmlReturn_t nvmlDeviceGetUUID(nvmlDevice_t d, char *uuid, unsigned int length) {
    gpu_t *g = gpu_of(d); mig_t *m = mig_of(d);
    if (!g) { trace("nvmlDeviceGetUUID"); return NVML_ERROR_INVALID_ARGUMENT; }
    nvmlReturn_t rc = rc_for("nvmlDeviceGetUUID", g, m);
    if (uuid && length) {
        if (m) snprintf(uuid, length, "MIG-00000000-0000-0000-%04x-%012x", g->index, m->slot);
        else snprintf(uuid, length, "GPU-00000000-0000-0000-0000-%012x", g->index);
    }
    return rc;
}

EXPORT nvmlReturn_t nvmlDeviceGetMemoryInfo(nvmlDevice_t d, nvmlMemory_t *mem) {
    gpu_t *g = gpu_of(d); mig_t *m = mig_of(d);
    if (!g) { trace("nvmlDeviceGetMemoryInfo"); return NVML_ERROR_INVALID_ARGUMENT; }
    nvmlReturn_t rc = rc_for("nvmlDeviceGetMemoryInfo", g, m);
    if (mem) { mem->total = m ? m->attrs[0] * 1024ULL * 1024ULL : g->memory_total; mem->free = mem->total; mem->used = 0; }
    return rc;
}

EXPORT nvmlReturn_t nvmlDeviceGetCudaComputeCapability(nvmlDevice_t d, int *major, int *minor) {
    gpu_t *g = gpu_of(d); mig_t *m = mig_of(d);
    if (!g) { trace("nvmlDeviceGetCudaComputeCapability"); return NVML_ERROR_INVALID_ARGUMENT; }
    nvmlReturn_t rc = rc_for("nvmlDeviceGetCudaComputeCapability", g, m);
    if (major) *major = g->cc_major;
    if (minor) *minor = g->cc_minor;
    return rc;
}

#ifndef FAKE_NO_MIG_SYMBOLS
EXPORT nvmlReturn_t nvmlDeviceGetMigMode(nvmlDevice_t d, unsigned int *cur, unsigned int *pend) {
    gpu_t *g = gpu_of(d); mig_t *m = mig_of(d);
    if (!g) { trace("nvmlDeviceGetMigMode"); return NVML_ERROR_INVALID_ARGUMENT; }
    nvmlReturn_t rc = rc_for("nvmlDeviceGetMigMode", g, m);
    if (rc != NVML_SUCCESS) return rc;
    if (m) return NVML_ERROR_INVALID_ARGUMENT;
    if (!g->has_mig_mode) return NVML_ERROR_NOT_SUPPORTED;
    if (cur) *cur = (unsigned)g->mig_cur;
    if (pend) *pend = (unsigned)g->mig_pend;
    return rc;
}
#endif

EXPORT nvmlReturn_t nvmlDeviceGetMaxMigDeviceCount(nvmlDevice_t d, unsigned int *count) {
    gpu_t *g = gpu_of(d); mig_t *m = mig_of(d);
    if (!g) { trace("nvmlDeviceGetMaxMigDeviceCount"); return NVML_ERROR_INVALID_ARGUMENT; }
    nvmlReturn_t rc = rc_for("nvmlDeviceGetMaxMigDeviceCount", g, m);
    if (count) *count = m ? 0u : (unsigned)g->max_mig;
    return rc;
}

EXPORT nvmlReturn_t nvmlDeviceGetMigDeviceHandleByIndex(nvmlDevice_t d, unsigned int index, nvmlDevice_t *mig) {
    gpu_t *g = gpu_of(d);
    if (!g || mig_of(d)) { trace("nvmlDeviceGetMigDeviceHandleByIndex"); return NVML_ERROR_INVALID_ARGUMENT; }
    for (int k = 0; k < g->nmigs; k++) {
        if (g->migs[k].slot == (int)index) {
            nvmlReturn_t rc = rc_for("nvmlDeviceGetMigDeviceHandleByIndex", g, &g->migs[k]);
            if (mig) *mig = (nvmlDevice_t)&g->migs[k];
            return rc;
        }
    }
    nvmlReturn_t rc = rc_for("nvmlDeviceGetMigDeviceHandleByIndex", g, NULL);
    if (rc != NVML_SUCCESS) return rc;
    if ((int)index >= g->max_mig) return NVML_ERROR_INVALID_ARGUMENT;
    return NVML_ERROR_NOT_FOUND;
}

EXPORT nvmlReturn_t nvmlDeviceIsMigDeviceHandle(nvmlDevice_t d, unsigned int *is_mig) {
    gpu_t *g = gpu_of(d); mig_t *m = mig_of(d);
    if (!g) { trace("nvmlDeviceIsMigDeviceHandle"); return NVML_ERROR_INVALID_ARGUMENT; }
    nvmlReturn_t rc = rc_for("nvmlDeviceIsMigDeviceHandle", g, m);
    if (is_mig) *is_mig = m ? 1u : 0u;
    return rc;
}

EXPORT nvmlReturn_t nvmlDeviceGetDeviceHandleFromMigDeviceHandle(nvmlDevice_t d, nvmlDevice_t *parent) {
    gpu_t *g = gpu_of(d); mig_t *m = mig_of(d);
    if (!g) { trace("nvmlDeviceGetDeviceHandleFromMigDeviceHandle"); return NVML_ERROR_INVALID_ARGUMENT; }
    nvmlReturn_t rc = rc_for("nvmlDeviceGetDeviceHandleFromMigDeviceHandle", g, m);
    if (rc != NVML_SUCCESS) return rc;
    if (!m) return NVML_ERROR_INVALID_ARGUMENT;
    if (parent) *parent = (nvmlDevice_t)g;
    return rc;
}

EXPORT nvmlReturn_t nvmlDeviceGetAttributes_v2(nvmlDevice_t d, nvmlDeviceAttributes_t *a) {
    gpu_t *g = gpu_of(d); mig_t *m = mig_of(d);
    if (!g) { trace("nvmlDeviceGetAttributes_v2"); return NVML_ERROR_INVALID_ARGUMENT; }
    nvmlReturn_t rc = rc_for("nvmlDeviceGetAttributes_v2", g, m);
    if (rc != NVML_SUCCESS) return rc;
    if (!m) return NVML_ERROR_INVALID_ARGUMENT;
    if (a) {
        a->memorySizeMB = m->attrs[0]; a->multiprocessorCount = (unsigned)m->attrs[1];
        a->gpuInstanceSliceCount = (unsigned)m->attrs[2]; a->computeInstanceSliceCount = (unsigned)m->attrs[3];
        a->sharedCopyEngineCount = (unsigned)m->attrs[4]; a->sharedDecoderCount = (unsigned)m->attrs[5];
        a->sharedEncoderCount = (unsigned)m->attrs[6]; a->sharedJpegCount = (unsigned)m->attrs[7];
        a->sharedOfaCount = (unsigned)m->attrs[8];
    }
    return rc;
}

EXPORT nvmlReturn_t nvmlDeviceGetGpuInstanceId(nvmlDevice_t d, unsigned int *id) {
    gpu_t *g = gpu_of(d); mig_t *m = mig_of(d);
    if (!g) { trace("nvmlDeviceGetGpuInstanceId"); return NVML_ERROR_INVALID_ARGUMENT; }
    nvmlReturn_t rc = rc_for("nvmlDeviceGetGpuInstanceId", g, m);
    if (rc != NVML_SUCCESS) return rc;
    if (!m) return NVML_ERROR_INVALID_ARGUMENT;
    if (id) *id = (unsigned)m->gi_id;
    return rc;
}

EXPORT nvmlReturn_t nvmlDeviceGetComputeInstanceId(nvmlDevice_t d, unsigned int *id) {
    gpu_t *g = gpu_of(d); mig_t *m = mig_of(d);
    if (!g) { trace("nvmlDeviceGetComputeInstanceId"); return NVML_ERROR_INVALID_ARGUMENT; }
    nvmlReturn_t rc = rc_for("nvmlDeviceGetComputeInstanceId", g, m);
    if (rc != NVML_SUCCESS) return rc;
    if (!m) return NVML_ERROR_INVALID_ARGUMENT;
    if (id) *id = (unsigned)m->ci_id;
    return rc;
}

EXPORT nvmlReturn_t nvmlDeviceGetGpuInstanceById(nvmlDevice_t d, unsigned int id, nvmlGpuInstance_t *gi) {
    gpu_t *g = gpu_of(d);
    if (!g) { trace("nvmlDeviceGetGpuInstanceById"); return NVML_ERROR_INVALID_ARGUMENT; }
    nvmlReturn_t rc = rc_for("nvmlDeviceGetGpuInstanceById", g, NULL);
    if (rc != NVML_SUCCESS) return rc;
    if (mig_of(d)) return NVML_ERROR_INVALID_ARGUMENT;
    for (int k = 0; k < g->nmigs; k++)
        if (g->migs[k].gi_id == (int)id) { if (gi) *gi = (nvmlGpuInstance_t)&g->migs[k].gi; return rc; }
    return NVML_ERROR_NOT_FOUND;
}

static mig_t *gi_mig(nvmlGpuInstance_t gi) { return (gi && ((inst_t *)gi)->kind == K_GI) ? ((inst_t *)gi)->m : NULL; }
static mig_t *ci_mig(nvmlComputeInstance_t ci) { return (ci && ((inst_t *)ci)->kind == K_CI) ? ((inst_t *)ci)->m : NULL; }

EXPORT nvmlReturn_t nvmlGpuInstanceGetInfo(nvmlGpuInstance_t gi, nvmlGpuInstanceInfo_t *info) {
    mig_t *m = gi_mig(gi);
    if (!m) { trace("nvmlGpuInstanceGetInfo"); return NVML_ERROR_INVALID_ARGUMENT; }
    nvmlReturn_t rc = rc_for("nvmlGpuInstanceGetInfo", m->gpu, NULL);
    if (rc != NVML_SUCCESS) return rc;
    if (info) { memset(info, 0, sizeof *info); info->device = (nvmlDevice_t)m->gpu; info->id = (unsigned)m->gi_id; info->profileId = (unsigned)m->gi_profile_id; }
    return rc;
}

EXPORT nvmlReturn_t nvmlGpuInstanceGetComputeInstanceById(nvmlGpuInstance_t gi, unsigned int id, nvmlComputeInstance_t *ci) {
    mig_t *m = gi_mig(gi);
    if (!m) { trace("nvmlGpuInstanceGetComputeInstanceById"); return NVML_ERROR_INVALID_ARGUMENT; }
    gpu_t *g = m->gpu;
    nvmlReturn_t rc = rc_for("nvmlGpuInstanceGetComputeInstanceById", g, NULL);
    if (rc != NVML_SUCCESS) return rc;
    for (int k = 0; k < g->nmigs; k++)
        if (g->migs[k].gi_id == m->gi_id && g->migs[k].ci_id == (int)id) { if (ci) *ci = (nvmlComputeInstance_t)&g->migs[k].ci; return rc; }
    return NVML_ERROR_NOT_FOUND;
}

EXPORT nvmlReturn_t nvmlComputeInstanceGetInfo_v2(nvmlComputeInstance_t ci, nvmlComputeInstanceInfo_t *info) {
    mig_t *m = ci_mig(ci);
    if (!m) { trace("nvmlComputeInstanceGetInfo_v2"); return NVML_ERROR_INVALID_ARGUMENT; }
    nvmlReturn_t rc = rc_for("nvmlComputeInstanceGetInfo_v2", m->gpu, NULL);
    if (rc != NVML_SUCCESS) return rc;
    if (info) { memset(info, 0, sizeof *info); info->device = (nvmlDevice_t)m->gpu; info->gpuInstance = (nvmlGpuInstance_t)&m->gi; info->id = (unsigned)m->ci_id; info->profileId = (unsigned)m->ci_profile_id; }
    return rc;
}

EXPORT nvmlReturn_t nvmlDeviceGetGpuInstanceProfileInfo(nvmlDevice_t d, unsigned int profile, nvmlGpuInstanceProfileInfo_t *info) {
    gpu_t *g = gpu_of(d);
    if (!g) { trace("nvmlDeviceGetGpuInstanceProfileInfo"); return NVML_ERROR_INVALID_ARGUMENT; }
    nvmlReturn_t rc = rc_for("nvmlDeviceGetGpuInstanceProfileInfo", g, NULL);
    if (rc != NVML_SUCCESS) return rc;
    if (mig_of(d)) return NVML_ERROR_INVALID_ARGUMENT;
    if (profile >= MAX_GI_IDX || !g->gi_prof[profile].present) return NVML_ERROR_NOT_SUPPORTED;
    if (info) { memset(info, 0, sizeof *info); info->id = (unsigned)g->gi_prof[profile].id; info->sliceCount = (unsigned)g->gi_prof[profile].slices; info->memorySizeMB = g->gi_prof[profile].mb; }
    return rc;
}

EXPORT nvmlReturn_t nvmlGpuInstanceGetComputeInstanceProfileInfo(nvmlGpuInstance_t gi, unsigned int profile, unsigned int eng, nvmlComputeInstanceProfileInfo_t *info) {
    mig_t *m = gi_mig(gi);
    if (!m) { trace("nvmlGpuInstanceGetComputeInstanceProfileInfo"); return NVML_ERROR_INVALID_ARGUMENT; }
    gpu_t *g = m->gpu;
    nvmlReturn_t rc = rc_for("nvmlGpuInstanceGetComputeInstanceProfileInfo", g, NULL);
    if (rc != NVML_SUCCESS) return rc;
    if (eng != 0) return NVML_ERROR_INVALID_ARGUMENT;
    /* the GI's profile id is that of the first MIG device carrying this gi_id (== RawNvml._gi_first_mig) */
    int gpid = m->gi_profile_id;
    for (int k = 0; k < g->nmigs; k++) if (g->migs[k].gi_id == m->gi_id) { gpid = g->migs[k].gi_profile_id; break; }
    if (gpid < 0 || gpid >= MAX_GI_IDX || profile >= MAX_CI_IDX || !g->ci_prof[gpid][profile].present) return NVML_ERROR_NOT_SUPPORTED;
    if (info) { memset(info, 0, sizeof *info); info->id = (unsigned)g->ci_prof[gpid][profile].id; info->sliceCount = (unsigned)g->ci_prof[gpid][profile].slices; }
    return rc;
}
