/*
 * fake libcuda.so.1 -- TEST DOUBLE for the Tegra (CUDA driver API) manager on hosts without a GPU.
 *
 * Exports the 7 driver entry points the reference binds (internal/cuda/cuda.go:103-109) and answers them from
 * the text file named by $GFD_FAKE_CUDA:
 *     version <int>                       cuDriverGetVersion
 *     device <total bytes> <cc major> <cc minor> <name ...>
 *     rc <symbol> <CUresult>              return-code override
 * cuDeviceTotalMem is the LEGACY unversioned entry point (the reference binds that name, not _v2): it writes
 * a 32-bit value, SATURATED at 0xFFFFFFFF for a device with more than 4 GiB -- observed on the real libcuda of a B200 box
 * (driver 580.x: 4294967295 for a 191 502 876 672-byte device; tests/test_gpu_cuda_backend.py pins it).  The oracle models
 * the same.  There is still no Tegra hardware and no reference test for this path.
 */
#define _GNU_SOURCE
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define EXPORT __attribute__((visibility("default")))
#define MAXD 16
static struct { unsigned long long total; int major, minor; char name[128]; } D[MAXD];
static int nd, version = 12040, loaded;
static struct { char sym[48]; int code; } RC[16];
static int nrc;

static void load(void) {
    if (loaded) return;
    loaded = 1;
    const char *p = getenv("GFD_FAKE_CUDA");
    FILE *f = p ? fopen(p, "r") : NULL;
    if (!f) return;
    char line[512];
    while (fgets(line, sizeof line, f)) {
        size_t L = strlen(line);
        while (L && (line[L - 1] == '\n' || line[L - 1] == '\r')) line[--L] = 0;
        int off = 0;
        if (!strncmp(line, "version ", 8)) version = atoi(line + 8);
        else if (!strncmp(line, "rc ", 3) && nrc < 16) { if (sscanf(line + 3, "%47s %d", RC[nrc].sym, &RC[nrc].code) == 2) nrc++; }
        else if (!strncmp(line, "device ", 7) && nd < MAXD) {
            if (sscanf(line + 7, "%llu %d %d %n", &D[nd].total, &D[nd].major, &D[nd].minor, &off) >= 3) {
                snprintf(D[nd].name, sizeof D[nd].name, "%s", line + 7 + off);
                nd++;
            }
        }
    }
    fclose(f);
}
static int rc(const char *sym) {
    load();
    for (int i = 0; i < nrc; i++) if (!strcmp(RC[i].sym, sym)) return RC[i].code;
    return 0;
}
EXPORT int cuInit(unsigned flags) { (void)flags; return rc("cuInit"); }
EXPORT int cuDriverGetVersion(int *v) { int r = rc("cuDriverGetVersion"); if (v) *v = version; return r; }
EXPORT int cuDeviceGetCount(int *c) { int r = rc("cuDeviceGetCount"); if (c) *c = nd; return r; }
EXPORT int cuDeviceGet(int *dev, int ordinal) { int r = rc("cuDeviceGet"); if (r) return r; if (ordinal < 0 || ordinal >= nd) return 101; *dev = ordinal; return 0; }
EXPORT int cuDeviceGetAttribute(int *pi, int attrib, int dev) {
    int r = rc("cuDeviceGetAttribute");
    if (r) return r;
    if (dev < 0 || dev >= nd) return 101;
    *pi = attrib == 75 ? D[dev].major : attrib == 76 ? D[dev].minor : 0;
    return 0;
}
EXPORT int cuDeviceTotalMem(unsigned int *bytes, int dev) {
    int r = rc("cuDeviceTotalMem");
    if (r) return r;
    if (dev < 0 || dev >= nd) return 101;
    *bytes = D[dev].total > 0xFFFFFFFFull ? 0xFFFFFFFFu : (unsigned int)D[dev].total;
    return 0;
}
EXPORT int cuDeviceGetName(char *name, int len, int dev) {
    int r = rc("cuDeviceGetName");
    if (r) return r;
    if (dev < 0 || dev >= nd) return 101;
    snprintf(name, (size_t)len, "%s", D[dev].name);
    return 0;
}
