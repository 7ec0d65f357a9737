/* Experiment helper (not product): does NVML_INIT_FLAG_NO_ATTACH + a per-GPU handle lookup by PCI bus id give a usable handle,
 * and what do init / lookup / shutdown cost then?   nvml_noattach_probe <iters> <flags> <threads: 0 = serial> */
#define _GNU_SOURCE
#include <dirent.h>
#include <dlfcn.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
static double now(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec * 1e3 + t.tv_nsec / 1e6; }
typedef struct { unsigned long long total, free, used; } mem_t;
static int (*bybus)(const char *, void **);
static int (*name)(void *, char *, unsigned);
static int (*meminfo)(void *, mem_t *);
static int (*cc)(void *, int *, int *);
static int (*migmode)(void *, unsigned *, unsigned *);
static char ids[64][32];
static int nids;
typedef struct { int i; int rc, rcn, rcm; double ms; char nm[96]; unsigned long long total; int maj, min; } job_t;
static void *work(void *p) {
    job_t *j = p;
    double t = now();
    void *h = NULL;
    j->rc = bybus(ids[j->i], &h);
    j->ms = now() - t;
    if (j->rc == 0) {
        mem_t m = {0};
        unsigned cur, pend;
        j->rcn = name(h, j->nm, 96);
        j->rcm = meminfo(h, &m);
        j->total = m.total;
        cc(h, &j->maj, &j->min);
        migmode(h, &cur, &pend);
    }
    return NULL;
}
int main(int argc, char **argv) {
    int iters = argc > 1 ? atoi(argv[1]) : 3, flags = argc > 2 ? atoi(argv[2]) : 2, threads = argc > 3 ? atoi(argv[3]) : 0;
    /* NVIDIA 3D controllers / VGA devices from sysfs (class 0x0302xx / 0x0300xx, vendor 0x10de) */
    DIR *d = opendir("/sys/bus/pci/devices");
    if (d) {
        struct dirent *e;
        while ((e = readdir(d)) && nids < 64) {
            if (e->d_name[0] == '.' || strlen(e->d_name) > 30) continue;
            char path[300], buf[32] = {0};
            snprintf(path, sizeof path, "/sys/bus/pci/devices/%s/vendor", e->d_name);
            FILE *f = fopen(path, "r");
            if (!f) continue;
            if (!fgets(buf, sizeof buf, f)) buf[0] = 0;
            fclose(f);
            if (strncmp(buf, "0x10de", 6)) continue;
            snprintf(path, sizeof path, "/sys/bus/pci/devices/%s/class", e->d_name);
            f = fopen(path, "r");
            if (!f) continue;
            if (!fgets(buf, sizeof buf, f)) buf[0] = 0;
            fclose(f);
            if (strncmp(buf, "0x0302", 6) && strncmp(buf, "0x0300", 6)) continue;
            memcpy(ids[nids++], e->d_name, strlen(e->d_name) + 1);
        }
        closedir(d);
    }
    printf("NVIDIA GPU functions in sysfs: %d\n", nids);
    for (int it = 0; it < iters; it++) {
        double t0 = now();
        void *lib = dlopen("libnvidia-ml.so.1", RTLD_LAZY | RTLD_GLOBAL);
        if (!lib) { printf("no NVML\n"); return 1; }
        int (*initf)(unsigned) = dlsym(lib, "nvmlInitWithFlags");
        int (*shut)(void) = dlsym(lib, "nvmlShutdown");
        int (*count)(unsigned *) = dlsym(lib, "nvmlDeviceGetCount_v2");
        bybus = dlsym(lib, "nvmlDeviceGetHandleByPciBusId_v2");
        name = dlsym(lib, "nvmlDeviceGetName");
        meminfo = dlsym(lib, "nvmlDeviceGetMemoryInfo");
        cc = dlsym(lib, "nvmlDeviceGetCudaComputeCapability");
        migmode = dlsym(lib, "nvmlDeviceGetMigMode");
        double t1 = now();
        int rc = initf((unsigned)flags);
        double t2 = now();
        unsigned c = 0;
        int rcc = count(&c);
        job_t jobs[64];
        pthread_t th[64];
        memset(jobs, 0, sizeof jobs);
        double t3 = now();
        for (int i = 0; i < nids; i++) {
            jobs[i].i = i;
            if (threads) pthread_create(&th[i], NULL, work, &jobs[i]); else work(&jobs[i]);
        }
        if (threads) for (int i = 0; i < nids; i++) pthread_join(th[i], NULL);
        double t4 = now();
        unsigned c2 = 0;
        count(&c2);
        int rs = shut();
        double t5 = now();
        dlclose(lib);
        printf("iter %d flags %d threads %d: init rc=%d %.3f ms, count rc=%d n=%u (after lookups n=%u), lookups %.3f ms, shutdown rc=%d %.3f ms, total %.3f ms\n",
               it, flags, threads, rc, t2 - t1, rcc, c, c2, t4 - t3, rs, t5 - t4, t5 - t0);
        for (int i = 0; i < nids && it == 0; i++)
            printf("   %s: bybus rc=%d %.3f ms name rc=%d '%s' mem rc=%d total=%llu cc %d.%d\n", ids[i], jobs[i].rc, jobs[i].ms, jobs[i].rcn, jobs[i].nm, jobs[i].rcm, jobs[i].total, jobs[i].maj, jobs[i].min);
    }
    return 0;
}
