/* Experiment helper (not product, not a bench): where does a labeling pass's NVML time go, and can it be spread?
 *   nvml_init_timing <iters> <init_flags> <threads>
 * init_flags: 0 = nvmlInit_v2 (what the reference does), 2 = NVML_INIT_FLAG_NO_ATTACH via nvmlInitWithFlags.
 * threads: 1 = serial per-GPU queries, N = one pthread per GPU.  Prints per-phase ms per iteration. */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
static double now(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec * 1e3 + t.tv_nsec / 1e6; }
typedef struct { unsigned long long total, free, used; } mem_t;
static int (*byidx)(unsigned, void **);
static int (*name)(void *, char *, unsigned);
static int (*meminfo)(void *, mem_t *);
static int (*cc)(void *, int *, int *);
static int (*migmode)(void *, unsigned *, unsigned *);
typedef struct { unsigned idx; double ms_handle, ms_rest; int rc; char name[96]; unsigned long long mem; int maj, min; unsigned mig; } job_t;
static void *work(void *p) {
    job_t *j = p;
    double t0 = now();
    void *h = NULL;
    j->rc = byidx(j->idx, &h);
    double t1 = now();
    if (!j->rc) {
        mem_t m = {0};
        unsigned cur = 0, pend = 0;
        j->rc |= name(h, j->name, 96);
        j->rc |= meminfo(h, &m);
        j->rc |= cc(h, &j->maj, &j->min);
        int r = migmode(h, &cur, &pend);
        j->mem = m.total; j->mig = cur; (void)r;
    }
    j->ms_handle = t1 - t0;
    j->ms_rest = now() - t1;
    return NULL;
}
int main(int argc, char **argv) {
    int n = argc > 1 ? atoi(argv[1]) : 6;
    unsigned flags = argc > 2 ? (unsigned)atoi(argv[2]) : 0;
    int threads = argc > 3 ? atoi(argv[3]) : 1;
    for (int i = 0; i < n; i++) {
        double t0 = now();
        void *lib = dlopen("libnvidia-ml.so.1", RTLD_LAZY | RTLD_GLOBAL);
        if (!lib) { printf("no NVML\n"); return 1; }
        int (*init)(void) = dlsym(lib, "nvmlInit_v2");
        int (*initf)(unsigned) = dlsym(lib, "nvmlInitWithFlags");
        int (*shut)(void) = dlsym(lib, "nvmlShutdown");
        int (*count)(unsigned *) = dlsym(lib, "nvmlDeviceGetCount_v2");
        byidx = dlsym(lib, "nvmlDeviceGetHandleByIndex_v2");
        name = dlsym(lib, "nvmlDeviceGetName");
        meminfo = dlsym(lib, "nvmlDeviceGetMemoryInfo");
        cc = dlsym(lib, "nvmlDeviceGetCudaComputeCapability");
        migmode = dlsym(lib, "nvmlDeviceGetMigMode");
        double t1 = now();
        int rc = flags ? initf(flags) : init();
        double t2 = now();
        unsigned c = 0;
        int rcc = count(&c);
        double t3 = now();
        job_t jobs[64];
        pthread_t th[64];
        memset(jobs, 0, sizeof jobs);
        if (c > 64) c = 64;
        for (unsigned g = 0; g < c; g++) jobs[g].idx = g;
        if (threads > 1) {
            for (unsigned g = 0; g < c; g++) pthread_create(&th[g], NULL, work, &jobs[g]);
            for (unsigned g = 0; g < c; g++) pthread_join(th[g], NULL);
        } else {
            for (unsigned g = 0; g < c; g++) work(&jobs[g]);
        }
        double t4 = now();
        shut();
        double t5 = now();
        dlclose(lib);
        double t6 = now();
        double hmax = 0, hsum = 0;
        int bad = 0;
        for (unsigned g = 0; g < c; g++) { if (jobs[g].ms_handle > hmax) hmax = jobs[g].ms_handle; hsum += jobs[g].ms_handle; bad |= jobs[g].rc; }
        printf("iter %d flags=%u threads=%d rc=%d/%d gpus=%u bad=%d: dlopen+dlsym %.2f init %.2f count %.2f perGPU %.2f (handle sum %.2f max %.2f) shutdown %.2f dlclose %.2f total %.2f ms | gpu0 '%s' %llu MiB cc %d.%d mig %u\n",
               i, flags, threads, rc, rcc, c, bad, t1 - t0, t2 - t1, t3 - t2, t4 - t3, hsum, hmax, t5 - t4, t6 - t5, t6 - t0,
               c ? jobs[0].name : "", c ? jobs[0].mem >> 20 : 0ull, c ? jobs[0].maj : 0, c ? jobs[0].min : 0, c ? jobs[0].mig : 0);
    }
    return 0;
}
