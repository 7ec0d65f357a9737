/* Experiment helper (not product, not a bench): where does a labeling pass's NVML time go?
 * Times dlopen / nvmlInit_v2 / a few calls / nvmlShutdown / dlclose separately, N iterations. */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <stdio.h>
#include <stdlib.h>
#include <signal.h>
#include <time.h>
static double now(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec * 1e3 + t.tv_nsec / 1e6; }
int main(int argc, char **argv) {
    int n = argc > 1 ? atoi(argv[1]) : 8;
    int block = argc > 2 ? atoi(argv[2]) : 0;
    const char *extra = argc > 3 ? argv[3] : NULL;
    if (block) { sigset_t s; sigemptyset(&s); sigaddset(&s, SIGHUP); sigaddset(&s, SIGINT); sigaddset(&s, SIGTERM); sigaddset(&s, SIGQUIT); sigprocmask(SIG_BLOCK, &s, NULL); }
    if (extra) { void *h = dlopen(extra, RTLD_NOW | RTLD_GLOBAL); printf("extra lib %s -> %p\n", extra, h); }
    for (int i = 0; i < n; i++) {
        double t0 = now();
        void *lib = dlopen("libnvidia-ml.so.1", RTLD_LAZY | RTLD_GLOBAL);
        double t1 = now();
        int (*init)(void) = dlsym(lib, "nvmlInit_v2");
        int (*shut)(void) = dlsym(lib, "nvmlShutdown");
        int (*count)(unsigned *) = dlsym(lib, "nvmlDeviceGetCount_v2");
        int (*byidx)(unsigned, void **) = dlsym(lib, "nvmlDeviceGetHandleByIndex_v2");
        int (*name)(void *, char *, unsigned) = dlsym(lib, "nvmlDeviceGetName");
        int rc = init();
        double t2 = now();
        unsigned c = 0; count(&c);
        void *h; byidx(0, &h); char nm[96]; name(h, nm, 96);
        double t3 = now();
        shut();
        double t4 = now();
        dlclose(lib);
        double t5 = now();
        printf("iter %d rc=%d dlopen %.3f init %.3f calls %.3f shutdown %.3f dlclose %.3f total %.3f ms\n", i, rc, t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4, t5 - t0);
    }
    return 0;
}
