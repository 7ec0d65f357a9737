/* Experiment helper (not product): WHEN does each GPU's first nvmlDeviceGetHandleByIndex return if all GPUs are asked at once?
 * One thread per GPU, started together after nvmlInit_v2; prints per-thread start / end (ms after the common start), sorted by end.
 * A staircase of ends means the driver serialises part of the attach.   nvml_attach_pattern <iters> [max threads] */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <time.h>
static double now(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec * 1e3 + t.tv_nsec / 1e6; }
static int (*byidx)(unsigned, void **);
static pthread_barrier_t bar;
static double T0, t_start[64], t_end[64];
static void *work(void *p) {
    long i = (long)p;
    void *h = NULL;
    pthread_barrier_wait(&bar);
    t_start[i] = now() - T0;
    byidx((unsigned)i, &h);
    t_end[i] = now() - T0;
    return NULL;
}
static int cmp(const void *a, const void *b) { double x = *(const double *)a, y = *(const double *)b; return x < y ? -1 : x > y; }
int main(int argc, char **argv) {
    int iters = argc > 1 ? atoi(argv[1]) : 5, cap = argc > 2 ? atoi(argv[2]) : 64;
    for (int it = 0; it < iters; it++) {
        void *lib = dlopen("libnvidia-ml.so.1", RTLD_LAZY | RTLD_GLOBAL);
        if (!lib) { printf("no NVML\n"); return 1; }
        int (*init)(void) = dlsym(lib, "nvmlInit_v2");
        int (*shut)(void) = dlsym(lib, "nvmlShutdown");
        int (*count)(unsigned *) = dlsym(lib, "nvmlDeviceGetCount_v2");
        byidx = dlsym(lib, "nvmlDeviceGetHandleByIndex_v2");
        double ti = now();
        init();
        ti = now() - ti;
        unsigned c = 0;
        count(&c);
        if ((int)c > cap) c = (unsigned)cap;
        pthread_t th[64];
        pthread_barrier_init(&bar, NULL, c + 1);
        for (long i = 0; i < (long)c; i++) pthread_create(&th[i], NULL, work, (void *)i);
        T0 = now();
        pthread_barrier_wait(&bar);
        for (unsigned i = 0; i < c; i++) pthread_join(th[i], NULL);
        double all = now() - T0;
        double ts = now();
        shut();
        ts = now() - ts;
        dlclose(lib);
        double ends[64];
        for (unsigned i = 0; i < c; i++) ends[i] = t_end[i];
        qsort(ends, c, sizeof(double), cmp);
        printf("iter %d: init %.2f ms, %u handles asked at once, all back after %.2f ms, shutdown %.2f ms; ends sorted:", it, ti, c, all, ts);
        for (unsigned i = 0; i < c; i++) printf(" %.2f", ends[i]);
        printf("\n");
    }
    return 0;
}
