/* Experiment helper (not product): does opening the /dev/nvidiaN nodes from parallel threads BEFORE nvmlInit_v2 move the
 * per-GPU cost out of nvmlInit_v2's serial loop?   nvml_preopen_probe <iters> <mode>
 *   mode 0: plain nvmlInit_v2 ... first handle per GPU (threads) ... nvmlShutdown      (what the product does today)
 *   mode 1: open every /dev/nvidia[0-9]* concurrently, keep the fds, then the same sequence, then close the fds
 *   mode 2: as 1, but the fds are closed again BEFORE nvmlInit_v2 (is it the open itself, or holding the node open?) */
#define _GNU_SOURCE
#include <dirent.h>
#include <dlfcn.h>
#include <fcntl.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>
static double now(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec * 1e3 + t.tv_nsec / 1e6; }
static int (*byidx)(unsigned, void **);
static int (*name)(void *, char *, unsigned);
static char nodes[64][64];
static int fds[64], nnodes;
static double open_ms[64];
static void *opener(void *p) {
    int i = (int)(long)p;
    double t = now();
    fds[i] = open(nodes[i], O_RDWR | O_CLOEXEC);
    open_ms[i] = now() - t;
    return NULL;
}
static void *attach(void *p) {
    void *h = NULL;
    char nm[96];
    if (byidx((unsigned)(long)p, &h) == 0) name(h, nm, 96);
    return NULL;
}
int main(int argc, char **argv) {
    int iters = argc > 1 ? atoi(argv[1]) : 5, mode = argc > 2 ? atoi(argv[2]) : 0;
    DIR *d = opendir("/dev");
    struct dirent *e;
    while (d && (e = readdir(d)) && nnodes < 64)
        if (!strncmp(e->d_name, "nvidia", 6) && e->d_name[6] >= '0' && e->d_name[6] <= '9') snprintf(nodes[nnodes++], 64, "/dev/%.40s", e->d_name);
    if (d) closedir(d);
    printf("mode %d, %d device nodes\n", mode, nnodes);
    for (int it = 0; it < iters; it++) {
        pthread_t th[64];
        double t0 = now(), topen = 0;
        if (mode) {
            for (long i = 0; i < nnodes; i++) pthread_create(&th[i], NULL, opener, (void *)i);
            for (int i = 0; i < nnodes; i++) pthread_join(th[i], NULL);
            topen = now() - t0;
            if (mode == 2) for (int i = 0; i < nnodes; i++) if (fds[i] >= 0) close(fds[i]);
        }
        void *lib = dlopen("libnvidia-ml.so.1", RTLD_LAZY | RTLD_GLOBAL);
        if (!lib) { printf("no NVML\n"); return 1; }
        int (*init)(void) = dlsym(lib, "nvmlInit_v2");
        int (*shut)(void) = dlsym(lib, "nvmlShutdown");
        int (*count)(unsigned *) = dlsym(lib, "nvmlDeviceGetCount_v2");
        byidx = dlsym(lib, "nvmlDeviceGetHandleByIndex_v2");
        name = dlsym(lib, "nvmlDeviceGetName");
        double t1 = now();
        int rc = init();
        double t2 = now();
        unsigned c = 0;
        count(&c);
        for (long i = 0; i < (long)c; i++) pthread_create(&th[i], NULL, attach, (void *)i);
        for (unsigned i = 0; i < c; i++) pthread_join(th[i], NULL);
        double t3 = now();
        int rs = shut();
        double t4 = now();
        dlclose(lib);
        if (mode == 1) for (int i = 0; i < nnodes; i++) if (fds[i] >= 0) close(fds[i]);
        double t5 = now();
        double mx = 0;
        for (int i = 0; i < nnodes; i++) if (open_ms[i] > mx) mx = open_ms[i];
        printf("iter %d: pre-open %.3f ms (slowest single open %.3f), init rc=%d %.3f ms, %u handles %.3f ms, shutdown rc=%d %.3f ms, close %.3f ms, TOTAL %.3f ms\n",
               it, topen, mx, rc, t2 - t1, c, t3 - t2, rs, t4 - t3, t5 - t4, t5 - t0);
    }
    return 0;
}
