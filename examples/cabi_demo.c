/*
 * Minimal C consumer of libgfd_b200.so: what a cgo / JNI / ctypes binding does, in plain C.
 * Walks the node through the resource.Manager / resource.Device surface and prints what it finds.
 *
 *   gcc -Iinclude examples/cabi_demo.c -o examples/cabi_demo -Lgpu-feature-discovery_b200 -lgfd_b200 -Wl,-rpath,'$ORIGIN/../gpu-feature-discovery_b200'
 *   GFD_FAKE_NVML=tests/fixtures/c5_mixed_hetero.nvmlfix LD_LIBRARY_PATH=gpu-feature-discovery_b200/fake_nvml/lib examples/cabi_demo
 */
#include <inttypes.h>
#include <stdio.h>

#include "gfd_b200.h"

#define CHECK(call) do { int rc_ = (call); if (rc_) { fprintf(stderr, "%s: %s\n", #call, gfd_strerror(rc_)); return 1; } } while (0)

int main(void) {
    char reason[512];
    int has_nvml = gfd_has_nvml(reason, sizeof reason);
    printf("nvml: %s (%s)\n", has_nvml ? "yes" : "no", reason);
    gfd_mgr *m = NULL;
    CHECK(gfd_mgr_new(has_nvml ? GFD_BACKEND_NVML : GFD_BACKEND_NULL, 0, &m));
    CHECK(gfd_init(m));
    CHECK(gfd_prefetch(m, 0)); /* optional: one host thread per GPU */
    int n = 0;
    CHECK(gfd_device_count(m, &n));
    char buf[256];
    if (n > 0) {
        unsigned maj, min;
        CHECK(gfd_driver_version(m, buf, sizeof buf));
        CHECK(gfd_cuda_driver_version(m, &maj, &min));
        printf("driver %s, cuda %u.%u, %d device(s)\n", buf, maj, min, n);
    }
    for (int i = 0; i < n; i++) {
        gfd_dev *d = NULL;
        uint64_t mb = 0;
        int cmaj = 0, cmin = 0, capable = 0, enabled = 0;
        CHECK(gfd_device_get(m, i, &d));
        CHECK(gfd_dev_name(d, buf, sizeof buf));
        CHECK(gfd_dev_total_memory_mb(d, &mb));
        CHECK(gfd_dev_cuda_compute_capability(d, &cmaj, &cmin));
        CHECK(gfd_dev_is_mig_capable(d, &capable));
        CHECK(gfd_dev_is_mig_enabled(d, &enabled));
        printf("gpu %d: %s, %" PRIu64 " MiB, cc %d.%d, mig capable=%d enabled=%d\n", i, buf, mb, cmaj, cmin, capable, enabled);
        if (!enabled) continue;
        int migs = 0;
        CHECK(gfd_dev_mig_count(d, &migs));
        for (int k = 0; k < migs; k++) {
            gfd_dev *mig = NULL;
            gfd_mig_attributes a;
            CHECK(gfd_dev_mig_get(d, k, &mig));
            int rc = gfd_dev_name(mig, buf, sizeof buf);
            if (rc) { printf("  mig %d: profile unresolved: %s\n", k, gfd_strerror(rc)); continue; }
            CHECK(gfd_dev_attributes(mig, &a));
            gfd_dev *parent = NULL;  /* Device.GetDeviceHandleFromMigDeviceHandle (nvml-mig-device.go:56-72) */
            char pname[96];
            CHECK(gfd_dev_parent(mig, &parent));
            CHECK(gfd_dev_name(parent, pname, sizeof pname));
            printf("  mig %d: %s, %" PRIu64 " MiB, %u SMs, slices gi=%u ci=%u, parent %s\n", k, buf, a.memory_mb, a.multiprocessors, a.slices_gi, a.slices_ci, pname);
        }
    }
    uint64_t calls = 0, probes = 0;
    gfd_call_stats(m, &calls, &probes);
    printf("%" PRIu64 " NVML calls\n", calls);
    CHECK(gfd_shutdown(m));
    gfd_mgr_free(m);
    return 0;
}
