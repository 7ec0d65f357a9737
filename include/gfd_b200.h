/*
 * gfd_b200.h -- C ABI of libgfd_b200.so, the B200-native backend for GPU Feature Discovery.
 *
 * This is the drop-in boundary: everything the reference's `resource.Manager` / `resource.Device`
 * interfaces (internal/resource/types.go:22-42) obtain through go-nvml + go-nvlib cgo bindings is
 * obtained here through plain C calls (pointers and sizes only), so a Go `resource.Manager` can bind
 * it with a ~100-line cgo stub and the C++ labeler in this repo binds it directly.
 *
 * Citations are reference paths (relative to the reference root); V/nvlib =
 * vendor/gitlab.com/nvidia/cloud-native/go-nvlib/pkg, V/go-nvml = vendor/github.com/NVIDIA/go-nvml/pkg.
 *
 * Return convention (every function returning `int` unless stated otherwise):
 *      0                      success
 *      1 .. 99999             an nvmlReturn_t, verbatim (e.g. 3 = NVML_ERROR_NOT_SUPPORTED, 999 = UNKNOWN)
 *      GFD_ERR_CUDA + e       a cudaError_t / CUresult e from the CUDA runtime/driver
 *      GFD_ERR_*              shim-level errors (fixed texts of the reference's own fmt.Errorf sites)
 * gfd_strerror() turns any of them into text.  No function throws, calls back, or retains caller memory.
 *
 * Threading: one thread may drive a manager start to finish (the reference is single-goroutine);
 * distinct managers may be used from distinct threads.  Handles (gfd_dev*) are owned by their manager
 * and are valid from gfd_init() until gfd_shutdown() -- the same lifetime NVML gives device handles
 * inside the reference's Init()...defer Shutdown() bracket (internal/lm/nvml.go:30-33).  A handle used after
 * gfd_shutdown() is answered with GFD_ERR_NOT_INITIALIZED; once the manager is initialised AGAIN (or freed) old
 * handles dangle and must not be used -- fetch new ones with gfd_device_get().
 *
 * The library has NO link-time dependency on libnvidia-ml or libcuda: both are dlopen()ed on first
 * use, so it loads on a CPU-only host (BASELINE config #1).  GPU queries have no CPU fallback: where
 * the driver is missing the call fails with an error code.
 */
#ifndef GFD_B200_H
#define GFD_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GFD_OK 0
#define GFD_ERR_CUDA 100000 /* + cudaError_t / CUresult */
#define GFD_ERR_SHIM 200000
enum {
    GFD_ERR_BAD_ARGUMENT = GFD_ERR_SHIM + 1,
    GFD_ERR_BUFFER_TOO_SMALL,
    GFD_ERR_NOT_INITIALIZED,      /* call outside the gfd_init()..gfd_shutdown() bracket */
    GFD_ERR_UNSUPPORTED_BACKEND,  /* null manager: "GetDriverVersion is unsupported" (internal/resource/null.go:50-57) */
    GFD_ERR_NOT_MIG_DEVICE,       /* "... is not supported for non-MIG devices" (internal/resource/nvml-device.go:63-70) */
    GFD_ERR_IS_MIG_DEVICE,        /* "... is not supported for MIG devices" (internal/resource/nvml-mig-device.go:75-92) */
    GFD_ERR_NOT_A_MIG_HANDLE,     /* "not a MIG device" (V/nvlib/nvlib/device/mig_device.go:45-47) */
    GFD_ERR_NO_MATCHING_PROFILE,  /* "no matching profile IDs found" (mig_device.go:147) */
    GFD_ERR_INVALID_GI_PROFILE,   /* "invalid GPU Instance Profile ID" (V/nvlib/nvlib/device/mig_profile.go:74-75) */
    GFD_ERR_INVALID_CI_PROFILE,   /* "invalid Compute Instance Profile ID" (mig_profile.go:94-95) */
    GFD_ERR_PCI_CONFIG_SHORT,     /* "entire PCI configuration is not read" (internal/vgpu/pciutil.go:116-118) */
    GFD_ERR_PCI_OUT_OF_RANGE,     /* where the reference's Go code would panic with index/slice out of range */
    GFD_ERR_VGPU_NO_VERSION_RECORD, /* "cannot find driver version record" (internal/vgpu/vgpu.go:143-145) */
    GFD_ERR_VGPU_RECORD_LOOP,     /* zero-length record: the reference spins forever (vgpu.go:121-126); we fail */
    GFD_ERR_NO_GPU,               /* a probe was asked for on a host without a usable CUDA device */
    GFD_ERR_PEER_ACCESS,          /* the two devices cannot map each other's memory */
    GFD_ERR_SIZE_LIMIT,           /* request larger than a stated maximum */
    GFD_ERR_ALIGNMENT,            /* device pointer not 16-byte aligned */
    GFD_ERR_DATA_MISMATCH         /* a probe read back something else than the pattern it had written */
};

/* Text for any status code.  For NVML codes: nvmlErrorString() while NVML is initialised, else the fixed
 * table of V/nvlib/nvml/return.go:38-93.  Returned pointer is valid until the next call on this thread. */
const char *gfd_strerror(int status);

/* ------------------------------------------------------------------------------------------------
 * Platform detection -- replaces info.HasNvml / info.IsTegraSystem (V/nvlib/nvlib/info/info.go:41-78)
 * used by resource.getManager (internal/resource/factory.go:42-74).
 * Return 1 / 0; `reason` receives the text the reference logs ("found NVML library", "could not load
 * NVML: <dlerror>", "<file> found", ...).  `root` is prepended to the Tegra probe paths ("" in prod).
 * ---------------------------------------------------------------------------------------------- */
int gfd_has_nvml(char *reason, size_t reason_len);
int gfd_is_tegra_system(const char *root, char *reason, size_t reason_len);

/* ------------------------------------------------------------------------------------------------
 * resource.Manager  (internal/resource/types.go:22-30)
 * ---------------------------------------------------------------------------------------------- */
typedef struct gfd_mgr gfd_mgr;
typedef struct gfd_dev gfd_dev;

enum gfd_backend {
    GFD_BACKEND_NVML = 0, /* NewNVMLManager  (internal/resource/nvml-lib.go:29-38)  */
    GFD_BACKEND_CUDA = 1, /* NewCudaManager  (internal/resource/cuda-lib.go:29-32; Tegra, libcuda only) */
    GFD_BACKEND_NULL = 2  /* NewNullManager  (internal/resource/null.go:30-32)      */
};

/* Flags for gfd_mgr_new */
#define GFD_MGR_KEEP_ATTACHED 0x1u /* hold one extra nvmlInit reference for the manager's lifetime so that
                                      every pass after the first skips the driver attach (new; opt-in) */

#define GFD_MGR_CURRENT_TABLES 0x2u /* resolve MIG GI/CI profile ids beyond the reference's NVML 11.x header (ids up to
                                      0x10 / 7 of the CUDA 12.9 nvml.h, e.g. B200 1g.45gb) instead of failing with
                                      "no matching profile IDs found" (new; opt-in; no parity claim) */

int gfd_mgr_new(int backend, unsigned flags, gfd_mgr **out); /* replaces resource.NewManager's backend pick */
void gfd_mgr_free(gfd_mgr *m);

/* Manager.Init / Manager.Shutdown (nvml-lib.go:82-97 -> V/nvlib/nvml/nvml.go:38-70 -> V/go-nvml/nvml/init.go:33-72):
 * dlopen("libnvidia-ml.so.1", RTLD_LAZY|RTLD_GLOBAL) + nvmlInit_v2 / nvmlShutdown + dlclose.
 * A missing library yields NVML_ERROR_LIBRARY_NOT_FOUND (12), as go-nvml does (init.go:36-38). */
int gfd_init(gfd_mgr *m);
int gfd_shutdown(gfd_mgr *m);

/* Manager.GetDriverVersion (nvml-lib.go:73-79): nvmlSystemGetDriverVersion, 80-byte buffer semantics. */
int gfd_driver_version(gfd_mgr *m, char *buf, size_t len);
/* Manager.GetCudaDriverVersion (nvml-lib.go:42-51): major = v/1000, minor = v%1000/10. */
int gfd_cuda_driver_version(gfd_mgr *m, unsigned *major, unsigned *minor);

/* Manager.GetDevices (nvml-lib.go:54-70 -> V/nvlib/nvlib/device/device.go:206-236,288-298): the walk
 * nvmlDeviceGetCount_v2 / nvmlDeviceGetHandleByIndex_v2 / nvmlDeviceGetName with the "NVIDIA DGX Display"
 * skip (V/nvlib/nvlib/device/api.go:53-55).  The walk runs ONCE per Init..Shutdown bracket and is cached
 * (the reference repeats it 4x per pass).  idx is the index AFTER the skip. */
int gfd_device_count(gfd_mgr *m, int *count);
int gfd_device_get(gfd_mgr *m, int idx, gfd_dev **dev);

/* Optional: fetch every per-GPU property the labeler may ask for (name, memory, compute capability,
 * MIG mode, MIG devices and their profiles) with `threads` host threads, one GPU per thread
 * (the path partitions per GPU, no collective).  Results AND error codes are cached per
 * (device, property); nothing is reported here -- an error surfaces only when the matching getter is
 * called, so the set of failing passes is exactly the reference's.  When called before the first
 * gfd_device_count of the bracket it also spreads the device walk itself (nvmlDeviceGetHandleByIndex is where
 * the driver attaches a GPU, ~8 ms each on B200).  threads == 1: no-op (serial); threads <= 0: one thread per
 * GPU, at most min(16, host cores).  Measured on 8 x B200: pass 152 ms serial -> 104 ms (profiles/). */
int gfd_prefetch(gfd_mgr *m, int threads);

/* How many NVML entry-point calls / dlopen probes this manager has made since gfd_init (instrumentation
 * for the labeling-pass report: the reference needs ~92 calls + 10 dlopen cycles for 8 GPUs). */
int gfd_call_stats(gfd_mgr *m, uint64_t *nvml_calls, uint64_t *dl_probes);

/* ------------------------------------------------------------------------------------------------
 * resource.Device  (internal/resource/types.go:32-42); impls nvml-device.go:26-88, nvml-mig-device.go:27-134
 * ---------------------------------------------------------------------------------------------- */
/* Device.GetName: full GPU -> nvmlDeviceGetName (96-byte buffer, V/go-nvml/nvml/device.go:60-64);
 * MIG device -> its profile string "<g>g.<gb>gb" | "<c>c.<g>g.<gb>gb" with '+' replaced by '.'
 * (nvml-mig-device.go:96-104 -> V/nvlib/nvlib/device/mig_device.go:52-148, mig_profile.go:55-115,200-209). */
int gfd_dev_name(gfd_dev *d, char *buf, size_t len);
/* Device.GetTotalMemoryMB: nvmlDeviceGetMemoryInfo (v1 struct).total / 2^20 (nvml-device.go:82-88);
 * MIG device -> attributes.memorySizeMB (nvml-mig-device.go:107-118). */
int gfd_dev_total_memory_mb(gfd_dev *d, uint64_t *mb);
/* Device.GetCudaComputeCapability (nvml-device.go:53-60); MIG device -> GFD_ERR_IS_MIG_DEVICE. */
int gfd_dev_cuda_compute_capability(gfd_dev *d, int *major, int *minor);
/* Device.IsMigCapable / IsMigEnabled (V/nvlib/nvlib/device/device.go:55-88): dlsym probe for
 * nvmlDeviceGetMigMode (missing => false), NOT_SUPPORTED => false, enabled iff current == 1. */
int gfd_dev_is_mig_capable(gfd_dev *d, int *capable);
int gfd_dev_is_mig_enabled(gfd_dev *d, int *enabled);
/* Device.GetMigDevices (nvml-device.go:34-50 -> device.go:91-118 -> mig_device.go:40-49): walk
 * 0..nvmlDeviceGetMaxMigDeviceCount skipping NOT_FOUND / INVALID_ARGUMENT; cached per bracket. */
int gfd_dev_mig_count(gfd_dev *d, int *count);
int gfd_dev_mig_get(gfd_dev *d, int idx, gfd_dev **mig);
/* Device.GetDeviceHandleFromMigDeviceHandle (nvml-mig-device.go:56-72). */
int gfd_dev_parent(gfd_dev *mig, gfd_dev **parent);

/* Device.GetAttributes (nvml-mig-device.go:35-53): nvmlDeviceGetAttributes_v2; the 9 label values. */
typedef struct gfd_mig_attributes {
    uint64_t memory_mb;        /* "memory"          */
    uint32_t multiprocessors;  /* "multiprocessors" */
    uint32_t slices_gi;        /* "slices.gi"       */
    uint32_t slices_ci;        /* "slices.ci"       */
    uint32_t engines_copy;     /* "engines.copy"    */
    uint32_t engines_decoder;  /* "engines.decoder" */
    uint32_t engines_encoder;  /* "engines.encoder" */
    uint32_t engines_jpeg;     /* "engines.jpeg"    */
    uint32_t engines_ofa;      /* "engines.ofa"     */
} gfd_mig_attributes;
int gfd_dev_attributes(gfd_dev *mig, gfd_mig_attributes *out);

/* New (no reference counterpart; used only by the opt-in probes to map an NVML device to a CUDA ordinal). */
int gfd_dev_uuid(gfd_dev *d, char *buf, size_t len);

/* ------------------------------------------------------------------------------------------------
 * The MIG arithmetic, exposed on its own so it can be tested against the oracle directly.
 * ---------------------------------------------------------------------------------------------- */
/* getMigMemorySizeGB (V/nvlib/nvlib/device/mig_profile.go:324-332), float64, Go math.Round semantics. */
uint64_t gfd_mig_memory_size_gb(uint64_t total_device_memory_bytes, uint64_t mig_memory_size_mb);
/* NewMigProfile + MigProfileInfo.String (mig_profile.go:55-115,200-209), '+me' NOT yet dotted. */
int gfd_mig_profile_string(int gi_profile_id, int ci_profile_id, uint64_t mig_memory_size_mb,
                           uint64_t device_memory_bytes, char *buf, size_t len);

/* ------------------------------------------------------------------------------------------------
 * vGPU host-driver detection from PCI config space -- replaces PCIDevice.GetVendorSpecificCapability,
 * Lib.IsVGPUDevice and Device.GetInfo (internal/vgpu/pciutil.go:115-149, vgpu.go:92-153).
 * `config` = the sysfs `config` file bytes; `cap` = len for exact Go slice semantics of a literal, or the
 * allocation os.ReadFile would have made (max(512, st_size+1)) for a sysfs read.
 * *is_vgpu = 0 and status 0 when the device has no vendor capability or it lacks the "VF" signature.
 * ---------------------------------------------------------------------------------------------- */
typedef struct gfd_vgpu_info {
    int has_vendor_capability;
    int vendor_capability_len;
    int is_vgpu;
    char host_driver_version[32]; /* 10 config bytes, each UTF-8 encoded as Go's string(uint8) does, NULs trimmed
                                   * at both ENDS only (strings.Trim): a NUL in the middle stays, so use the lengths */
    char host_driver_branch[32];
    int host_driver_version_len;  /* bytes in host_driver_version (a terminating NUL follows them) */
    int host_driver_branch_len;
} gfd_vgpu_info;
int gfd_vgpu_parse(const uint8_t *config, size_t len, size_t cap, gfd_vgpu_info *out);

/* ------------------------------------------------------------------------------------------------
 * NEW, OPT-IN capability probes (sm_100a kernels; the reference launches no GPU work and has no
 * counterpart -- they produce extra labels only when asked for, never alter the default label set).
 *
 * Low level: caller owns device memory and the stream (`stream` is a cudaStream_t passed as void*;
 * NULL = the legacy default stream).  All device pointers must be 16-byte aligned.  Launches are
 * asynchronous; the device is taken from the pointer, not from the caller's current device.
 * ---------------------------------------------------------------------------------------------- */
typedef struct gfd_checksum {
    uint64_t xor64; /* XOR of all little-endian 64-bit words (tail zero-padded to 8 bytes) */
    uint64_t sum64; /* their sum mod 2^64 */
} gfd_checksum;

/* Fill `bytes` at d_buf with word[i] = 0x9E3779B97F4A7C15 * (i + seed) (SURVEY.md 8(d) probe input). */
int gfd_hbm_fill_pattern(void *d_buf, size_t bytes, uint64_t seed, void *stream);

/* The checksum a buffer of `bytes` filled by gfd_hbm_fill_pattern(…, seed) must have, computed on the GPU from the
 * pattern's definition without reading any buffer (same scratch rules as gfd_hbm_read_checksum -- header zeroed once with
 * gfd_hbm_scratch_init, declared below; the device is the one d_out lives on).  fill -> read_checksum == pattern_checksum is a write/read verification of the memory in between;
 * the self-contained probes below do exactly that and fail with GFD_ERR_DATA_MISMATCH otherwise. */
int gfd_hbm_pattern_checksum(size_t bytes, uint64_t seed, gfd_checksum *d_out, void *d_scratch, size_t scratch_bytes,
                             void *stream);
/* Bytes of device scratch gfd_hbm_read_checksum needs, for any input size: a 16-byte header (the tile counter the
 * CTAs draw work from + the count of finished CTAs) + one 16-byte partial sum per CTA.
 * The header must be ZERO before the first call that uses a scratch: call gfd_hbm_scratch_init() once after allocating it
 * (or zero its first 16 bytes yourself).  After that nothing is cleared per call: a checksum is exactly one kernel launch,
 * whose last CTA folds the partials, writes the result and puts the header back to zero.  Two calls sharing one scratch
 * must not overlap on the GPU (they would share the counters, and could leave them non-zero: re-init the scratch then). */
size_t gfd_hbm_scratch_bytes(void);
/* Zero the scratch header on `stream` (asynchronous).  Once per scratch allocation; also the way to recover a scratch
 * whose last user was killed mid-launch or raced with another. */
int gfd_hbm_scratch_init(void *d_scratch, size_t scratch_bytes, void *stream);
/* hbm_read_probe: stream `bytes` from HBM with 128-bit coalesced loads and fold them into *d_out
 * (device memory; the kernel runs on d_out's GPU -- a d_buf on another GPU is pulled over NVLink, the peer mapping being
 * enabled on first use, GFD_ERR_PEER_ACCESS if the pair has none).  Algorithmic traffic: `bytes` read.  d_scratch: gfd_hbm_scratch_bytes() bytes, initialised once
 * (above).  One kernel launch on `stream`, nothing else. */
int gfd_hbm_read_checksum(const void *d_buf, size_t bytes, gfd_checksum *d_out, void *d_scratch,
                          size_t scratch_bytes, void *stream);
/* Copy variant: d_dst[i] = d_src[i].  Algorithmic traffic: 2 * bytes.  Either pointer may be a
 * peer-mapped address on another GPU (that is the NVLink probe's kernel). */
int gfd_hbm_copy(void *d_dst, const void *d_src, size_t bytes, void *stream);

/* gfd_hbm_read_checksum without any allocation, by the caller or by the library.  No counterpart in the
 * reference.  Meant for callers that hold the data on the GPU but must not (or cannot) allocate device memory
 * when they need the checksum: an agent next to a job that already owns almost all of HBM (its cudaMalloc can fail,
 * or tip the job over), code living under a framework's allocator rules.  The per-CTA partial sums go to a fixed
 * 16 KiB array inside the library's device image (each GPU that loads the library's kernels holds it until exit).
 *   - bytes <= GFD_NOALLOC_MAX_BYTES, else GFD_ERR_SIZE_LIMIT (split the buffer at 8-byte multiples and combine
 *     the pieces: xor of xor64, sum of sum64); also GFD_ERR_SIZE_LIMIT on a GPU whose launch would need more than
 *     1023 CTAs (the array holds a 16-byte header + 1023 partials; a B200 uses 296).
 *   - the first call on each GPU resolves the array's address and may load the kernels there: do it once at
 *     start-up, before any stream capture begins.
 *   - calls on the same GPU share the array; ones that overlap in time on the GPU corrupt each other (and may leave
 *     the header non-zero, which spoils later calls too).  Issue them on one stream or serialise them.  Nothing needs
 *     clearing between calls: the header is zero when the library is loaded and each launch zeroes it again as it ends.
 *   - d_buf is validated before anything is launched: it must be device (or managed) memory ON THE SAME GPU as d_out.  A host
 *     pointer, or a buffer on another GPU (reading it would need a peer mapping, which this entry point will not set up),
 *     is refused with GFD_ERR_BAD_ARGUMENT instead of faulting the caller's context; gfd_hbm_read_checksum takes peer input.
 * Same result, bit for bit, as gfd_hbm_read_checksum.  d_buf 16-byte aligned, d_out device memory.  One kernel launch. */
#define GFD_NOALLOC_MAX_BYTES (16ull << 30)
int gfd_hbm_read_checksum_noalloc(const void *d_buf, size_t bytes, gfd_checksum *d_out, void *stream);

/* Host-buffer form (the end-to-end call): copy `bytes` from host memory to GPU `cuda_device`, checksum
 * them there, bring the 16-byte result back.  Staging memory is allocated on first use and cached. */
int gfd_checksum_host(int cuda_device, const void *host_buf, size_t bytes, gfd_checksum *out);

/* Self-contained probes: allocate, fill, warm up, time `iters` launches with CUDA events, verify the bytes that arrived
 * against the pattern (GFD_ERR_DATA_MISMATCH otherwise), free.  gfd_probe_hbm takes READ, COPY or WRITE. */
enum gfd_probe_mode {
    GFD_PROBE_READ = 0,
    GFD_PROBE_COPY = 1,
    GFD_PROBE_PUSH = 2, /* p2p only: the source GPU stores into the peer */
    GFD_PROBE_WRITE = 3, /* hbm only: the fill kernel is what gets timed (bytes written / time) */
    GFD_PROBE_DMA = 4   /* p2p only: the destination GPU's stream pulls with cudaMemcpyAsync (the copy engine, no kernel of this
                           library) -- the link's own limit beside what SM-issued loads and stores get; verified like COPY */
};
typedef struct gfd_probe_result {
    double gbps;           /* algorithmic bytes (bytes, or 2*bytes for copy over HBM; bytes per direction for p2p) / time */
    double ms_per_launch;
    uint64_t bytes;        /* buffer size used */
    int iters;
    int mode;
    int sm_count;
    int grid, block;
    gfd_checksum checksum; /* of the bytes that arrived: the source as read by the last timed launch (read mode), the
                            * destination re-read after the last launch (copy, push) */
    gfd_checksum expected; /* what the fill pattern must fold to (gfd_hbm_pattern_checksum) */
    int verified;          /* 1 = checksum == expected (otherwise the call returns GFD_ERR_DATA_MISMATCH) */
} gfd_probe_result;
int gfd_probe_hbm(int cuda_device, size_t bytes, int warmup, int iters, int mode, gfd_probe_result *out);
/* NVLink peer probe: a kernel running on `dst_device` pulls `bytes` from a buffer on `src_device`
 * through a peer mapping (read = fold, copy = store locally).  No NCCL: nothing is reduced. */
int gfd_probe_p2p(int src_device, int dst_device, size_t bytes, int warmup, int iters, int mode,
                  gfd_probe_result *out);
/* All directed pairs of the first n_devices GPUs, measured CONCURRENTLY: n-1 rounds of n/2 disjoint pairs (round-robin
 * tournament), both directions of every pair at once, so each round keeps all n GPUs pulling from a distinct peer.
 * gbps_matrix[runner * n + peer] receives GB/s (diagonal 0); runner = the GPU whose kernel moved the bytes (the puller for
 * READ/COPY, the pusher for PUSH).  *ring_aggregate_gbps (optional) = sum over GPUs when
 * every GPU pulls from its right-hand neighbour simultaneously.  One kernel per direction, no NCCL, no collective. */
int gfd_probe_p2p_matrix(int n_devices, size_t bytes, int warmup, int iters, int mode, double *gbps_matrix,
                         double *ring_aggregate_gbps);
/* CUDA ordinal whose UUID equals `uuid` ("GPU-xxxxxxxx-...."), or an error. */
int gfd_cuda_device_by_uuid(const char *uuid, int *ordinal);
int gfd_cuda_device_count(int *count);

#ifdef __cplusplus
}
#endif
#endif /* GFD_B200_H */
