// MIG profile arithmetic, status text, and the vGPU PCI config-space parser for libgfd_b200.so.
//
// Replaces: V/nvlib/nvlib/device/mig_profile.go:55-115,200-209,324-332 (NewMigProfile, String,
// getMigMemorySizeGB), V/nvlib/nvml/return.go:26-93 (status text), internal/vgpu/pciutil.go:115-149 and
// internal/vgpu/vgpu.go:92-153 (vendor capability walk, "VF" signature, host-driver record).
#include <cmath>
#include <cstdio>
#include <cstring>

#include "gfd_internal.h"

namespace gfd {

// nvml.GPU_INSTANCE_PROFILE_* / COMPUTE_INSTANCE_PROFILE_* of the reference's header (const.go:538-571)
static int gi_slices(int id) {
    switch (id) {
    case 0: return 1;  // 1_SLICE
    case 7: return 1;  // 1_SLICE_REV1 (+me)
    case 1: return 2;
    case 2: return 3;
    case 3: return 4;
    case 6: return 6;
    case 4: return 7;
    case 5: return 8;
    default: return 0;
    }
}
static int ci_slices(int id) {
    switch (id) {
    case 0: return 1;
    case 1: return 2;
    case 2: return 3;
    case 3: return 4;
    case 6: return 6;
    case 4: return 7;
    case 5: return 8;
    default: return 0;
    }
}

// NEW, opt-in (GFD_MGR_CURRENT_TABLES): the GI/CI profile ids the CUDA 12.9 nvml.h adds after the reference's header
// was cut (nvml.h: NVML_GPU_INSTANCE_PROFILE_2_SLICE_REV1 .. _2_SLICE_ALL_ME, NVML_COMPUTE_INSTANCE_PROFILE_1_SLICE_REV1).
// Suffixes follow nvidia-smi's profile names: +me, +gfx, -me, +me.all.  No reference counterpart, no parity claim.
static int gi_slices_current(int id, const char **suffix) {
    *suffix = "";
    switch (id) {
    case 0x8: *suffix = "+me"; return 2;
    case 0x9: return 1;  // 1 slice with twice the memory: the GB figure already tells them apart (B200 1g.45gb)
    case 0xA: *suffix = "+gfx"; return 1;
    case 0xB: *suffix = "+gfx"; return 2;
    case 0xC: *suffix = "+gfx"; return 4;
    case 0xD: *suffix = "-me"; return 1;
    case 0xE: *suffix = "-me"; return 2;
    case 0xF: *suffix = "+me.all"; return 1;
    case 0x10: *suffix = "+me.all"; return 2;
    default: return 0;
    }
}

int mig_profile_string(int gi_profile_id, int ci_profile_id, uint64_t mig_mb, uint64_t dev_bytes, std::string *out, bool current_tables) {
    const char *suffix = (gi_profile_id == 7) ? "+me" : "";
    int g = gi_slices(gi_profile_id);
    if (!g && current_tables) g = gi_slices_current(gi_profile_id, &suffix);
    if (!g) return GFD_ERR_INVALID_GI_PROFILE;
    int c = ci_slices(ci_profile_id);
    if (!c && current_tables && ci_profile_id == 7) c = 1;  // NVML_COMPUTE_INSTANCE_PROFILE_1_SLICE_REV1
    if (!c) return GFD_ERR_INVALID_CI_PROFILE;
    // Go: GB: int(getMigMemorySizeGB(...)) printed with %d
    long long gb = (long long)gfd_mig_memory_size_gb(dev_bytes, mig_mb);
    char buf[96];
    if (c == g)
        snprintf(buf, sizeof buf, "%dg.%lldgb%s", g, gb, suffix);
    else
        snprintf(buf, sizeof buf, "%dc.%dg.%lldgb%s", c, g, gb, suffix);
    *out = buf;
    return 0;
}

}  // namespace gfd

GFD_API uint64_t gfd_mig_memory_size_gb(uint64_t total_device_memory, uint64_t mig_memory_size_mb) {
    // mig_profile.go:324-332 -- float64 throughout; math.Round is half-away-from-zero == C round().
    // Compiled without -ffast-math / FMA contraction so every step rounds as Go's does.
    const double frac_denominator = 8.0;
    const double one_mb = 1024.0 * 1024.0;
    const uint64_t one_gb = 1024ull * 1024ull * 1024ull;
    volatile double fractional = ((double)mig_memory_size_mb * one_mb) / (double)total_device_memory;
    fractional = std::ceil(fractional * frac_denominator) / frac_denominator;
    volatile double total_mem_gb = (double)((total_device_memory + one_gb - 1) / one_gb);
    volatile double prod = fractional * total_mem_gb;
    double r = std::round(prod);
    // Go's uint64(float) of NaN/negative/huge is implementation-defined; NVML never yields those for a live device.
    if (!(r >= 0.0)) return 0;
    if (r >= 18446744073709551615.0) return UINT64_MAX;
    return (uint64_t)r;
}

GFD_API int gfd_mig_profile_string(int gi_profile_id, int ci_profile_id, uint64_t mig_mb, uint64_t dev_bytes, char *buf,
                                   size_t len) {
    std::string s;
    int rc = gfd::mig_profile_string(gi_profile_id, ci_profile_id, mig_mb, dev_bytes, &s);
    if (rc) return rc;
    return gfd::copy_out(s, buf, len);
}

// ---- status text -----------------------------------------------------------------------------------
static const char *nvml_default_text(int r) {  // defaultErrorStringFunc (V/nvlib/nvml/return.go:38-93)
    switch (r) {
    case 0: return "SUCCESS";
    case 1: return "ERROR_UNINITIALIZED";
    case 2: return "ERROR_INVALID_ARGUMENT";
    case 3: return "ERROR_NOT_SUPPORTED";
    case 4: return "ERROR_NO_PERMISSION";
    case 5: return "ERROR_ALREADY_INITIALIZED";
    case 6: return "ERROR_NOT_FOUND";
    case 7: return "ERROR_INSUFFICIENT_SIZE";
    case 8: return "ERROR_INSUFFICIENT_POWER";
    case 9: return "ERROR_DRIVER_NOT_LOADED";
    case 10: return "ERROR_TIMEOUT";
    case 11: return "ERROR_IRQ_ISSUE";
    case 12: return "ERROR_LIBRARY_NOT_FOUND";
    case 13: return "ERROR_FUNCTION_NOT_FOUND";
    case 14: return "ERROR_CORRUPTED_INFOROM";
    case 15: return "ERROR_GPU_IS_LOST";
    case 16: return "ERROR_RESET_REQUIRED";
    case 17: return "ERROR_OPERATING_SYSTEM";
    case 18: return "ERROR_LIB_RM_VERSION_MISMATCH";
    case 19: return "ERROR_IN_USE";
    case 20: return "ERROR_MEMORY";
    case 21: return "ERROR_NO_DATA";
    case 22: return "ERROR_VGPU_ECC_NOT_SUPPORTED";
    case 23: return "ERROR_INSUFFICIENT_RESOURCES";
    case 999: return "ERROR_UNKNOWN";
    default: return nullptr;
    }
}

GFD_API const char *gfd_strerror(int status) {
    static thread_local char buf[160];
    if (status == 0) return "SUCCESS";
    if (status > 0 && status < GFD_ERR_CUDA) {
        const char *s = gfd::nvml_error_string_if_loaded(status);
        if (!s) s = nvml_default_text(status);
        if (s) return s;
        snprintf(buf, sizeof buf, "Unknown return value: %d", status);
        return buf;
    }
    if (status >= GFD_ERR_CUDA && status < GFD_ERR_SHIM) {
        snprintf(buf, sizeof buf, "CUDA error %d", status - GFD_ERR_CUDA);
        return buf;
    }
    switch (status) {
    case GFD_ERR_BAD_ARGUMENT: return "bad argument";
    case GFD_ERR_BUFFER_TOO_SMALL: return "buffer too small";
    case GFD_ERR_NOT_INITIALIZED: return "manager is not initialized";
    case GFD_ERR_UNSUPPORTED_BACKEND: return "operation is unsupported by this manager";
    case GFD_ERR_NOT_MIG_DEVICE: return "operation is not supported for non-MIG devices";
    case GFD_ERR_IS_MIG_DEVICE: return "operation is not supported for MIG devices";
    case GFD_ERR_NOT_A_MIG_HANDLE: return "not a MIG device";
    case GFD_ERR_NO_MATCHING_PROFILE: return "no matching profile IDs found";
    case GFD_ERR_INVALID_GI_PROFILE: return "invalid GPU Instance Profile ID";
    case GFD_ERR_INVALID_CI_PROFILE: return "invalid Compute Instance Profile ID";
    case GFD_ERR_PCI_CONFIG_SHORT:
        return "entire PCI configuration is not read. Please run GFD with privileged mode to read complete PCI "
               "configuration data";
    case GFD_ERR_PCI_OUT_OF_RANGE: return "PCI configuration index out of range";
    case GFD_ERR_VGPU_NO_VERSION_RECORD: return "cannot find driver version record in vendor specific capability";
    case GFD_ERR_VGPU_RECORD_LOOP: return "zero-length record in vGPU vendor specific capability";
    case GFD_ERR_NO_GPU: return "no usable CUDA device (the probes have no CPU fallback)";
    case GFD_ERR_PEER_ACCESS: return "peer access between the two devices is not available";
    case GFD_ERR_SIZE_LIMIT: return "size exceeds the supported maximum";
    case GFD_ERR_ALIGNMENT: return "device pointer is not 16-byte aligned";
    case GFD_ERR_DATA_MISMATCH: return "probe data mismatch: the bytes read back are not the pattern that was written";
    default:
        snprintf(buf, sizeof buf, "unknown status %d", status);
        return buf;
    }
}

// ---- vGPU -------------------------------------------------------------------------------------------
namespace {
// Go's string(uint8): the UTF-8 encoding of code point U+00XX
size_t put_go_byte(char *dst, uint8_t c) {
    if (c < 0x80) { dst[0] = (char)c; return 1; }
    dst[0] = (char)(0xC0 | (c >> 6));
    dst[1] = (char)(0x80 | (c & 0x3F));
    return 2;
}
// 10 config bytes -> Go string -> strings.Trim(s, "\x00")
int go_field(const uint8_t *p, char *out, size_t out_len) {
    char tmp[24];
    size_t n = 0;
    for (int i = 0; i < 10; i++) n += put_go_byte(tmp + n, p[i]);
    size_t b = 0, e = n;
    while (b < e && tmp[b] == 0) b++;
    while (e > b && tmp[e - 1] == 0) e--;
    size_t L = e - b;
    if (L >= out_len) L = out_len - 1;
    memcpy(out, tmp + b, L);
    out[L] = 0;
    return (int)L;
}
}  // namespace

GFD_API int gfd_vgpu_parse(const uint8_t *cfg, size_t len, size_t cap, gfd_vgpu_info *out) {
    if (!cfg || !out) return GFD_ERR_BAD_ARGUMENT;
    memset(out, 0, sizeof *out);
    if (cap < len) cap = len;
    // GetVendorSpecificCapability (pciutil.go:115-149)
    if (len < 256) return GFD_ERR_PCI_CONFIG_SHORT;
    if ((cfg[0x06] & 0x10) == 0) return 0;
    uint8_t visited[256] = {0};
    size_t pos = cfg[0x34];
    size_t cap_pos = 0, cap_len = 0;
    bool found = false;
    while (pos != 0) {
        if (pos + 2 >= len) return GFD_ERR_PCI_OUT_OF_RANGE;  // GetByte(pos+2) would panic
        uint8_t id = cfg[pos], next = cfg[pos + 1], length = cfg[pos + 2];
        if (visited[pos]) break;   // chain looped
        if (id == 0xff) break;     // chain broken
        if (id == 0x09) {
            if (pos + length > cap) return GFD_ERR_PCI_OUT_OF_RANGE;  // slice bounds out of range
            cap_pos = pos;
            cap_len = length;
            found = true;
            break;
        }
        visited[pos]++;
        pos = next;
    }
    if (!found) return 0;
    out->has_vendor_capability = 1;
    out->vendor_capability_len = (int)cap_len;
    // capability = cfg[pos : pos+length], zero-extended past len (Go slices may run to cap; fresh allocation is zero)
    uint8_t c[256];
    for (size_t i = 0; i < cap_len; i++) c[i] = (cap_pos + i < len) ? cfg[cap_pos + i] : 0;
    // IsVGPUDevice (vgpu.go:92-105)
    if (cap_len < 5 || c[3] != 0x56 || c[4] != 0x46) return 0;
    out->is_vgpu = 1;
    // GetInfo (vgpu.go:108-153)
    if (cap_len <= 5) return GFD_ERR_PCI_OUT_OF_RANGE;  // GetByte(capability, 5) panics
    size_t p = 5;
    uint8_t record = c[5];
    while (record != 0 && p < cap_len) {
        if (p + 1 >= cap_len) return GFD_ERR_PCI_OUT_OF_RANGE;
        uint8_t rl = c[p + 1];
        if (rl == 0) return GFD_ERR_VGPU_RECORD_LOOP;  // the reference never terminates here
        p += rl;
        if (p >= cap_len) return GFD_ERR_PCI_OUT_OF_RANGE;
        record = c[p];
    }
    if (record == 0 && p + 2 + 10 + 10 <= cap_len) {
        out->host_driver_version_len = go_field(c + p + 2, out->host_driver_version, sizeof out->host_driver_version);
        out->host_driver_branch_len = go_field(c + p + 12, out->host_driver_branch, sizeof out->host_driver_branch);
        return 0;
    }
    return GFD_ERR_VGPU_NO_VERSION_RECORD;
}
