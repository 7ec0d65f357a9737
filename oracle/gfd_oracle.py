"""CPU oracle for the GFD labeling pass -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this.  The product (libgfd_b200.so + the gpu-feature-discovery
binary) never routes through it.

It is a plain-Python restatement of the reference's label semantics
(Telemaco019/gpu-feature-discovery == NVIDIA GFD v0.7.0 fork; Go, cannot be
built here: no Go toolchain).  Every function cites the reference file:line it
follows; paths are relative to the reference root, vendored code abbreviated

    V/nvlib = vendor/gitlab.com/nvidia/cloud-native/go-nvlib/pkg
    V/go-nvml = vendor/github.com/NVIDIA/go-nvml/pkg
    V/spec = vendor/github.com/NVIDIA/k8s-device-plugin/api/config/v1

PARITY PINNING
  * pinned: the label-builder layer (internal/lm, internal/mig, the
    resource.Device/Manager contract, internal/vgpu) is checked against the
    reference's own golden vectors -- 32 exact label-map table cases, the
    TestFailOnNVMLInitError matrix, the two vGPU PCI config-space known-answer
    blobs and the run()-level regex goldens -- extracted mechanically by
    tests/golden/extract_reference_vectors.py (see tests/test_oracle_golden.py).
  * PARITY UNPINNED at the NVML boundary: no reference test exercises
    internal/resource/nvml-*.go or go-nvlib pkg/nvlib/device/* (profile
    resolution, getMigMemorySizeGB, the DGX-Display skip, '+' -> '.').  For
    those this file is a line-by-line restatement only.

Layering mirrors the reference so that the tests read like its tests:
    Manager / Device        internal/resource/types.go:22-42
    MockManager/MockDevice  internal/resource/testing/resource-testing.go:31-134
    NvmlManager/...         internal/resource/nvml-*.go over a "raw NVML" object
    label builders          internal/lm/*.go
    run_pass                cmd/gpu-feature-discovery/main.go:148-187
"""
import itertools
import math
import os

FULL_GPU_RESOURCE_NAME = "nvidia.com/gpu"  # internal/lm/resource.go:27

# NVML return codes (nvml.h / V/go-nvml/nvml/const.go)
NVML_SUCCESS = 0
NVML_ERROR_UNINITIALIZED = 1
NVML_ERROR_INVALID_ARGUMENT = 2
NVML_ERROR_NOT_SUPPORTED = 3
NVML_ERROR_NO_PERMISSION = 4
NVML_ERROR_NOT_FOUND = 6
NVML_ERROR_LIBRARY_NOT_FOUND = 12
NVML_ERROR_FUNCTION_NOT_FOUND = 13
NVML_ERROR_UNKNOWN = 999

NVML_DEVICE_MIG_ENABLE = 1               # V/go-nvml/nvml/const.go:535
GPU_INSTANCE_PROFILE_COUNT = 8           # V/go-nvml/nvml/const.go:555 (reference's NVML 11.x header)
COMPUTE_INSTANCE_PROFILE_COUNT = 7       # const.go:571
COMPUTE_INSTANCE_ENGINE_PROFILE_COUNT = 1  # const.go:575

# GI/CI profile id -> slice count (V/nvlib/nvlib/device/mig_profile.go:56-96, ids from const.go:538-571)
_GI_SLICES = {0: 1, 7: 1, 1: 2, 2: 3, 3: 4, 6: 6, 4: 7, 5: 8}
_CI_SLICES = {0: 1, 1: 2, 2: 3, 3: 4, 6: 6, 4: 7, 5: 8}
GPU_INSTANCE_PROFILE_1_SLICE_REV1 = 7

SKIPPED_DEVICE_NAMES = ("NVIDIA DGX Display",)  # V/nvlib/nvlib/device/api.go:53-55


class GfdError(Exception):
    """A Go `error` returned up the call chain (the pass fails, exit 1)."""


class GoPanic(GfdError):
    """A Go runtime panic (index/slice out of range); the reference process dies -- still 'no file'."""


class GoHang(GfdError):
    """The reference would loop forever here (vgpu.go:121-126 with a zero-length record).  The new labeler
    cannot hang a DaemonSet, so it fails the pass instead; the oracle flags the case so tests can tell."""


# ----------------------------------------------------------------------------
# Go standard-library behaviours the label text depends on
# ----------------------------------------------------------------------------
def go_round(x):
    """math.Round: half away from zero (V/nvlib/nvlib/device/mig_profile.go:331). Python's round() is half-even."""
    t = math.trunc(x)
    if abs(x - t) >= 0.5:
        return t + math.copysign(1.0, x)
    return float(t)


_GO_SPACE = {0x09, 0x0A, 0x0B, 0x0C, 0x0D, 0x20, 0x85, 0xA0, 0x1680, 0x2028, 0x2029, 0x202F, 0x205F, 0x3000} | set(
    range(0x2000, 0x200B))


def _decode_rune(b, i):
    """utf8.DecodeRune: (rune, width); invalid -> (0xFFFD, 1)."""
    c = b[i]
    if c < 0x80:
        return c, 1
    for n, lo in ((2, 0x80), (3, 0x800), (4, 0x10000)):
        lead_mask = (0xFF << (7 - n)) & 0xFF
        if c & lead_mask == (lead_mask << 1) & 0xFF:
            if i + n > len(b):
                return 0xFFFD, 1
            r = c & (0xFF >> (n + 1))
            for k in range(1, n):
                cc = b[i + k]
                if cc & 0xC0 != 0x80:
                    return 0xFFFD, 1
                r = (r << 6) | (cc & 0x3F)
            if r < lo or r > 0x10FFFF or 0xD800 <= r <= 0xDFFF:
                return 0xFFFD, 1
            return r, n
    return 0xFFFD, 1


def go_trim_space(b):
    """strings.TrimSpace on a Go string holding arbitrary bytes (internal/lm/machine-type.go:50)."""
    b = bytes(b)
    start = 0
    while start < len(b):
        r, w = _decode_rune(b, start)
        if r not in _GO_SPACE:
            break
        start += w
    end = len(b)
    while end > start:
        # utf8.DecodeLastRune: back up over at most 3 continuation bytes
        lim = max(start, end - 4)
        k = end - 1
        while k > lim and b[k] & 0xC0 == 0x80:
            k -= 1
        r, w = _decode_rune(b, k)
        if k + w != end:
            r, w, k = 0xFFFD, 1, end - 1
        if r not in _GO_SPACE:
            break
        end = k
    return b[start:end]


def _s(b):
    return b.decode("utf-8", "surrogateescape") if isinstance(b, (bytes, bytearray)) else b


def go_byte_to_string(c):
    """Go `string(uint8)`: the UTF-8 encoding of the code point, NOT the raw byte (internal/vgpu/vgpu.go:131,136)."""
    return chr(c).encode("utf-8")


def fmt_v(value):
    """fmt.Sprintf("%v") for the value kinds labels carry: ints, uints, strings, bools (internal/lm/resource.go:139)."""
    if isinstance(value, bool):
        return "true" if value else "false"
    return str(value)


# ----------------------------------------------------------------------------
# Config (the subset of V/spec Config the labeling pass reads)
# ----------------------------------------------------------------------------
class ReplicatedResource:
    """V/spec/replicas.go:36-41"""

    def __init__(self, name, replicas, rename=""):
        self.name, self.replicas, self.rename = name, replicas, rename


class Config:
    """V/spec/config.go:33-38 + flags.go:48-72, flattened."""

    def __init__(self, mig_strategy="none", fail_on_init_error=True, oneshot=True, no_timestamp=False,
                 output_file="", machine_type_file="/sys/class/dmi/id/product_name", sleep_interval=60.0,
                 time_slicing=()):
        self.mig_strategy = mig_strategy
        self.fail_on_init_error = fail_on_init_error
        self.oneshot = oneshot
        self.no_timestamp = no_timestamp
        self.output_file = output_file
        self.machine_type_file = machine_type_file
        self.sleep_interval = sleep_interval
        self.time_slicing = list(time_slicing)  # [ReplicatedResource]


# ----------------------------------------------------------------------------
# resource.Manager / resource.Device  (internal/resource/types.go:22-42)
# ----------------------------------------------------------------------------
class Device:
    def is_mig_enabled(self): raise NotImplementedError
    def is_mig_capable(self): raise NotImplementedError
    def get_mig_devices(self): raise NotImplementedError
    def get_attributes(self): raise NotImplementedError
    def get_name(self): raise NotImplementedError
    def get_total_memory_mb(self): raise NotImplementedError
    def get_device_handle_from_mig_device_handle(self): raise NotImplementedError
    def get_cuda_compute_capability(self): raise NotImplementedError


class Manager:
    def init(self): raise NotImplementedError
    def shutdown(self): raise NotImplementedError
    def get_devices(self): raise NotImplementedError
    def get_driver_version(self): raise NotImplementedError
    def get_cuda_driver_version(self): raise NotImplementedError


class NullManager(Manager):
    """internal/resource/null.go:23-57"""

    def init(self): return None
    def shutdown(self): return None
    def get_devices(self): return []
    def get_cuda_driver_version(self): raise GfdError("GetCudaDriverVersion is unsupported")
    def get_driver_version(self): raise GfdError("GetDriverVersion is unsupported")


class FallbackManager(Manager):
    """internal/resource/fallback.go:23-64 -- becomes a Null manager on the first Init error."""

    def __init__(self, wraps):
        self.wraps, self.fallback = wraps, NullManager()
        self.log = []

    def init(self):
        try:
            self.wraps.init()
        except GfdError as e:
            self.log.append("Warning: failed to initialize resource manager: %s" % e)
            self.wraps = self.fallback
        return None

    def shutdown(self): return self.wraps.shutdown()
    def get_devices(self): return self.wraps.get_devices()
    def get_cuda_driver_version(self): return self.wraps.get_cuda_driver_version()
    def get_driver_version(self): return self.wraps.get_driver_version()


class CudaDevice(Device):
    """internal/resource/cuda-device.go:25-98 over internal/cuda/api.go.  `raw` = a raw CUDA-driver object whose methods
    return (CUresult, value): oracle/nvml_model.py:RawCuda."""

    def __init__(self, raw, ordinal):
        self.raw, self.d = raw, ordinal

    def get_attributes(self): raise GfdError("GetAttributes is not supported for CUDA devices")
    def get_device_handle_from_mig_device_handle(self): raise GfdError("GetDeviceHandleFromMigDeviceHandle is unsupported for CUDA devices")
    def get_mig_devices(self): raise GfdError("GetMigDevices is unsupported for CUDA devices")
    def is_mig_capable(self): return False
    def is_mig_enabled(self): return False

    def get_cuda_compute_capability(self):
        r, major = self.raw.device_get_attribute(75, self.d)
        if r:
            raise GfdError("failed to get CUDA compute capability major for device: result=%d" % r)
        r, minor = self.raw.device_get_attribute(76, self.d)
        if r:
            raise GfdError("failed to get CUDA compute capability minor for device: result=%d" % r)
        return major, minor

    def get_total_memory_mb(self):
        r, total = self.raw.device_total_mem(self.d)
        if r:
            raise GfdError("failed to get memory info for device: %d" % r)
        return total // (1024 * 1024)

    def get_name(self):
        r, name = self.raw.device_get_name(self.d)
        if r:
            raise GfdError("failed to get device name: %d" % r)
        return name


class CudaManager(Manager):
    """internal/resource/cuda-lib.go:25-88 (Tegra: no NVML, libcuda only)."""

    def __init__(self, raw):
        self.raw = raw

    def init(self):
        r, _ = self.raw.init()
        if r:
            raise GfdError("%d" % r)

    def shutdown(self): return None

    def get_devices(self):
        r, count = self.raw.device_get_count()
        if r:
            raise GfdError("failed to get number of CUDA devices: %d" % r)
        out = []
        for i in range(count):
            r, d = self.raw.device_get(i)
            if r:
                raise GfdError("failed to get CUDA device %d: %d" % (i, r))
            out.append(CudaDevice(self.raw, d))
        return out

    def get_cuda_driver_version(self):
        r, v = self.raw.driver_get_version()
        if r:
            raise GfdError("failed to get driver version: %d" % r)
        return v // 1000, v % 100 // 10  # cuda-lib.go:59-60: `% 100`, unlike the NVML manager's `% 1000`

    def get_driver_version(self):
        return "unknown.unknown.unknown"  # cuda-lib.go:66-68


def with_config(manager, config):
    """internal/resource/factory.go:33-39"""
    if config.fail_on_init_error:
        return manager
    return FallbackManager(manager)


# ---- the reference's test doubles (internal/resource/testing/resource-testing.go) -------------
class MockDevice(Device):
    """rt.NewDeviceMock / NewFullGPU / NewMigEnabledDevice (resource-testing.go:31-58, 88-104)."""

    def __init__(self, mig_enabled, migs=(), name="MOCKMODEL", total_memory_mb=300):
        self.mig_enabled = mig_enabled
        self.name, self.total_memory_mb = name, total_memory_mb
        self.migs = list(migs)
        for m in self.migs:
            m.parent = self

    def get_name(self): return self.name
    def get_cuda_compute_capability(self): return (0, 0) if self.mig_enabled else (8, 0)
    def get_total_memory_mb(self): return self.total_memory_mb
    def is_mig_enabled(self): return self.mig_enabled
    def is_mig_capable(self): return self.mig_enabled
    def get_mig_devices(self): return list(self.migs)
    # moq panics on unset funcs; the tables never reach these
    def get_attributes(self): raise GoPanic("DeviceMock.GetAttributesFunc: method is nil")
    def get_device_handle_from_mig_device_handle(self): raise GoPanic("DeviceMock.GetDeviceHandleFromMigDeviceHandleFunc: method is nil")


class MockMigDevice(Device):
    """rt.NewMigDevice(gi, ci, gb, attributes...) (resource-testing.go:60-86)."""

    def __init__(self, gi, ci, gb, attributes=None):
        self.attrs = {
            "memory": gb, "multiprocessors": 0, "slices.gi": gi, "slices.ci": ci,
            "engines.copy": 0, "engines.decoder": 0, "engines.encoder": 0, "engines.jpeg": 0, "engines.ofa": 0,
        }
        self.attrs.update(attributes or {})
        self.name = "%dg.%dgb" % (gi, gb)
        self.parent = None

    def get_name(self): return self.name
    def get_attributes(self): return dict(self.attrs)

    def get_device_handle_from_mig_device_handle(self):
        if self.parent is None:
            raise GoPanic("DeviceMock.GetDeviceHandleFromMigDeviceHandleFunc: method is nil")
        return self.parent


class MockManager(Manager):
    """rt.NewManagerMockWithDevices / WithErrorOnInit (resource-testing.go:106-134)."""

    def __init__(self, devices=(), error_on_init=None):
        self.devices, self.error_on_init = list(devices), error_on_init

    def init(self):
        if self.error_on_init:
            raise GfdError(self.error_on_init)

    def shutdown(self): return None
    def get_driver_version(self): return "400.300"
    def get_devices(self): return list(self.devices)
    def get_cuda_driver_version(self): return (8, 0)


# ---- NVML-backed manager ----------------------------------------------------------------------
# `raw` is any object with the raw-NVML method set of oracle/nvml_model.py:RawNvml (fixture driven)
# or oracle/nvml_model.py:PynvmlRaw (the real library on a GPU box).  Every method returns (rc, value).
def _nvml_err(raw, rc):
    """nvml.Return.Error() (V/nvlib/nvml/return.go:26-33)."""
    return raw.error_string(rc)


class NvmlManager(Manager):
    """internal/resource/nvml-lib.go:24-97 over V/nvlib/nvlib/device/device.go."""

    def __init__(self, raw):
        self.raw = raw

    def init(self):
        rc, _ = self.raw.init()
        if rc != NVML_SUCCESS:
            raise GfdError(_nvml_err(self.raw, rc))

    def shutdown(self):
        rc, _ = self.raw.shutdown()
        if rc != NVML_SUCCESS:
            raise GfdError(_nvml_err(self.raw, rc))

    def get_cuda_driver_version(self):
        # nvml-lib.go:42-51: major = v/1000, minor = v%1000/10
        rc, v = self.raw.system_get_cuda_driver_version()
        if rc != NVML_SUCCESS:
            raise GfdError(_nvml_err(self.raw, rc))
        return v // 1000, v % 1000 // 10

    def get_driver_version(self):
        rc, v = self.raw.system_get_driver_version()
        if rc != NVML_SUCCESS:
            raise GfdError(_nvml_err(self.raw, rc))
        return v

    def get_devices(self):
        # nvml-lib.go:54-70 -> devicelib.GetDevices -> VisitDevices (device.go:206-236, 288-298)
        raw = self.raw
        rc, count = raw.device_get_count()
        if rc != NVML_SUCCESS:
            raise GfdError("error getting device count: %s" % _nvml_err(raw, rc))
        devs = []
        for i in range(count):
            rc, h = raw.device_get_handle_by_index(i)
            if rc != NVML_SUCCESS:
                raise GfdError("error getting device handle for index '%d': %s" % (i, _nvml_err(raw, rc)))
            # isSkipped (device.go:192-203)
            rc, name = raw.device_get_name(h)
            if rc != NVML_SUCCESS:
                raise GfdError("error checking whether device is skipped: error getting device name: %s" % _nvml_err(raw, rc))
            if name in SKIPPED_DEVICE_NAMES:
                continue
            devs.append(NvmlDevice(raw, h))
        return devs


class NvmlDevice(Device):
    """internal/resource/nvml-device.go:26-88 over go-nvlib device (device.go:55-118)."""

    def __init__(self, raw, handle):
        self.raw, self.h = raw, handle

    def _mig_mode(self):
        # device.go:55-88: dlopen+dlsym probe; missing symbol or NOT_SUPPORTED => false
        if not self.raw.lookup_symbol("nvmlDeviceGetMigMode"):
            return None
        rc, mode = self.raw.device_get_mig_mode(self.h)
        if rc == NVML_ERROR_NOT_SUPPORTED:
            return None
        if rc != NVML_SUCCESS:
            raise GfdError("error getting MIG mode: %s" % _nvml_err(self.raw, rc))
        return mode  # (current, pending)

    def is_mig_capable(self):
        return self._mig_mode() is not None

    def is_mig_enabled(self):
        mode = self._mig_mode()
        return mode is not None and mode[0] == NVML_DEVICE_MIG_ENABLE

    def get_mig_devices(self):
        # nvml-device.go:34-50 -> device.go:91-118 (VisitMigDevices) -> mig_device.go:40-49 (NewMigDevice)
        raw = self.raw
        rc, count = raw.device_get_max_mig_device_count(self.h)
        if rc != NVML_SUCCESS:
            raise GfdError("error getting max MIG device count: %s" % _nvml_err(raw, rc))
        migs = []
        for i in range(count):
            rc, mh = raw.device_get_mig_device_handle_by_index(self.h, i)
            if rc in (NVML_ERROR_NOT_FOUND, NVML_ERROR_INVALID_ARGUMENT):
                continue
            if rc != NVML_SUCCESS:
                raise GfdError("error getting MIG device handle at index '%d': %s" % (i, _nvml_err(raw, rc)))
            rc, is_mig = raw.device_is_mig_device_handle(mh)
            if rc != NVML_SUCCESS:
                raise GfdError("error creating new MIG device wrapper: error checking if device is a MIG device: %s" % _nvml_err(raw, rc))
            if not is_mig:
                raise GfdError("error creating new MIG device wrapper: not a MIG device")
            migs.append(NvmlMigDevice(raw, mh))
        return migs

    def get_cuda_compute_capability(self):
        rc, cc = self.raw.device_get_cuda_compute_capability(self.h)
        if rc != NVML_SUCCESS:
            raise GfdError(_nvml_err(self.raw, rc))
        return cc

    def get_attributes(self):
        raise GfdError("GetAttributes is not supported for non-MIG devices")

    def get_device_handle_from_mig_device_handle(self):
        raise GfdError("GetDeviceHandleFromMigDeviceHandle is not supported for non-MIG devices")

    def get_name(self):
        rc, name = self.raw.device_get_name(self.h)
        if rc != NVML_SUCCESS:
            raise GfdError(_nvml_err(self.raw, rc))
        return name

    def get_total_memory_mb(self):
        # nvml-device.go:82-88: nvmlDeviceGetMemoryInfo (v1 struct) .Total / (1024*1024)
        rc, total = self.raw.device_get_memory_info(self.h)
        if rc != NVML_SUCCESS:
            raise GfdError(_nvml_err(self.raw, rc))
        return total // (1024 * 1024)


def get_mig_memory_size_gb(total_device_memory, mig_memory_size_mb):
    """getMigMemorySizeGB (V/nvlib/nvlib/device/mig_profile.go:324-332), float64 arithmetic as in Go."""
    frac_denominator = 8
    one_mb = 1024 * 1024
    one_gb = 1024 * 1024 * 1024
    fractional = (float(mig_memory_size_mb) * one_mb) / float(total_device_memory)
    fractional = math.ceil(fractional * frac_denominator) / frac_denominator
    total_mem_gb = float((total_device_memory + one_gb - 1) // one_gb)
    return int(go_round(fractional * total_mem_gb))


def new_mig_profile_string(gi_profile_id, ci_profile_id, mig_memory_size_mb, device_memory_bytes):
    """NewMigProfile + MigProfileInfo.String (mig_profile.go:55-115, 200-209)."""
    if gi_profile_id not in _GI_SLICES:
        raise GfdError("invalid GPU Instance Profile ID: %d" % gi_profile_id)
    if ci_profile_id not in _CI_SLICES:
        raise GfdError("invalid Compute Instance Profile ID: %d" % ci_profile_id)
    g, c = _GI_SLICES[gi_profile_id], _CI_SLICES[ci_profile_id]
    gb = get_mig_memory_size_gb(device_memory_bytes, mig_memory_size_mb)
    suffix = "+me" if gi_profile_id == GPU_INSTANCE_PROFILE_1_SLICE_REV1 else ""
    if c == g:
        return "%dg.%dgb%s" % (g, gb, suffix)
    return "%dc.%dg.%dgb%s" % (c, g, gb, suffix)


class NvmlMigDevice(Device):
    """internal/resource/nvml-mig-device.go:27-134 over go-nvlib migdevice (mig_device.go:52-148)."""

    def __init__(self, raw, handle):
        self.raw, self.h = raw, handle
        self._profile = None

    def get_attributes(self):
        rc, a = self.raw.device_get_attributes(self.h)
        if rc != NVML_SUCCESS:
            raise GfdError(_nvml_err(self.raw, rc))
        # nvml-mig-device.go:40-50
        return {
            "memory": a["memory_size_mb"], "multiprocessors": a["multiprocessor_count"],
            "slices.gi": a["gpu_instance_slice_count"], "slices.ci": a["compute_instance_slice_count"],
            "engines.copy": a["shared_copy_engine_count"], "engines.decoder": a["shared_decoder_count"],
            "engines.encoder": a["shared_encoder_count"], "engines.jpeg": a["shared_jpeg_count"],
            "engines.ofa": a["shared_ofa_count"],
        }

    def get_device_handle_from_mig_device_handle(self):
        rc, p = self.raw.device_get_device_handle_from_mig_device_handle(self.h)
        if rc != NVML_SUCCESS:
            raise GfdError(_nvml_err(self.raw, rc))
        return NvmlDevice(self.raw, p)

    def is_mig_capable(self): raise GfdError("IsMigCapable is not supported for MIG devices")
    def is_mig_enabled(self): raise GfdError("IsMigEnabled is not supported for MIG devices")
    def get_mig_devices(self): raise GfdError("GetMigDevices is not implemented for MIG devices")
    def get_cuda_compute_capability(self): raise GfdError("GetCudaComputeCapability is not supported for MIG devices")

    def get_profile(self):
        """migdevice.GetProfile (mig_device.go:52-148); cached per wrapper object (:53-55, :141)."""
        if self._profile is not None:
            return self._profile
        raw = self.raw

        def must(rc_val, what):
            rc, val = rc_val
            if rc != NVML_SUCCESS:
                raise GfdError("%s: %s" % (what, _nvml_err(raw, rc)))
            return val

        parent = must(raw.device_get_device_handle_from_mig_device_handle(self.h), "error getting parent device handle")
        parent_total = must(raw.device_get_memory_info(parent), "error getting parent memory info")
        attrs = must(raw.device_get_attributes(self.h), "error getting MIG device attributes")
        gi_id = must(raw.device_get_gpu_instance_id(self.h), "error getting MIG device GPU Instance ID")
        ci_id = must(raw.device_get_compute_instance_id(self.h), "error getting MIG device Compute Instance ID")
        gi = must(raw.device_get_gpu_instance_by_id(parent, gi_id), "error getting GPU Instance")
        ci = must(raw.gpu_instance_get_compute_instance_by_id(gi, ci_id), "error getting Compute Instance")
        gi_info = must(raw.gpu_instance_get_info(gi), "error getting GPU Instance info")
        ci_info = must(raw.compute_instance_get_info(ci), "error getting Compute Instance info")

        for i in range(GPU_INSTANCE_PROFILE_COUNT):
            rc, gi_prof = raw.device_get_gpu_instance_profile_info(parent, i)
            if rc in (NVML_ERROR_NOT_SUPPORTED, NVML_ERROR_INVALID_ARGUMENT):
                continue
            if rc != NVML_SUCCESS:
                raise GfdError("error getting GPU Instance profile info: %s" % _nvml_err(raw, rc))
            if gi_prof["id"] != gi_info["profile_id"]:
                continue
            for j in range(COMPUTE_INSTANCE_PROFILE_COUNT):
                for k in range(COMPUTE_INSTANCE_ENGINE_PROFILE_COUNT):
                    rc, ci_prof = raw.gpu_instance_get_compute_instance_profile_info(gi, j, k)
                    if rc in (NVML_ERROR_NOT_SUPPORTED, NVML_ERROR_INVALID_ARGUMENT):
                        continue
                    if rc != NVML_SUCCESS:
                        raise GfdError("error getting Compute Instance profile info: %s" % _nvml_err(raw, rc))
                    if ci_prof["id"] != ci_info["profile_id"]:
                        continue
                    # NB: the *loop indices* i, j go into NewMigProfile, not the ids (mig_device.go:136)
                    try:
                        self._profile = new_mig_profile_string(i, j, attrs["memory_size_mb"], parent_total)
                    except GfdError as e:
                        raise GfdError("error creating MIG profile: %s" % e)
                    return self._profile
        raise GfdError("no matching profile IDs found")

    def get_name(self):
        # nvml-mig-device.go:96-104: profile string with '+' -> '.'
        try:
            p = self.get_profile()
        except GfdError as e:
            raise GfdError("failed to get MIG profile: %s" % e)
        return p.replace("+", ".")

    def get_total_memory_mb(self):
        return self.get_attributes()["memory"]


# ----------------------------------------------------------------------------
# Go map iteration order.  The reference ranges over Go maps in three places
# where order changes the RESULT (internal/lm/mig-strategy.go:157,169): which
# device model's labels win on a heterogeneous node.  `order` is a callable
# keys -> ordered keys; the default (sorted) is what the new labeler emits.
# possible_label_sets() enumerates every order the reference could take.
# ----------------------------------------------------------------------------
def _sorted_order(keys):
    return sorted(keys)


# ----------------------------------------------------------------------------
# internal/mig/mig.go:24-124
# ----------------------------------------------------------------------------
class DeviceInfo:
    def __init__(self, manager):
        self.manager, self.devices_map = manager, None

    def get_devices_map(self):
        if self.devices_map is not None:
            return self.devices_map
        m = {}
        for d in self.manager.get_devices():
            m.setdefault(d.is_mig_enabled(), []).append(d)
        self.devices_map = m
        return m

    def get_devices_with_mig_enabled(self): return self.get_devices_map().get(True, [])
    def get_devices_with_mig_disabled(self): return self.get_devices_map().get(False, [])

    def any_mig_enabled_device_is_empty(self):
        enabled = self.get_devices_map().get(True, [])
        if len(enabled) == 0:
            return True
        for d in enabled:
            if len(d.get_mig_devices()) == 0:
                return True
        return False

    def get_all_mig_devices(self):
        migs = []
        for d in self.get_devices_map().get(True, []):
            migs.extend(d.get_mig_devices())
        return migs


# ----------------------------------------------------------------------------
# internal/lm/resource.go
# ----------------------------------------------------------------------------
class ResourceLabeler:
    """resourceLabeler (resource.go:113-226). config None == sharing disabled (:31-33, :194-196)."""

    def __init__(self, resource_name, config):
        self.resource_name, self.config = resource_name, config

    def key(self, suffix): return self.resource_name + "." + suffix
    def single(self, suffix, value): return {self.key(suffix): fmt_v(value)}
    def labels(self, suffix_values): return {self.key(k): fmt_v(v) for k, v in suffix_values.items()}

    def replication_info(self):
        if self.config is None:
            return None
        for r in self.config.time_slicing:
            if r.name == self.resource_name:
                return r
        return None

    def is_shared(self):
        r = self.replication_info()
        return r is not None and r.replicas > 1

    def is_renamed(self):
        r = self.replication_info()
        return r is not None and r.rename != ""

    def product_label(self, *parts):
        stripped = [p.replace(" ", "-") for p in parts if p != ""]
        if not stripped:
            return {}
        if self.is_shared() and not self.is_renamed():
            stripped.append("SHARED")
        return self.single("product", "-".join(stripped))

    def replicas_label(self):
        replicas = 1
        if self.config is None:
            replicas = 0
        else:
            r = self.replication_info()
            if r is not None and r.replicas > 1:
                replicas = r.replicas
        return self.single("replicas", replicas)

    def base_labeler(self, count, *parts):
        out = {}
        out.update(self.product_label(*parts))
        out.update(self.single("count", count))
        out.update(self.replicas_label())
        return out


def get_arch_family(major, minor):
    """getArchFamily (resource.go:261-284) -- stops at hopper; compute 10.x => 'undefined'."""
    table = {1: "tesla", 2: "fermi", 3: "kepler", 5: "maxwell", 6: "pascal", 8: "ampere", 9: "hopper"}
    if major == 7:
        return "volta" if minor < 5 else "turing"
    return table.get(major, "undefined")


def new_gpu_resource_labeler(config, device, count):
    """NewGPUResourceLabeler (resource.go:36-73)."""
    if count == 0:
        return {}
    try:
        model = device.get_name()
    except GfdError as e:
        raise GfdError("failed to get device model: %s" % e)
    try:
        total_mb = device.get_total_memory_mb()
    except GfdError as e:
        raise GfdError("failed to get memory info for device: %s" % e)
    rl = ResourceLabeler(FULL_GPU_RESOURCE_NAME, config)
    try:
        major, minor = device.get_cuda_compute_capability()
    except GfdError as e:
        raise GfdError("failed to create architecture labels: failed to determine CUDA compute capability: %s" % e)
    out = rl.base_labeler(count, model)
    if total_mb != 0:
        out.update(rl.single("memory", total_mb))
    if major != 0:  # newArchitectureLabels (resource.go:239-258)
        out.update(rl.labels({"family": get_arch_family(major, minor), "compute.major": major, "compute.minor": minor}))
    return out


def new_mig_resource_labeler(resource_name, config, device, count):
    """NewMIGResourceLabeler (resource.go:76-111)."""
    if count == 0:
        return {}
    try:
        parent = device.get_device_handle_from_mig_device_handle()
    except GfdError as e:
        raise GfdError("failed to get parent of MIG device: %s" % e)
    model = parent.get_name()
    mig_profile = device.get_name()
    rl = ResourceLabeler(resource_name, config)
    try:
        attrs = device.get_attributes()
    except GfdError as e:
        raise GfdError("faled to get MIG attribute labels: unable to get attributes of MIG device: %s" % e)
    out = rl.base_labeler(count, model, "MIG", mig_profile)
    out.update(rl.labels(attrs))
    return out


# ----------------------------------------------------------------------------
# internal/lm/mig-strategy.go
# ----------------------------------------------------------------------------
def new_gpu_labelers(manager, config, order=_sorted_order):
    """newGPULabelers (mig-strategy.go:113-179)."""
    by_mig = DeviceInfo(manager).get_devices_map()
    if len(by_mig) == 0:
        raise GfdError("no GPU devices detected")
    counts, mig_enabled, full = {}, {}, {}
    for d in by_mig.get(True, []):
        n = d.get_name()
        mig_enabled[n] = d
        counts[n] = counts.get(n, 0) + 1
    for d in by_mig.get(False, []):
        n = d.get_name()
        full[n] = d
        counts[n] = counts.get(n, 0) + 1
    out = {}
    for n in order(list(mig_enabled)):
        out.update(new_gpu_resource_labeler(None, mig_enabled[n], counts[n]))  # WithoutSharing: config=nil
    for n in order(list(full)):
        out.update(new_gpu_resource_labeler(config, full[n], counts[n]))
    return out


def new_invalid_mig_strategy_labeler(device, reason):
    """newInvalidMigStrategyLabeler (mig-strategy.go:243-262). No config => never '-SHARED'."""
    model = device.get_name()
    rl = ResourceLabeler(FULL_GPU_RESOURCE_NAME, None)
    out = rl.product_label(model, "MIG", "INVALID")
    out[rl.key("count")] = "0"
    out[rl.key("replicas")] = "0"
    out[rl.key("memory")] = "0"
    return out


def _new_mig_device_labelers(resources, config, order):
    """newMIGDeviceLabelers (mig-strategy.go:297-309); resources: name -> [resource_name, device, count]."""
    out = {}
    for n in order(list(resources)):
        rn, dev, cnt = resources[n]
        out.update(new_mig_resource_labeler(rn, config, dev, cnt))
    return out


def new_mig_strategy_single_labeler(manager, config, order=_sorted_order):
    """newMigStrategySingleLabeler (mig-strategy.go:181-241)."""
    di = DeviceInfo(manager)
    enabled = di.get_devices_with_mig_enabled()
    if len(enabled) == 0:
        return {}
    if di.any_mig_enabled_device_is_empty():
        return new_invalid_mig_strategy_labeler(enabled[0], "at least one MIG device is enabled but empty")
    if len(di.get_devices_with_mig_disabled()) != 0:
        return new_invalid_mig_strategy_labeler(enabled[0], "devices with MIG enabled and disable detected")
    resources = {}
    for mig in di.get_all_mig_devices():
        try:
            n = mig.get_name()
        except GfdError as e:
            raise GfdError("unable to get MIG device name: %s" % e)
        if n not in resources:
            resources[n] = [FULL_GPU_RESOURCE_NAME, mig, 0]
        resources[n][2] += 1
    if len(resources) != 1:
        return new_invalid_mig_strategy_labeler(enabled[0], "more than one MIG device type present on node")
    return _new_mig_device_labelers(resources, config, order)


def new_mig_strategy_mixed_labeler(manager, config, order=_sorted_order):
    """newMigStrategyMixedLabeler (mig-strategy.go:264-295)."""
    di = DeviceInfo(manager)
    resources = {}
    for mig in di.get_all_mig_devices():
        try:
            n = mig.get_name()
        except GfdError as e:
            raise GfdError("unable to get MIG device name: %s" % e)
        if n not in resources:
            resources[n] = ["nvidia.com/mig-" + n, mig, 0]
        resources[n][2] += 1
    return _new_mig_device_labelers(resources, config, order)


def new_mig_labeler(manager, config, order=_sorted_order):
    """newMigLabeler (mig-strategy.go:84-110) + migStrategyLabeler (strategy.go:20-28)."""
    s = config.mig_strategy
    if s == "none":
        labels = {}
    elif s == "single":
        labels = new_mig_strategy_single_labeler(manager, config, order)
    elif s == "mixed":
        labels = new_mig_strategy_mixed_labeler(manager, config, order)
    else:
        raise GfdError("unknown strategy: %s" % s)
    out = {} if s == "none" else {"nvidia.com/mig.strategy": s}
    out.update(labels)
    return out


def new_resource_labeler(manager, config, order=_sorted_order):
    """NewResourceLabeler (mig-strategy.go:45-77)."""
    if len(manager.get_devices()) == 0:
        return {}
    out = new_gpu_labelers(manager, config, order)
    if config.mig_strategy == "none":
        return out
    out.update(new_mig_labeler(manager, config, order))
    return out


# ----------------------------------------------------------------------------
# internal/lm/nvml.go, machine-type.go, timestamp.go, vgpu.go
# ----------------------------------------------------------------------------
def get_machine_type(path, read_file=None):
    """newMachineTypeLabeler/getMachineType (machine-type.go:30-51)."""
    if path == "":
        mt = "unknown"
    else:
        try:
            if read_file is not None:
                data = read_file(path)
            else:
                with open(path, "rb") as f:
                    data = f.read()
            mt = _s(go_trim_space(data))
        except OSError:
            mt = "unknown"
    return {"nvidia.com/gpu.machine": mt.replace(" ", "-")}


def new_version_labeler(manager):
    """newVersionLabeler (nvml.go:75-106)."""
    try:
        dv = manager.get_driver_version()
    except GfdError as e:
        raise GfdError("error getting driver version: %s" % e)
    parts = dv.split(".")
    if len(parts) > 3 or len(parts) < 2:
        raise GfdError('error getting driver version: Version "%s" does not match format "X.Y[.Z]"' % dv)
    rev = parts[2] if len(parts) > 2 else ""
    try:
        cmaj, cmin = manager.get_cuda_driver_version()
    except GfdError as e:
        raise GfdError("error getting cuda driver version: %s" % e)
    return {
        "nvidia.com/cuda.driver.major": parts[0], "nvidia.com/cuda.driver.minor": parts[1],
        "nvidia.com/cuda.driver.rev": rev,
        "nvidia.com/cuda.runtime.major": "%d" % cmaj, "nvidia.com/cuda.runtime.minor": "%d" % cmin,
    }


def new_mig_capability_labeler(manager):
    """newMigCapabilityLabeler (nvml.go:110-137)."""
    devices = manager.get_devices()
    if len(devices) == 0:
        return {}
    capable = False
    for d in devices:
        try:
            capable = d.is_mig_capable()
        except GfdError as e:
            raise GfdError("error getting mig capability: %s" % e)
        if capable:
            break
    return {"nvidia.com/mig.capable": "true" if capable else "false"}


def new_nvml_labeler(manager, config, order=_sorted_order, read_file=None):
    """NewNVMLLabeler (nvml.go:29-72): everything between Init and the deferred Shutdown."""
    try:
        manager.init()
    except GfdError as e:
        raise GfdError("failed to initialize NVML: %s" % e)
    try:
        try:
            devices = manager.get_devices()
        except GfdError as e:
            raise GfdError("error getting devices: %s" % e)
        if len(devices) == 0:
            return {}
        out = {}
        out.update(get_machine_type(config.machine_type_file, read_file))
        out.update(new_version_labeler(manager))
        out.update(new_mig_capability_labeler(manager))
        out.update(new_resource_labeler(manager, config, order))
        return out
    finally:
        try:
            manager.shutdown()  # `defer manager.Shutdown()`: the return value is dropped
        except GfdError:
            pass


# ---- vGPU (internal/vgpu/pciutil.go, vgpu.go; internal/lm/vgpu.go) ----------------------------
class PCIDevice:
    def __init__(self, address, config, vendor="0x10de", class_="", cap=None):
        self.address, self.config, self.vendor, self.class_ = address, bytes(config), vendor, class_
        # Go slices have len AND cap: config[pos:pos+length] may legally run past len up to cap.
        # os.ReadFile allocates max(512, statsize+1) so sysfs reads have slack; literals have cap == len.
        self.cap = len(self.config) if cap is None else cap


def _get_byte(buf, pos):
    if pos < 0 or pos >= len(buf):
        raise GoPanic("index out of range [%d] with length %d" % (pos, len(buf)))
    return buf[pos]


def nvidia_pci_devices(root="/sys/bus/pci/devices"):
    """NvidiaPCILib.Devices (pciutil.go:70-112). ReadDir order is by filename."""
    try:
        names = sorted(os.listdir(root))
    except OSError as e:
        raise GfdError("unable to read PCI bus devices: %s" % e)
    out = []
    for name in names:
        p = os.path.join(root, name)
        try:
            with open(os.path.join(p, "vendor"), "rb") as f:
                vendor = f.read()
        except OSError as e:
            raise GfdError("unable to read PCI device vendor id for %s: %s" % (name, e))
        if _s(go_trim_space(vendor)) != "0x10de":
            continue
        try:
            with open(os.path.join(p, "class"), "rb") as f:
                class_ = f.read()
        except OSError as e:
            raise GfdError("unable to read PCI device class for %s: %s" % (name, e))
        try:
            with open(os.path.join(p, "config"), "rb") as f:
                config = f.read()
        except OSError as e:
            raise GfdError("unable to read PCI configuration space for %s: %s" % (name, e))
        if len(class_) < 4:
            raise GoPanic("slice bounds out of range [:4] with length %d" % len(class_))  # string(class)[0:4]
        size = os.stat(os.path.join(p, "config")).st_size
        out.append(PCIDevice(name, config, "0x10de", _s(class_[0:4]), cap=max(512, size + 1, len(config))))
    return out


def get_vendor_specific_capability(dev):
    """PCIDevice.GetVendorSpecificCapability (pciutil.go:115-149). Returns bytes or None."""
    cfg = dev.config
    if len(cfg) < 256:
        raise GfdError("entire PCI configuration is not read for device %s. Please run GFD with privileged mode "
                       "to read complete PCI configuration data" % dev.address)
    if cfg[0x06] & 0x10 == 0:
        return None
    visited = [0] * 256
    pos = _get_byte(cfg, 0x34)
    while pos != 0:
        cap_id = _get_byte(cfg, pos + 0)
        nxt = _get_byte(cfg, pos + 1)
        length = _get_byte(cfg, pos + 2)
        if visited[pos] != 0:
            break
        if cap_id == 0xFF:
            break
        if cap_id == 0x09:
            if pos + length > dev.cap:
                raise GoPanic("slice bounds out of range [:%d] with capacity %d" % (pos + length, dev.cap))
            padded = cfg + b"\x00" * max(0, pos + length - len(cfg))
            return padded[pos:pos + length]
        visited[pos] += 1
        pos = nxt
    return None


def is_vgpu_device(capability):
    """Lib.IsVGPUDevice (vgpu.go:92-105): signature 'VF' at bytes 3..4."""
    return len(capability) >= 5 and capability[3] == 0x56 and capability[4] == 0x46


def vgpu_get_info(capability, address=""):
    """Device.GetInfo (vgpu.go:108-153) -> (host_driver_version, host_driver_branch) as bytes."""
    cap = capability
    if len(cap) == 0:
        raise GfdError("vendor capability record is not populated for device %s" % address)
    pos = 5
    record = _get_byte(cap, 5)
    while record != 0 and pos < len(cap):
        record_length = _get_byte(cap, pos + 1)
        if record_length == 0:
            raise GoHang("zero-length vGPU capability record at offset %d for device %s" % (pos, address))
        pos = pos + record_length
        record = _get_byte(cap, pos)
    if record == 0 and pos + 2 + 10 + 10 <= len(cap):
        i = pos + 2
        ver = b"".join(go_byte_to_string(_get_byte(cap, k)) for k in range(i, i + 10))
        br = b"".join(go_byte_to_string(_get_byte(cap, k)) for k in range(i + 10, i + 20))
        return ver.strip(b"\x00"), br.strip(b"\x00")
    raise GfdError("cannot find driver version record in vendor specific capability for device %s" % address)


def vgpu_labels(pci_devices):
    """Lib.Devices (vgpu.go:66-90) + vgpuLabeler.Labels (internal/lm/vgpu.go:37-55)."""
    vgpus = []
    for dev in pci_devices:
        try:
            cap = get_vendor_specific_capability(dev)
        except GoPanic:
            raise
        except GfdError as e:
            raise GfdError("unable to get vGPU devices: unable to read vendor specific capability for %s: %s" % (dev.address, e))
        if cap is None:
            continue
        if is_vgpu_device(cap):
            vgpus.append((dev, cap))
    labels = {}
    if vgpus:
        labels["nvidia.com/vgpu.present"] = "true"
    for dev, cap in vgpus:
        try:
            ver, br = vgpu_get_info(cap, dev.address)
        except GoPanic:
            raise
        except GfdError as e:
            raise GfdError("error getting vGPU device info: %s" % e)
        labels["nvidia.com/vgpu.host-driver-version"] = _s(ver)
        labels["nvidia.com/vgpu.host-driver-branch"] = _s(br)
    return labels


# ----------------------------------------------------------------------------
# cmd/gpu-feature-discovery/main.go:148-187 -- one labeling pass
# ----------------------------------------------------------------------------
def run_pass(manager, pci_devices, config, timestamp=None, order=_sorted_order, read_file=None):
    """One iteration of run(): returns (labels dict, warnings list).  Raises GfdError when the reference
    would return an error (exit 1, no file written).  `pci_devices` is a list of PCIDevice or a callable
    returning one (sysfs is read lazily, AFTER the NVML bracket: internal/lm/vgpu.go:37)."""
    labels = {}
    if not config.no_timestamp:
        labels["nvidia.com/gfd.timestamp"] = "%d" % (0 if timestamp is None else timestamp)
    try:
        nvml_labels = new_nvml_labeler(manager, config, order, read_file)
    except GoPanic:
        raise
    except GfdError as e:
        raise GfdError("error creating NVML labeler: %s" % e)
    devs = pci_devices() if callable(pci_devices) else pci_devices
    try:
        v = vgpu_labels(devs)
    except GoPanic:
        raise
    except GfdError as e:
        raise GfdError("error generating labels: error generating labels: %s" % e)
    labels.update(nvml_labels)
    labels.update(v)
    warnings = []
    if len(labels) <= 1:
        warnings.append("Warning: no labels generated from any source")
    return labels, warnings


def render(labels):
    """Labels.WriteTo (internal/lm/labels.go:55-66) with keys sorted (the reference's order is Go-map random;
    parity is defined after LC_ALL=C sort -- SURVEY.md Fact 3)."""
    return "".join("%s=%s\n" % (k, labels[k]) for k in sorted(labels, key=lambda s: s.encode("utf-8", "surrogateescape")))


def possible_label_sets(manager_factory, pci_devices, config, timestamp=None, read_file=None, limit=720):
    """Every label set the reference could emit given Go's random map order (only differs on heterogeneous nodes).
    manager_factory() must return a fresh manager each call."""
    seen, outs = set(), []
    probe = manager_factory()
    probe.init()
    names = sorted({d.get_name() for d in probe.get_devices()})
    probe.shutdown()
    for n, perm in enumerate(itertools.permutations(names)):
        if n >= limit:
            break
        rank = {name: i for i, name in enumerate(perm)}
        order = lambda keys, rank=rank: sorted(keys, key=lambda k: rank.get(k, 0))
        text = render(run_pass(manager_factory(), pci_devices, config, timestamp, order, read_file)[0])
        if text not in seen:
            seen.add(text)
            outs.append(text)
    return outs
