"""ctypes view of libgfd_b200.so -- names, argument meaning and error behaviour mirror the reference's
`resource.Manager` / `resource.Device` interfaces (internal/resource/types.go:22-42) so tests read like
the reference's own.  Every call goes through the C ABI declared in include/gfd_b200.h.

No fallback: a missing library, a missing symbol or a non-zero status raises.
"""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libgfd_b200.so")
BIN_PATH = os.path.join(HERE, "gpu-feature-discovery")
FAKE_NVML_DIR = os.path.join(HERE, "fake_nvml", "lib")
FAKE_NVML_NOMIG_DIR = os.path.join(HERE, "fake_nvml", "lib_nomig")

if not os.path.exists(LIB_PATH):
    raise ImportError("libgfd_b200.so is not built (run `make` or __graft_entry__.build()); there is no fallback path")

lib = ctypes.CDLL(LIB_PATH)

GFD_ERR_CUDA = 100000
GFD_ERR_SHIM = 200000
(GFD_ERR_BAD_ARGUMENT, GFD_ERR_BUFFER_TOO_SMALL, GFD_ERR_NOT_INITIALIZED, GFD_ERR_UNSUPPORTED_BACKEND,
 GFD_ERR_NOT_MIG_DEVICE, GFD_ERR_IS_MIG_DEVICE, GFD_ERR_NOT_A_MIG_HANDLE, GFD_ERR_NO_MATCHING_PROFILE,
 GFD_ERR_INVALID_GI_PROFILE, GFD_ERR_INVALID_CI_PROFILE, GFD_ERR_PCI_CONFIG_SHORT, GFD_ERR_PCI_OUT_OF_RANGE,
 GFD_ERR_VGPU_NO_VERSION_RECORD, GFD_ERR_VGPU_RECORD_LOOP, GFD_ERR_NO_GPU, GFD_ERR_PEER_ACCESS,
 GFD_ERR_SIZE_LIMIT, GFD_ERR_ALIGNMENT, GFD_ERR_DATA_MISMATCH) = range(GFD_ERR_SHIM + 1, GFD_ERR_SHIM + 20)

BACKEND_NVML, BACKEND_CUDA, BACKEND_NULL = 0, 1, 2
MGR_KEEP_ATTACHED = 1
MGR_CURRENT_TABLES = 2
PROBE_READ, PROBE_COPY, PROBE_PUSH, PROBE_WRITE, PROBE_DMA = 0, 1, 2, 3, 4

# every symbol include/gfd_b200.h declares (tests/test_cabi_symbols.py checks the header against this list)
SYMBOLS = [
    "gfd_strerror", "gfd_has_nvml", "gfd_is_tegra_system", "gfd_mgr_new", "gfd_mgr_free", "gfd_init", "gfd_shutdown",
    "gfd_driver_version", "gfd_cuda_driver_version", "gfd_device_count", "gfd_device_get", "gfd_prefetch",
    "gfd_call_stats", "gfd_dev_name", "gfd_dev_total_memory_mb", "gfd_dev_cuda_compute_capability",
    "gfd_dev_is_mig_capable", "gfd_dev_is_mig_enabled", "gfd_dev_mig_count", "gfd_dev_mig_get", "gfd_dev_parent",
    "gfd_dev_attributes", "gfd_dev_uuid", "gfd_mig_memory_size_gb", "gfd_mig_profile_string", "gfd_vgpu_parse",
    "gfd_hbm_fill_pattern", "gfd_hbm_pattern_checksum", "gfd_hbm_scratch_bytes", "gfd_hbm_scratch_init", "gfd_hbm_read_checksum", "gfd_hbm_read_checksum_noalloc", "gfd_hbm_copy", "gfd_checksum_host",
    "gfd_probe_hbm", "gfd_probe_p2p", "gfd_probe_p2p_matrix", "gfd_cuda_device_by_uuid", "gfd_cuda_device_count",
]


class GfdError(RuntimeError):
    def __init__(self, status, where=""):
        self.status = status
        RuntimeError.__init__(self, "%s%s (status %d)" % (where + ": " if where else "", strerror(status), status))


class MigAttributes(ctypes.Structure):
    _fields_ = [("memory_mb", ctypes.c_uint64), ("multiprocessors", ctypes.c_uint32), ("slices_gi", ctypes.c_uint32),
                ("slices_ci", ctypes.c_uint32), ("engines_copy", ctypes.c_uint32), ("engines_decoder", ctypes.c_uint32),
                ("engines_encoder", ctypes.c_uint32), ("engines_jpeg", ctypes.c_uint32), ("engines_ofa", ctypes.c_uint32)]


class VgpuInfo(ctypes.Structure):
    _fields_ = [("has_vendor_capability", ctypes.c_int), ("vendor_capability_len", ctypes.c_int), ("is_vgpu", ctypes.c_int),
                ("host_driver_version", ctypes.c_char * 32), ("host_driver_branch", ctypes.c_char * 32),
                ("host_driver_version_len", ctypes.c_int), ("host_driver_branch_len", ctypes.c_int)]

    def version(self):  # bytes, interior NULs kept (c_char arrays read as C strings would stop at the first one)
        return ctypes.string_at(ctypes.addressof(self) + VgpuInfo.host_driver_version.offset, self.host_driver_version_len)

    def branch(self):
        return ctypes.string_at(ctypes.addressof(self) + VgpuInfo.host_driver_branch.offset, self.host_driver_branch_len)


class Checksum(ctypes.Structure):
    _fields_ = [("xor64", ctypes.c_uint64), ("sum64", ctypes.c_uint64)]

    def tuple(self):
        return (int(self.xor64), int(self.sum64))


class ProbeResult(ctypes.Structure):
    _fields_ = [("gbps", ctypes.c_double), ("ms_per_launch", ctypes.c_double), ("bytes", ctypes.c_uint64),
                ("iters", ctypes.c_int), ("mode", ctypes.c_int), ("sm_count", ctypes.c_int), ("grid", ctypes.c_int),
                ("block", ctypes.c_int), ("checksum", Checksum), ("expected", Checksum), ("verified", ctypes.c_int)]


_vp, _sz, _u64, _int = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_uint64, ctypes.c_int
_P = ctypes.POINTER
_SIG = {
    "gfd_strerror": (ctypes.c_char_p, [_int]),
    "gfd_has_nvml": (_int, [ctypes.c_char_p, _sz]),
    "gfd_is_tegra_system": (_int, [ctypes.c_char_p, ctypes.c_char_p, _sz]),
    "gfd_mgr_new": (_int, [_int, ctypes.c_uint, _P(_vp)]),
    "gfd_mgr_free": (None, [_vp]),
    "gfd_init": (_int, [_vp]),
    "gfd_shutdown": (_int, [_vp]),
    "gfd_driver_version": (_int, [_vp, ctypes.c_char_p, _sz]),
    "gfd_cuda_driver_version": (_int, [_vp, _P(ctypes.c_uint), _P(ctypes.c_uint)]),
    "gfd_device_count": (_int, [_vp, _P(_int)]),
    "gfd_device_get": (_int, [_vp, _int, _P(_vp)]),
    "gfd_prefetch": (_int, [_vp, _int]),
    "gfd_call_stats": (_int, [_vp, _P(_u64), _P(_u64)]),
    "gfd_dev_name": (_int, [_vp, ctypes.c_char_p, _sz]),
    "gfd_dev_total_memory_mb": (_int, [_vp, _P(_u64)]),
    "gfd_dev_cuda_compute_capability": (_int, [_vp, _P(_int), _P(_int)]),
    "gfd_dev_is_mig_capable": (_int, [_vp, _P(_int)]),
    "gfd_dev_is_mig_enabled": (_int, [_vp, _P(_int)]),
    "gfd_dev_mig_count": (_int, [_vp, _P(_int)]),
    "gfd_dev_mig_get": (_int, [_vp, _int, _P(_vp)]),
    "gfd_dev_parent": (_int, [_vp, _P(_vp)]),
    "gfd_dev_attributes": (_int, [_vp, _P(MigAttributes)]),
    "gfd_dev_uuid": (_int, [_vp, ctypes.c_char_p, _sz]),
    "gfd_mig_memory_size_gb": (_u64, [_u64, _u64]),
    "gfd_mig_profile_string": (_int, [_int, _int, _u64, _u64, ctypes.c_char_p, _sz]),
    "gfd_vgpu_parse": (_int, [ctypes.c_char_p, _sz, _sz, _P(VgpuInfo)]),
    "gfd_hbm_fill_pattern": (_int, [_vp, _sz, _u64, _vp]),
    "gfd_hbm_pattern_checksum": (_int, [_sz, _u64, _vp, _vp, _sz, _vp]),
    "gfd_hbm_scratch_bytes": (_sz, []),
    "gfd_hbm_scratch_init": (_int, [_vp, _sz, _vp]),
    "gfd_hbm_read_checksum": (_int, [_vp, _sz, _vp, _vp, _sz, _vp]),
    "gfd_hbm_read_checksum_noalloc": (_int, [_vp, _sz, _vp, _vp]),
    "gfd_hbm_copy": (_int, [_vp, _vp, _sz, _vp]),
    "gfd_checksum_host": (_int, [_int, _vp, _sz, _P(Checksum)]),
    "gfd_probe_hbm": (_int, [_int, _sz, _int, _int, _int, _P(ProbeResult)]),
    "gfd_probe_p2p": (_int, [_int, _int, _sz, _int, _int, _int, _P(ProbeResult)]),
    "gfd_probe_p2p_matrix": (_int, [_int, _sz, _int, _int, _int, _P(ctypes.c_double), _P(ctypes.c_double)]),
    "gfd_cuda_device_by_uuid": (_int, [ctypes.c_char_p, _P(_int)]),
    "gfd_cuda_device_count": (_int, [_P(_int)]),
}
for _name, (_res, _args) in _SIG.items():
    _f = getattr(lib, _name)  # AttributeError here == the library does not export what the header declares
    _f.restype, _f.argtypes = _res, _args


def strerror(status):
    return lib.gfd_strerror(status).decode("utf-8", "replace")


def _check(status, where=""):
    if status != 0:
        raise GfdError(status, where)


def _str_call(fn, handle, n=256):
    buf = ctypes.create_string_buffer(n)
    _check(fn(handle, buf, n), fn.__name__)
    return buf.value.decode("utf-8", "surrogateescape")


def has_nvml():
    buf = ctypes.create_string_buffer(512)
    return bool(lib.gfd_has_nvml(buf, 512)), buf.value.decode("utf-8", "replace")


def is_tegra_system(root=""):
    buf = ctypes.create_string_buffer(512)
    return bool(lib.gfd_is_tegra_system(root.encode(), buf, 512)), buf.value.decode("utf-8", "replace")


class Device:
    """resource.Device (internal/resource/types.go:32-42)."""

    def __init__(self, handle):
        self.h = handle

    def get_name(self): return _str_call(lib.gfd_dev_name, self.h)
    def get_uuid(self): return _str_call(lib.gfd_dev_uuid, self.h)

    def get_total_memory_mb(self):
        v = _u64()
        _check(lib.gfd_dev_total_memory_mb(self.h, ctypes.byref(v)), "GetTotalMemoryMB")
        return int(v.value)

    def get_cuda_compute_capability(self):
        a, b = _int(), _int()
        _check(lib.gfd_dev_cuda_compute_capability(self.h, ctypes.byref(a), ctypes.byref(b)), "GetCudaComputeCapability")
        return a.value, b.value

    def is_mig_capable(self):
        v = _int()
        _check(lib.gfd_dev_is_mig_capable(self.h, ctypes.byref(v)), "IsMigCapable")
        return bool(v.value)

    def is_mig_enabled(self):
        v = _int()
        _check(lib.gfd_dev_is_mig_enabled(self.h, ctypes.byref(v)), "IsMigEnabled")
        return bool(v.value)

    def get_mig_devices(self):
        n = _int()
        _check(lib.gfd_dev_mig_count(self.h, ctypes.byref(n)), "GetMigDevices")
        out = []
        for i in range(n.value):
            h = _vp()
            _check(lib.gfd_dev_mig_get(self.h, i, ctypes.byref(h)), "GetMigDevices")
            out.append(Device(h))
        return out

    def get_device_handle_from_mig_device_handle(self):
        h = _vp()
        _check(lib.gfd_dev_parent(self.h, ctypes.byref(h)), "GetDeviceHandleFromMigDeviceHandle")
        return Device(h)

    def get_attributes(self):
        a = MigAttributes()
        _check(lib.gfd_dev_attributes(self.h, ctypes.byref(a)), "GetAttributes")
        return {"memory": a.memory_mb, "multiprocessors": a.multiprocessors, "slices.gi": a.slices_gi,
                "slices.ci": a.slices_ci, "engines.copy": a.engines_copy, "engines.decoder": a.engines_decoder,
                "engines.encoder": a.engines_encoder, "engines.jpeg": a.engines_jpeg, "engines.ofa": a.engines_ofa}


class Manager:
    """resource.Manager (internal/resource/types.go:22-30) over the C ABI."""

    def __init__(self, backend=BACKEND_NVML, flags=0):
        self.h = _vp()
        _check(lib.gfd_mgr_new(backend, flags, ctypes.byref(self.h)), "gfd_mgr_new")

    def close(self):
        if self.h:
            lib.gfd_mgr_free(self.h)
            self.h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def init(self): _check(lib.gfd_init(self.h), "Init")
    def shutdown(self): _check(lib.gfd_shutdown(self.h), "Shutdown")
    def get_driver_version(self): return _str_call(lib.gfd_driver_version, self.h)
    def prefetch(self, threads): _check(lib.gfd_prefetch(self.h, threads), "gfd_prefetch")

    def get_cuda_driver_version(self):
        a, b = ctypes.c_uint(), ctypes.c_uint()
        _check(lib.gfd_cuda_driver_version(self.h, ctypes.byref(a), ctypes.byref(b)), "GetCudaDriverVersion")
        return a.value, b.value

    def get_devices(self):
        n = _int()
        _check(lib.gfd_device_count(self.h, ctypes.byref(n)), "GetDevices")
        out = []
        for i in range(n.value):
            h = _vp()
            _check(lib.gfd_device_get(self.h, i, ctypes.byref(h)), "GetDevices")
            out.append(Device(h))
        return out

    def call_stats(self):
        a, b = _u64(), _u64()
        _check(lib.gfd_call_stats(self.h, ctypes.byref(a), ctypes.byref(b)))
        return int(a.value), int(b.value)


def mig_memory_size_gb(total_bytes, mig_mb):
    return int(lib.gfd_mig_memory_size_gb(total_bytes, mig_mb))


def mig_profile_string(gi_profile_id, ci_profile_id, mig_mb, total_bytes):
    buf = ctypes.create_string_buffer(96)
    _check(lib.gfd_mig_profile_string(gi_profile_id, ci_profile_id, mig_mb, total_bytes, buf, 96), "NewMigProfile")
    return buf.value.decode()


def vgpu_parse(config, cap=None):
    info = VgpuInfo()
    config = bytes(config)
    _check(lib.gfd_vgpu_parse(config, len(config), len(config) if cap is None else cap, ctypes.byref(info)), "vgpu")
    return info


# ---- probes (device pointers are plain ints, e.g. torch.Tensor.data_ptr(); streams torch's .cuda_stream) ----
def hbm_scratch_bytes():
    return int(lib.gfd_hbm_scratch_bytes())


def hbm_scratch_init(d_scratch, scratch_bytes, stream=0):
    _check(lib.gfd_hbm_scratch_init(d_scratch, scratch_bytes, stream), "gfd_hbm_scratch_init")


def hbm_fill_pattern(d_buf, nbytes, seed=0, stream=0):
    _check(lib.gfd_hbm_fill_pattern(d_buf, nbytes, seed, stream), "gfd_hbm_fill_pattern")


def hbm_pattern_checksum(nbytes, seed, d_out, d_scratch, scratch_bytes, stream=0):
    _check(lib.gfd_hbm_pattern_checksum(nbytes, seed, d_out, d_scratch, scratch_bytes, stream), "gfd_hbm_pattern_checksum")


def hbm_read_checksum(d_buf, nbytes, d_out, d_scratch, scratch_bytes, stream=0):
    _check(lib.gfd_hbm_read_checksum(d_buf, nbytes, d_out, d_scratch, scratch_bytes, stream), "gfd_hbm_read_checksum")


def hbm_read_checksum_noalloc(d_buf, nbytes, d_out, stream=0):
    _check(lib.gfd_hbm_read_checksum_noalloc(d_buf, nbytes, d_out, stream), "gfd_hbm_read_checksum_noalloc")


NOALLOC_MAX_BYTES = 16 << 30


def hbm_copy(d_dst, d_src, nbytes, stream=0):
    _check(lib.gfd_hbm_copy(d_dst, d_src, nbytes, stream), "gfd_hbm_copy")


def checksum_host(device, host_ptr, nbytes):
    out = Checksum()
    _check(lib.gfd_checksum_host(device, host_ptr, nbytes, ctypes.byref(out)), "gfd_checksum_host")
    return out.tuple()


def probe_hbm(device, nbytes, warmup=3, iters=20, mode=PROBE_READ):
    r = ProbeResult()
    _check(lib.gfd_probe_hbm(device, nbytes, warmup, iters, mode, ctypes.byref(r)), "gfd_probe_hbm")
    return r


def probe_p2p(src, dst, nbytes, warmup=3, iters=20, mode=PROBE_READ):
    r = ProbeResult()
    _check(lib.gfd_probe_p2p(src, dst, nbytes, warmup, iters, mode, ctypes.byref(r)), "gfd_probe_p2p")
    return r


def probe_p2p_matrix(n, nbytes, warmup=2, iters=10, mode=PROBE_READ):
    m = (ctypes.c_double * (n * n))()
    ring = ctypes.c_double()
    _check(lib.gfd_probe_p2p_matrix(n, nbytes, warmup, iters, mode, m, ctypes.byref(ring)), "gfd_probe_p2p_matrix")
    return [[m[r * n + c] for c in range(n)] for r in range(n)], ring.value


def cuda_device_count():
    n = _int()
    _check(lib.gfd_cuda_device_count(ctypes.byref(n)), "gfd_cuda_device_count")
    return n.value
