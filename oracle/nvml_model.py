"""Raw-NVML models for the oracle -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Two interchangeable back-ends for oracle/gfd_oracle.py:NvmlManager, each exposing the ~24 NVML entry
points the reference actually calls (V/go-nvml/nvml/nvml.go) as methods that return
`(rc, value)`:

  * RawNvml    -- interprets a `.nvmlfix` fixture, the same text file the C fake
                  `libnvidia-ml.so.1` (gpu-feature-discovery_b200/fake_nvml/fake_nvml.c) interprets, so
                  the oracle and the product see the *same* simulated node.
  * PynvmlRaw  -- the real NVML on a GPU box through `pynvml` (harvests the raw values; no label logic).

Fixture grammar (one directive per line, '#' comments, indentation ignored):

    driver_version <str>            cuda_driver_version <int>
    hide_symbol <nvml symbol>       (dlsym fails: models a pre-MIG driver; C side = the *_nomig build)
    rc <nvml symbol> <code>         (return-code override; scope = the block it appears in)
    gpu                             (starts a GPU block; index = order of appearance)
      name <rest of line>           memory_total <bytes>          cc <major> <minor>
      mig_mode <current> <pending>  (absent => nvmlDeviceGetMigMode returns NOT_SUPPORTED)
      max_mig_count <n>
      gi_profile <idx> id <id> slices <n> memory_mb <mb>     (absent idx => NOT_SUPPORTED)
      ci_profile <gi profile id> <idx> id <id> slices <n>    (absent => NOT_SUPPORTED)
      mig <slot>                    (starts a MIG-device block in that MIG slot; absent slot => NOT_FOUND)
        gi_id <n>  ci_id <n>  gi_profile_id <n>  ci_profile_id <n>
        attrs <memMB> <SMs> <giSlices> <ciSlices> <copy> <dec> <enc> <jpeg> <ofa>
"""
from .gfd_oracle import (NVML_SUCCESS, NVML_ERROR_INVALID_ARGUMENT, NVML_ERROR_NOT_SUPPORTED,
                         NVML_ERROR_NOT_FOUND, NVML_ERROR_UNINITIALIZED, NVML_ERROR_UNKNOWN)

_ERR_NAMES = {
    0: "SUCCESS", 1: "ERROR_UNINITIALIZED", 2: "ERROR_INVALID_ARGUMENT", 3: "ERROR_NOT_SUPPORTED",
    4: "ERROR_NO_PERMISSION", 5: "ERROR_ALREADY_INITIALIZED", 6: "ERROR_NOT_FOUND", 7: "ERROR_INSUFFICIENT_SIZE",
    8: "ERROR_INSUFFICIENT_POWER", 9: "ERROR_DRIVER_NOT_LOADED", 10: "ERROR_TIMEOUT", 11: "ERROR_IRQ_ISSUE",
    12: "ERROR_LIBRARY_NOT_FOUND", 13: "ERROR_FUNCTION_NOT_FOUND", 14: "ERROR_CORRUPTED_INFOROM",
    15: "ERROR_GPU_IS_LOST", 16: "ERROR_RESET_REQUIRED", 17: "ERROR_OPERATING_SYSTEM",
    18: "ERROR_LIB_RM_VERSION_MISMATCH", 19: "ERROR_IN_USE", 20: "ERROR_MEMORY", 21: "ERROR_NO_DATA",
    22: "ERROR_VGPU_ECC_NOT_SUPPORTED", 23: "ERROR_INSUFFICIENT_RESOURCES", 999: "ERROR_UNKNOWN",
}

ATTR_FIELDS = ("memory_size_mb", "multiprocessor_count", "gpu_instance_slice_count", "compute_instance_slice_count",
               "shared_copy_engine_count", "shared_decoder_count", "shared_encoder_count", "shared_jpeg_count",
               "shared_ofa_count")


# ----------------------------------------------------------------------------
# fixture model + (de)serialisation
# ----------------------------------------------------------------------------
class MigFix:
    def __init__(self, slot, gi_id=0, ci_id=0, gi_profile_id=0, ci_profile_id=0, attrs=(0,) * 9, rc=None):
        self.slot, self.gi_id, self.ci_id = slot, gi_id, ci_id
        self.gi_profile_id, self.ci_profile_id = gi_profile_id, ci_profile_id
        self.attrs = tuple(attrs)
        self.rc = dict(rc or {})


class GpuFix:
    def __init__(self, name="NVIDIA B200", memory_total=0, cc=(0, 0), mig_mode=None, max_mig_count=0,
                 gi_profiles=None, ci_profiles=None, migs=(), rc=None):
        self.name, self.memory_total, self.cc = name, memory_total, tuple(cc)
        self.mig_mode = None if mig_mode is None else tuple(mig_mode)
        self.max_mig_count = max_mig_count
        self.gi_profiles = dict(gi_profiles or {})  # idx -> (id, slices, memory_mb)
        self.ci_profiles = dict(ci_profiles or {})  # (gi_profile_id, idx) -> (id, slices)
        self.migs = list(migs)
        self.rc = dict(rc or {})


class NodeFix:
    def __init__(self, driver_version="575.57.08", cuda_driver_version=12090, gpus=(), rc=None, hidden=()):
        self.driver_version, self.cuda_driver_version = driver_version, cuda_driver_version
        self.gpus = list(gpus)
        self.rc = dict(rc or {})
        self.hidden = set(hidden)

    def dumps(self):
        out = ["driver_version %s" % self.driver_version, "cuda_driver_version %d" % self.cuda_driver_version]
        out += ["hide_symbol %s" % s for s in sorted(self.hidden)]
        out += ["rc %s %d" % kv for kv in sorted(self.rc.items())]
        for g in self.gpus:
            out.append("gpu")
            out.append("  name %s" % g.name)
            out.append("  memory_total %d" % g.memory_total)
            out.append("  cc %d %d" % g.cc)
            if g.mig_mode is not None:
                out.append("  mig_mode %d %d" % g.mig_mode)
            out.append("  max_mig_count %d" % g.max_mig_count)
            for idx, (pid, sl, mb) in sorted(g.gi_profiles.items()):
                out.append("  gi_profile %d id %d slices %d memory_mb %d" % (idx, pid, sl, mb))
            for (gp, idx), (pid, sl) in sorted(g.ci_profiles.items()):
                out.append("  ci_profile %d %d id %d slices %d" % (gp, idx, pid, sl))
            out += ["  rc %s %d" % kv for kv in sorted(g.rc.items())]
            for m in g.migs:
                out.append("  mig %d" % m.slot)
                out.append("    gi_id %d" % m.gi_id)
                out.append("    ci_id %d" % m.ci_id)
                out.append("    gi_profile_id %d" % m.gi_profile_id)
                out.append("    ci_profile_id %d" % m.ci_profile_id)
                out.append("    attrs " + " ".join("%d" % a for a in m.attrs))
                out += ["    rc %s %d" % kv for kv in sorted(m.rc.items())]
        return "\n".join(out) + "\n"


def loads(text):
    node, gpu, mig = NodeFix(gpus=[]), None, None
    for line in text.splitlines():
        line = line.split("#", 1)[0].strip()
        if not line:
            continue
        tok = line.split()
        k = tok[0]
        if k == "driver_version":
            node.driver_version = tok[1] if len(tok) > 1 else ""
        elif k == "cuda_driver_version":
            node.cuda_driver_version = int(tok[1])
        elif k == "hide_symbol":
            node.hidden.add(tok[1])
        elif k == "rc":
            scope = mig if mig is not None else gpu if gpu is not None else node
            scope.rc[tok[1]] = int(tok[2])
        elif k == "gpu":
            gpu, mig = GpuFix(name=""), None
            node.gpus.append(gpu)
        elif k == "name":
            gpu.name = line[len("name"):].strip()
        elif k == "memory_total":
            gpu.memory_total = int(tok[1])
        elif k == "cc":
            gpu.cc = (int(tok[1]), int(tok[2]))
        elif k == "mig_mode":
            gpu.mig_mode = (int(tok[1]), int(tok[2]))
        elif k == "max_mig_count":
            gpu.max_mig_count = int(tok[1])
        elif k == "gi_profile":
            kv = dict(zip(tok[2::2], map(int, tok[3::2])))
            gpu.gi_profiles[int(tok[1])] = (kv["id"], kv["slices"], kv["memory_mb"])
        elif k == "ci_profile":
            kv = dict(zip(tok[3::2], map(int, tok[4::2])))
            gpu.ci_profiles[(int(tok[1]), int(tok[2]))] = (kv["id"], kv["slices"])
        elif k == "mig":
            mig = MigFix(int(tok[1]))
            gpu.migs.append(mig)
        elif k in ("gi_id", "ci_id", "gi_profile_id", "ci_profile_id"):
            setattr(mig, k, int(tok[1]))
        elif k == "attrs":
            mig.attrs = tuple(int(x) for x in tok[1:10])
        else:
            raise ValueError("unknown fixture directive: %r" % line)
    return node


def load(path):
    with open(path) as f:
        return loads(f.read())


# ----------------------------------------------------------------------------
# RawNvml: the fixture interpreter (must agree, call for call, with fake_nvml.c)
# ----------------------------------------------------------------------------
class RawNvml:
    def __init__(self, node):
        self.node = node if isinstance(node, NodeFix) else loads(node)
        self.initialized = 0
        self.trace = []  # NVML symbols in call order; "dlsym:<sym>" for lookup probes

    # handles are tuples: ("gpu", i) / ("mig", i, slot_list_index) / ("gi", i, gi_id) / ("ci", i, gi_id, ci_id)
    def _gpu(self, h): return self.node.gpus[h[1]]
    def _mig(self, h): return self.node.gpus[h[1]].migs[h[2]]

    def _rc(self, sym, h=None):
        self.trace.append(sym)
        if sym not in ("nvmlInit_v2", "nvmlErrorString") and not self.initialized:
            return NVML_ERROR_UNINITIALIZED
        if h is not None:
            if h[0] == "mig" and sym in self._mig(h).rc:
                return self._mig(h).rc[sym]
            if sym in self._gpu(h).rc:
                return self._gpu(h).rc[sym]
        return self.node.rc.get(sym, NVML_SUCCESS)

    def error_string(self, rc):
        return _ERR_NAMES.get(rc, "Unknown return value: %d" % rc)

    def lookup_symbol(self, sym):
        self.trace.append("dlsym:" + sym)
        return sym not in self.node.hidden

    def init(self):
        rc = self._rc("nvmlInit_v2")
        if rc == NVML_SUCCESS:
            self.initialized += 1
        return rc, None

    def shutdown(self):
        rc = self._rc("nvmlShutdown")
        if rc != NVML_ERROR_UNINITIALIZED and self.initialized > 0:
            self.initialized -= 1  # an injected error still tears down (same rule as fake_nvml.c)
        return rc, None

    def system_get_driver_version(self):
        return self._rc("nvmlSystemGetDriverVersion"), self.node.driver_version

    def system_get_cuda_driver_version(self):
        return self._rc("nvmlSystemGetCudaDriverVersion"), self.node.cuda_driver_version

    def device_get_count(self):
        return self._rc("nvmlDeviceGetCount_v2"), len(self.node.gpus)

    def device_get_handle_by_index(self, i):
        if not (0 <= i < len(self.node.gpus)):
            self.trace.append("nvmlDeviceGetHandleByIndex_v2")
            return NVML_ERROR_INVALID_ARGUMENT, None
        h = ("gpu", i)
        return self._rc("nvmlDeviceGetHandleByIndex_v2", h), h

    def device_get_name(self, h):
        rc = self._rc("nvmlDeviceGetName", h)
        if h[0] == "mig":
            return rc, self._gpu(h).name + " MIG"
        return rc, self._gpu(h).name

    def device_get_memory_info(self, h):
        rc = self._rc("nvmlDeviceGetMemoryInfo", h)
        if h[0] == "mig":
            return rc, self._mig(h).attrs[0] * 1024 * 1024
        return rc, self._gpu(h).memory_total

    def device_get_cuda_compute_capability(self, h):
        return self._rc("nvmlDeviceGetCudaComputeCapability", h), self._gpu(h).cc

    def device_get_mig_mode(self, h):
        rc = self._rc("nvmlDeviceGetMigMode", h)
        if rc != NVML_SUCCESS:
            return rc, None
        if h[0] != "gpu":
            return NVML_ERROR_INVALID_ARGUMENT, None
        if self._gpu(h).mig_mode is None:
            return NVML_ERROR_NOT_SUPPORTED, None
        return rc, self._gpu(h).mig_mode

    def device_get_max_mig_device_count(self, h):
        return self._rc("nvmlDeviceGetMaxMigDeviceCount", h), (self._gpu(h).max_mig_count if h[0] == "gpu" else 0)

    def device_get_mig_device_handle_by_index(self, h, idx):
        g = self._gpu(h)
        for k, m in enumerate(g.migs):
            if m.slot == idx:
                mh = ("mig", h[1], k)
                return self._rc("nvmlDeviceGetMigDeviceHandleByIndex", mh), mh
        rc = self._rc("nvmlDeviceGetMigDeviceHandleByIndex", h)
        if rc != NVML_SUCCESS:
            return rc, None
        if idx >= g.max_mig_count:
            return NVML_ERROR_INVALID_ARGUMENT, None
        return NVML_ERROR_NOT_FOUND, None

    def device_is_mig_device_handle(self, h):
        return self._rc("nvmlDeviceIsMigDeviceHandle", h), h[0] == "mig"

    def device_get_device_handle_from_mig_device_handle(self, h):
        rc = self._rc("nvmlDeviceGetDeviceHandleFromMigDeviceHandle", h)
        if rc != NVML_SUCCESS:
            return rc, None
        if h[0] != "mig":
            return NVML_ERROR_INVALID_ARGUMENT, None
        return rc, ("gpu", h[1])

    def device_get_attributes(self, h):
        rc = self._rc("nvmlDeviceGetAttributes_v2", h)
        if rc != NVML_SUCCESS:
            return rc, None
        if h[0] != "mig":
            return NVML_ERROR_INVALID_ARGUMENT, None
        return rc, dict(zip(ATTR_FIELDS, self._mig(h).attrs))

    def device_get_gpu_instance_id(self, h):
        rc = self._rc("nvmlDeviceGetGpuInstanceId", h)
        if rc != NVML_SUCCESS:
            return rc, None
        if h[0] != "mig":
            return NVML_ERROR_INVALID_ARGUMENT, None
        return rc, self._mig(h).gi_id

    def device_get_compute_instance_id(self, h):
        rc = self._rc("nvmlDeviceGetComputeInstanceId", h)
        if rc != NVML_SUCCESS:
            return rc, None
        if h[0] != "mig":
            return NVML_ERROR_INVALID_ARGUMENT, None
        return rc, self._mig(h).ci_id

    def device_get_gpu_instance_by_id(self, h, gi_id):
        rc = self._rc("nvmlDeviceGetGpuInstanceById", h)
        if rc != NVML_SUCCESS:
            return rc, None
        if h[0] != "gpu":
            return NVML_ERROR_INVALID_ARGUMENT, None
        for m in self._gpu(h).migs:
            if m.gi_id == gi_id:
                return rc, ("gi", h[1], gi_id)
        return NVML_ERROR_NOT_FOUND, None

    def _gi_first_mig(self, gi):
        for m in self.node.gpus[gi[1]].migs:
            if m.gi_id == gi[2]:
                return m
        return None

    def gpu_instance_get_info(self, gi):
        rc = self._rc("nvmlGpuInstanceGetInfo", ("gpu", gi[1]))
        if rc != NVML_SUCCESS:
            return rc, None
        return rc, {"id": gi[2], "profile_id": self._gi_first_mig(gi).gi_profile_id}

    def gpu_instance_get_compute_instance_by_id(self, gi, ci_id):
        rc = self._rc("nvmlGpuInstanceGetComputeInstanceById", ("gpu", gi[1]))
        if rc != NVML_SUCCESS:
            return rc, None
        for m in self.node.gpus[gi[1]].migs:
            if m.gi_id == gi[2] and m.ci_id == ci_id:
                return rc, ("ci", gi[1], gi[2], ci_id)
        return NVML_ERROR_NOT_FOUND, None

    def compute_instance_get_info(self, ci):
        rc = self._rc("nvmlComputeInstanceGetInfo_v2", ("gpu", ci[1]))
        if rc != NVML_SUCCESS:
            return rc, None
        for m in self.node.gpus[ci[1]].migs:
            if m.gi_id == ci[2] and m.ci_id == ci[3]:
                return rc, {"id": ci[3], "profile_id": m.ci_profile_id}
        return NVML_ERROR_NOT_FOUND, None

    def device_get_gpu_instance_profile_info(self, h, idx):
        rc = self._rc("nvmlDeviceGetGpuInstanceProfileInfo", h)
        if rc != NVML_SUCCESS:
            return rc, None
        if h[0] != "gpu":
            return NVML_ERROR_INVALID_ARGUMENT, None
        p = self._gpu(h).gi_profiles.get(idx)
        if p is None:
            return NVML_ERROR_NOT_SUPPORTED, None
        return rc, {"id": p[0], "slice_count": p[1], "memory_size_mb": p[2]}

    def gpu_instance_get_compute_instance_profile_info(self, gi, idx, eng):
        rc = self._rc("nvmlGpuInstanceGetComputeInstanceProfileInfo", ("gpu", gi[1]))
        if rc != NVML_SUCCESS:
            return rc, None
        if eng != 0:
            return NVML_ERROR_INVALID_ARGUMENT, None
        gpid = self._gi_first_mig(gi).gi_profile_id
        p = self.node.gpus[gi[1]].ci_profiles.get((gpid, idx))
        if p is None:
            return NVML_ERROR_NOT_SUPPORTED, None
        return rc, {"id": p[0], "slice_count": p[1]}


# ----------------------------------------------------------------------------
# Node presets used by tests/fixtures (values from SURVEY.md 8(d): 183359 MiB, cc 10.0, driver 575.57.08)
# ----------------------------------------------------------------------------
B200_TOTAL_BYTES = 183359 * 1024 * 1024
# profile idx == profile id on real NVML.  id -> (slices, memory_mb)
B200_GI_PROFILES = {0: (1, 22912), 1: (2, 45824), 2: (3, 91648), 3: (4, 91648), 4: (7, 182272),
                    7: (1, 22912), 9: (1, 45824)}
# attrs for a <slices>-slice instance: SMs, copy, dec, enc, jpeg, ofa (illustrative; passed through verbatim)
B200_SLICE_ATTRS = {1: (18, 1, 1, 0, 1, 0), 2: (36, 2, 2, 0, 2, 0), 3: (54, 3, 3, 0, 3, 0),
                    4: (72, 4, 4, 0, 4, 0), 7: (132, 7, 7, 0, 7, 1)}
_CI_IDX_SLICES = {0: 1, 1: 2, 2: 3, 3: 4, 4: 7}  # CI profile idx/id -> slices


def b200_gpu(mig=None, name="NVIDIA B200", total=B200_TOTAL_BYTES, cc=(10, 0), mig_mode="auto"):
    """A B200 GpuFix.  mig = None (MIG disabled) or a list of (gi_profile_id, ci_profile_id) per MIG device
    (one CI per GI; an empty list = MIG enabled but empty)."""
    g = GpuFix(name=name, memory_total=total, cc=cc, max_mig_count=7)
    g.gi_profiles = {idx: (idx, sl, mb) for idx, (sl, mb) in B200_GI_PROFILES.items()}
    for gp, (gsl, _) in B200_GI_PROFILES.items():
        for cidx, csl in _CI_IDX_SLICES.items():
            if csl <= gsl:
                g.ci_profiles[(gp, cidx)] = (cidx, csl)
    if mig_mode == "auto":
        g.mig_mode = (0, 0) if mig is None else (1, 1)
    else:
        g.mig_mode = mig_mode
    for slot, (gp, cp) in enumerate(mig or []):
        gsl, mb = B200_GI_PROFILES[gp]
        csl = _CI_IDX_SLICES[cp]
        sm, ce, dec, enc, jpg, ofa = B200_SLICE_ATTRS[gsl]
        if gp == 7:  # +me keeps the media engines
            enc, ofa = 0, 1
        sm = sm * csl // gsl
        g.migs.append(MigFix(slot, gi_id=slot + 1, ci_id=0, gi_profile_id=gp, ci_profile_id=cp,
                             attrs=(mb, sm, gsl, csl, ce, dec, enc, jpg, ofa)))
    return g


# ----------------------------------------------------------------------------
# PynvmlRaw: real NVML on the GPU box (harvest only)
# ----------------------------------------------------------------------------
class PynvmlRaw:
    """Same method set over pynvml.  Handles are pynvml handles."""

    def __init__(self):
        import pynvml
        self.nv = pynvml
        self.trace = []

    def _call(self, sym, fn, *a):
        self.trace.append(sym)
        try:
            return NVML_SUCCESS, fn(*a)
        except self.nv.NVMLError as e:
            return int(e.value), None

    def error_string(self, rc):
        try:
            s = self.nv.nvmlErrorString(rc)
            return s.decode() if isinstance(s, bytes) else s
        except Exception:
            return _ERR_NAMES.get(rc, "Unknown return value: %d" % rc)

    def lookup_symbol(self, sym):
        self.trace.append("dlsym:" + sym)
        import ctypes
        try:
            lib = ctypes.CDLL("libnvidia-ml.so.1")
            return hasattr(lib, sym)
        except OSError:
            return False

    @staticmethod
    def _str(v):
        return v.decode() if isinstance(v, bytes) else v

    @staticmethod
    def _v1(fn):
        """The reference binds the v1-struct symbols (V/go-nvml/nvml/nvml.go:2354,2364); current pynvml defaults to the `...V`
        (v2 struct) entry points, so ask for version 1 where the wrapper takes a version."""
        import inspect
        if "version" in inspect.signature(fn).parameters:
            return lambda *a: fn(*a, version=1)
        return fn

    def init(self): return self._call("nvmlInit_v2", self.nv.nvmlInit)
    def shutdown(self): return self._call("nvmlShutdown", self.nv.nvmlShutdown)

    def system_get_driver_version(self):
        rc, v = self._call("nvmlSystemGetDriverVersion", self.nv.nvmlSystemGetDriverVersion)
        return rc, (self._str(v) if rc == 0 else None)

    def system_get_cuda_driver_version(self):
        return self._call("nvmlSystemGetCudaDriverVersion", self.nv.nvmlSystemGetCudaDriverVersion)

    def device_get_count(self): return self._call("nvmlDeviceGetCount_v2", self.nv.nvmlDeviceGetCount)
    def device_get_handle_by_index(self, i): return self._call("nvmlDeviceGetHandleByIndex_v2", self.nv.nvmlDeviceGetHandleByIndex, i)

    def device_get_name(self, h):
        rc, v = self._call("nvmlDeviceGetName", self.nv.nvmlDeviceGetName, h)
        return rc, (self._str(v) if rc == 0 else None)

    def device_get_memory_info(self, h):
        rc, v = self._call("nvmlDeviceGetMemoryInfo", self.nv.nvmlDeviceGetMemoryInfo, h)  # v1 struct
        return rc, (int(v.total) if rc == 0 else None)

    def device_get_cuda_compute_capability(self, h):
        rc, v = self._call("nvmlDeviceGetCudaComputeCapability", self.nv.nvmlDeviceGetCudaComputeCapability, h)
        return rc, (tuple(int(x) for x in v) if rc == 0 else None)

    def device_get_mig_mode(self, h):
        rc, v = self._call("nvmlDeviceGetMigMode", self.nv.nvmlDeviceGetMigMode, h)
        return rc, (tuple(int(x) for x in v) if rc == 0 else None)

    def device_get_max_mig_device_count(self, h): return self._call("nvmlDeviceGetMaxMigDeviceCount", self.nv.nvmlDeviceGetMaxMigDeviceCount, h)
    def device_get_mig_device_handle_by_index(self, h, i): return self._call("nvmlDeviceGetMigDeviceHandleByIndex", self.nv.nvmlDeviceGetMigDeviceHandleByIndex, h, i)

    def device_is_mig_device_handle(self, h):
        rc, v = self._call("nvmlDeviceIsMigDeviceHandle", self.nv.nvmlDeviceIsMigDeviceHandle, h)
        return rc, bool(v) if rc == 0 else None

    def device_get_device_handle_from_mig_device_handle(self, h):
        return self._call("nvmlDeviceGetDeviceHandleFromMigDeviceHandle", self.nv.nvmlDeviceGetDeviceHandleFromMigDeviceHandle, h)

    def device_get_attributes(self, h):
        rc, a = self._call("nvmlDeviceGetAttributes_v2", self.nv.nvmlDeviceGetAttributes_v2, h)
        if rc != 0:
            return rc, None
        return rc, {
            "memory_size_mb": int(a.memorySizeMB), "multiprocessor_count": int(a.multiprocessorCount),
            "gpu_instance_slice_count": int(a.gpuInstanceSliceCount),
            "compute_instance_slice_count": int(a.computeInstanceSliceCount),
            "shared_copy_engine_count": int(a.sharedCopyEngineCount), "shared_decoder_count": int(a.sharedDecoderCount),
            "shared_encoder_count": int(a.sharedEncoderCount), "shared_jpeg_count": int(a.sharedJpegCount),
            "shared_ofa_count": int(a.sharedOfaCount)}

    def device_get_gpu_instance_id(self, h): return self._call("nvmlDeviceGetGpuInstanceId", self.nv.nvmlDeviceGetGpuInstanceId, h)
    def device_get_compute_instance_id(self, h): return self._call("nvmlDeviceGetComputeInstanceId", self.nv.nvmlDeviceGetComputeInstanceId, h)
    def device_get_gpu_instance_by_id(self, h, i): return self._call("nvmlDeviceGetGpuInstanceById", self.nv.nvmlDeviceGetGpuInstanceById, h, i)

    def gpu_instance_get_info(self, gi):
        rc, v = self._call("nvmlGpuInstanceGetInfo", self.nv.nvmlGpuInstanceGetInfo, gi)
        return rc, ({"id": int(v.id), "profile_id": int(v.profileId)} if rc == 0 else None)

    def gpu_instance_get_compute_instance_by_id(self, gi, i):
        return self._call("nvmlGpuInstanceGetComputeInstanceById", self.nv.nvmlGpuInstanceGetComputeInstanceById, gi, i)

    def compute_instance_get_info(self, ci):
        fn = getattr(self.nv, "nvmlComputeInstanceGetInfo_v2", None) or self.nv.nvmlComputeInstanceGetInfo
        rc, v = self._call("nvmlComputeInstanceGetInfo_v2", fn, ci)
        return rc, ({"id": int(v.id), "profile_id": int(v.profileId)} if rc == 0 else None)

    def device_get_gpu_instance_profile_info(self, h, idx):
        rc, v = self._call("nvmlDeviceGetGpuInstanceProfileInfo", self._v1(self.nv.nvmlDeviceGetGpuInstanceProfileInfo), h, idx)
        return rc, ({"id": int(v.id), "slice_count": int(v.sliceCount), "memory_size_mb": int(v.memorySizeMB)} if rc == 0 else None)

    def gpu_instance_get_compute_instance_profile_info(self, gi, idx, eng):
        rc, v = self._call("nvmlGpuInstanceGetComputeInstanceProfileInfo",
                           self._v1(self.nv.nvmlGpuInstanceGetComputeInstanceProfileInfo), gi, idx, eng)
        return rc, ({"id": int(v.id), "slice_count": int(v.sliceCount)} if rc == 0 else None)


# ----------------------------------------------------------------------------
# RawCuda: the fixture interpreter for the Tegra path (must agree with fake_cuda/fake_cuda.c)
# ----------------------------------------------------------------------------
class RawCuda:
    """devices: list of (total_bytes, cc_major, cc_minor, name).  Every method returns (CUresult, value)."""

    def __init__(self, devices=(), version=12040, rc=None):
        self.devices, self.version, self.rc = list(devices), version, dict(rc or {})

    def dumps(self):
        out = ["version %d" % self.version] + ["rc %s %d" % kv for kv in sorted(self.rc.items())]
        out += ["device %d %d %d %s" % d for d in self.devices]
        return "\n".join(out) + "\n"

    def _rc(self, sym): return self.rc.get(sym, 0)
    def init(self): return self._rc("cuInit"), None
    def driver_get_version(self): return self._rc("cuDriverGetVersion"), self.version
    def device_get_count(self): return self._rc("cuDeviceGetCount"), len(self.devices)

    def device_get(self, i):
        r = self._rc("cuDeviceGet")
        if r:
            return r, None
        return (0, i) if 0 <= i < len(self.devices) else (101, None)

    def device_get_attribute(self, attr, d):
        r = self._rc("cuDeviceGetAttribute")
        if r:
            return r, None
        return 0, {75: self.devices[d][1], 76: self.devices[d][2]}.get(attr, 0)

    def device_total_mem(self, d):
        # the reference binds the legacy unversioned cuDeviceTotalMem (32-bit out-parameter) into a zeroed uint64
        # (internal/cuda/cuda.go:108, api.go:107-112).  The real driver SATURATES that value at 0xFFFFFFFF for devices above
        # 4 GiB (measured through the same symbol on a B200 box, tests/test_gpu_cuda_backend.py); it does not wrap.
        r = self._rc("cuDeviceTotalMem")
        return (r, None) if r else (0, min(self.devices[d][0], 0xFFFFFFFF))

    def device_get_name(self, d):
        r = self._rc("cuDeviceGetName")
        return (r, None) if r else (0, self.devices[d][3][:95])
