"""Helpers for CPU tests that drive libgfd_b200.so against the fake libnvidia-ml.so.1.

The fake is pulled into THIS process by absolute path with RTLD_GLOBAL; because it carries the SONAME
libnvidia-ml.so.1, the shim's own dlopen("libnvidia-ml.so.1") then resolves to it.  The fixture file is
re-read at every nvmlInit_v2, so tests switch nodes by rewriting $GFD_FAKE_NVML."""
import ctypes
import importlib
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
B = importlib.import_module("gpu-feature-discovery_b200.binding")
from oracle import gfd_oracle as O  # noqa: E402

_fake = None


def load_fake():
    global _fake
    if _fake is None:
        _fake = ctypes.CDLL(os.path.join(B.FAKE_NVML_DIR, "libnvidia-ml.so.1"), mode=ctypes.RTLD_GLOBAL)
    return _fake


def use_node(node, tmp_path, name="node.nvmlfix"):
    p = os.path.join(str(tmp_path), name)
    with open(p, "w") as f:
        f.write(node.dumps())
    os.environ["GFD_FAKE_NVML"] = p
    return p


class ShimDevice(O.Device):
    """Adapts binding.Device to the oracle's Device interface (errors become O.GfdError)."""

    def __init__(self, d):
        self.d = d

    def _w(self, fn, *a):
        try:
            return fn(*a)
        except B.GfdError as e:
            raise O.GfdError(str(e))

    def is_mig_enabled(self): return self._w(self.d.is_mig_enabled)
    def is_mig_capable(self): return self._w(self.d.is_mig_capable)
    def get_mig_devices(self): return [ShimDevice(x) for x in self._w(self.d.get_mig_devices)]
    def get_attributes(self): return self._w(self.d.get_attributes)
    def get_name(self): return self._w(self.d.get_name)
    def get_total_memory_mb(self): return self._w(self.d.get_total_memory_mb)
    def get_device_handle_from_mig_device_handle(self): return ShimDevice(self._w(self.d.get_device_handle_from_mig_device_handle))
    def get_cuda_compute_capability(self): return self._w(self.d.get_cuda_compute_capability)


class ShimManager(O.Manager):
    def __init__(self, backend=B.BACKEND_NVML, flags=0, prefetch_threads=0):
        self.m = B.Manager(backend, flags)
        self.prefetch_threads = prefetch_threads

    def _w(self, fn, *a):
        try:
            return fn(*a)
        except B.GfdError as e:
            raise O.GfdError(str(e))

    def init(self):
        self._w(self.m.init)
        if self.prefetch_threads:  # > 1: that many threads; < 0: auto (one per GPU); 0: no prefetch at all
            self.m.prefetch(self.prefetch_threads if self.prefetch_threads > 0 else 0)

    def shutdown(self): return self._w(self.m.shutdown)
    def get_devices(self): return [ShimDevice(x) for x in self._w(self.m.get_devices)]
    def get_driver_version(self): return self._w(self.m.get_driver_version)
    def get_cuda_driver_version(self): return self._w(self.m.get_cuda_driver_version)


def binary_path():
    """The labeler binary under test; GFD_TEST_BINDIR=build/asan selects the ASan/UBSan build (`make asan`)."""
    d = os.environ.get("GFD_TEST_BINDIR")
    return os.path.join(ROOT, d, "gpu-feature-discovery") if d else B.BIN_PATH


def run_cli(args, env=None, fake=True, nomig=False, timeout=60):
    """Run the gpu-feature-discovery binary; returns (returncode, stdout, stderr)."""
    e = dict(os.environ)
    for k in list(e):
        if k.startswith("GFD_") or k in ("MIG_STRATEGY", "FAIL_ON_INIT_ERROR", "CONFIG_FILE"):
            del e[k]
    if fake:
        e["LD_LIBRARY_PATH"] = (B.FAKE_NVML_NOMIG_DIR if nomig else B.FAKE_NVML_DIR) + ":" + e.get("LD_LIBRARY_PATH", "")
    e.update(env or {})
    p = subprocess.run([binary_path()] + list(args), env=e, capture_output=True, timeout=timeout)
    return p.returncode, p.stdout.decode("utf-8", "surrogateescape"), p.stderr.decode("utf-8", "surrogateescape")
