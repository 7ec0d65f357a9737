"""-m gpu: the CUDA-driver manager (the reference's Tegra backend: internal/resource/cuda-lib.go, cuda-device.go,
internal/cuda/*) against the REAL libcuda.so.1 of the box.

There is no Tegra hardware here, but the seven driver entry points the backend binds (internal/cuda/cuda.go:103-109) exist in every
libcuda, so on a B200 box the backend can be driven for real: product (C ABI, GFD_BACKEND_CUDA; and the labeler binary with NVML
hidden and a Tegra marker file) vs the oracle's CudaManager fed through ctypes from the same library.  This also pins the one
modelling assumption of the fake libcuda: what the LEGACY unversioned cuDeviceTotalMem -- the symbol the reference binds -- writes
for a device with more than 4 GiB."""
import ctypes
import importlib
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class RealCuda:
    """The RawCuda method set (oracle/nvml_model.py) over the real driver library.  Every method returns (CUresult, value)."""

    def __init__(self):
        self.lib = ctypes.CDLL("libcuda.so.1")

    def init(self):
        return int(self.lib.cuInit(0)), None

    def driver_get_version(self):
        v = ctypes.c_int(0)
        return int(self.lib.cuDriverGetVersion(ctypes.byref(v))), v.value

    def device_get_count(self):
        n = ctypes.c_int(0)
        return int(self.lib.cuDeviceGetCount(ctypes.byref(n))), n.value

    def device_get(self, i):
        d = ctypes.c_int(0)
        r = int(self.lib.cuDeviceGet(ctypes.byref(d), i))
        return r, (d.value if r == 0 else None)

    def device_get_attribute(self, attr, d):
        v = ctypes.c_int(0)
        r = int(self.lib.cuDeviceGetAttribute(ctypes.byref(v), attr, d))
        return r, (v.value if r == 0 else None)

    def device_total_mem(self, d):
        # exactly what the reference's cgo call does: the UNVERSIONED symbol, a zeroed 64-bit out-parameter (internal/cuda/api.go:107-112)
        v = ctypes.c_uint64(0)
        r = int(self.lib.cuDeviceTotalMem(ctypes.byref(v), d))
        return r, (int(v.value) if r == 0 else None)

    def device_total_mem_v2(self, d):
        v = ctypes.c_size_t(0)
        r = int(self.lib.cuDeviceTotalMem_v2(ctypes.byref(v), d))
        return r, int(v.value)

    def device_get_name(self, d):
        buf = ctypes.create_string_buffer(96)
        r = int(self.lib.cuDeviceGetName(buf, 96, d))
        return r, (buf.value.decode() if r == 0 else None)


@pytest.fixture(scope="module")
def B():
    return importlib.import_module("gpu-feature-discovery_b200.binding")


def oracle_labels(strategy="none"):
    from oracle import gfd_oracle as O
    cfg = O.Config(mig_strategy=strategy, no_timestamp=True, machine_type_file="")
    labels, _ = O.run_pass(O.CudaManager(RealCuda()), [], cfg)
    return labels


def test_cabi_cuda_backend_matches_the_driver(B):
    raw = RealCuda()
    assert raw.init()[0] == 0
    m = B.Manager(backend=B.BACKEND_CUDA)
    m.init()
    try:
        devs = m.get_devices()
        assert len(devs) == raw.device_get_count()[1] >= 1
        assert m.get_driver_version() == "unknown.unknown.unknown"                       # cuda-lib.go:66-68
        v = raw.driver_get_version()[1]
        assert m.get_cuda_driver_version() == (v // 1000, v % 100 // 10)                 # cuda-lib.go:59-60 (`% 100`)
        for i, d in enumerate(devs):
            assert d.get_name() == raw.device_get_name(i)[1] == "NVIDIA B200"
            assert d.get_cuda_compute_capability() == (raw.device_get_attribute(75, i)[1], raw.device_get_attribute(76, i)[1]) == (10, 0)
            legacy, full = raw.device_total_mem(i)[1], raw.device_total_mem_v2(i)[1]
            assert d.get_total_memory_mb() == legacy // (1024 * 1024)
            assert full > (1 << 32)
            # what the legacy symbol does with > 4 GiB: recorded so that fake_cuda.c / RawCuda model the real driver
            assert legacy in (full, full & 0xFFFFFFFF, min(full, 0xFFFFFFFF)), (legacy, full)
            assert d.is_mig_capable() is False and d.is_mig_enabled() is False           # cuda-device.go:60-68
    finally:
        m.shutdown()
        m.close()


def test_legacy_total_mem_behaviour_is_the_one_the_fake_models():
    """The fake libcuda and RawCuda model the legacy cuDeviceTotalMem as `what the real driver does`; this is that check."""
    from oracle import nvml_model as M
    raw = RealCuda()
    assert raw.init()[0] == 0
    legacy, full = raw.device_total_mem(0)[1], raw.device_total_mem_v2(0)[1]
    model = M.RawCuda([(full, 10, 0, "NVIDIA B200")]).device_total_mem(0)[1]
    assert model == legacy, "real driver: legacy=%d for total=%d; model says %d" % (legacy, full, model)


@pytest.mark.parametrize("strategy", ["none", "single"])
def test_labeler_binary_on_the_cuda_backend_matches_oracle(B, strategy, tmp_path):
    """NVML hidden (an unloadable libnvidia-ml.so.1 first on the library path) + a Tegra marker file: the factory picks the CUDA
    manager (factory.go:59-70) and the label file must be the oracle's."""
    from oracle import gfd_oracle as O
    hide = os.path.join(str(tmp_path), "hide")
    os.makedirs(hide)
    open(os.path.join(hide, "libnvidia-ml.so.1"), "w").write("not an ELF file\n")
    root = os.path.join(str(tmp_path), "root")
    os.makedirs(os.path.join(root, "etc"))
    open(os.path.join(root, "etc", "nv_tegra_release"), "w").write("# R35 (release), REVISION: 4.1\n")
    empty = os.path.join(str(tmp_path), "pci")
    os.makedirs(empty)
    out = os.path.join(str(tmp_path), "gfd")
    env = dict(os.environ, LD_LIBRARY_PATH=hide + ":" + os.environ.get("LD_LIBRARY_PATH", ""), GFD_B200_TEGRA_ROOT=root,
               GFD_B200_PCI_DEVICES_ROOT=empty)
    p = subprocess.run([B.BIN_PATH, "--oneshot", "--no-timestamp", "--machine-type-file=", "--mig-strategy=" + strategy, "-o", out],
                       capture_output=True, text=True, env=env, timeout=120)
    if "Detected NVML platform" in p.stderr:
        pytest.skip("the loader fell through to the system NVML: cannot hide it on this box")
    assert "Using CUDA manager" in p.stderr, p.stderr
    assert p.returncode == 0, p.stderr
    assert open(out).read() == O.render(oracle_labels(strategy))
    kv = dict(l.split("=", 1) for l in open(out).read().splitlines())
    assert kv["nvidia.com/gpu.product"] == "NVIDIA-B200" and kv["nvidia.com/cuda.driver.major"] == "unknown"
    assert kv["nvidia.com/gpu.family"] == "undefined" and kv["nvidia.com/mig.capable"] == "false"
