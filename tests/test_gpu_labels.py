"""-m gpu: the labeling pass on the REAL NVML of the box.

  * product CLI file == the Python oracle fed with raw values harvested through pynvml + the real sysfs
  * product CLI file == the CPU baseline's file (oracle/ref_replay, the call-sequence-faithful C replay)
  * C-ABI getters == pynvml raw values
  * the opt-in flags add ONLY their own labels; the default label set never changes
(MIG cannot be enabled on the shared box, so single/mixed run on a MIG-disabled GPU here; MIG-enabled nodes
are covered by the fake-NVML fixtures in the CPU suite.)"""
import importlib
import json
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def B():
    return importlib.import_module("gpu-feature-discovery_b200.binding")


def run_cli(B, tmp_path, *args):
    out = os.path.join(str(tmp_path), "gfd")
    p = subprocess.run([B.BIN_PATH, "--oneshot", "--no-timestamp", "-o", out] + list(args), capture_output=True, text=True, timeout=300)
    return p.returncode, (open(out).read() if os.path.exists(out) else None), p.stderr


def oracle_text(strategy):
    from oracle import gfd_oracle as O, nvml_model as M
    cfg = O.Config(mig_strategy=strategy, no_timestamp=True)
    try:
        labels, _ = O.run_pass(O.NvmlManager(M.PynvmlRaw()), lambda: O.nvidia_pci_devices(), cfg)
        return O.render(labels)
    except O.GfdError:
        return None


@pytest.mark.parametrize("strategy", ["none", "single", "mixed"])
def test_cli_matches_oracle_on_real_nvml(B, strategy, tmp_path):
    want = oracle_text(strategy)
    rc, got, err = run_cli(B, tmp_path, "--mig-strategy=" + strategy)
    if want is None:
        assert rc == 1 and got is None, err
    else:
        assert rc == 0, err
        assert got == want
        assert "nvidia.com/gpu.product=NVIDIA-B200\n" in got
        assert "nvidia.com/gpu.family=undefined\n" in got          # the reference's table stops at hopper: parity, not a bug here
        assert "nvidia.com/gpu.compute.major=10\n" in got


@pytest.mark.parametrize("strategy", ["none", "single", "mixed"])
def test_cli_matches_cpu_baseline_replay(B, strategy, tmp_path):
    out = os.path.join(str(tmp_path), "replay")
    p = subprocess.run([os.path.join(ROOT, "oracle", "ref_replay"), "--mig-strategy", strategy, "--no-timestamp", "-o", out],
                       capture_output=True, text=True, timeout=300)
    rc, got, err = run_cli(B, tmp_path, "--mig-strategy=" + strategy)
    if p.returncode != 0:
        assert rc == 1, err
        return
    assert rc == 0, err
    assert got == open(out).read()
    stats = json.loads(p.stdout.strip().splitlines()[-1])
    assert stats["nvml_calls"] >= 20


def test_cabi_getters_match_pynvml(B):
    import pynvml
    pynvml.nvmlInit()
    try:
        m = B.Manager()
        m.init()
        devs = m.get_devices()
        assert len(devs) == pynvml.nvmlDeviceGetCount()
        drv = pynvml.nvmlSystemGetDriverVersion()
        assert m.get_driver_version() == (drv.decode() if isinstance(drv, bytes) else drv)
        v = pynvml.nvmlSystemGetCudaDriverVersion()
        assert m.get_cuda_driver_version() == (v // 1000, v % 1000 // 10)
        for i, d in enumerate(devs):
            h = pynvml.nvmlDeviceGetHandleByIndex(i)
            name = pynvml.nvmlDeviceGetName(h)
            assert d.get_name() == (name.decode() if isinstance(name, bytes) else name)
            assert d.get_total_memory_mb() == pynvml.nvmlDeviceGetMemoryInfo(h).total // (1024 * 1024)
            assert d.get_cuda_compute_capability() == tuple(pynvml.nvmlDeviceGetCudaComputeCapability(h))
            cur, _ = pynvml.nvmlDeviceGetMigMode(h)
            assert d.is_mig_capable() is True and d.is_mig_enabled() == (cur == 1)
            uuid = pynvml.nvmlDeviceGetUUID(h)
            assert d.get_uuid() == (uuid.decode() if isinstance(uuid, bytes) else uuid)
        calls, probes = m.call_stats()
        m.shutdown()
        m.close()
        assert calls <= 4 + 7 * len(devs)
    finally:
        pynvml.nvmlShutdown()


def test_opt_in_flags_only_add_their_own_labels(B, tmp_path):
    rc, base, err = run_cli(B, tmp_path, "--mig-strategy=none")
    assert rc == 0, err
    rc, got, err = run_cli(B, tmp_path, "--mig-strategy=none", "--probe-hbm", "--probe-size-mib=1024", "--prefetch-threads=4",
                           "--keep-nvml-attached", "--timing")
    assert rc == 0, err
    base_lines, got_lines = set(base.splitlines()), set(got.splitlines())
    assert base_lines <= got_lines
    extra = sorted(got_lines - base_lines)
    assert [l.split("=")[0] for l in extra] == ["nvidia.com/gpu.probe.hbm.copy-gbps", "nvidia.com/gpu.probe.hbm.read-gbps",
                                                "nvidia.com/gpu.probe.hbm.size-mib", "nvidia.com/gpu.probe.hbm.write-gbps"]
    vals = dict(l.split("=") for l in extra)
    assert 1000 < int(vals["nvidia.com/gpu.probe.hbm.read-gbps"]) < 20000
    assert "timing: pass" in err


def test_probe_size_backs_off_when_hbm_is_short(B, tmp_path):
    """--probe-size-mib larger than HBM: the labeler halves the buffer until it fits (down to 256 MiB) and says what it used."""
    rc, got, err = run_cli(B, tmp_path, "--mig-strategy=none", "--probe-hbm", "--probe-size-mib=262144")  # 256 GiB > 180 GB
    assert rc == 0, err
    vals = dict(l.split("=") for l in got.splitlines())
    assert 256 <= int(vals["nvidia.com/gpu.probe.hbm.size-mib"]) < 262144
    assert 1000 < int(vals["nvidia.com/gpu.probe.hbm.read-gbps"]) < 20000
    assert 1000 < int(vals["nvidia.com/gpu.probe.hbm.copy-gbps"]) < 20000
    assert 1000 < int(vals["nvidia.com/gpu.probe.hbm.write-gbps"]) < 20000


def test_keep_attached_multi_pass_same_file(B, tmp_path):
    rc, base, err = run_cli(B, tmp_path, "--mig-strategy=none")
    out = os.path.join(str(tmp_path), "multi")
    p = subprocess.run([B.BIN_PATH, "--passes", "4", "--sleep-interval", "0s", "--no-timestamp", "--keep-nvml-attached", "--timing", "-o", out],
                       capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr
    assert open(out).read() == base
    assert p.stderr.count("timing: pass") == 4


def test_uuid_maps_to_cuda_ordinal(B):
    import ctypes
    m = B.Manager()
    m.init()
    uuid = m.get_devices()[0].get_uuid()
    m.shutdown()
    m.close()
    o = ctypes.c_int(-1)
    assert B.lib.gfd_cuda_device_by_uuid(uuid.encode(), ctypes.byref(o)) == 0 and o.value >= 0
