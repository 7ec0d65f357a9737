"""-m gpu: the labeling pass on the REAL NVML of the box.

  * product CLI file == the Python oracle fed with raw values harvested through pynvml + the real sysfs
  * product CLI file == the CPU baseline's file (oracle/ref_replay, the call-sequence-faithful C replay)
  * C-ABI getters == pynvml raw values
  * the opt-in flags add ONLY their own labels; the default label set never changes
Every comparison is against whatever state the node is in, so the same tests are the MIG-enabled parity tests when the box's
GPU is partitioned (tools/mig_hw_check.sh does that under a restore trap and runs this file; the `mig_*` tests below skip
unless a MIG device is visible).  MIG-enabled nodes are also covered by the fake-NVML fixtures in the CPU suite."""
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


def mig_state():
    """(some GPU has MIG mode enabled, number of MIG devices NVML shows) -- asked of pynvml, not of the product."""
    import pynvml
    pynvml.nvmlInit()
    try:
        enabled, n = False, 0
        for i in range(pynvml.nvmlDeviceGetCount()):
            h = pynvml.nvmlDeviceGetHandleByIndex(i)
            try:
                cur, _ = pynvml.nvmlDeviceGetMigMode(h)
            except pynvml.NVMLError:
                continue
            if cur != 1:
                continue
            enabled = True
            for j in range(pynvml.nvmlDeviceGetMaxMigDeviceCount(h)):
                try:
                    pynvml.nvmlDeviceGetMigDeviceHandleByIndex(h, j)
                    n += 1
                except pynvml.NVMLError:
                    pass
        return enabled, n
    finally:
        pynvml.nvmlShutdown()


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
        if not mig_state()[0]:
            assert "nvidia.com/gpu.product=NVIDIA-B200\n" in got
            assert "nvidia.com/gpu.family=undefined\n" in got      # the reference's table stops at hopper: parity, not a bug here
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
            if cur == 1:
                _check_mig_devices_against_pynvml(B, pynvml, h, d)
            assert d.is_mig_capable() is True and d.is_mig_enabled() == (cur == 1)
            uuid = pynvml.nvmlDeviceGetUUID(h)
            assert d.get_uuid() == (uuid.decode() if isinstance(uuid, bytes) else uuid)
        calls, probes = m.call_stats()
        m.shutdown()
        m.close()
        if not mig_state()[0]:
            assert calls <= 4 + 7 * len(devs)
    finally:
        pynvml.nvmlShutdown()


def _check_mig_devices_against_pynvml(B, nv, h, d):
    """Every MIG device the C ABI lists under GPU `d`: attributes, GI/CI-derived profile name, memory GB and parent handle
    against values read through pynvml and the oracle's arithmetic (mig_device.go:52-148, mig_profile.go:55-115,200-209,324-332,
    nvml-mig-device.go:35-72,96-118)."""
    from oracle import gfd_oracle as O
    total = int(nv.nvmlDeviceGetMemoryInfo(h).total)
    migs = d.get_mig_devices()
    want = []
    for j in range(nv.nvmlDeviceGetMaxMigDeviceCount(h)):
        try:
            want.append(nv.nvmlDeviceGetMigDeviceHandleByIndex(h, j))
        except nv.NVMLError:
            pass
    assert len(migs) == len(want)
    for md, mh in zip(migs, want):
        a = nv.nvmlDeviceGetAttributes_v2(mh)
        assert md.get_attributes() == {
            "memory": a.memorySizeMB, "multiprocessors": a.multiprocessorCount, "slices.gi": a.gpuInstanceSliceCount,
            "slices.ci": a.computeInstanceSliceCount, "engines.copy": a.sharedCopyEngineCount,
            "engines.decoder": a.sharedDecoderCount, "engines.encoder": a.sharedEncoderCount,
            "engines.jpeg": a.sharedJpegCount, "engines.ofa": a.sharedOfaCount}
        assert md.get_total_memory_mb() == a.memorySizeMB
        gb = O.get_mig_memory_size_gb(total, int(a.memorySizeMB))
        assert B.mig_memory_size_gb(total, int(a.memorySizeMB)) == gb
        gi = nv.nvmlDeviceGetGpuInstanceById(h, nv.nvmlDeviceGetGpuInstanceId(mh))
        gi_pid = int(nv.nvmlGpuInstanceGetInfo(gi).profileId)
        ci = nv.nvmlGpuInstanceGetComputeInstanceById(gi, nv.nvmlDeviceGetComputeInstanceId(mh))
        ci_pid = int((getattr(nv, "nvmlComputeInstanceGetInfo_v2", None) or nv.nvmlComputeInstanceGetInfo)(ci).profileId)
        if gi_pid <= 7 and ci_pid <= 6:  # what the reference's NVML 11.x header can name (V/go-nvml/nvml/const.go:538-575)
            name = md.get_name()
            assert name.endswith("%dgb" % gb) or name.endswith("%dgb.me" % gb), (name, gb)
            assert "+" not in name
            pname0 = nv.nvmlDeviceGetName(h)
            if a.gpuInstanceSliceCount == 1 and gi_pid == 0 and ci_pid == 0 and "B200" in (pname0.decode() if isinstance(pname0, bytes) else pname0):
                assert name == "1g.23gb" and gb == 23    # 1/8 of 180 GiB = 22.5, Go math.Round -> 23 (SURVEY Fact 4)
        else:
            with pytest.raises(B.GfdError):
                md.get_name()
        parent = md.get_device_handle_from_mig_device_handle()
        pname = nv.nvmlDeviceGetName(h)
        assert parent.get_name() == (pname.decode() if isinstance(pname, bytes) else pname)


def test_mig_enabled_single_and_mixed_label_sets(B, tmp_path):
    """BASELINE configs[3]/[4] on a real partitioned B200: the label SHAPES the reference's own MIG tests pin
    (cmd/gpu-feature-discovery/mig_test.go:70-72,122-125,240-243,300-305; tests/expected-output-mig-*.txt) on top of the
    exact diff against the oracle that test_cli_matches_oracle_on_real_nvml already makes."""
    import re
    enabled, n_mig = mig_state()
    if n_mig == 0:
        pytest.skip("no MIG device visible on this node")
    regs = json.load(open(os.path.join(ROOT, "tests", "golden", "expected_output_regexes.json")))["files"]
    for strategy, key in (("single", "expected-output-mig-single.txt"), ("mixed", "expected-output-mig-mixed.txt")):
        rc, got, err = run_cli(B, tmp_path, "--mig-strategy=" + strategy)
        want = oracle_text(strategy)
        if want is None:
            assert rc == 1, err
            continue
        assert rc == 0 and got == want, err
        kv = dict(l.split("=", 1) for l in got.splitlines())
        assert kv["nvidia.com/mig.strategy"] == strategy
        pats = [re.compile(p.strip()) for p in regs[key]]
        for line in got.splitlines():       # checkResult (main_test.go:403-435): every non-vgpu line matches one of the file's regexes
            if "vgpu" in line or re.match(r"nvidia\.com/mig-(\d+c\.)?\d+g\.\d+gb\.me\.|nvidia\.com/mig-\d+c\.", line):
                continue                    # the reference's regex files only know plain "<g>g.<gb>gb" resources
            assert any(p.search(line) for p in pats), line
        if strategy == "single" and not kv["nvidia.com/gpu.product"].endswith("MIG-INVALID"):
            assert int(kv["nvidia.com/gpu.count"]) == n_mig
            assert re.fullmatch(r"NVIDIA-B200-MIG-\d+(c\.\d+)?g\.\d+gb(\.me)?", kv["nvidia.com/gpu.product"]), kv["nvidia.com/gpu.product"]
            for k in ("memory", "multiprocessors", "slices.gi", "slices.ci", "engines.copy", "engines.decoder", "engines.encoder",
                      "engines.jpeg", "engines.ofa"):
                assert "nvidia.com/gpu." + k in kv
        if strategy == "mixed":
            counts = {k: int(v) for k, v in kv.items() if re.fullmatch(r"nvidia\.com/mig-.*\.count", k)}
            assert sum(counts.values()) == n_mig and counts


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


def test_probe_nvlink_adds_read_and_write_labels(B, tmp_path):
    """--probe-nvlink: with peers, the minimum over all directed pairs for peer loads AND peer stores (both verified inside the
    probe); on a single-GPU box the flag adds nothing.  The default label set never changes."""
    import torch
    rc, base, err = run_cli(B, tmp_path, "--mig-strategy=none")
    assert rc == 0, err
    rc, got, err = run_cli(B, tmp_path, "--mig-strategy=none", "--probe-nvlink", "--probe-size-mib=256")
    assert rc == 0, err
    extra = sorted(set(got.splitlines()) - set(base.splitlines()))
    assert set(base.splitlines()) <= set(got.splitlines())
    n = torch.cuda.device_count()
    if n < 2:
        assert extra == []
        return
    kv = dict(l.split("=") for l in extra)
    assert sorted(kv) == ["nvidia.com/gpu.probe.nvlink.p2p-dma-gbps", "nvidia.com/gpu.probe.nvlink.p2p-read-gbps",
                          "nvidia.com/gpu.probe.nvlink.p2p-write-gbps", "nvidia.com/gpu.probe.nvlink.pairs",
                          "nvidia.com/gpu.probe.nvlink.ring-gbps"]
    assert 50 < int(kv["nvidia.com/gpu.probe.nvlink.p2p-dma-gbps"]) < 2000
    assert int(kv["nvidia.com/gpu.probe.nvlink.pairs"]) == n * (n - 1)
    assert 50 < int(kv["nvidia.com/gpu.probe.nvlink.p2p-read-gbps"]) < 2000
    assert 50 < int(kv["nvidia.com/gpu.probe.nvlink.p2p-write-gbps"]) < 2000
