"""-m gpu: the SURVEY section-8(f) rows that are built, on the REAL NVML of the box.

  (i)   --config-file with sharing.timeSlicing (V/spec/replicas.go:121-224, internal/lm/resource.go:159-226,
        cmd/gpu-feature-discovery/main.go:236-270): file == oracle with the same time-slicing entries
  (ii)  --current-tables: exactly one line differs from the default file on a B200 (gpu.family), the stale table being
        internal/lm/resource.go:261-284
  (iii) the daemon loop on real NVML (main.go:189-232; TestRunSleep, main_test.go:184-271): rewrites keep the timestamp,
        SIGHUP restarts with a new one, SIGTERM removes the file and gfd-tmp"""
import importlib
import json
import os
import signal
import subprocess
import time

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def B():
    return importlib.import_module("gpu-feature-discovery_b200.binding")


def run_cli(B, tmp_path, *args):
    out = os.path.join(str(tmp_path), "gfd")
    if os.path.exists(out):
        os.unlink(out)
    p = subprocess.run([B.BIN_PATH, "--oneshot", "--no-timestamp", "-o", out] + list(args), capture_output=True, text=True, timeout=300)
    return p.returncode, (open(out).read() if os.path.exists(out) else None), p.stderr


def oracle_text(strategy, time_slicing=()):
    from oracle import gfd_oracle as O, nvml_model as M
    cfg = O.Config(mig_strategy=strategy, no_timestamp=True,
                   time_slicing=[O.ReplicatedResource(n, r, rn) for n, r, rn in time_slicing])
    labels, _ = O.run_pass(O.NvmlManager(M.PynvmlRaw()), lambda: O.nvidia_pci_devices(), cfg)
    return O.render(labels)


def mig_enabled():
    import pynvml
    pynvml.nvmlInit()
    try:
        for i in range(pynvml.nvmlDeviceGetCount()):
            try:
                if pynvml.nvmlDeviceGetMigMode(pynvml.nvmlDeviceGetHandleByIndex(i))[0] == 1:
                    return True
            except pynvml.NVMLError:
                pass
        return False
    finally:
        pynvml.nvmlShutdown()


# (document, what disableResourceRenamingInConfig leaves of it, expected replicas, product gets the -SHARED suffix)
TS = [
    ("replicas2", "version: v1\nsharing:\n  timeSlicing:\n    resources:\n    - name: nvidia.com/gpu\n      replicas: 2\n",
     [("nvidia.com/gpu", 2, "")], 2, True),
    # an explicit rename is neutralised by main.go:246-249: the resource is NOT renamed, so the product keeps -SHARED
    ("rename_neutralised", "sharing:\n  timeSlicing:\n    resources:\n    - name: gpu\n      rename: gpu-shared\n      replicas: 4\n",
     [("nvidia.com/gpu", 4, "")], 4, True),
    # renameByDefault survives (main.go:250-253 forces the DEFAULT rename): a renamed resource carries no -SHARED suffix
    ("rename_by_default", "sharing:\n  timeSlicing:\n    renameByDefault: true\n    resources:\n    - name: nvidia.com/gpu\n      replicas: 3\n",
     [("nvidia.com/gpu", 3, "nvidia.com/gpu.shared")], 3, False),
    ("other_resource", "sharing:\n  timeSlicing:\n    resources:\n    - name: nvidia.com/not-gpu\n      replicas: 2\n",
     [("nvidia.com/not-gpu", 2, "")], 1, False),
    ("json", json.dumps({"version": "v1", "sharing": {"timeSlicing": {"resources": [{"name": "nvidia.com/gpu", "devices": "all", "replicas": 2}]}}}),
     [("nvidia.com/gpu", 2, "")], 2, True),
]


@pytest.mark.parametrize("name,doc,ts,replicas,shared", TS, ids=[t[0] for t in TS])
def test_config_file_time_slicing_on_real_nvml(B, tmp_path, name, doc, ts, replicas, shared):
    cfg = os.path.join(str(tmp_path), "config.json" if name == "json" else "config.yaml")
    open(cfg, "w").write(doc)
    rc, got, err = run_cli(B, tmp_path, "--mig-strategy=none", "--config-file", cfg)
    assert rc == 0, err
    assert got == oracle_text("none", ts)
    kv = dict(l.split("=", 1) for l in got.splitlines())
    if not mig_enabled():   # a MIG-enabled GPU is labelled without sharing (mig-strategy.go:157-165)
        assert kv["nvidia.com/gpu.replicas"] == str(replicas)
        assert kv["nvidia.com/gpu.product"] == ("NVIDIA-B200-SHARED" if shared else "NVIDIA-B200")
    # every other line is the one the default pass writes
    rc, base, err = run_cli(B, tmp_path, "--mig-strategy=none")
    assert rc == 0, err
    changed = {l.split("=")[0] for l in set(got.splitlines()) ^ set(base.splitlines())}
    assert changed <= {"nvidia.com/gpu.replicas", "nvidia.com/gpu.product"}


def test_config_file_replicas_below_two_is_rejected(B, tmp_path):
    cfg = os.path.join(str(tmp_path), "config.yaml")
    open(cfg, "w").write("sharing:\n  timeSlicing:\n    resources:\n    - name: nvidia.com/gpu\n      replicas: 1\n")
    rc, got, err = run_cli(B, tmp_path, "--config-file", cfg)
    assert rc == 1 and got is None
    assert "number of replicas must be >= 2" in err      # V/spec/replicas.go:209-211


def test_config_file_flags_section_and_cli_precedence(B, tmp_path):
    """flags.migStrategy from the file, overridden by the command line (V/spec/config.go:41-57)."""
    cfg = os.path.join(str(tmp_path), "config.yaml")
    open(cfg, "w").write("version: v1\nflags:\n  migStrategy: mixed\n  gfd:\n    noTimestamp: true\n    oneshot: true\n")
    out = os.path.join(str(tmp_path), "o")
    p = subprocess.run([B.BIN_PATH, "--config-file", cfg, "-o", out], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr
    assert open(out).read() == oracle_text("mixed")
    p = subprocess.run([B.BIN_PATH, "--config-file", cfg, "--mig-strategy=none", "-o", out], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr
    assert open(out).read() == oracle_text("none")


def test_current_tables_changes_only_the_family_line_on_b200(B, tmp_path):
    rc, base, err = run_cli(B, tmp_path, "--mig-strategy=none")
    assert rc == 0, err
    rc, cur, err = run_cli(B, tmp_path, "--mig-strategy=none", "--current-tables")
    assert rc == 0, err
    if mig_enabled():
        pytest.skip("MIG-enabled GPU carries no architecture labels")
    diff = sorted(set(base.splitlines()) ^ set(cur.splitlines()))
    assert diff == ["nvidia.com/gpu.family=blackwell", "nvidia.com/gpu.family=undefined"]
    assert "nvidia.com/gpu.family=undefined" in base.splitlines()    # default = the reference's stale table (parity)
    assert "nvidia.com/gpu.family=blackwell" in cur.splitlines()
    assert len(base.splitlines()) == len(cur.splitlines())


def _timestamp(path):
    for l in open(path).read().splitlines():
        if l.startswith("nvidia.com/gfd.timestamp="):
            return l.split("=")[1]
    return None


def _wait(pred, timeout=60.0):
    t0 = time.time()
    while time.time() - t0 < timeout:
        try:
            if pred():
                return True
        except (OSError, ValueError):
            pass
        time.sleep(0.02)
    return False


def test_daemon_sleep_loop_sighup_and_sigterm_on_real_nvml(B, tmp_path):
    d = os.path.join(str(tmp_path), "features.d")
    os.makedirs(d)
    out = os.path.join(d, "gfd")
    p = subprocess.Popen([B.BIN_PATH, "--sleep-interval=1s", "--mig-strategy=none", "-o", out], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    try:
        assert _wait(lambda: os.path.exists(out) and _timestamp(out))
        ts1, ino1 = _timestamp(out), os.stat(out).st_ino
        # the loop rewrites the file every interval (a new inode: write-temp-then-rename) with the SAME timestamp
        assert _wait(lambda: os.stat(out).st_ino != ino1, timeout=30), "no rewrite within 30 s"
        assert _timestamp(out) == ts1                                     # main.go:158 / main_test.go:266-267
        want = oracle_text("none")
        body = "".join(l + "\n" for l in open(out).read().splitlines() if not l.startswith("nvidia.com/gfd.timestamp="))
        assert body == want
        # SIGHUP: run() returns restart=true, the file is removed, start() rebuilds the manager and a NEW timestamp appears
        time.sleep(1.1)                                                   # the timestamp has one-second resolution
        p.send_signal(signal.SIGHUP)
        assert _wait(lambda: os.path.exists(out) and _timestamp(out) not in (None, ts1), timeout=60), "no new timestamp after SIGHUP"
        assert int(_timestamp(out)) > int(ts1)
        # SIGTERM: exit 0, output file and gfd-tmp removed (main.go:149-156,212-232)
        p.send_signal(signal.SIGTERM)
        _, err = p.communicate(timeout=60)
    finally:
        if p.poll() is None:
            p.kill()
            p.communicate()
    assert p.returncode == 0, err
    assert not os.path.exists(out) and not os.path.exists(os.path.join(d, "gfd-tmp"))
    assert "restarting" in err.lower() or "sighup" in err.lower(), err[-1500:]
    assert "Exiting" in err
