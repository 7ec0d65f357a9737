"""oracle/ref_replay (the CPU-baseline binary) must be call-sequence-faithful to the reference, i.e. to the
Python oracle whose raw-NVML trace IS the reference's call sequence: same NVML calls in the same order, same
number of dlopen probes, same label file -- on every fixture (CPU only, fake NVML)."""
import json
import os
import subprocess

import pytest

import nodes as N
from oracle import gfd_oracle as O, nvml_model as M
from shimutil import B, ROOT, use_node

NODES = N.nodes()
REPLAY = os.path.join(ROOT, "oracle", "ref_replay")
CASES = [(n, s) for n in sorted(NODES) for s in ("none", "single", "mixed") if not NODES[n].hidden]


def run_replay(tmp_path, node, strategy, passes=1):
    fix = use_node(node, tmp_path)
    out = os.path.join(str(tmp_path), "replay.out")
    trace = os.path.join(str(tmp_path), "trace.txt")
    empty = os.path.join(str(tmp_path), "pci-empty")
    os.makedirs(empty, exist_ok=True)
    mt = os.path.join(str(tmp_path), "product_name")
    open(mt, "w").write("DGX B200\n")
    env = dict(os.environ, GFD_FAKE_NVML=fix, GFD_FAKE_NVML_TRACE=trace,
               LD_LIBRARY_PATH=B.FAKE_NVML_DIR + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
    p = subprocess.run([REPLAY, "--mig-strategy", strategy, "--no-timestamp", "--machine-type-file", mt, "-o", out, "--pci-root", empty,
                        "--passes", str(passes)], env=env, capture_output=True, timeout=60)
    calls = [l for l in open(trace).read().split() if l != "nvmlErrorString"] if os.path.exists(trace) else []
    text = open(out).read() if os.path.exists(out) else None
    return p.returncode, text, calls, p.stdout.decode(), p.stderr.decode()


@pytest.mark.parametrize("name,strategy", CASES, ids=["%s-%s" % c for c in CASES])
def test_replay_matches_oracle_trace_and_labels(name, strategy, tmp_path):
    node = NODES[name]
    raw = M.RawNvml(node)
    cfg = O.Config(mig_strategy=strategy, no_timestamp=True, machine_type_file="/x")
    try:
        labels, _ = O.run_pass(O.NvmlManager(raw), [], cfg, read_file=lambda p: b"DGX B200\n")
        want = O.render(labels)
    except O.GfdError:
        want = None
    rc, got, calls, so, se = run_replay(tmp_path, node, strategy)
    ref_calls = [t for t in raw.trace if not t.startswith("dlsym:") and t != "nvmlErrorString"]
    ref_probes = sum(1 for t in raw.trace if t.startswith("dlsym:"))
    if want is None:
        assert rc == 1 and got is None, se
        return
    assert rc == 0, se
    assert got == want
    assert calls == ref_calls
    stats = json.loads(so)
    assert stats["nvml_calls"] == len(ref_calls) and stats["dl_cycles"] == ref_probes


def test_replay_multiple_passes_reinit_every_pass(tmp_path):
    rc, got, calls, so, se = run_replay(tmp_path, NODES["c3_8xB200"], "none", passes=3)
    assert rc == 0, se
    assert calls.count("nvmlInit_v2") == 3 and calls.count("nvmlShutdown") == 3 and len(calls) == 3 * 92
    assert len(json.loads(so)["ms"]) == 3
