"""The plain-C consumer of the C ABI (examples/cabi_demo.c) runs against the committed `.nvmlfix` fixtures, and those
fixtures are what tests/nodes.py renders today."""
import os
import subprocess

import nodes as N
from shimutil import B, ROOT

FIX = os.path.join(ROOT, "tests", "fixtures")


def test_committed_fixtures_are_current():
    ns = N.nodes()
    for fn in sorted(os.listdir(FIX)):
        if fn.endswith(".nvmlfix"):
            body = open(os.path.join(FIX, fn)).read().split("\n", 1)[1]
            assert body == ns[fn[:-len(".nvmlfix")]].dumps(), "%s is stale: run tests/fixtures/generate.py" % fn


def test_c_consumer_walks_a_mixed_mig_node():
    env = dict(os.environ, GFD_FAKE_NVML=os.path.join(FIX, "c5_mixed_hetero.nvmlfix"),
               LD_LIBRARY_PATH=B.FAKE_NVML_DIR + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
    p = subprocess.run([os.path.join(ROOT, "examples", "cabi_demo")], env=env, capture_output=True, text=True, timeout=60)
    assert p.returncode == 0, p.stderr
    assert "driver 575.57.08, cuda 12.9, 8 device(s)" in p.stdout
    assert "  mig 0: 3g.90gb, 91648 MiB, 54 SMs, slices gi=3 ci=3, parent NVIDIA B200" in p.stdout   # gfd_dev_parent
    assert "  mig 0: 1g.23gb.me, 22912 MiB" in p.stdout and "  mig 0: 7g.180gb, 182272 MiB" in p.stdout
    assert p.stdout.count("mig capable=1 enabled=0") == 4


def test_c_consumer_reports_unresolvable_profile():
    env = dict(os.environ, GFD_FAKE_NVML=os.path.join(FIX, "gi_profile_id_9_fails.nvmlfix"),
               LD_LIBRARY_PATH=B.FAKE_NVML_DIR + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
    p = subprocess.run([os.path.join(ROOT, "examples", "cabi_demo")], env=env, capture_output=True, text=True, timeout=60)
    assert p.returncode == 0 and "profile unresolved: no matching profile IDs found" in p.stdout
