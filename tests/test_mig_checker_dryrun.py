"""CPU: the MIG-enabled branches of the `-m gpu` label tests, run against SIMULATED partitioned nodes.

The pool's B200s refuse `nvidia-smi -mig 1` (profiles/r2_mig_refused/), so on the GPU box tests/test_gpu_labels.py only ever sees a
MIG-disabled GPU and its MIG-enabled assertions skip.  Here the same test functions run in a child pytest whose pynvml (NVIDIA's own
Python binding, with its own ctypes struct layouts) and whose product binary both load the fake libnvidia-ml.so.1: product file vs
oracle(pynvml), vs the C replay, C-ABI MIG getters vs pynvml, and the reference's regex goldens -- so the checker that waits for a
partitioned B200 is exercised code, and the product's reading of the NVML structs is cross-checked against pynvml's."""
import importlib
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FIX = os.path.join(ROOT, "tests", "fixtures")
B = importlib.import_module("gpu-feature-discovery_b200.binding")


@pytest.mark.parametrize("fixture,expect_mig_test", [
    ("c4_single_8x7x1g23", "passed"),        # BASELINE configs[3]
    ("c5_mixed_hetero", "passed"),           # BASELINE configs[4]
    ("c4_invalid_one_empty", "passed"),
    ("gi_profile_id_9_fails", "passed"),     # reference fails the pass: product must too
    ("a100_mixed", "passed"),
    ("c2_1xB200", "skipped"),                # no MIG device: exactly what the GPU box shows
])
def test_gpu_label_tests_over_simulated_mig_node(fixture, expect_mig_test, tmp_path):
    env = dict(os.environ, GFD_FAKE_NVML=os.path.join(FIX, fixture + ".nvmlfix"),
               LD_LIBRARY_PATH=B.FAKE_NVML_DIR + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
    p = subprocess.run([sys.executable, "-m", "pytest", os.path.join(ROOT, "tests", "test_gpu_labels.py"), "-q", "-x", "-rs",
                        "-p", "no:cacheprovider", "-k", "matches or getters or mig_enabled"],
                       capture_output=True, text=True, env=env, cwd=ROOT, timeout=600)
    assert p.returncode == 0, p.stdout[-4000:] + p.stderr[-2000:]
    tail = p.stdout.strip().splitlines()[-1]
    if expect_mig_test == "skipped":
        assert "1 skipped" in tail and "7 passed" in tail, tail
    else:
        assert "8 passed" in tail and "skipped" not in tail, tail


def test_gpu_f_row_tests_over_simulated_b200(tmp_path):
    """tests/test_gpu_f_rows.py (config file + time-slicing, --current-tables, daemon loop / SIGHUP / SIGTERM) dry-run the same way."""
    env = dict(os.environ, GFD_FAKE_NVML=os.path.join(FIX, "c2_1xB200.nvmlfix"),
               LD_LIBRARY_PATH=B.FAKE_NVML_DIR + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
    p = subprocess.run([sys.executable, "-m", "pytest", os.path.join(ROOT, "tests", "test_gpu_f_rows.py"), "-q", "-x", "-p", "no:cacheprovider"],
                       capture_output=True, text=True, env=env, cwd=ROOT, timeout=600)
    assert p.returncode == 0, p.stdout[-4000:] + p.stderr[-2000:]
    assert "9 passed" in p.stdout.strip().splitlines()[-1]
