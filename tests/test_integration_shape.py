"""The reference's docker integration test (tests/integration-tests.py:12-33,36-78) without docker: run the daemon with
its default life-cycle (no --oneshot), wait for the features.d file, then apply the SAME bijective check -- every
expected regex consumed by exactly one label (re.match, i.e. anchored at the start), no label left over -- against
the reference's tests/expected-output.txt (15 labels, MIG strategy none).  CPU: fake NVML; GPU: real NVML."""
import os
import re
import signal

import pytest

from shimutil import B
import test_cli_flags_and_lifecycle as L

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def check_labels(expected_regexs, labels):
    """check_labels (tests/integration-tests.py:19-33), same algorithm."""
    expected_regexs, labels = list(expected_regexs), list(labels)
    for label in labels[:]:
        for rx in expected_regexs[:]:
            if rx.match(label):
                expected_regexs.remove(rx)
                labels.remove(label)
                break
    return [r.pattern for r in expected_regexs], labels


def expected(name="expected-output.txt"):
    import json
    return [re.compile(l.strip()) for l in json.load(open(os.path.join(GOLDEN, "expected_output_regexes.json")))["files"][name]]


def run_daemon_and_collect(tmp_path, spawn):
    out = os.path.join(str(tmp_path), "features.d", "gfd")
    os.makedirs(os.path.dirname(out))
    p = spawn(out)
    try:
        assert L.wait_for(out, timeout=60)
        content = [x.strip() for x in open(out).readlines()]
    finally:
        p.send_signal(signal.SIGTERM)
        p.communicate(timeout=30)
    return content


def test_bijective_match_on_simulated_b200(tmp_path):
    content = run_daemon_and_collect(tmp_path, lambda out: L.spawn(tmp_path, ["-o", out, "--machine-type-file", "/dev/null"], node="c3_8xB200"))
    missing, unexpected = check_labels(expected(), content)
    assert missing == [] and unexpected == [], (missing, unexpected)
    assert len(content) == 15


@pytest.mark.gpu
def test_bijective_match_on_real_nvml(tmp_path):
    import subprocess

    def spawn(out):
        return subprocess.Popen([B.BIN_PATH, "-o", out], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    content = run_daemon_and_collect(tmp_path, spawn)
    missing, unexpected = check_labels(expected(), content)
    assert missing == [] and unexpected == [], (missing, unexpected)
