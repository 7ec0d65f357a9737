"""The multi-GPU property of the labeling pass, checked without GPUs: the fake NVML can charge a latency per
device-scoped call (GFD_FAKE_NVML_LATENCY_US -- on real B200s the first handle per GPU costs ~8 ms of driver attach), and
with it the per-GPU host threads must (a) shorten the pass roughly by the number of GPUs and (b) change no byte of output.
The real numbers are in profiles/ (8 x B200: 157 ms serial -> 103 ms with threads; nvmlInit_v2 itself cannot be spread)."""
import os
import re

import nodes as N
from shimutil import run_cli, use_node


def timed_passes(tmp_path, node, threads, strategy="none", latency_us=3000):
    fix = use_node(N.nodes()[node], str(tmp_path))
    empty = os.path.join(str(tmp_path), "pci")
    os.makedirs(empty, exist_ok=True)
    out = os.path.join(str(tmp_path), "out-%d" % threads)
    rc, _, se = run_cli(["--passes", "4", "--sleep-interval", "0s", "--timing", "--no-timestamp", "--mig-strategy=" + strategy,
                         "--prefetch-threads=%d" % threads, "-o", out],
                        env={"GFD_FAKE_NVML": fix, "GFD_B200_PCI_DEVICES_ROOT": empty, "GFD_FAKE_NVML_LATENCY_US": str(latency_us)})
    assert rc == 0, se
    ms = [float(x) for x in re.findall(r"timing: pass ([0-9.]+) ms", se)]
    assert len(ms) == 4
    return min(ms), open(out).read()


def test_threads_spread_the_per_gpu_latency_and_keep_the_output(tmp_path):
    serial, text1 = timed_passes(tmp_path, "c3_8xB200", 1)
    threaded, text8 = timed_passes(tmp_path, "c3_8xB200", 0)
    assert text1 == text8
    # serial: 8 GPUs x (handle, name, MIG mode) + memory and cc of the one representative GPU = 26 calls x 3 ms = 78 ms;
    # one thread per GPU: 5 calls x 3 ms = 15 ms (+ whatever the host adds)
    assert serial > 70, serial
    assert threaded < serial / 3, (serial, threaded)


def test_threads_also_spread_the_mig_walk(tmp_path):
    serial, text1 = timed_passes(tmp_path, "c4_single_8x7x1g23", 1, strategy="single", latency_us=300)
    threaded, text8 = timed_passes(tmp_path, "c4_single_8x7x1g23", 0, strategy="single", latency_us=300)
    assert text1 == text8 and "nvidia.com/gpu.count=56\n" in text1
    assert threaded < serial / 2.5, (serial, threaded)


def test_single_gpu_node_takes_the_serial_path(tmp_path):
    serial, text1 = timed_passes(tmp_path, "c2_1xB200", 1, latency_us=1000)
    auto, text0 = timed_passes(tmp_path, "c2_1xB200", 0, latency_us=1000)
    assert text1 == text0
    assert abs(auto - serial) < max(3.0, 0.5 * serial), (serial, auto)  # nothing to spread: same cost, no thread overhead to speak of


def test_pass_survives_a_host_that_refuses_new_threads(tmp_path):
    """A pod with a pids limit can refuse pthread_create (EAGAIN).  The per-GPU helpers, the node-level query thread and the vGPU
    scan thread are all optional: with every thread creation failing the pass must still produce the same file, serially."""
    import subprocess
    from test_cli_parity import NODES, cli_pass, oracle_pass
    shim_c = os.path.join(str(tmp_path), "nothreads.c")
    shim_so = os.path.join(str(tmp_path), "nothreads.so")
    open(shim_c, "w").write("#include <errno.h>\n#include <pthread.h>\n"
                            "int pthread_create(pthread_t *t, const pthread_attr_t *a, void *(*f)(void *), void *arg) {\n"
                            "    (void)t; (void)a; (void)f; (void)arg; return EAGAIN;\n}\n")
    subprocess.run(["gcc", "-shared", "-fPIC", "-o", shim_so, shim_c], check=True)
    for name, strategy in (("c3_8xB200", "none"), ("c4_single_8x7x1g23", "single"), ("c5_mixed_hetero", "mixed")):
        want, _ = oracle_pass(NODES[name], strategy)
        rc, got, stderr, _ = cli_pass(tmp_path, NODES[name], strategy, extra_env={"LD_PRELOAD": shim_so})
        assert rc == 0, stderr
        assert got == want
