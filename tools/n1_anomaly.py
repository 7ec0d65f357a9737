"""Experiment helper: on a multi-GPU box, does the labeling pass time depend on what the PARENT holds (nothing / a CUDA
context on GPU 0, as bench.py --gpus 1 does) or is the spread host noise?  Prints per-pass ms, phases and /proc/loadavg."""
import os, statistics, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "gpu-feature-discovery_b200", "gpu-feature-discovery")
def passes(tag, extra=()):
    p = subprocess.run([BIN, "--passes", "13", "--sleep-interval", "0s", "--timing", "--no-timestamp", "-o", "/tmp/n1_o"] + list(extra), capture_output=True, text=True)
    ms, ph = [], []
    for l in p.stderr.splitlines():
        if "timing: pass" in l:
            ms.append(float(l.split("timing: pass ")[1].split(" ms")[0]))
            ph.append(l.split("[")[1].rstrip("]"))
    ms = ms[3:]
    print("%-34s load %s affinity %d | median %.1f min %.1f max %.1f | last: %s" % (tag, open("/proc/loadavg").read().split()[0], len(os.sched_getaffinity(0)),
          statistics.median(ms), min(ms), max(ms), ph[-1]), flush=True)
for rep in range(2):
    passes("bare parent, threads auto")
    passes("bare parent, serial", ["--prefetch-threads=1"])
import torch
torch.cuda.set_device(0); torch.zeros(1, device="cuda"); torch.cuda.synchronize()
for rep in range(3):
    passes("parent holds ctx on GPU0, auto")
    passes("parent holds ctx on GPU0, serial", ["--prefetch-threads=1"])
ctx = [torch.zeros(1, device="cuda:%d" % d) for d in range(torch.cuda.device_count())]
torch.cuda.synchronize()
for rep in range(2):
    passes("parent holds ctx on ALL GPUs, auto")
    passes("parent holds ctx on ALL GPUs, serial", ["--prefetch-threads=1"])
