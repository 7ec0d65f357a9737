"""Experiment helper: one-shot exec->exit wall-clock of the labeler (default and serial walk) vs ref_replay.

profiles/r1c_e2e_breakdown_1gpu.txt was taken with an earlier build that still had the _exit() shortcut and an env knob
to disable it ("fast exit" / "with shutdown" rows); the shortcut measured no gain and is gone, so is the knob."""
import os, statistics, subprocess, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "gpu-feature-discovery_b200", "gpu-feature-discovery")
REPLAY = os.path.join(ROOT, "oracle", "ref_replay")
def run(cmd, env=None, n=25):
    v, inner = [], []
    for i in range(n + 3):
        t0 = time.perf_counter()
        p = subprocess.run(cmd, capture_output=True, text=True, env=env)
        dt = (time.perf_counter() - t0) * 1e3
        if i >= 3:
            v.append(dt)
            for l in p.stderr.splitlines():
                if "timing: pass" in l:
                    inner.append(float(l.split("timing: pass ")[1].split(" ms")[0]))
    return {"median": round(statistics.median(v), 2), "min": round(min(v), 2), "inner_median": round(statistics.median(inner), 2) if inner else None}
args = ["--oneshot", "--no-timestamp", "--timing", "-o", "/tmp/e2e_o"]
for rep in range(2):
    print("default     ", run([BIN] + args))
    print("ref_replay  ", run([REPLAY, "--no-timestamp", "-o", "/tmp/e2e_r"]))
    print("serial      ", run([BIN] + args + ["--prefetch-threads=1"]))
