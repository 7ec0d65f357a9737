"""Experiment helper: does the one-shot exec->exit time depend on what the PARENT process holds (torch imported, a CUDA
context, a 4 GiB allocation, a just-terminated nvidia-smi poller)?  Prints mean / median / max of 30 runs per case."""
import os, statistics, subprocess, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "gpu-feature-discovery_b200", "gpu-feature-discovery")
REPLAY = os.path.join(ROOT, "oracle", "ref_replay")
CASES = {"product": [BIN, "--oneshot", "--no-timestamp", "--mig-strategy=none", "-o", "/tmp/ctx_o"],
         "replay": [REPLAY, "--mig-strategy", "none", "--no-timestamp", "-o", "/tmp/ctx_r"]}
def run(tag):
    for name, cmd in CASES.items():
        v = []
        for i in range(35):
            t0 = time.perf_counter()
            subprocess.run(cmd, capture_output=True)
            if i >= 5:
                v.append((time.perf_counter() - t0) * 1e3)
        print("%-28s %-8s mean %.2f median %.2f max %.2f" % (tag, name, sum(v) / len(v), statistics.median(v), max(v)), flush=True)
run("bare python")
import torch
run("torch imported")
torch.cuda.init(); torch.zeros(1, device="cuda"); torch.cuda.synchronize()
run("CUDA context")
buf = torch.empty(4 << 30, dtype=torch.uint8, device="cuda"); buf.zero_(); torch.cuda.synchronize()
run("CUDA context + 4 GiB")
p = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm", "--format=csv,noheader", "-lms", "100"], stdout=subprocess.DEVNULL)
time.sleep(1.0); p.terminate(); p.wait()
run("after nvidia-smi poller")
del buf; torch.cuda.empty_cache()
run("buffer freed")
