"""Experiment helper (not product): does a labeling pass cost the same when a binary's processes run back to back as when
two different NVML clients alternate?  Prints medians + phase medians for: replay alone, product alone, alternating chunks,
replay alternating with a trivial third NVML client (pynvml init/shutdown)."""
import os, statistics, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
BIN = os.path.join(ROOT, "gpu-feature-discovery_b200", "gpu-feature-discovery")
env = bench.bench_env(1)
out_p, out_r = "/tmp/exp_p", "/tmp/exp_r"
prod = lambda n: bench.product_passes(BIN, n, out_p, env)
repl = lambda n: bench.replay_passes(n, out_r, env)[0]
def third(n):
    subprocess.run([sys.executable, "-c", "import pynvml\nfor _ in range(%d):\n pynvml.nvmlInit(); h=pynvml.nvmlDeviceGetHandleByIndex(0); pynvml.nvmlDeviceGetName(h); pynvml.nvmlShutdown()" % n], check=True)
    return [(0.0, {})] * n
def show(tag, rows):
    st, ph = bench.stats(rows)
    print("%-44s median %.2f  min %.2f  max %.2f  phases %s" % (tag, st["median"], st["min"], st["max"], ph), flush=True)
for rep in range(2):
    a, _ = bench.series(3, 20, repl); show("replay alone", a)
    a, _ = bench.series(3, 20, prod); show("product alone", a)
    a, b = bench.series(3, 20, prod, repl); show("product, alternating with replay", a); show("replay, alternating with product", b)
    a, b = bench.series(3, 20, repl, third); show("replay, alternating with a pynvml client", a)
    a, b = bench.series(3, 20, prod, third); show("product, alternating with a pynvml client", a)
    a, b = bench.series(3, 20, repl, repl); show("replay, alternating with replay (A)", a); show("replay, alternating with replay (B)", b)
