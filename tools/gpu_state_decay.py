"""Experiment helper (not product): how long after GPU activity does a labeling pass stay cheaper?  Runs a burst of the HBM probe
(in a child process that exits, so no context stays behind), then times replay passes at increasing delays."""
import os, statistics, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
env = bench.bench_env(1)
def passes(n=6):
    rows, _ = bench.replay_passes(n, "/tmp/decay_r", env)
    return statistics.median(r[0] for r in rows[1:])
def burst():
    subprocess.run([sys.executable, os.path.join(ROOT, "tools", "probe_once.py"), str(1 << 30)], check=True, capture_output=True)
print("idle box, before any GPU work: %.2f %.2f %.2f" % (passes(), passes(), passes()), flush=True)
for rep in range(2):
    burst()
    t0 = time.time()
    line = []
    for delay in (0, 1, 2, 4, 6, 9, 12, 16, 20, 25, 30, 40):
        while time.time() - t0 < delay:
            time.sleep(0.05)
        line.append("t+%ds %.2f" % (delay, passes()))
    print("after a burst: " + "  ".join(line), flush=True)
