"""Experiment helper (not product, not a bench): print probe bandwidth for a few sizes/modes.
Variant knobs are read by the library from the environment once per process, so run one process per variant."""
import importlib, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
B = importlib.import_module("gpu-feature-discovery_b200.binding")
out = {"env": {k: v for k, v in os.environ.items() if k.startswith("GFD_B200_")}}
for mode, name in ((B.PROBE_READ, "read"), (B.PROBE_COPY, "copy")):
    for gib in (1, 4):
        r = B.probe_hbm(0, gib << 30, warmup=3, iters=20, mode=mode)
        out["%s_%dGiB" % (name, gib)] = {"gbps": round(r.gbps, 1), "ms": round(r.ms_per_launch, 4), "grid": r.grid}
print(json.dumps(out))
