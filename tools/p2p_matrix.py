"""Experiment helper: NVLink P2P probe for every directed GPU pair (read and copy), sequentially."""
import importlib, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
B = importlib.import_module("gpu-feature-discovery_b200.binding")
n = B.cuda_device_count()
out = {"n": n, "read": {}, "copy": {}}
for s in range(n):
    for d in range(n):
        if s == d:
            continue
        for mode, name in ((B.PROBE_READ, "read"), (B.PROBE_COPY, "copy")):
            try:
                r = B.probe_p2p(s, d, 1 << 30, warmup=2, iters=10, mode=mode)
                out[name]["%d->%d" % (s, d)] = round(r.gbps, 1)
            except B.GfdError as e:
                out[name]["%d->%d" % (s, d)] = "ERR %d" % e.status
print(json.dumps(out))
