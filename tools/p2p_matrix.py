"""Experiment helper: NVLink P2P probes -- sequential per directed pair, then the concurrent tournament matrix."""
import importlib, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
B = importlib.import_module("gpu-feature-discovery_b200.binding")
n = B.cuda_device_count()
out = {"n": n, "read": {}, "copy": {}, "push": {}}
t0 = time.time()
for s in range(0 if "--concurrent-only" not in sys.argv else n, n):
    for d in range(n):
        if s == d:
            continue
        for mode, name in ((B.PROBE_READ, "read"), (B.PROBE_COPY, "copy"), (B.PROBE_PUSH, "push")):
            try:
                r = B.probe_p2p(s, d, 1 << 30, warmup=2, iters=10, mode=mode)
                out[name]["%d->%d" % (s, d)] = round(r.gbps, 1)
            except B.GfdError as e:
                out[name]["%d->%d" % (s, d)] = "ERR %d" % e.status
out["sequential_seconds"] = round(time.time() - t0, 2)
for mode, name in ((B.PROBE_READ, "read"), (B.PROBE_COPY, "copy"), (B.PROBE_PUSH, "push")):
    t0 = time.time()
    m, ring = B.probe_p2p_matrix(n, 1 << 30, warmup=2, iters=10, mode=mode)
    out["concurrent_" + name] = {"matrix_row_puller_col_source": [[round(v, 1) for v in row] for row in m], "ring_aggregate_gbps": round(ring, 1),
                                 "seconds": round(time.time() - t0, 2)}
print(json.dumps(out))
