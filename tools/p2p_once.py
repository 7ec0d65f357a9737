"""One directed GPU pair, one mode, two timed launches: the command ncu profiles for the NVLink byte counters
(profiles/r2_ncu_p2p_summary.md).  usage: python tools/p2p_once.py read|copy|push [MiB]"""
import importlib, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
B = importlib.import_module("gpu-feature-discovery_b200.binding")
mode = {"read": B.PROBE_READ, "copy": B.PROBE_COPY, "push": B.PROBE_PUSH}[sys.argv[1]]
mib = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
r = B.probe_p2p(0, 1, mib << 20, warmup=1, iters=2, mode=mode)  # src = GPU 0, dst = GPU 1 (READ/COPY run on 1, PUSH runs on 0)
print(json.dumps({"mode": sys.argv[1], "bytes": r.bytes, "gbps": round(r.gbps, 1), "verified": r.verified, "grid": r.grid}))
