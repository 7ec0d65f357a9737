"""us per gfd_hbm_read_checksum call at small and large sizes, for two builds of the library side by side:
the round-1 shape (cudaMemsetAsync + fold kernel + 1-CTA finishing kernel; build/r1/libgfd_b200_r1.so, built from
`git show <r1>:.../gfd_probe.cu`) and the current one (ONE kernel, last CTA finishes).  CUDA events on the launching stream,
CALLS back-to-back calls after warm-up; the result of every size is also compared between the two builds.

usage: python tools/checksum_latency.py [--old build/r1/libgfd_b200_r1.so] [--calls 200]"""
import argparse
import ctypes
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load(path):
    lib = ctypes.CDLL(path)
    vp, sz = ctypes.c_void_p, ctypes.c_size_t
    lib.gfd_hbm_read_checksum.argtypes = [vp, sz, vp, vp, sz, vp]
    lib.gfd_hbm_read_checksum.restype = ctypes.c_int
    lib.gfd_hbm_read_checksum_noalloc.argtypes = [vp, sz, vp, vp]
    lib.gfd_hbm_read_checksum_noalloc.restype = ctypes.c_int
    lib.gfd_hbm_scratch_bytes.restype = sz
    return lib


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--old", default=os.path.join(ROOT, "build", "r1", "libgfd_b200_r1.so"))
    ap.add_argument("--new", default=os.path.join(ROOT, "gpu-feature-discovery_b200", "libgfd_b200.so"))
    ap.add_argument("--calls", type=int, default=200)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    libs = [("r1_memset+fold+finish", load(a.old)), ("r2_single_launch", load(a.new))] if os.path.exists(a.old) else [("r2_single_launch", load(a.new))]
    sizes = [64 << 10, 256 << 10, 1 << 20, 4 << 20, 8 << 20, 16 << 20, 64 << 20, 1 << 30, 4 << 30]
    big = torch.empty(max(sizes), dtype=torch.uint8, device=dev)
    new = libs[-1][1]
    new.gfd_hbm_fill_pattern.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_uint64, ctypes.c_void_p]
    assert new.gfd_hbm_fill_pattern(big.data_ptr(), big.numel(), 3, torch.cuda.current_stream().cuda_stream) == 0
    torch.cuda.synchronize()
    out = torch.zeros(2, dtype=torch.int64, device=dev)
    st = torch.cuda.current_stream()
    rows = []
    for n in sizes:
        row = {"bytes": n}
        sums = []
        for name, lib in libs:
            scratch = torch.zeros(int(lib.gfd_hbm_scratch_bytes()), dtype=torch.uint8, device=dev)  # zeroed: valid for both builds
            for entry in ("scratch", "noalloc"):
                def call():
                    if entry == "scratch":
                        rc = lib.gfd_hbm_read_checksum(big.data_ptr(), n, out.data_ptr(), scratch.data_ptr(), scratch.numel(), st.cuda_stream)
                    else:
                        rc = lib.gfd_hbm_read_checksum_noalloc(big.data_ptr(), n, out.data_ptr(), st.cuda_stream)
                    assert rc == 0, rc
                calls = a.calls if n <= (64 << 20) else 20
                for _ in range(10):
                    call()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(st)
                for _ in range(calls):
                    call()
                e1.record(st)
                torch.cuda.synchronize()
                us = e0.elapsed_time(e1) * 1e3 / calls
                row["%s/%s_us" % (name, entry)] = round(us, 2)
                sums.append(tuple(int(v) for v in out.cpu().numpy().view(np.uint64)))
        row["all_builds_agree"] = len(set(sums)) == 1
        row["gbps_new"] = round(n / (row["r2_single_launch/scratch_us"] * 1e-6) / 1e9, 1)
        rows.append(row)
        print(json.dumps(row), flush=True)
    ok = all(r["all_builds_agree"] for r in rows)
    print(json.dumps({"ok": ok}))
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
