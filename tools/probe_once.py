"""Profiling helper: a few launches of the probe kernels through the C ABI (what ncu wraps; not a bench)."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
B = importlib.import_module("gpu-feature-discovery_b200.binding")
n = int(sys.argv[1]) if len(sys.argv) > 1 else (4 << 30)
dev = torch.device("cuda:0")
buf = torch.empty(n, dtype=torch.uint8, device=dev)
dst = torch.empty(n, dtype=torch.uint8, device=dev)
scratch = torch.zeros(B.hbm_scratch_bytes(), dtype=torch.uint8, device=dev)  # header must start at zero (gfd_hbm_scratch_init)
out = torch.zeros(2, dtype=torch.int64, device=dev)
st = torch.cuda.current_stream().cuda_stream
B.hbm_fill_pattern(buf.data_ptr(), n, 0, st)
for _ in range(6):
    B.hbm_read_checksum(buf.data_ptr(), n, out.data_ptr(), scratch.data_ptr(), scratch.numel(), st)
for _ in range(4):
    B.hbm_copy(dst.data_ptr(), buf.data_ptr(), n, st)
torch.cuda.synchronize()
print("checksum %016x %016x" % tuple(int(v) & (2**64 - 1) for v in out.cpu().tolist()))
