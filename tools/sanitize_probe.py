"""compute-sanitizer target: every kernel of libgfd_b200.so on small, ragged sizes (tails, partial tiles, empty)."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
B = importlib.import_module("gpu-feature-discovery_b200.binding")
dev = torch.device("cuda:0")
scratch = torch.empty(B.hbm_scratch_bytes(), dtype=torch.uint8, device=dev)
out = torch.zeros(2, dtype=torch.int64, device=dev)
st = torch.cuda.current_stream().cuda_stream
for n in (0, 1, 15, 16, 17, 4097, 65535, 65536, 65537, 3 * 65536 + 5, (8 << 20) + 13):
    buf = torch.empty(max(n, 1), dtype=torch.uint8, device=dev)   # exactly n bytes: any overrun is out of bounds
    dst = torch.empty(max(n, 1), dtype=torch.uint8, device=dev)
    B.hbm_fill_pattern(buf.data_ptr(), n, 7, st)
    B.hbm_read_checksum(buf.data_ptr(), n, out.data_ptr(), scratch.data_ptr(), scratch.numel(), st)
    a = tuple(out.cpu().tolist())
    B.hbm_copy(dst.data_ptr(), buf.data_ptr(), n, st)
    B.hbm_read_checksum(dst.data_ptr(), n, out.data_ptr(), scratch.data_ptr(), scratch.numel(), st)
    b = tuple(out.cpu().tolist())
    B.hbm_pattern_checksum(n, 7, out.data_ptr(), scratch.data_ptr(), scratch.numel(), st)
    c = tuple(out.cpu().tolist())
    lib = B.lib
    assert lib.gfd_hbm_read_checksum_noalloc(buf.data_ptr(), n, out.data_ptr(), st) == 0
    d = tuple(out.cpu().tolist())
    assert a == b == c == d, (n, a, b, c, d)
for mode in (B.PROBE_READ, B.PROBE_COPY, B.PROBE_WRITE):
    r = B.probe_hbm(0, (4 << 20) + 24, warmup=1, iters=2, mode=mode)
    assert r.verified == 1
print("sanitize target ok")
