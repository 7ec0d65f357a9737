"""Small-case self-check of every kernel in libgfd_b200.so on ragged sizes (tails, partial tiles, empty), with guard
bands around every written buffer: a store past either end flips a canary.  (compute-sanitizer is closed on this pool, so
out-of-bounds WRITES are caught this way; out-of-bounds reads only by the four-way checksum agreement below.)"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
B = importlib.import_module("gpu-feature-discovery_b200.binding")
dev = torch.device("cuda:0")
scratch = torch.empty(B.hbm_scratch_bytes(), dtype=torch.uint8, device=dev)
out = torch.zeros(2, dtype=torch.int64, device=dev)
st = torch.cuda.current_stream().cuda_stream
G = 4096  # guard bytes on each side (keeps the 16-byte alignment of the payload)


def guarded(n):
    t = torch.full((n + 2 * G,), 0xA5, dtype=torch.uint8, device=dev)
    return t, t[G:G + n] if n else t[G:G + 1]


def guards_intact(t, n):
    return bool((t[:G] == 0xA5).all()) and bool((t[G + n:] == 0xA5).all())


for n in (0, 1, 15, 16, 17, 4097, 32767, 32768, 32769, 65535, 65536, 65537, 3 * 65536 + 5, (8 << 20) + 13):
    tb, buf = guarded(n)
    td, dst = guarded(n)
    B.hbm_fill_pattern(buf.data_ptr(), n, 7, st)
    B.hbm_read_checksum(buf.data_ptr(), n, out.data_ptr(), scratch.data_ptr(), scratch.numel(), st)
    a = tuple(out.cpu().tolist())
    B.hbm_copy(dst.data_ptr(), buf.data_ptr(), n, st)
    B.hbm_read_checksum(dst.data_ptr(), n, out.data_ptr(), scratch.data_ptr(), scratch.numel(), st)
    b = tuple(out.cpu().tolist())
    B.hbm_pattern_checksum(n, 7, out.data_ptr(), scratch.data_ptr(), scratch.numel(), st)
    c = tuple(out.cpu().tolist())
    assert B.lib.gfd_hbm_read_checksum_noalloc(buf.data_ptr(), n, out.data_ptr(), st) == 0
    d = tuple(out.cpu().tolist())
    assert a == b == c == d, (n, a, b, c, d)
    assert guards_intact(tb, n) and guards_intact(td, n), "a kernel stored outside its buffer at n=%d" % n
for mode in (B.PROBE_READ, B.PROBE_COPY, B.PROBE_WRITE):
    r = B.probe_hbm(0, (4 << 20) + 24, warmup=1, iters=2, mode=mode)
    assert r.verified == 1
print("self-check ok")
