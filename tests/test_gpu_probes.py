"""-m gpu parity tests for the probe kernels: every call goes through the C ABI of libgfd_b200.so; the
checker is the plain-C oracle (oracle/checksum_oracle.c) on the same bytes.  Bar: bit-exact (integer work).

torch is used only as plumbing: device buffers, streams, host<->device copies."""
import ctypes
import importlib
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def env():
    import torch
    assert torch.cuda.is_available(), "these tests need a B200"
    B = importlib.import_module("gpu-feature-discovery_b200.binding")
    so = os.path.join(ROOT, "oracle", "libgfd_oracle_c.so")
    assert os.path.exists(so), "oracle/libgfd_oracle_c.so not built (make / __graft_entry__.build())"
    orc = ctypes.CDLL(so)
    orc.oracle_checksum.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.POINTER(ctypes.c_uint64)]
    orc.oracle_fill_pattern.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_uint64]
    orc.oracle_pattern_checksum.argtypes = [ctypes.c_size_t, ctypes.c_uint64, ctypes.POINTER(ctypes.c_uint64)]
    dev = torch.device("cuda:0")
    # garbage on purpose: the one-time gfd_hbm_scratch_init is what makes the header usable, nothing is cleared per call
    scratch = torch.full((B.hbm_scratch_bytes(),), 0xA5, dtype=torch.uint8, device=dev)
    B.hbm_scratch_init(scratch.data_ptr(), scratch.numel(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    out = torch.zeros(2, dtype=torch.int64, device=dev)

    class Env:
        pass
    e = Env()
    e.torch, e.B, e.orc, e.dev, e.scratch, e.out = torch, B, orc, dev, scratch, out
    return e


def oracle_checksum(env, arr):
    arr = np.ascontiguousarray(arr, dtype=np.uint8)
    o = (ctypes.c_uint64 * 2)()
    env.orc.oracle_checksum(arr.ctypes.data, arr.size, o)
    return int(o[0]), int(o[1])


def gpu_checksum(env, t, nbytes=None, stream=None):
    torch = env.torch
    nbytes = t.numel() * t.element_size() if nbytes is None else nbytes
    st = torch.cuda.current_stream() if stream is None else stream
    env.out.fill_(-1)
    env.B.hbm_read_checksum(t.data_ptr(), nbytes, env.out.data_ptr(), env.scratch.data_ptr(), env.scratch.numel(), st.cuda_stream)
    st.synchronize()
    x, s = env.out.cpu().numpy().view(np.uint64)
    return int(x), int(s)


# sizes: empty, sub-word, sub-vector, ragged everywhere, exactly one 64 KiB tile, tile+-1, many tiles + tail
SIZES = [0, 1, 7, 8, 15, 16, 17, 31, 32, 33, 4095, 4096, 65535, 65536, 65537, 65536 * 3 + 5, (1 << 20) + 13,
         296 * 65536, 296 * 65536 + 65536 - 16, (64 << 20) + 9]


@pytest.mark.parametrize("n", SIZES)
def test_read_checksum_matches_oracle_random_bytes(env, n):
    rng = np.random.default_rng(1234 + n)
    host = rng.integers(0, 256, size=n, dtype=np.uint8)
    t = env.torch.from_numpy(host).to(env.dev) if n else env.torch.empty(0, dtype=env.torch.uint8, device=env.dev)
    assert gpu_checksum(env, t, n) == oracle_checksum(env, host)


def test_read_checksum_unaligned_to_32_bytes(env):
    """a 16-byte aligned (not 32-byte aligned) base must take the 128-bit path and still be exact"""
    rng = np.random.default_rng(7)
    n = (3 << 20) + 48
    host = rng.integers(0, 256, size=n + 16, dtype=np.uint8)
    t = env.torch.from_numpy(host).to(env.dev)
    view = t[16:]
    assert view.data_ptr() % 32 == 16
    assert gpu_checksum(env, view, n) == oracle_checksum(env, host[16:])


def test_misaligned_pointer_is_rejected(env):
    t = env.torch.zeros(1024, dtype=env.torch.uint8, device=env.dev)
    with pytest.raises(env.B.GfdError) as ei:
        env.B.hbm_read_checksum(t.data_ptr() + 4, 512, env.out.data_ptr(), env.scratch.data_ptr(), env.scratch.numel(), 0)
    assert ei.value.status == env.B.GFD_ERR_ALIGNMENT
    with pytest.raises(env.B.GfdError) as ei:
        env.B.hbm_read_checksum(t.data_ptr(), 512, env.out.data_ptr(), env.scratch.data_ptr(), 16, 0)
    assert ei.value.status == env.B.GFD_ERR_BUFFER_TOO_SMALL


@pytest.mark.parametrize("n", [0, 5, 16, 1000, 65536, 65536 * 5 + 7, (8 << 20) + 3])
@pytest.mark.parametrize("seed", [0, 12345])
def test_fill_pattern_matches_oracle(env, n, seed):
    t = env.torch.zeros(max(n, 1) + 64, dtype=env.torch.uint8, device=env.dev)
    env.B.hbm_fill_pattern(t.data_ptr(), n, seed, env.torch.cuda.current_stream().cuda_stream)
    env.torch.cuda.synchronize()
    got = t.cpu().numpy()
    want = np.zeros(n, dtype=np.uint8)
    env.orc.oracle_fill_pattern(want.ctypes.data, n, seed)
    assert np.array_equal(got[:n], want)
    assert not got[n:].any(), "fill wrote past the end"


@pytest.mark.parametrize("n", [0, 1, 7, 8, 9, 15, 16, 17, 24, 31, 4097, 65536, 65536 * 7 + 9, (32 << 20) + 13])
@pytest.mark.parametrize("seed", [0, 12345, 7 << 40])
def test_pattern_checksum_is_the_checksum_of_the_filled_buffer(env, n, seed):
    """gfd_hbm_pattern_checksum (no memory read) == oracle checksum of the oracle's pattern bytes == what the GPU reads back
    from a buffer it filled: the three-way agreement the probes' write -> read verification rests on."""
    torch, B = env.torch, env.B
    st = torch.cuda.current_stream().cuda_stream
    B.hbm_pattern_checksum(n, seed, env.out.data_ptr(), env.scratch.data_ptr(), env.scratch.numel(), st)
    torch.cuda.synchronize()
    expected = tuple(int(v) for v in env.out.cpu().numpy().view(np.uint64))
    want = np.zeros(n, dtype=np.uint8)
    env.orc.oracle_fill_pattern(want.ctypes.data, n, seed)
    assert expected == oracle_checksum(env, want)
    t = torch.zeros(max(n, 1) + 64, dtype=torch.uint8, device=env.dev)
    B.hbm_fill_pattern(t.data_ptr(), n, seed, st)
    assert gpu_checksum(env, t, nbytes=n) == expected
    if n >= 8:  # one flipped bit anywhere must break the agreement
        t[n // 2] ^= 0x10
        assert gpu_checksum(env, t, nbytes=n) != expected


@pytest.mark.parametrize("n", [0, 1, 15, 16, 4097, 65536, 65536 * 7 + 9, (32 << 20) + 1])
def test_copy_is_bit_exact(env, n):
    rng = np.random.default_rng(99 + n)
    host = rng.integers(0, 256, size=n, dtype=np.uint8)
    src = env.torch.from_numpy(host).to(env.dev) if n else env.torch.empty(0, dtype=env.torch.uint8, device=env.dev)
    dst = env.torch.full((n + 32,), 0xAB, dtype=env.torch.uint8, device=env.dev)
    env.B.hbm_copy(dst.data_ptr(), src.data_ptr(), n, env.torch.cuda.current_stream().cuda_stream)
    env.torch.cuda.synchronize()
    got = dst.cpu().numpy()
    assert np.array_equal(got[:n], host)
    assert (got[n:] == 0xAB).all(), "copy wrote past the end"


def test_full_size_properties_4gib(env):
    """BASELINE probe size (4 GiB) through size-independent properties:
    fill -> checksum equals the closed-form pattern checksum; copy -> same checksum; checksum of two halves
    combines (xor / +) to the checksum of the whole; idempotent on repeat."""
    torch, B = env.torch, env.B
    n = 1 << 32
    st = torch.cuda.current_stream().cuda_stream
    buf = torch.empty(n, dtype=torch.uint8, device=env.dev)
    B.hbm_fill_pattern(buf.data_ptr(), n, 0, st)
    whole = gpu_checksum(env, buf)
    o = (ctypes.c_uint64 * 2)()
    env.orc.oracle_pattern_checksum(n // 8, 0, o)
    assert whole == (int(o[0]), int(o[1]))
    assert gpu_checksum(env, buf) == whole
    half = n // 2
    a = gpu_checksum(env, buf[:half], half)
    b = gpu_checksum(env, buf[half:], half)
    assert (a[0] ^ b[0], (a[1] + b[1]) % (1 << 64)) == whole
    dst = torch.empty(n, dtype=torch.uint8, device=env.dev)
    B.hbm_copy(dst.data_ptr(), buf.data_ptr(), n, st)
    assert gpu_checksum(env, dst) == whole
    # a single flipped byte anywhere must change the result
    dst[n - 3] ^= 0x40
    assert gpu_checksum(env, dst) != whole


def test_checksum_on_side_stream(env):
    torch = env.torch
    rng = np.random.default_rng(5)
    host = rng.integers(0, 256, size=(4 << 20) + 3, dtype=np.uint8)
    t = torch.from_numpy(host).to(env.dev)
    torch.cuda.synchronize()
    s = torch.cuda.Stream()
    assert gpu_checksum(env, t, host.size, stream=s) == oracle_checksum(env, host)


def test_checksum_host_end_to_end(env):
    """gfd_checksum_host: host buffer in, 16 bytes out; pinned and pageable sources."""
    torch, B = env.torch, env.B
    rng = np.random.default_rng(11)
    for n in (0, 3, 1 << 20, (16 << 20) + 11):
        host = rng.integers(0, 256, size=n, dtype=np.uint8)
        want = oracle_checksum(env, host)
        assert B.checksum_host(0, host.ctypes.data if n else 0, n) == want
        if n:
            pinned = torch.from_numpy(host).pin_memory()
            assert B.checksum_host(0, pinned.data_ptr(), n) == want


def test_probe_hbm_self_contained(env):
    B = env.B
    n = 1 << 30
    r = B.probe_hbm(0, n, warmup=2, iters=5, mode=B.PROBE_READ)
    o = (ctypes.c_uint64 * 2)()
    env.orc.oracle_pattern_checksum(n // 8, 0, o)
    assert r.checksum.tuple() == (int(o[0]), int(o[1]))
    assert r.verified == 1 and r.expected.tuple() == r.checksum.tuple()  # the probe checked what it read against the pattern
    assert r.bytes == n and r.iters == 5 and r.sm_count >= 100 and r.grid == r.sm_count * 2
    assert 500 < r.gbps < 20000, r.gbps
    c = B.probe_hbm(0, n, warmup=2, iters=5, mode=B.PROBE_COPY)
    assert 500 < c.gbps < 20000, c.gbps
    assert c.verified == 1 and c.checksum.tuple() == r.checksum.tuple()  # the copy's destination was re-read and checked
    w = B.probe_hbm(0, n, warmup=2, iters=5, mode=B.PROBE_WRITE)  # the fill kernel timed; what it wrote is re-read and checked
    assert 500 < w.gbps < 20000, w.gbps
    assert w.verified == 1 and w.checksum.tuple() == r.checksum.tuple()
    w = B.probe_hbm(0, (64 << 20) + 40, warmup=1, iters=2, mode=B.PROBE_WRITE)  # ragged: partial last tile + 8 tail bytes
    assert w.verified == 1


def test_probe_rejects_bad_arguments(env):
    B = env.B
    for args in ((99, 1 << 20, 1, 1, 0), (0, 0, 1, 1, 0), (0, 1 << 20, 1, 0, 0), (0, 1 << 20, 1, 1, 7)):
        with pytest.raises(B.GfdError):
            B.probe_hbm(*args[:2], warmup=args[2], iters=args[3], mode=args[4])


@pytest.mark.parametrize("n", [0, 1, 15, 16, 17, 4097, 32767, 32768, 32769, 65535, 65536, 65537, 3 * 65536 + 5, (8 << 20) + 13])
def test_no_store_outside_the_buffer_and_all_paths_agree(env, n):
    """Guard bands around every written buffer (compute-sanitizer is not available on the pool): fill and copy must leave the
    canaries on both sides alone, and fill -> read, copy -> read, the pattern's definition and the no-alloc read must agree."""
    torch, B, G = env.torch, env.B, 4096
    st = torch.cuda.current_stream().cuda_stream

    def guarded():
        t = torch.full((n + 2 * G,), 0xA5, dtype=torch.uint8, device=env.dev)
        return t, t[G:G + max(n, 1)]

    tb, buf = guarded()
    td, dst = guarded()
    B.hbm_fill_pattern(buf.data_ptr(), n, 7, st)
    a = gpu_checksum(env, buf, nbytes=n)
    B.hbm_copy(dst.data_ptr(), buf.data_ptr(), n, st)
    b = gpu_checksum(env, dst, nbytes=n)
    B.hbm_pattern_checksum(n, 7, env.out.data_ptr(), env.scratch.data_ptr(), env.scratch.numel(), st)
    torch.cuda.synchronize()
    c = tuple(int(v) for v in env.out.cpu().numpy().view(np.uint64))
    assert B.lib.gfd_hbm_read_checksum_noalloc(buf.data_ptr(), n, env.out.data_ptr(), st) == 0
    torch.cuda.synchronize()
    d = tuple(int(v) for v in env.out.cpu().numpy().view(np.uint64))
    assert a == b == c == d
    for t in (tb, td):
        assert bool((t[:G] == 0xA5).all()) and bool((t[G + n:] == 0xA5).all()), "a kernel stored outside its buffer"


def test_out_of_memory_is_an_error_and_leaves_no_stale_state(env):
    """A probe size that cannot be allocated must come back as cudaErrorMemoryAllocation (status 100002) -- and the NEXT
    call must work: a failed cudaMalloc also parks its code in the runtime's last-error slot, which the launch checks read."""
    B = env.B
    with pytest.raises(B.GfdError) as e:
        B.probe_hbm(0, 1 << 40, warmup=1, iters=1, mode=B.PROBE_READ)  # 1 TiB
    assert e.value.status == 100000 + 2
    r = B.probe_hbm(0, 256 << 20, warmup=1, iters=2, mode=B.PROBE_READ)
    o = (ctypes.c_uint64 * 2)()
    env.orc.oracle_pattern_checksum((256 << 20) // 8, 0, o)
    assert r.checksum.tuple() == (int(o[0]), int(o[1]))
    with pytest.raises(B.GfdError) as e:
        B.probe_hbm(0, 1 << 40, warmup=1, iters=1, mode=B.PROBE_COPY)
    assert e.value.status == 100000 + 2
    host = np.arange(1 << 16, dtype=np.uint64)
    assert B.checksum_host(0, host.ctypes.data, host.nbytes) == oracle_checksum(env, host.view(np.uint8))


def test_p2p_probe(env):
    """NVLink peer probe: needs >= 2 GPUs (run with gpurun --gpus 2); the pulled bytes' checksum must equal
    the pattern's closed form, i.e. every byte crossed the link intact."""
    B = env.B
    if env.torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    n = 1 << 28
    r = B.probe_p2p(0, 1, n, warmup=2, iters=5, mode=B.PROBE_READ)
    o = (ctypes.c_uint64 * 2)()
    env.orc.oracle_pattern_checksum(n // 8, 0, o)
    assert r.checksum.tuple() == (int(o[0]), int(o[1])) and r.verified == 1
    assert 50 < r.gbps < 2000, r.gbps
    c = B.probe_p2p(1, 0, n, warmup=2, iters=5, mode=B.PROBE_COPY)
    assert 50 < c.gbps < 2000, c.gbps
    p = B.probe_p2p(1, 0, n, warmup=2, iters=5, mode=B.PROBE_PUSH)
    assert c.verified == 1 and p.verified == 1 and c.checksum.tuple() == p.checksum.tuple() == r.checksum.tuple()
    d = B.probe_p2p(1, 0, n, warmup=2, iters=5, mode=B.PROBE_DMA)     # the copy engine over the same link, verified the same way
    assert d.verified == 1 and d.checksum.tuple() == r.checksum.tuple() and 50 < d.gbps < 2000, d.gbps


def test_p2p_matrix_concurrent_tournament(env):
    """gfd_probe_p2p_matrix: every directed pair measured, all GPUs busy in every round; needs >= 2 GPUs."""
    B = env.B
    n = env.torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs 2 GPUs")
    m, ring = B.probe_p2p_matrix(n, 1 << 28, warmup=1, iters=5, mode=B.PROBE_READ)
    for a in range(n):
        for b in range(n):
            if a == b:
                assert m[a][b] == 0.0
            else:
                assert 50 < m[a][b] < 2000, (a, b, m[a][b])
    assert ring > 50 * n
    md, ring_d = B.probe_p2p_matrix(n, 1 << 28, warmup=1, iters=5, mode=B.PROBE_DMA)
    assert all(50 < md[a][b] < 2000 for a in range(n) for b in range(n) if a != b) and ring_d > 50 * n


def test_random_sizes_and_offsets(env):
    """40 random (offset, size) windows of one random buffer: read checksum and copy vs the C oracle / numpy, bit-exact."""
    torch = env.torch
    rng = np.random.default_rng(20261106)
    total = (96 << 20) + 4096
    host = rng.integers(0, 256, size=total, dtype=np.uint8)
    buf = torch.from_numpy(host).to(env.dev)
    dst = torch.empty(total + 64, dtype=torch.uint8, device=env.dev)
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(40):
        off = int(rng.integers(0, 4096)) * 16
        n = int(rng.choice([rng.integers(0, 70000), rng.integers(0, 1 << 20), rng.integers(0, total - off)]))
        n = min(n, total - off)
        view = buf[off:off + n] if n else buf[off:off]
        assert gpu_checksum(env, view, n) == oracle_checksum(env, host[off:off + n]), (off, n)
        dst.fill_(0x5A)
        env.B.hbm_copy(dst.data_ptr(), buf.data_ptr() + off, n, st)
        torch.cuda.synchronize()
        got = dst.cpu().numpy()
        assert np.array_equal(got[:n], host[off:off + n]) and (got[n:n + 64] == 0x5A).all(), (off, n)


def test_copy_and_checksum_take_peer_pointers_without_prior_setup(env):
    """gfd_hbm_copy / gfd_hbm_read_checksum with the source on another GPU: the library enables the peer mapping itself (the kernel
    runs on the destination's GPU and pulls over NVLink); bytes arrive intact.  Needs >= 2 GPUs."""
    torch, B = env.torch, env.B
    if torch.cuda.device_count() < 2 or not torch.cuda.can_device_access_peer(0, 1):
        pytest.skip("needs 2 peer-capable GPUs")
    n = (6 << 20) + 40
    host = np.random.default_rng(77).integers(0, 256, size=n, dtype=np.uint8)
    far = torch.from_numpy(host).to("cuda:1")
    near = torch.zeros(n, dtype=torch.uint8, device=env.dev)
    torch.cuda.synchronize()
    B.hbm_copy(near.data_ptr(), far.data_ptr(), n, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert bool((near.cpu() == torch.from_numpy(host)).all())
    assert gpu_checksum(env, far, nbytes=n) == oracle_checksum(env, host) == gpu_checksum(env, near, nbytes=n)
