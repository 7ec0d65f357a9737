"""gfd_hbm_read_checksum_noalloc -- the checksum entry point for callers that may not allocate GPU memory
(a monitoring agent beside a job that owns nearly all of HBM, code bound to a framework's allocator).  Its scratch is a fixed array inside the library, so the test
hands it only an input buffer and a 16-byte output, and checks, through the C ABI, that the answer is bit-for-bit
the one gfd_hbm_read_checksum gives with caller-provided scratch (integer work: the tolerance is zero).

Every case makes two no-alloc calls back to back -- the array is shared between calls and keeps the previous
call's partial sums, so the second call is the one that would expose stale-scratch bugs.

This file sorts last on purpose: if it fails under `pytest -x`, everything else has already run.  It does not
depend on state left by other tests and passes in a fresh process on its own."""
import ctypes
import importlib
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    import torch
    assert torch.cuda.is_available(), "needs a B200"
    B = importlib.import_module("gpu-feature-discovery_b200.binding")
    dev = torch.device("cuda:0")

    class Ctx:
        pass
    c = Ctx()
    c.torch, c.B, c.dev = torch, B, dev
    c.scratch = torch.empty(B.hbm_scratch_bytes(), dtype=torch.uint8, device=dev)
    B.hbm_scratch_init(c.scratch.data_ptr(), c.scratch.numel(), torch.cuda.current_stream().cuda_stream)  # the allocating sibling's one-time set-up
    c.out_ref = torch.zeros(2, dtype=torch.int64, device=dev)
    c.out_new = torch.zeros(2, dtype=torch.int64, device=dev)
    # the CPU statement of the checksum (test infrastructure): the no-alloc path is also held against it directly
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    c.orc = ctypes.CDLL(os.path.join(root, "oracle", "libgfd_oracle_c.so"))
    c.orc.oracle_checksum.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.POINTER(ctypes.c_uint64)]
    return c


def oracle_entry_point(c, host):
    o = (ctypes.c_uint64 * 2)()
    host = np.ascontiguousarray(host, dtype=np.uint8)
    c.orc.oracle_checksum(host.ctypes.data, host.size, o)
    return int(o[0]), int(o[1])


def reference_entry_point(c, t, n):
    st = c.torch.cuda.current_stream()
    c.out_ref.fill_(-1)
    c.B.hbm_read_checksum(t.data_ptr(), n, c.out_ref.data_ptr(), c.scratch.data_ptr(), c.scratch.numel(), st.cuda_stream)
    st.synchronize()
    return tuple(int(v) for v in c.out_ref.cpu().numpy().view(np.uint64))


def noalloc_entry_point(c, t, n):
    st = c.torch.cuda.current_stream()
    c.out_new.fill_(-1)
    c.B.hbm_read_checksum_noalloc(t.data_ptr(), n, c.out_new.data_ptr(), st.cuda_stream)
    st.synchronize()
    return tuple(int(v) for v in c.out_new.cpu().numpy().view(np.uint64))


def device_bytes(c, n, seed):
    host = np.random.default_rng(seed).integers(0, 256, size=n, dtype=np.uint8)
    if n == 0:
        return c.torch.empty(0, dtype=c.torch.uint8, device=c.dev)
    return c.torch.from_numpy(host).to(c.dev)


# (first call, second call): a large launch followed by a small one leaves most scratch slots stale, and vice versa
PAIRS = [((64 << 20) + 9, 5), (5, (64 << 20) + 9), (0, 65536 * 3 + 7), (65536 * 296, 65536 * 296), (1 << 20, 0), (4097, 4097)]


@pytest.mark.parametrize("first,second", PAIRS)
def test_two_calls_in_a_row_match_the_allocating_entry_point(ctx, first, second):
    a = device_bytes(ctx, first, 1000 + first)
    b = device_bytes(ctx, second, 2000 + second)
    want_a = reference_entry_point(ctx, a, first)
    want_b = reference_entry_point(ctx, b, second)
    got_a = noalloc_entry_point(ctx, a, first)
    got_b = noalloc_entry_point(ctx, b, second)
    assert got_a == want_a
    assert got_b == want_b


def test_same_input_twice_is_stable(ctx):
    n = (32 << 20) + 24
    t = device_bytes(ctx, n, 7)
    want = reference_entry_point(ctx, t, n)
    assert noalloc_entry_point(ctx, t, n) == want
    assert noalloc_entry_point(ctx, t, n) == want


def test_oversized_request_is_refused_with_size_limit(ctx):
    B = ctx.B
    t = ctx.torch.zeros(1024, dtype=ctx.torch.uint8, device=ctx.dev)
    with pytest.raises(B.GfdError) as ei:
        B.hbm_read_checksum_noalloc(t.data_ptr(), B.NOALLOC_MAX_BYTES + 16, ctx.out_new.data_ptr(), 0)
    assert ei.value.status == B.GFD_ERR_SIZE_LIMIT
    assert "maximum" in str(ei.value)
    # the limit itself is a size check only: nothing is read when it trips, and a small request still works afterwards
    assert noalloc_entry_point(ctx, t, 1024) == (0, 0)


def test_bad_pointers_are_refused(ctx):
    B = ctx.B
    t = ctx.torch.zeros(1024, dtype=ctx.torch.uint8, device=ctx.dev)
    with pytest.raises(B.GfdError) as ei:
        B.hbm_read_checksum_noalloc(t.data_ptr() + 8, 256, ctx.out_new.data_ptr(), 0)
    assert ei.value.status == B.GFD_ERR_ALIGNMENT
    with pytest.raises(B.GfdError) as ei:
        B.hbm_read_checksum_noalloc(t.data_ptr(), 256, 0, 0)
    assert ei.value.status == B.GFD_ERR_BAD_ARGUMENT


@pytest.mark.parametrize("n", [0, 1, 15, 16, 65536 + 3, (48 << 20) + 11])
def test_matches_the_cpu_oracle_directly_twice(ctx, n):
    """Not only equal to the other GPU entry point: equal to the byte-at-a-time C statement of the checksum, on random bytes,
    on two calls in a row (the second call meets the array as the first one left it)."""
    host = np.random.default_rng(31337 + n).integers(0, 256, size=n, dtype=np.uint8)
    t = ctx.torch.from_numpy(host).to(ctx.dev) if n else ctx.torch.empty(0, dtype=ctx.torch.uint8, device=ctx.dev)
    want = oracle_entry_point(ctx, host)
    assert noalloc_entry_point(ctx, t, n) == want
    assert noalloc_entry_point(ctx, t, n) == want


def test_host_pointer_is_refused_not_faulted(ctx):
    """A pageable host address as the input must come back as BAD_ARGUMENT (or the runtime's invalid-value code), exactly as
    the allocating entry point answers -- not as an illegal-address fault that kills the caller's CUDA context.  The GPU must
    still work afterwards."""
    B = ctx.B
    host = np.zeros(4096 + 16, dtype=np.uint8)
    addr = (host.ctypes.data + 15) // 16 * 16
    st_new = B.lib.gfd_hbm_read_checksum_noalloc(addr, 4096, ctx.out_new.data_ptr(), 0)
    st_ref = B.lib.gfd_hbm_read_checksum(addr, 4096, ctx.out_ref.data_ptr(), ctx.scratch.data_ptr(), ctx.scratch.numel(), 0)
    assert st_new != 0 and st_new == st_ref
    assert st_new in (B.GFD_ERR_BAD_ARGUMENT, 100000 + 1)
    ctx.torch.cuda.synchronize()                    # no sticky error
    t = device_bytes(ctx, 70001, 5)
    assert noalloc_entry_point(ctx, t, 70001) == reference_entry_point(ctx, t, 70001)


def test_second_gpu_gets_its_own_array_address(ctx):
    """The array's device address differs per GPU and is looked up on first use per GPU: call on cuda:0 first, then cuda:1,
    then cuda:0 again, each twice, each against the allocating entry point on the same GPU."""
    torch, B = ctx.torch, ctx.B
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    n = (8 << 20) + 40
    host = np.random.default_rng(99).integers(0, 256, size=n, dtype=np.uint8)
    want = oracle_entry_point(ctx, host)
    for dev_index in (0, 1, 0, 1):
        dev = torch.device("cuda:%d" % dev_index)
        with torch.cuda.device(dev):
            t = torch.from_numpy(host).to(dev)
            out = torch.zeros(2, dtype=torch.int64, device=dev)
            scratch = torch.zeros(B.hbm_scratch_bytes(), dtype=torch.uint8, device=dev)
            st = torch.cuda.current_stream(dev)
            for _ in range(2):
                out.fill_(-1)
                B.hbm_read_checksum_noalloc(t.data_ptr(), n, out.data_ptr(), st.cuda_stream)
                st.synchronize()
                assert tuple(int(v) for v in out.cpu().numpy().view(np.uint64)) == want
            out.fill_(-1)
            B.hbm_read_checksum(t.data_ptr(), n, out.data_ptr(), scratch.data_ptr(), scratch.numel(), st.cuda_stream)
            st.synchronize()
            assert tuple(int(v) for v in out.cpu().numpy().view(np.uint64)) == want


def test_input_on_another_gpu_is_refused_but_the_allocating_sibling_reads_it(ctx):
    """Input on cuda:1, result on cuda:0.  The no-alloc entry point sets no peer mapping up, so it must refuse (BAD_ARGUMENT)
    rather than launch a kernel that faults; the allocating entry point enables the mapping itself and pulls the bytes over
    NVLink.  Runs in a fresh process too: nothing else has enabled peer access before."""
    torch, B = ctx.torch, ctx.B
    if torch.cuda.device_count() < 2 or not torch.cuda.can_device_access_peer(0, 1):
        pytest.skip("needs 2 peer-capable GPUs")
    n = (4 << 20) + 24
    host = np.random.default_rng(123).integers(0, 256, size=n, dtype=np.uint8)
    far = torch.from_numpy(host).to("cuda:1")
    torch.cuda.synchronize()
    with pytest.raises(B.GfdError) as ei:
        B.hbm_read_checksum_noalloc(far.data_ptr(), n, ctx.out_new.data_ptr(), 0)
    assert ei.value.status == B.GFD_ERR_BAD_ARGUMENT
    assert reference_entry_point(ctx, far, n) == oracle_entry_point(ctx, host)     # peer loads, mapping enabled on demand
    near = far.to("cuda:0")
    assert noalloc_entry_point(ctx, near, n) == oracle_entry_point(ctx, host)      # and the GPU is fine afterwards
