"""gfd_hbm_read_checksum_noalloc -- the checksum entry point for callers that may not allocate GPU memory
(a monitoring agent beside a job that owns nearly all of HBM, code bound to a framework's allocator).  Its scratch is a fixed array inside the library, so the test
hands it only an input buffer and a 16-byte output, and checks, through the C ABI, that the answer is bit-for-bit
the one gfd_hbm_read_checksum gives with caller-provided scratch (integer work: the tolerance is zero).

Every case makes two no-alloc calls back to back -- the array is shared between calls and keeps the previous
call's partial sums, so the second call is the one that would expose stale-scratch bugs.

This file sorts last on purpose: if it fails under `pytest -x`, everything else has already run.  It does not
depend on state left by other tests and passes in a fresh process on its own."""
import importlib

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
    c.out_ref = torch.zeros(2, dtype=torch.int64, device=dev)
    c.out_new = torch.zeros(2, dtype=torch.int64, device=dev)
    return c


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
