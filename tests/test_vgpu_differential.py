"""gfd_vgpu_parse (C ABI) vs the oracle's restatement of internal/vgpu/{pciutil,vgpu}.go on GENERATED PCI config spaces.

The reference pins this code with two 256-byte blobs (tests/golden/vgpu_blobs.json); everything around them -- chains
that loop or break, capabilities that run past the buffer, records of length zero, fields with NULs or bytes >= 0x80
in them -- is covered here by construction: hypothesis draws a config space with a capability chain, a vendor
capability with a record list and a version/branch tail, then damages some bytes, and the shim must agree with the
oracle on EVERYTHING: which of {no capability, not a vGPU, parsed, error / Go panic / Go hang} happens, the capability
length, and the two strings byte for byte (Go's string(uint8) turns a byte >= 0x80 into two UTF-8 bytes, and
strings.Trim removes NULs at the two ends only)."""
import importlib

from hypothesis import HealthCheck, given, settings, strategies as st

from oracle import gfd_oracle as O

B = importlib.import_module("gpu-feature-discovery_b200.binding")

FIELD = st.lists(st.one_of(st.sampled_from([0, 0, 0x20, 0x30, 0x2E, 0x5F, 0x72, 0x80, 0xFF, 0xC3]), st.integers(0, 255)), min_size=10, max_size=10)


@st.composite
def config_space(draw):
    size = draw(st.sampled_from([64, 255, 256, 256, 256, 257, 300, 512, 4096]))
    cfg = bytearray(draw(st.binary(min_size=size, max_size=size))) if draw(st.booleans()) else bytearray(size)
    if size < 0x50:
        return bytes(cfg), size
    if draw(st.integers(0, 9)):
        cfg[0x06] |= 0x10
    # a capability chain: a few ordinary capabilities, then (usually) the vendor-specific one
    n_before = draw(st.integers(0, 3))
    positions = draw(st.lists(st.integers(0x40, min(size, 256) - 1), min_size=n_before + 1, max_size=n_before + 1, unique=True))
    cfg[0x34] = positions[0]
    for i, pos in enumerate(positions):
        last = i == len(positions) - 1
        if pos + 2 >= size:
            continue
        if not last:
            cfg[pos] = draw(st.sampled_from([0x01, 0x05, 0x10, 0x11, 0xFF, 0x09]))
            cfg[pos + 1] = positions[i + 1] if draw(st.integers(0, 7)) else draw(st.sampled_from([0, pos, positions[0]]))
            cfg[pos + 2] = draw(st.integers(0, 255))
            continue
        cfg[pos] = 0x09 if draw(st.integers(0, 7)) else draw(st.integers(0, 255))
        cfg[pos + 1] = draw(st.sampled_from([0, 0, pos]))
        records = draw(st.lists(st.tuples(st.integers(1, 255), st.integers(0, 6)), max_size=3))  # (record id, record length)
        body = bytearray([0, 0, 0, 0x56, 0x46]) if draw(st.integers(0, 7)) else bytearray(draw(st.binary(min_size=5, max_size=5)))
        for rid, rlen in records:
            rec = bytearray([rid, rlen]) + bytearray(draw(st.binary(min_size=max(0, rlen - 2), max_size=max(0, rlen - 2))))
            body += rec
        body += bytearray([0, draw(st.integers(0, 255))]) + bytearray(draw(FIELD)) + bytearray(draw(FIELD))
        length = draw(st.one_of(st.just(len(body)), st.integers(0, 255)))
        cfg[pos + 2] = length & 0xFF
        end = min(size, pos + len(body))
        cfg[pos + 3:end] = body[3:end - pos]
    for _ in range(draw(st.integers(0, 3))):  # damage
        cfg[draw(st.integers(0, size - 1))] = draw(st.integers(0, 255))
    cap = draw(st.sampled_from([size, size, max(512, size + 1)]))
    return bytes(cfg), cap


def oracle_outcome(cfg, cap):
    dev = O.PCIDevice("0000:00:00.0", cfg, cap=cap)
    try:
        c = O.get_vendor_specific_capability(dev)
    except O.GoPanic:
        return ("status", B.GFD_ERR_PCI_OUT_OF_RANGE)
    except O.GfdError:
        return ("status", B.GFD_ERR_PCI_CONFIG_SHORT)
    if c is None:
        return ("none",)
    if not O.is_vgpu_device(c):
        return ("not-vgpu", len(c))
    try:
        ver, br = O.vgpu_get_info(c)
    except O.GoPanic:
        return ("status", B.GFD_ERR_PCI_OUT_OF_RANGE)
    except O.GoHang:
        return ("status", B.GFD_ERR_VGPU_RECORD_LOOP)
    except O.GfdError:
        return ("status", B.GFD_ERR_VGPU_NO_VERSION_RECORD)
    return ("vgpu", len(c), ver, br)


def shim_outcome(cfg, cap):
    try:
        info = B.vgpu_parse(cfg, cap)
    except B.GfdError as e:
        return ("status", e.status)
    if not info.has_vendor_capability:
        return ("none",)
    if not info.is_vgpu:
        return ("not-vgpu", info.vendor_capability_len)
    return ("vgpu", info.vendor_capability_len, info.version(), info.branch())


@settings(max_examples=1500, deadline=None, suppress_health_check=[HealthCheck.too_slow, HealthCheck.data_too_large])
@given(config_space())
def test_vgpu_parse_equals_oracle(case):
    cfg, cap = case
    assert shim_outcome(cfg, cap) == oracle_outcome(cfg, cap)


def test_generator_reaches_every_outcome():
    """The differential test is only as good as its generator: make sure it really produces parsed vGPUs, every error
    kind and fields with interior NULs / high bytes (fixed seed, so this is deterministic)."""
    from hypothesis import seed

    seen, interesting = set(), set()

    @seed(20221222)
    @settings(max_examples=2000, deadline=None, database=None, suppress_health_check=list(HealthCheck))
    @given(config_space())
    def collect(case):
        o = oracle_outcome(*case)
        seen.add(o[0] if o[0] != "status" else o[1])
        if o[0] == "vgpu":
            for f in o[2:]:
                if b"\x00" in f:
                    interesting.add("interior-nul")
                if any(b >= 0x80 for b in f):
                    interesting.add("utf8")
    collect()
    assert {"none", "not-vgpu", "vgpu", B.GFD_ERR_PCI_OUT_OF_RANGE, B.GFD_ERR_PCI_CONFIG_SHORT, B.GFD_ERR_VGPU_RECORD_LOOP,
            B.GFD_ERR_VGPU_NO_VERSION_RECORD} <= seen, seen
    assert interesting == {"interior-nul", "utf8"}, interesting


def test_interior_nul_is_kept_like_go_does():
    cfg = bytearray(256)
    cfg[0x06] = 0x10
    cfg[0x34] = 0x40
    body = bytearray([0x09, 0, 0, 0x56, 0x46, 0, 0]) + bytearray(b"\x005\x0025.1\x00\x00\x00") + bytearray(b"r\xe9v\x00\x00\x00\x00\x00\x00\x00")
    body[2] = len(body)
    cfg[0x40:0x40 + len(body)] = body
    info = B.vgpu_parse(bytes(cfg))
    assert info.is_vgpu and info.version() == b"5\x0025.1" and info.branch() == "rév".encode()
    assert oracle_outcome(bytes(cfg), 256) == ("vgpu", len(body), b"5\x0025.1", "rév".encode())
