"""The gpu-feature-discovery binary vs the oracle: label FILE parity on simulated nodes (CPU only).

For every named node (tests/nodes.py) x MIG strategy the binary runs --oneshot against the fake
libnvidia-ml.so.1 and its output file must equal the oracle's rendering byte for byte (the binary writes keys
sorted; the oracle renders sorted -- SURVEY.md Fact 3).  Where the oracle says the reference fails, the binary
must exit 1 and write nothing."""
import json
import os
import stat

import pytest

import nodes as N
from oracle import gfd_oracle as O, nvml_model as M
from shimutil import B, run_cli, use_node

NODES = N.nodes()
CASES = [(n, s) for n in sorted(NODES) for s in N.STRATEGIES]
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def cli_pass(tmp_path, node, strategy, extra_args=(), extra_env=None, machine=b"DGX B200\n", pci_root=None, nomig=False):
    fix = use_node(node, tmp_path)
    out = os.path.join(str(tmp_path), "features.d", "gfd")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    mt = os.path.join(str(tmp_path), "product_name")
    if machine is not None:
        with open(mt, "wb") as f:
            f.write(machine)
    if pci_root is None:
        pci_root = os.path.join(str(tmp_path), "pci-empty")
        os.makedirs(pci_root, exist_ok=True)
    env = {"GFD_FAKE_NVML": fix, "GFD_B200_PCI_DEVICES_ROOT": pci_root}
    env.update(extra_env or {})
    args = ["--oneshot", "--no-timestamp", "--mig-strategy=" + strategy, "--machine-type-file", mt, "-o", out] + list(extra_args)
    rc, so, se = run_cli(args, env=env, nomig=nomig)
    text = open(out).read() if os.path.exists(out) else None
    return rc, text, se, out


def oracle_pass(node, strategy, machine=b"DGX B200\n", time_slicing=(), pci=()):
    cfg = O.Config(mig_strategy=strategy, no_timestamp=True, machine_type_file="/x", time_slicing=time_slicing)

    def rf(_):
        if machine is None:
            raise OSError("no such file")
        return machine
    try:
        labels, warnings = O.run_pass(O.NvmlManager(M.RawNvml(node)), list(pci), cfg, read_file=rf)
        return O.render(labels), warnings
    except O.GfdError:
        return None, []


@pytest.mark.parametrize("name,strategy", CASES, ids=["%s-%s" % c for c in CASES])
def test_label_file_matches_oracle(name, strategy, tmp_path):
    node = NODES[name]
    want, _ = oracle_pass(node, strategy)
    rc, got, stderr, out = cli_pass(tmp_path, node, strategy, nomig=bool(node.hidden))
    if want is None:
        assert rc == 1 and got is None, stderr
        assert "Error: " in stderr
    else:
        assert rc == 0, stderr
        assert got == want
        assert stat.S_IMODE(os.stat(out).st_mode) == 0o644
        assert os.path.isdir(os.path.join(os.path.dirname(out), "gfd-tmp"))        # oneshot leaves it behind
        assert os.listdir(os.path.join(os.path.dirname(out), "gfd-tmp")) == []


@pytest.mark.parametrize("name", ["c3_8xB200", "c4_single_8x7x1g23", "c5_mixed_hetero", "err_memory_on_non_representative_gpu", "gi_id9_but_other_gpu_empty"])
@pytest.mark.parametrize("strategy", ["none", "single", "mixed"])
def test_prefetch_threads_flag_same_output(name, strategy, tmp_path):
    want, _ = oracle_pass(NODES[name], strategy)
    rc, got, stderr, _ = cli_pass(tmp_path, NODES[name], strategy, extra_args=["--prefetch-threads=8"])
    assert (got == want) and (rc == (1 if want is None else 0)), stderr


def test_heterogeneous_node_output_is_one_the_reference_could_emit(tmp_path):
    """Go map order makes the reference's output random on mixed-model nodes; ours must be one of its possibilities."""
    for name in ("hetero_models", "hetero_mig_and_full"):
        node = NODES[name]
        cfg = O.Config(mig_strategy="mixed", no_timestamp=True, machine_type_file="/x")
        possible = O.possible_label_sets(lambda: O.NvmlManager(M.RawNvml(node)), [], cfg, read_file=lambda p: b"DGX B200\n")
        rc, got, stderr, _ = cli_pass(tmp_path, node, "mixed")
        assert rc == 0 and got in possible, (stderr, len(possible))
        assert len(possible) >= 2 or name != "hetero_models"


@pytest.mark.parametrize("machine,want", [(b"NVIDIA DGX B200\n", "NVIDIA-DGX-B200"), (b"  PowerEdge XE9680 \r\n", "PowerEdge-XE9680"),
                                          (b"", ""), (b"\xc2\xa0Odd\xe2\x80\x83Box\xc2\x85", "Odd Box"), (None, "unknown")])
def test_machine_type(machine, want, tmp_path):
    rc, got, stderr, _ = cli_pass(tmp_path, NODES["c2_1xB200"], "none", machine=machine)
    assert rc == 0, stderr
    oracle_text, _ = oracle_pass(NODES["c2_1xB200"], "none", machine=machine)
    assert got == oracle_text
    assert "nvidia.com/gpu.machine=%s\n" % want in got


def test_machine_type_empty_path(tmp_path):
    fix = use_node(NODES["c2_1xB200"], tmp_path)
    out = os.path.join(str(tmp_path), "o")
    empty = os.path.join(str(tmp_path), "pci-empty")
    os.makedirs(empty)
    rc, _, se = run_cli(["--oneshot", "--no-timestamp", "--machine-type-file=", "-o", out],
                        env={"GFD_FAKE_NVML": fix, "GFD_B200_PCI_DEVICES_ROOT": empty})
    assert rc == 0, se
    assert "nvidia.com/gpu.machine=unknown\n" in open(out).read()


# ---- time slicing through the config file -----------------------------------------------------------------
def write_config(tmp_path, doc, name="config.yaml"):
    p = os.path.join(str(tmp_path), name)
    with open(p, "w") as f:
        f.write(doc)
    return p


TS_CASES = [
    ("gpu x2", "sharing:\n  timeSlicing:\n    resources:\n    - name: nvidia.com/gpu\n      replicas: 2\n",
     [("nvidia.com/gpu", 2, "")]),
    ("not matching", "sharing:\n  timeSlicing:\n    resources:\n    - name: nvidia.com/not-gpu\n      replicas: 2\n",
     [("nvidia.com/not-gpu", 2, "")]),
    ("rename is neutralised", "sharing:\n  timeSlicing:\n    resources:\n    - name: gpu\n      rename: gpu-shared\n      replicas: 4\n",
     [("nvidia.com/gpu", 4, "")]),
    ("renameByDefault keeps .shared", "sharing:\n  timeSlicing:\n    renameByDefault: true\n    resources:\n    - name: nvidia.com/gpu\n      replicas: 3\n",
     [("nvidia.com/gpu", 3, "nvidia.com/gpu.shared")]),
    ("mig resource", "version: v1\nsharing:\n  timeSlicing:\n    resources:\n    - name: nvidia.com/mig-1g.23gb\n      devices: [0, 1]\n      replicas: 5\n",
     [("nvidia.com/mig-1g.23gb", 5, "")]),
    ("json form", json.dumps({"version": "v1", "sharing": {"timeSlicing": {"resources": [{"name": "nvidia.com/gpu", "devices": "all", "replicas": 2}]}}}),
     [("nvidia.com/gpu", 2, "")]),
]


@pytest.mark.parametrize("label,doc,ts", TS_CASES, ids=[c[0] for c in TS_CASES])
@pytest.mark.parametrize("node,strategy", [("c2_1xB200", "none"), ("c4_single_8x7x1g23", "single"), ("c5_mixed_hetero", "mixed"),
                                           ("c4_invalid_one_empty", "single")])
def test_time_slicing(label, doc, ts, node, strategy, tmp_path):
    cfg = write_config(tmp_path, doc)
    want, _ = oracle_pass(NODES[node], strategy, time_slicing=[O.ReplicatedResource(n, r, rn) for n, r, rn in ts])
    rc, got, stderr, _ = cli_pass(tmp_path, NODES[node], strategy, extra_args=["--config-file", cfg])
    assert rc == 0, stderr
    assert got == want


# ---- vGPU through a sysfs tree -----------------------------------------------------------------------------
VG = json.load(open(os.path.join(GOLDEN, "vgpu_blobs.json")))


def make_pci_tree(tmp_path, devices):
    root = os.path.join(str(tmp_path), "pci")
    for addr, (vendor, klass, config) in devices.items():
        d = os.path.join(root, addr)
        os.makedirs(d)
        open(os.path.join(d, "vendor"), "w").write(vendor + "\n")
        if klass is not None:
            open(os.path.join(d, "class"), "w").write(klass + "\n")
        if config is not None:
            open(os.path.join(d, "config"), "wb").write(config)
    os.makedirs(root, exist_ok=True)
    return root


def oracle_pci(root):
    return lambda: O.nvidia_pci_devices(root)


def vgpu_case(tmp_path, devices, node="c2_1xB200"):
    root = make_pci_tree(tmp_path, devices)
    cfg = O.Config(no_timestamp=True, machine_type_file="/x")
    try:
        labels, _ = O.run_pass(O.NvmlManager(M.RawNvml(NODES[node])), oracle_pci(root), cfg, read_file=lambda p: b"DGX B200\n")
        want = O.render(labels)
    except O.GfdError as e:
        want = None
    rc, got, stderr, _ = cli_pass(tmp_path, NODES[node], "none", pci_root=root)
    return want, rc, got, stderr


def test_vgpu_reference_blobs_in_sysfs(tmp_path):
    blobs = VG["blobs"]
    want, rc, got, stderr = vgpu_case(tmp_path, {
        "0000:00:01.0": ("0x8086", "0x060000", b"\x86\x80" + bytes(62)),                      # not NVIDIA: never read further
        "0000:3b:00.0": ("0x10de", "0x030200", bytes.fromhex(blobs["gpuPassThroughConfig"])),
        "0000:5e:00.0": ("0x10de", "0x030200", bytes.fromhex(blobs["vgpuConfig"])),
    })
    assert rc == 0 and got == want, stderr
    assert "nvidia.com/vgpu.present=true\n" in got
    assert "nvidia.com/vgpu.host-driver-version=%s\n" % VG["expect"]["host_driver_version"] in got
    assert "nvidia.com/vgpu.host-driver-branch=%s\n" % VG["expect"]["host_driver_branch"] in got
    _check_result(got, "expected-output-vgpu.txt", is_vgpu=True)  # the vGPU half of TestRunOneshot's checkResult


def test_vgpu_unprivileged_short_config_fails_the_pass(tmp_path):
    """pciutil.go:116-118: < 256 bytes of config space (unprivileged container) is an error for ANY 0x10de function."""
    want, rc, got, stderr = vgpu_case(tmp_path, {"0000:c0:00.0": ("0x10de", "0x030200", bytes.fromhex(VG["blobs"]["vgpuConfig"])[:64])})
    assert want is None and rc == 1 and got is None
    assert "privileged" in stderr


def test_vgpu_with_no_nvml_devices_still_labels(tmp_path):
    want, rc, got, stderr = vgpu_case(tmp_path, {"0000:5e:00.0": ("0x10de", "0x030200", bytes.fromhex(VG["blobs"]["vgpuConfig"]))}, node="zero_gpus")
    assert rc == 0 and got == want and got.count("\n") == 3, stderr


def _cfg_with_cap(cap_bytes, cap_pos=0xC8):
    cfg = bytearray(256)
    cfg[0:2] = b"\xde\x10"
    cfg[0x06] = 0x10
    cfg[0x34] = cap_pos
    cfg[cap_pos:cap_pos + len(cap_bytes)] = cap_bytes
    return bytes(cfg[:256]) if cap_pos + len(cap_bytes) <= 256 else bytes(cfg)


VGPU_EDGE = {
    "no capability list bit": bytes(6) + b"\x00" + bytes(249),
    "chain loops": (lambda c: bytes(c))(bytearray(_cfg_with_cap(bytes([0x01, 0xC8, 0x08])))),
    "chain broken 0xff": _cfg_with_cap(bytes([0xFF, 0x00, 0x08])),
    "vendor cap, not VF": _cfg_with_cap(bytes([0x09, 0x00, 0x14, 0x01]) + bytes(16)),
    "VF, high bytes in version": _cfg_with_cap(bytes([0x09, 0x00, 0x1B, 0x56, 0x46, 0x00, 0x16]) + b"5\xe9\xff.1\x80\x00\x00\x00\x00" + b"r\xc3\xa9_00\x00\x00\x00\x00"),
    "VF, NUL inside the version (Go keeps it; only the ends are trimmed)": _cfg_with_cap(bytes([0x09, 0x00, 0x1B, 0x56, 0x46, 0x00, 0x16]) + b"\x005\x0025.1\x00\x00\x00" + b"r5\x0025\x00\x00\x00\x00\x00"),
    "VF, two records before version": _cfg_with_cap(bytes([0x09, 0x00, 0x23, 0x56, 0x46, 0x01, 0x04, 0xAA, 0xBB, 0x02, 0x04, 0xCC, 0xDD, 0x00, 0x16]) + b"535.54.03\x00" + b"r535_00\x00\x00\x00"),
    "VF, version record truncated": _cfg_with_cap(bytes([0x09, 0x00, 0x10, 0x56, 0x46, 0x00, 0x16]) + b"535.54.03"),
    "VF, zero-length record (reference would spin)": _cfg_with_cap(bytes([0x09, 0x00, 0x1B, 0x56, 0x46, 0x05, 0x00]) + bytes(20)),
    "VF, record chain runs off the end": _cfg_with_cap(bytes([0x09, 0x00, 0x08, 0x56, 0x46, 0x05, 0x40, 0x00])),
    "VF, capability exactly 5 bytes": _cfg_with_cap(bytes([0x09, 0x00, 0x05, 0x56, 0x46])),
    "cap pointer at 0xFE": (lambda c: bytes(c))(bytearray(_cfg_with_cap(b"", 0xFE))),
    "cap slice past len but within ReadFile cap": _cfg_with_cap(bytes([0x09, 0x00, 0x40, 0x56, 0x46, 0x00, 0x16]) + b"470.82.01\x00" + b"r470_00\x00\x00\x00", 0xE0),
}


@pytest.mark.parametrize("label", sorted(VGPU_EDGE), ids=sorted(VGPU_EDGE))
def test_vgpu_edge_cases(label, tmp_path):
    want, rc, got, stderr = vgpu_case(tmp_path, {"0000:5e:00.0": ("0x10de", "0x030200", VGPU_EDGE[label])})
    if want is None:
        assert rc == 1 and got is None, stderr
    else:
        assert rc == 0 and got == want, stderr


def test_missing_pci_root_fails(tmp_path):
    rc, got, stderr, _ = cli_pass(tmp_path, NODES["c2_1xB200"], "none", pci_root=os.path.join(str(tmp_path), "nope"))
    assert rc == 1 and got is None and "unable to read PCI bus devices" in stderr


def test_nvidia_function_without_class_file_fails(tmp_path):
    want, rc, got, stderr = vgpu_case(tmp_path, {"0000:5e:00.0": ("0x10de", None, bytes(256))})
    assert want is None and rc == 1


# ---- --current-tables (NEW, opt-in): the only flag that changes reference labels, and only when given ---------------
def test_current_tables_is_opt_in_and_resolves_new_mig_profiles(tmp_path):
    node = NODES["gi_profile_id_9_fails"]  # B200 1g.45gb: GI profile id 9, beyond the reference's NVML 11.x header
    rc, got, stderr, _ = cli_pass(tmp_path, node, "mixed")
    assert rc == 1 and got is None and "no matching profile IDs found" in stderr          # default = the reference's failure
    rc, got, stderr, _ = cli_pass(tmp_path, node, "mixed", extra_args=["--current-tables"])
    assert rc == 0, stderr
    assert "nvidia.com/mig-1g.45gb.count=1\n" in got and "nvidia.com/mig-1g.45gb.product=NVIDIA-B200-MIG-1g.45gb\n" in got
    assert "nvidia.com/gpu.family=blackwell\n" in got
    # on a node the reference handles, the flag changes nothing but the family label
    want, _ = oracle_pass(NODES["c5_mixed_hetero"], "mixed")
    rc, got, stderr, _ = cli_pass(tmp_path, NODES["c5_mixed_hetero"], "mixed", extra_args=["--current-tables"])
    assert rc == 0 and got == want.replace("gpu.family=undefined", "gpu.family=blackwell"), stderr


def test_current_tables_profile_names():
    MIBS = 183359 * 1024 * 1024
    cases = {(8, 1): "2g.45gb+me", (9, 0): "1g.45gb", (0xA, 0): "1g.23gb+gfx", (0xB, 1): "2g.45gb+gfx", (0xC, 3): "4g.90gb+gfx",
             (0xD, 0): "1g.23gb-me", (0xE, 0): "1c.2g.45gb-me", (0xF, 7): "1g.23gb+me.all", (0x10, 1): "2g.45gb+me.all"}
    for (gi, ci), want in cases.items():
        mb = {1: 22912, 2: 45824, 4: 91648}[int(want.split("g.")[0][-1])] if "1g.45gb" not in want else 45824
        with pytest.raises(B.GfdError):
            B.mig_profile_string(gi, ci, mb, MIBS)  # the plain C-ABI call keeps the reference's tables


# ---- the node the reference's README shows (README.md:138-150) ----------------------------------------------------
def test_readme_node_example(tmp_path):
    """The only VALUES the reference documents for the NVML boundary: an A100 node's labels in its README.  A node with that
    GPU (device name, 39538 MiB, cc 8.0), that two-part driver version and CUDA 11.1 must give every label shown there --
    from the oracle and from the binary.  Differences that are the README's age, not ours: the listing predates
    `gpu.replicas` / `mig.capable` / `mig.strategy` (extra labels are fine) and shows the machine name before
    machine-type.go:37 turned spaces into dashes."""
    ex = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "readme_node_example.json")))["labels"]
    node = M.NodeFix(driver_version="%s.%s" % (ex["nvidia.com/cuda.driver.major"], ex["nvidia.com/cuda.driver.minor"]),
                     cuda_driver_version=int(ex["nvidia.com/cuda.runtime.major"]) * 1000 + int(ex["nvidia.com/cuda.runtime.minor"]) * 10,
                     gpus=[M.GpuFix(name=ex["nvidia.com/gpu.product"], memory_total=int(ex["nvidia.com/gpu.memory"]) << 20,
                                    cc=(int(ex["nvidia.com/gpu.compute.major"]), int(ex["nvidia.com/gpu.compute.minor"])), mig_mode=(0, 0))])
    machine = ex["nvidia.com/gpu.machine"].encode() + b"\n"
    want, _ = oracle_pass(node, "none", machine=machine)
    rc, got, stderr, _ = cli_pass(tmp_path, node, "none", machine=machine)
    assert rc == 0 and got == want, stderr
    lines = dict(l.split("=", 1) for l in got.splitlines())
    for k, v in ex.items():
        if k == "nvidia.com/gfd.timestamp":
            continue
        if k == "nvidia.com/gpu.machine":
            v = v.replace(" ", "-")
        assert lines[k] == v, (k, lines.get(k), v)


# ---- the reference's run()-level tests (cmd/gpu-feature-discovery/main_test.go, mig_test.go) through the BINARY ------
# The reference drives run() with mock devices; the binary has no mock seam, so each scenario is rebuilt as a simulated NVML
# node with the same shape (an A100-like GPU called MOCKMODEL; MIG devices 3g.20gb / 1g.5gb) and checked the way
# checkResult (main_test.go:403-435) checks: every output line must match >= 1 regex of the scenario's expected-output file,
# plus the scenario's own spot assertions wherever they do not depend on the mocks' made-up memory sizes.
_REGEX = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "expected_output_regexes.json")))["files"]


def _check_result(text, expected_file, is_vgpu=False):
    import re
    regexps = [re.compile(l) for l in _REGEX[expected_file]]
    for line in text.rstrip("\n").split("\n"):
        if ("vgpu" in line) != is_vgpu:
            continue
        assert any(r.search(line) for r in regexps), "line does not match any regexp of %s: %s" % (expected_file, line)


def _mock_like(mig=None, enabled=None):
    g = N.a100_gpu(mig=mig)
    g.name = "MOCKMODEL"
    if enabled is not None:
        g.mig_mode = (1, 1) if enabled else (0, 0)
    return M.NodeFix(gpus=[g])


def _run_binary(tmp_path, node, strategy, timestamp=False):
    rc, got, stderr, out = cli_pass(tmp_path, node, strategy, machine=b"product-name\n",
                                    extra_args=["--no-timestamp=false"] if timestamp else [])
    assert rc == 0, stderr
    return got, dict(l.split("=", 1) for l in got.splitlines())


def test_binary_run_oneshot(tmp_path):  # TestRunOneshot, main_test.go:91-134 (its vGPU half: test_vgpu_reference_blobs_in_sysfs)
    got, labels = _run_binary(tmp_path, _mock_like(), "none", timestamp=True)
    _check_result(got, "expected-output.txt")
    assert len(labels) == 15 and labels["nvidia.com/gpu.machine"] == "product-name" and "nvidia.com/gfd.timestamp" in labels


def test_binary_run_with_no_timestamp(tmp_path):  # TestRunWithNoTimestamp, main_test.go:136-180
    got, labels = _run_binary(tmp_path, _mock_like(), "none")
    _check_result(got, "expected-output.txt")
    assert "nvidia.com/gfd.timestamp" not in labels


def test_binary_mig_strategy_none(tmp_path):  # TestMigStrategyNone, mig_test.go:17-72
    got, labels = _run_binary(tmp_path, _mock_like(mig=[(2, 2), (2, 2)]), "none")
    _check_result(got, "expected-output-mig-none.txt")
    assert labels["nvidia.com/gpu.count"] == "1" and labels["nvidia.com/gpu.product"] == "MOCKMODEL"


def test_binary_mig_strategy_single_no_mig_devices(tmp_path):  # TestMigStrategySingleForNoMigDevices, mig_test.go:74-126
    got, labels = _run_binary(tmp_path, _mock_like(), "single")
    _check_result(got, "expected-output-mig-single.txt")
    assert labels["nvidia.com/mig.strategy"] == "single" and labels["nvidia.com/gpu.count"] == "1"
    assert labels["nvidia.com/gpu.product"] == "MOCKMODEL"


def test_binary_mig_strategy_single_mig_disabled(tmp_path):  # TestMigStrategySingleForMigDeviceMigDisabled, mig_test.go:128-186
    got, labels = _run_binary(tmp_path, _mock_like(mig=[(2, 2), (2, 2)], enabled=False), "single")
    _check_result(got, "expected-output-mig-single.txt")
    assert labels["nvidia.com/gpu.count"] == "1" and labels["nvidia.com/gpu.product"] == "MOCKMODEL"


def test_binary_mig_strategy_single(tmp_path):  # TestMigStrategySingle, mig_test.go:188-245
    got, labels = _run_binary(tmp_path, _mock_like(mig=[(2, 2), (2, 2)]), "single")
    _check_result(got, "expected-output-mig-single.txt")
    assert labels["nvidia.com/mig.strategy"] == "single" and labels["nvidia.com/gpu.count"] == "2"
    assert labels["nvidia.com/gpu.product"] == "MOCKMODEL-MIG-3g.20gb"


def test_binary_mig_strategy_mixed(tmp_path):  # TestMigStrategyMixed, mig_test.go:247-306
    got, labels = _run_binary(tmp_path, _mock_like(mig=[(2, 2), (0, 0)]), "mixed")
    _check_result(got, "expected-output-mig-mixed.txt")
    assert labels["nvidia.com/mig.strategy"] == "mixed" and labels["nvidia.com/gpu.count"] == "1"
    assert labels["nvidia.com/gpu.product"] == "MOCKMODEL"
    assert "nvidia.com/mig-3g.20gb.count" in labels and "nvidia.com/mig-1g.5gb.count" in labels


# ---- the logged error chain follows the reference's fmt.Errorf wrappers (internal/lm level) -----------------------------
ERROR_CHAINS = [
    # (node, strategy, chain as cmd/gpu-feature-discovery/main.go:84-89 logs it)
    ("err_count", "none", "error creating NVML labeler: error getting devices: "),                                   # labeler.go:36, nvml.go:37
    ("err_driver_version", "none", "error creating NVML labeler: failed to construct version labeler: error getting driver version: "),  # nvml.go:51,78
    ("err_memory_on_representative_gpu", "none", "error creating NVML labeler: error creating resource labeler: failed to construct GPU labeler: "
     "failed to construct labeler: failed to get memory info for device: "),                # nvml.go:61, mig-strategy.go:58,173, resource.go:48
    ("err_mig_mode_gpu1", "none", "error creating NVML labeler: error creating resource labeler: failed to construct GPU labeler: "
     "error getting map of devices: "),                                                     # mig-strategy.go:58,119
    ("err_cc", "none", "error creating NVML labeler: error creating resource labeler: failed to construct GPU labeler: failed to construct labeler: "
     "failed to create architecture labels: failed to determine CUDA compute capability: "),  # resource.go:58,242
    ("err_max_mig_count", "single", "error creating NVML labeler: error creating resource labeler: failed to construct MIG resource labeler: "
     "failed to create labeler for mig-strategy=single: failed to check for empty MIG-enabled devices: "),  # mig-strategy.go:69,95,194
    ("err_mig_handle", "mixed", "error creating NVML labeler: error creating resource labeler: failed to construct MIG resource labeler: "
     "failed to create labeler for mig-strategy=mixed: unable to retrieve list of MIG devices: "),           # mig-strategy.go:69,100,272
    ("err_gi_profile_info", "single", "error creating NVML labeler: error creating resource labeler: failed to construct MIG resource labeler: "
     "failed to create labeler for mig-strategy=single: unable to get MIG device name: failed to get MIG profile: "),  # :218, nvml-mig-device.go:99
    ("c2_1xB200", "bogus", "error creating NVML labeler: error creating resource labeler: failed to construct MIG resource labeler: "
     "unknown strategy: bogus"),                                                            # mig-strategy.go:69,102
]


@pytest.mark.parametrize("name,strategy,chain", ERROR_CHAINS, ids=["%s-%s" % c[:2] for c in ERROR_CHAINS])
def test_error_chain_reads_like_the_reference(name, strategy, chain, tmp_path):
    rc, got, stderr, _ = cli_pass(tmp_path, NODES[name], strategy)
    assert rc == 1 and got is None
    errors = [l for l in stderr.splitlines() if " Error: " in l]
    assert errors and ("Error: " + chain) in errors[-1], errors
