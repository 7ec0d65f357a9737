"""Pin the oracle against the reference's OWN golden vectors (CPU only).

The vectors under tests/golden/ are extracted mechanically from the reference's Go test tables by
tests/golden/extract_reference_vectors.py; each test below mirrors the Go test it is named after.
"""
import json
import os
import re

import pytest

from oracle import gfd_oracle as O

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
V = json.load(open(os.path.join(GOLDEN, "lm_vectors.json")))


def mk_device(d):
    migs = [O.MockMigDevice(m["gi"], m["ci"], m["gb"], m["attributes"]) for m in d["migs"]]
    return O.MockDevice(d["mig_enabled"], migs)


def mk_config(ts=None, **kw):
    res = [O.ReplicatedResource(r["name"], r["replicas"], r.get("rename", "")) for r in (ts or {}).get("resources", [])]
    return O.Config(time_slicing=res, **kw)


def ids(cases):
    return [c["description"] for c in cases]


# internal/lm/resource_test.go:27-132
@pytest.mark.parametrize("case", V["TestGPUResourceLabeler"]["cases"], ids=ids(V["TestGPUResourceLabeler"]["cases"]))
def test_gpu_resource_labeler(case):
    device = mk_device(V["TestGPUResourceLabeler"]["device"])
    labels = O.new_gpu_resource_labeler(mk_config(case["timeSlicing"]), device, case["count"])
    assert labels == (case["expectedLabels"] or {})


# internal/lm/resource_test.go:134-304
@pytest.mark.parametrize("case", V["TestMigResourceLabeler"]["cases"], ids=ids(V["TestMigResourceLabeler"]["cases"]))
def test_mig_resource_labeler(case):
    d = V["TestMigResourceLabeler"]["device"]
    mig = O.MockMigDevice(d["gi"], d["ci"], d["gb"], d["attributes"])
    O.MockDevice(True, [mig])  # rt.NewMigEnabledDevice(device) wires the parent
    labels = O.new_mig_resource_labeler(case["resourceName"], mk_config(case["timeSlicing"]), mig, case["count"])
    assert labels == (case["expectedLabels"] or {})


# internal/lm/mig-strategy_test.go:28-195
@pytest.mark.parametrize("case", V["TestMigStrategyNoneLabels"]["cases"], ids=ids(V["TestMigStrategyNoneLabels"]["cases"]))
def test_mig_strategy_none_labels(case):
    mgr = O.MockManager([mk_device(d) for d in case["devices"]])
    labels = O.new_resource_labeler(mgr, mk_config(case["timeSlicing"], mig_strategy="none"))
    assert labels == (case["expectedLabels"] or {})


# internal/lm/mig-strategy_test.go:197-402
@pytest.mark.parametrize("case", V["TestMigStrategySingleLabels"]["cases"], ids=ids(V["TestMigStrategySingleLabels"]["cases"]))
def test_mig_strategy_single_labels(case):
    mgr = O.MockManager([mk_device(d) for d in case["devices"]])
    labels = O.new_resource_labeler(mgr, mk_config(mig_strategy="single"))
    assert labels == (case["expectedLabels"] or {})


# internal/lm/nvml_test.go:11-77
@pytest.mark.parametrize("case", V["TestMigCapabilityLabeler"]["cases"], ids=ids(V["TestMigCapabilityLabeler"]["cases"]))
def test_mig_capability_labeler(case):
    mgr = O.MockManager([mk_device(d) for d in case["devices"]])
    assert O.new_mig_capability_labeler(mgr) == (case["expectedLabels"] or {})


def test_mock_constants_match_reference_fixture():
    m = V["mock"]
    d = O.MockDevice(False)
    assert d.get_name() == m["name"] and d.get_total_memory_mb() == m["total_memory_mb"]
    assert list(d.get_cuda_compute_capability()) == m["cc_full"]
    assert list(O.MockDevice(True).get_cuda_compute_capability()) == m["cc_mig_enabled"]
    mgr = O.MockManager()
    assert mgr.get_driver_version() == m["driver_version"]
    assert list(mgr.get_cuda_driver_version()) == [m["cuda_major"], m["cuda_minor"]]


# ---- vGPU known-answer blobs (internal/vgpu/vgpu_test.go:36-74, pciutil_test.go:26-42) -------------------
VG = json.load(open(os.path.join(GOLDEN, "vgpu_blobs.json")))


def _mock_pci():
    b = VG["blobs"]
    return [O.PCIDevice("passthrough", bytes.fromhex(b["gpuPassThroughConfig"]), class_="300"),
            O.PCIDevice("vgpu", bytes.fromhex(b["vgpuConfig"]), class_="300")]


def test_vgpu_capability_kats():
    e = VG["expect"]
    pt, vg = _mock_pci()
    assert len(pt.config) == len(vg.config) == e["config_len"]
    assert pt.config[0] | pt.config[1] << 8 == 0x10DE and vg.config[0] | vg.config[1] << 8 == 0x10DE
    cap_pt = O.get_vendor_specific_capability(pt)
    cap_vg = O.get_vendor_specific_capability(vg)
    assert len(cap_pt) == e["passthrough_cap_len"] and not O.is_vgpu_device(cap_pt)
    assert len(cap_vg) == e["vgpu_cap_len"] and cap_vg[0] == e["vgpu_cap_id"] and O.is_vgpu_device(cap_vg)
    ver, br = O.vgpu_get_info(cap_vg, "vgpu")
    assert ver.decode() == e["host_driver_version"] and br.decode() == e["host_driver_branch"]


def test_vgpu_labels_from_mock():
    labels = O.vgpu_labels(_mock_pci())
    assert labels == {"nvidia.com/vgpu.present": "true",
                      "nvidia.com/vgpu.host-driver-version": VG["expect"]["host_driver_version"],
                      "nvidia.com/vgpu.host-driver-branch": VG["expect"]["host_driver_branch"]}


# ---- run()-level goldens (cmd/gpu-feature-discovery/main_test.go, mig_test.go) ---------------------------
def check_result(text, expected_file, is_vgpu):
    """checkResult (main_test.go:403-435): every (non-)vgpu line must match >= 1 regex (unanchored)."""
    regexps = [re.compile(l) for l in json.load(open(os.path.join(GOLDEN, "expected_output_regexes.json")))["files"][expected_file]]
    for line in text.rstrip("\n").split("\n"):
        if ("vgpu" in line) != is_vgpu:
            continue
        assert any(r.search(line) for r in regexps), "line does not match any regexp: %s" % line


def machine_file(_path):
    return b"product-name\n"  # setupMachineFile (main_test.go:452-456)


def run(devices, strategy="none", no_timestamp=False, fail_on_init_error=True, error_on_init=None):
    cfg = O.Config(mig_strategy=strategy, no_timestamp=no_timestamp, fail_on_init_error=fail_on_init_error,
                   machine_type_file="/tmp/machine-type")
    mgr = O.with_config(O.MockManager(devices, error_on_init), cfg)
    labels, _ = O.run_pass(mgr, _mock_pci(), cfg, timestamp=1700000000, read_file=machine_file)
    return labels, O.render(labels)


def test_run_oneshot():  # main_test.go:91-134
    labels, text = run([O.MockDevice(False)])
    check_result(text, "expected-output.txt", False)
    check_result(text, "expected-output-vgpu.txt", True)
    assert len(labels) == 15 + 3 and labels["nvidia.com/gpu.machine"] == "product-name"


def test_run_with_no_timestamp():  # main_test.go:136-180
    labels, text = run([O.MockDevice(False)], no_timestamp=True)
    check_result(text, "expected-output.txt", False)
    assert "nvidia.com/gfd.timestamp=" not in text


def test_mig_strategy_none_run():  # mig_test.go:17-72
    dev = O.MockDevice(True, [O.MockMigDevice(3, 0, 20000), O.MockMigDevice(3, 0, 20000)])
    labels, text = run([dev])
    check_result(text, "expected-output-mig-none.txt", False)
    assert labels["nvidia.com/gpu.count"] == "1" and labels["nvidia.com/gpu.product"] == "MOCKMODEL"
    assert labels["nvidia.com/gpu.memory"] == "300"


def test_mig_strategy_single_no_mig_devices():  # mig_test.go:74-126
    labels, text = run([O.MockDevice(False)], "single")
    check_result(text, "expected-output-mig-single.txt", False)
    assert labels["nvidia.com/mig.strategy"] == "single" and labels["nvidia.com/gpu.count"] == "1"
    assert labels["nvidia.com/gpu.product"] == "MOCKMODEL" and labels["nvidia.com/gpu.memory"] == "300"


def test_mig_strategy_single_mig_disabled():  # mig_test.go:128-186
    dev = O.MockDevice(False, [O.MockMigDevice(3, 0, 20000), O.MockMigDevice(3, 0, 20000)])
    labels, text = run([dev], "single")
    check_result(text, "expected-output-mig-single.txt", False)
    assert labels["nvidia.com/gpu.count"] == "1" and labels["nvidia.com/gpu.product"] == "MOCKMODEL"
    assert labels["nvidia.com/gpu.memory"] == "300"


def test_mig_strategy_single_run():  # mig_test.go:188-245
    dev = O.MockDevice(True, [O.MockMigDevice(3, 0, 20), O.MockMigDevice(3, 0, 20)])
    labels, text = run([dev], "single")
    check_result(text, "expected-output-mig-single.txt", False)
    assert labels["nvidia.com/mig.strategy"] == "single" and labels["nvidia.com/gpu.count"] == "2"
    assert labels["nvidia.com/gpu.product"] == "MOCKMODEL-MIG-3g.20gb" and labels["nvidia.com/gpu.memory"] == "20"


def test_mig_strategy_mixed_run():  # mig_test.go:247-306
    dev = O.MockDevice(True, [O.MockMigDevice(3, 0, 20), O.MockMigDevice(1, 0, 5)])
    labels, text = run([dev], "mixed")
    check_result(text, "expected-output-mig-mixed.txt", False)
    assert labels["nvidia.com/mig.strategy"] == "mixed" and labels["nvidia.com/gpu.count"] == "1"
    assert labels["nvidia.com/gpu.product"] == "MOCKMODEL" and labels["nvidia.com/gpu.memory"] == "300"
    assert "nvidia.com/mig-3g.20gb.count" in labels and "nvidia.com/mig-1g.5gb.count" in labels


INIT = json.load(open(os.path.join(GOLDEN, "init_error_matrix.json")))


@pytest.mark.parametrize("case", INIT["cases"], ids=ids(INIT["cases"]))
def test_fail_on_nvml_init_error(case):  # main_test.go:273-380
    err = "manager.Init error" if case["errorOnInit"] else None
    if case["expectError"]:
        with pytest.raises(O.GfdError):
            run([O.MockDevice(False)], case["migStrategy"], fail_on_init_error=case["failOnInitError"], error_on_init=err)
    else:
        run([O.MockDevice(False)], case["migStrategy"], fail_on_init_error=case["failOnInitError"], error_on_init=err)


def test_fallback_manager():  # internal/resource/fallback_test.go:26-62
    f = O.FallbackManager(O.MockManager([], error_on_init="init failed"))
    assert f.init() is None and f.shutdown() is None and f.get_devices() == []

    class BadShutdown(O.MockManager):
        def shutdown(self):
            raise O.GfdError("should not be called")

    f = O.FallbackManager(BadShutdown())
    f.init()
    with pytest.raises(O.GfdError, match="should not be called"):
        f.shutdown()


# ---- hand-derived vectors for the unpinned arithmetic -------------------------------------
MIB = 1024 * 1024


@pytest.mark.parametrize("total_mib,mig_mb,gb", [
    (40536, 4864, 5), (40536, 19968, 20), (40536, 40192, 40), (81920, 9728, 10), (81559, 20096, 20),
    (183359, 22912, 23), (183359, 45824, 45), (183359, 91648, 90), (183359, 182272, 180)])
def test_get_mig_memory_size_gb(total_mib, mig_mb, gb):
    assert O.get_mig_memory_size_gb(total_mib * MIB, mig_mb) == gb


def test_go_round_is_half_away_from_zero():
    assert [O.go_round(x) for x in (22.5, 23.5, -22.5, 0.49999999999999994, 2.4999, 0.5)] == [23, 24, -23, 0, 2, 1]


def test_go_trim_space_unicode():
    assert O.go_trim_space(b" \t\xc2\xa0DGX B200\xe2\x80\x83\r\n") == b"DGX B200"
    assert O.go_trim_space(b"\xa0x\xa0") == b"\xa0x\xa0"  # lone 0xA0 is invalid UTF-8 => RuneError, not space
    assert O.get_machine_type("/x", lambda p: b"NVIDIA DGX B200\n") == {"nvidia.com/gpu.machine": "NVIDIA-DGX-B200"}
    assert O.get_machine_type("") == {"nvidia.com/gpu.machine": "unknown"}
    assert O.get_machine_type("/definitely/not/here") == {"nvidia.com/gpu.machine": "unknown"}


def test_arch_family_is_stale_on_purpose():
    assert O.get_arch_family(10, 0) == "undefined" and O.get_arch_family(9, 0) == "hopper"
    assert O.get_arch_family(7, 0) == "volta" and O.get_arch_family(7, 5) == "turing" and O.get_arch_family(4, 0) == "undefined"
