"""The Tegra path: no NVML, /etc/nv_tegra_release present -> CUDA-driver manager (internal/resource/factory.go:59-70,
cuda-lib.go, cuda-device.go, internal/cuda/*).  Product = gpu-feature-discovery over a fake libcuda.so.1; checker = the
oracle's CudaManager over the same fixture.  (Parity for this path is unpinned: no reference test touches it and there
is no Tegra hardware here; the fake's one modelling choice, what the legacy cuDeviceTotalMem stores above 4 GiB, follows the
real libcuda of a B200 box -- tests/test_gpu_cuda_backend.py.)"""
import os

import pytest

from oracle import gfd_oracle as O, nvml_model as M
from shimutil import B, run_cli

CUDA_DIR = os.path.join(os.path.dirname(B.LIB_PATH), "fake_cuda", "lib")
GIB = 1 << 30

CASES = {
    "orin": M.RawCuda([(3 * GIB + 123456789, 8, 7, "Orin (nvgpu)")], version=12020),
    "xavier two devices": M.RawCuda([(2 * GIB, 7, 2, "Xavier"), (2 * GIB, 7, 2, "Xavier")], version=11040),
    "memory above 4 GiB saturates in the legacy call": M.RawCuda([(61 * GIB + 5 * (1 << 20), 8, 7, "Orin (nvgpu)")], version=12060),
    "no devices": M.RawCuda([], version=12020),
    "init fails": M.RawCuda([(GIB, 8, 7, "Orin")], rc={"cuInit": 100}),
    "count fails": M.RawCuda([(GIB, 8, 7, "Orin")], rc={"cuDeviceGetCount": 3}),
    "name fails": M.RawCuda([(GIB, 8, 7, "Orin")], rc={"cuDeviceGetName": 999}),
    "attribute fails": M.RawCuda([(GIB, 8, 7, "Orin")], rc={"cuDeviceGetAttribute": 1}),
    "memory fails": M.RawCuda([(GIB, 8, 7, "Orin")], rc={"cuDeviceTotalMem": 2}),
    "version fails": M.RawCuda([(GIB, 8, 7, "Orin")], rc={"cuDriverGetVersion": 999}),
}


def tegra_root(tmp_path):
    root = os.path.join(str(tmp_path), "root")
    os.makedirs(os.path.join(root, "etc"))
    open(os.path.join(root, "etc", "nv_tegra_release"), "w").write("# R35 (release), REVISION: 4.1\n")
    return root


@pytest.mark.parametrize("name", sorted(CASES))
@pytest.mark.parametrize("strategy,foie", [("none", True), ("single", True), ("mixed", False)])
def test_tegra_label_file_matches_oracle(name, strategy, foie, tmp_path):
    raw = CASES[name]
    cfg = O.Config(mig_strategy=strategy, no_timestamp=True, machine_type_file="", fail_on_init_error=foie)
    try:
        labels, _ = O.run_pass(O.with_config(O.CudaManager(raw), cfg), [], cfg)
        want = O.render(labels)
    except O.GfdError:
        want = None
    fix = os.path.join(str(tmp_path), "cuda.fix")
    open(fix, "w").write(raw.dumps())
    empty = os.path.join(str(tmp_path), "pci")
    os.makedirs(empty)
    out = os.path.join(str(tmp_path), "gfd")
    rc, _, se = run_cli(["--oneshot", "--no-timestamp", "--machine-type-file=", "--mig-strategy=" + strategy,
                         "--fail-on-init-error=%s" % str(foie).lower(), "-o", out], fake=False,
                        env={"GFD_FAKE_CUDA": fix, "GFD_B200_TEGRA_ROOT": tegra_root(tmp_path), "GFD_B200_PCI_DEVICES_ROOT": empty,
                             "LD_LIBRARY_PATH": CUDA_DIR})
    got = open(out).read() if os.path.exists(out) else None
    assert "Detected non-NVML platform" in se and "Detected Tegra platform" in se and "Using CUDA manager" in se
    if want is None:
        assert rc == 1 and got is None, se
    else:
        assert rc == 0 and got == want, se


def test_tegra_labels_spot_values(tmp_path):
    raw = CASES["orin"]
    cfg = O.Config(no_timestamp=True, machine_type_file="")
    labels, _ = O.run_pass(O.CudaManager(raw), [], cfg)
    assert labels["nvidia.com/cuda.driver.major"] == "unknown" and labels["nvidia.com/cuda.driver.rev"] == "unknown"
    assert labels["nvidia.com/cuda.runtime.major"] == "12" and labels["nvidia.com/cuda.runtime.minor"] == "2"
    assert labels["nvidia.com/gpu.product"] == "Orin-(nvgpu)" and labels["nvidia.com/mig.capable"] == "false"
    assert labels["nvidia.com/gpu.family"] == "ampere" and labels["nvidia.com/gpu.memory"] == str((3 * GIB + 123456789) // (1 << 20))
