"""The C-ABI library loads on a CPU-only host and exports every symbol include/gfd_b200.h declares (no compute)."""
import importlib
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "gfd_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(gfd_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_exactly_the_header():
    B = importlib.import_module("gpu-feature-discovery_b200.binding")
    declared = header_symbols()
    assert declared == sorted(B.SYMBOLS)
    nm = subprocess.run(["nm", "-D", "--defined-only", B.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = sorted(set(re.findall(r" T (gfd_[a-z0-9_]+)", nm)))
    assert exported == declared


def test_no_link_time_dependency_on_nvml_or_libcuda():
    B = importlib.import_module("gpu-feature-discovery_b200.binding")
    needed = subprocess.run(["readelf", "-d", B.LIB_PATH], capture_output=True, text=True, check=True).stdout
    assert "libnvidia-ml" not in needed and "libcuda.so" not in needed and "libcudart" not in needed


def test_binary_has_no_link_to_the_oracle_or_the_driver_libs():
    """The labeler carries the host-only C-ABI sources statically and dlopen()s libgfd_b200.so only for the probes."""
    B = importlib.import_module("gpu-feature-discovery_b200.binding")
    needed = subprocess.run(["readelf", "-d", B.BIN_PATH], capture_output=True, text=True, check=True).stdout
    for lib in ("oracle", "libnvidia-ml", "libcuda", "libcudart", "libstdc++"):
        assert lib not in needed
    strings = subprocess.run(["strings", "-a", B.BIN_PATH], capture_output=True, text=True, check=True).stdout
    assert "libgfd_b200.so" in strings and "ref_replay" not in strings and "gfd_oracle" not in strings


def test_probe_flag_without_a_gpu_fails_the_pass_loudly(tmp_path):
    """No CPU fallback: --probe-hbm on a host without a usable CUDA device is an error, not a silent skip."""
    import torch
    if torch.cuda.is_available():
        import pytest
        pytest.skip("GPU present")
    B = importlib.import_module("gpu-feature-discovery_b200.binding")
    out = os.path.join(str(tmp_path), "gfd")
    p = subprocess.run([B.BIN_PATH, "--oneshot", "--fail-on-init-error=false", "--probe-hbm", "-o", out], capture_output=True, text=True,
                       env={"PATH": os.environ.get("PATH", "")})
    assert p.returncode == 1 and not os.path.exists(out)
    assert "probe requested but no CUDA device is usable" in p.stderr


def test_gpu_entry_points_fail_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        import pytest
        pytest.skip("GPU present")
    B = importlib.import_module("gpu-feature-discovery_b200.binding")
    import pytest
    for call in (lambda: B.probe_hbm(0, 1 << 20), lambda: B.cuda_device_count(), lambda: B.checksum_host(0, 0, 0),
                 lambda: B.hbm_copy(16, 32, 16), lambda: B.hbm_fill_pattern(16, 16)):
        with pytest.raises(B.GfdError) as ei:
            call()
        assert ei.value.status == B.GFD_ERR_NO_GPU


def test_mig_arithmetic_matches_oracle_exhaustively_on_a_grid():
    from oracle import gfd_oracle as O
    B = importlib.import_module("gpu-feature-discovery_b200.binding")
    MIB = 1 << 20
    totals = [40536, 81920, 81559, 183359, 143771, 24576, 15360, 97871, 196608, 1, 1023, 1025]
    for t in totals:
        for mb in list(range(0, 200000, 977)) + [4864, 9728, 19968, 20096, 22912, 45824, 91648, 182272]:
            assert B.mig_memory_size_gb(t * MIB, mb) == O.get_mig_memory_size_gb(t * MIB, mb), (t, mb)
    for gi in range(-1, 12):
        for ci in range(-1, 9):
            try:
                want = O.new_mig_profile_string(gi, ci, 22912, 183359 * MIB)
            except O.GfdError:
                want = None
            try:
                got = B.mig_profile_string(gi, ci, 22912, 183359 * MIB)
            except B.GfdError:
                got = None
            assert got == want, (gi, ci)


def test_platform_detection_on_this_host(tmp_path):
    B = importlib.import_module("gpu-feature-discovery_b200.binding")
    ok, reason = B.is_tegra_system(str(tmp_path))
    assert not ok and reason.endswith("/sys/devices/soc0/family file not found")
    os.makedirs(os.path.join(str(tmp_path), "etc"))
    open(os.path.join(str(tmp_path), "etc", "nv_tegra_release"), "w").write("# R35\n")
    ok, reason = B.is_tegra_system(str(tmp_path))
    assert ok and reason.endswith("/etc/nv_tegra_release found")
    os.remove(os.path.join(str(tmp_path), "etc", "nv_tegra_release"))
    os.makedirs(os.path.join(str(tmp_path), "sys", "devices", "soc0"))
    fam = os.path.join(str(tmp_path), "sys", "devices", "soc0", "family")
    open(fam, "w").write("Tegra\n")
    assert B.is_tegra_system(str(tmp_path)) == (True, fam + " has 'tegra' prefix")
    open(fam, "w").write("jep106:0426\n")
    assert B.is_tegra_system(str(tmp_path)) == (False, fam + " has no 'tegra' prefix")
