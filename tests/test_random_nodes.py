"""Differential testing on RANDOM simulated nodes (hypothesis): product == oracle, byte for byte.

Each example draws a node (0-8 GPUs of 1-3 models, MIG off / enabled-empty / enabled with 1-7 instances of random
GI+CI profiles incl. ids the reference cannot resolve, random per-call NVML failures), a MIG strategy and a
time-slicing config; the oracle (Python, RawNvml) and the product (C ABI over the C fake NVML, driven in-process
with per-GPU threads ON) must agree on the rendered label file or on failing.  The CLI binary is spot-checked on
a subset to keep the CPU suite fast."""
import os

from hypothesis import HealthCheck, given, settings, strategies as st

from oracle import gfd_oracle as O, nvml_model as M
from shimutil import ShimManager, load_fake, run_cli, use_node

MIB = 1024 * 1024
MODELS = [("NVIDIA B200", 183359 * MIB, (10, 0)), ("NVIDIA A100-SXM4-40GB", 40536 * MIB, (8, 0)), ("NVIDIA H100 80GB HBM3", 81559 * MIB, (9, 0)),
          ("Tesla T4", 15360 * MIB, (7, 5)), ("NVIDIA DGX Display", 4096 * MIB, (7, 5)), ("Odd  Name", 0, (0, 0))]
FAILABLE = ["nvmlDeviceGetName", "nvmlDeviceGetMemoryInfo", "nvmlDeviceGetCudaComputeCapability", "nvmlDeviceGetMigMode",
            "nvmlDeviceGetMaxMigDeviceCount", "nvmlDeviceGetGpuInstanceProfileInfo", "nvmlGpuInstanceGetComputeInstanceProfileInfo",
            "nvmlDeviceGetGpuInstanceById", "nvmlGpuInstanceGetInfo", "nvmlGpuInstanceGetComputeInstanceById",
            "nvmlComputeInstanceGetInfo_v2", "nvmlDeviceGetHandleByIndex_v2"]
MIG_FAILABLE = ["nvmlDeviceGetAttributes_v2", "nvmlDeviceGetGpuInstanceId", "nvmlDeviceGetComputeInstanceId",
                "nvmlDeviceGetDeviceHandleFromMigDeviceHandle", "nvmlDeviceIsMigDeviceHandle", "nvmlDeviceGetMigDeviceHandleByIndex"]
CODES = [999, 3, 2, 6, 15]


@st.composite
def gpu_fix(draw):
    name, total, cc = draw(st.sampled_from(MODELS))
    state = draw(st.sampled_from(["off", "off", "unsupported", "empty", "mig", "mig"]))
    if state == "unsupported":
        g = M.GpuFix(name=name, memory_total=total, cc=cc, mig_mode=None)
    elif state == "off":
        g = M.b200_gpu(None, name=name, total=total or 1, cc=cc)
        g.memory_total = total
    else:
        n = 0 if state == "empty" else draw(st.integers(1, 7))
        prof = []
        for _ in range(n):
            gp = draw(st.sampled_from([0, 0, 1, 2, 3, 4, 7, 9]))
            gsl = M.B200_GI_PROFILES[gp][0]
            cp = draw(st.sampled_from([c for c, s in ((0, 1), (1, 2), (2, 3), (3, 4), (4, 7)) if s <= gsl]))
            prof.append((gp, cp))
        g = M.b200_gpu(prof, name=name, total=total or (1 << 30), cc=cc)
        if draw(st.booleans()) and len(g.migs) >= 2:   # two compute instances inside one GPU instance
            g.migs[1].gi_id = g.migs[0].gi_id
            g.migs[1].gi_profile_id = g.migs[0].gi_profile_id
            g.migs[1].ci_id = 1
    if draw(st.integers(0, 9)) == 0:
        g.rc[draw(st.sampled_from(FAILABLE))] = draw(st.sampled_from(CODES))
    if g.migs and draw(st.integers(0, 9)) == 0:
        m = draw(st.sampled_from(g.migs))
        m.rc[draw(st.sampled_from(MIG_FAILABLE))] = draw(st.sampled_from(CODES))
    return g


@st.composite
def node_fix(draw):
    gpus = draw(st.lists(gpu_fix(), min_size=0, max_size=8))
    node = M.NodeFix(gpus=gpus, driver_version=draw(st.sampled_from(["575.57.08", "470.57", "535", "1.2.3.4", "", "550.54.15"])),
                     cuda_driver_version=draw(st.sampled_from([12090, 11080, 13000, 0, 12345])))
    if draw(st.integers(0, 14)) == 0:
        node.rc[draw(st.sampled_from(["nvmlInit_v2", "nvmlDeviceGetCount_v2", "nvmlSystemGetDriverVersion",
                                      "nvmlSystemGetCudaDriverVersion", "nvmlShutdown"]))] = draw(st.sampled_from(CODES))
    return node


TS = st.lists(st.tuples(st.sampled_from(["nvidia.com/gpu", "nvidia.com/mig-1g.23gb", "nvidia.com/mig-3g.90gb", "nvidia.com/other"]),
                        st.integers(2, 9), st.sampled_from(["", "", "x.shared"])), max_size=3)


def both(node, strategy, ts, fail_on_init_error, tmp_dir):
    cfg = O.Config(mig_strategy=strategy, no_timestamp=True, machine_type_file="", fail_on_init_error=fail_on_init_error,
                   time_slicing=[O.ReplicatedResource(n, r, rn) for n, r, rn in ts])
    try:
        labels, _ = O.run_pass(O.with_config(O.NvmlManager(M.RawNvml(node)), cfg), [], cfg)
        want = O.render(labels)
    except O.GfdError:
        want = None
    use_node(node, tmp_dir)
    mgr = ShimManager(prefetch_threads=-1)  # auto: one thread per GPU
    try:
        labels, _ = O.run_pass(O.with_config(mgr, cfg), [], cfg)
        got = O.render(labels)
    except O.GfdError:
        got = None
    finally:
        mgr.m.close()
    return want, got


@settings(max_examples=400, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow])
@given(node=node_fix(), strategy=st.sampled_from(["none", "single", "mixed", "bogus"]), ts=TS, foie=st.booleans())
def test_shim_equals_oracle_on_random_nodes(node, strategy, ts, foie, tmp_path_factory):
    load_fake()
    tmp = tmp_path_factory.mktemp("rnd")
    want, got = both(node, strategy, ts, foie, tmp)
    assert got == want, node.dumps()


@settings(max_examples=60, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow])
@given(node=node_fix(), strategy=st.sampled_from(["none", "single", "mixed"]), foie=st.booleans())
def test_cli_equals_oracle_on_random_nodes(node, strategy, foie, tmp_path_factory):
    tmp = str(tmp_path_factory.mktemp("rndcli"))
    cfg = O.Config(mig_strategy=strategy, no_timestamp=True, machine_type_file="", fail_on_init_error=foie)
    try:
        labels, _ = O.run_pass(O.with_config(O.NvmlManager(M.RawNvml(node)), cfg), [], cfg)
        want = O.render(labels)
    except O.GfdError:
        want = None
    fix = use_node(node, tmp)
    empty = os.path.join(tmp, "pci")
    os.makedirs(empty, exist_ok=True)
    out = os.path.join(tmp, "gfd")
    rc, _, se = run_cli(["--oneshot", "--no-timestamp", "--machine-type-file=", "--mig-strategy=" + strategy,
                         "--fail-on-init-error=%s" % ("true" if foie else "false"), "-o", out],
                        env={"GFD_FAKE_NVML": fix, "GFD_B200_PCI_DEVICES_ROOT": empty})
    got = open(out).read() if os.path.exists(out) else None
    if want is None:
        assert rc == 1 and got is None, (se, node.dumps())
    else:
        assert rc == 0 and got == want, (se, node.dumps())
