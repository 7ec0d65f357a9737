"""C-ABI Manager/Device surface of libgfd_b200.so vs the oracle, on simulated nodes (CPU only).

Both sides see the same `.nvmlfix` node: the oracle through RawNvml (Python), the shim through the fake
libnvidia-ml.so.1 (C).  The oracle's label builders are run over BOTH managers, so any difference in a name,
a profile string, an attribute, a count or in *whether an error occurs* shows up as a label diff."""

import pytest

import nodes as N
from oracle import gfd_oracle as O, nvml_model as M
from shimutil import B, ShimManager, load_fake, use_node

NODES = N.nodes()
CASES = [(n, s) for n in sorted(NODES) for s in N.STRATEGIES]


def oracle_result(node, strategy, time_slicing=()):
    cfg = O.Config(mig_strategy=strategy, no_timestamp=True, machine_type_file="", time_slicing=time_slicing)
    raw = M.RawNvml(node)
    try:
        labels, _ = O.run_pass(O.NvmlManager(raw), [], cfg)
        return O.render(labels), raw
    except O.GfdError:
        return None, raw


def shim_result(strategy, time_slicing=(), **kw):
    cfg = O.Config(mig_strategy=strategy, no_timestamp=True, machine_type_file="", time_slicing=time_slicing)
    mgr = ShimManager(**kw)
    try:
        labels, _ = O.run_pass(mgr, [], cfg)
        return O.render(labels), mgr
    except O.GfdError:
        return None, mgr
    finally:
        pass


@pytest.mark.parametrize("name,strategy", CASES, ids=["%s-%s" % c for c in CASES])
def test_shim_matches_oracle(name, strategy, tmp_path):
    node = NODES[name]
    if node.hidden:
        pytest.skip("needs the *_nomig fake build; covered by the CLI tests")
    load_fake()
    use_node(node, tmp_path)
    want, _ = oracle_result(node, strategy)
    got, mgr = shim_result(strategy)
    mgr.m.close()
    assert got == want


@pytest.mark.parametrize("name", ["c3_8xB200", "c4_single_8x7x1g23", "c5_mixed_hetero", "err_memory_on_non_representative_gpu",
                                  "err_mig_attrs_second", "gi_id9_but_other_gpu_empty", "hetero_models"])
@pytest.mark.parametrize("strategy", ["none", "single", "mixed"])
def test_prefetch_threads_do_not_change_results(name, strategy, tmp_path):
    """gfd_prefetch caches errors instead of raising them, so the failing set is still the reference's."""
    load_fake()
    node = NODES[name]
    use_node(node, tmp_path)
    want, _ = oracle_result(node, strategy)
    got, mgr = shim_result(strategy, prefetch_threads=8)
    mgr.m.close()
    assert got == want


def test_keep_attached_manager_sees_node_changes(tmp_path):
    """GFD_MGR_KEEP_ATTACHED holds NVML open across passes; every pass still re-enumerates."""
    load_fake()
    use_node(NODES["c2_1xB200"], tmp_path)
    mgr = ShimManager(flags=B.MGR_KEEP_ATTACHED)
    cfg = O.Config(no_timestamp=True, machine_type_file="")
    a, _ = O.run_pass(mgr, [], cfg)
    b, _ = O.run_pass(mgr, [], cfg)
    assert a == b and a["nvidia.com/gpu.count"] == "1"
    mgr.m.close()


def test_call_counts_vs_reference_model(tmp_path):
    """SURVEY.md 3.2 call-count model: the reference needs ~92 NVML calls + 10 dlopen probes for 8 GPUs, MIG none;
    the shim needs one walk.  The oracle's trace IS the reference's call sequence."""
    load_fake()
    node = NODES["c3_8xB200"]
    use_node(node, tmp_path)
    _, raw = oracle_result(node, "none")
    ref_calls = [t for t in raw.trace if not t.startswith("dlsym:")]
    ref_probes = [t for t in raw.trace if t.startswith("dlsym:")]
    mgr = ShimManager()
    cfg = O.Config(no_timestamp=True, machine_type_file="")
    mgr.init()
    O.new_resource_labeler(mgr, cfg)
    O.new_version_labeler(mgr)
    O.new_mig_capability_labeler(mgr)
    calls, probes = mgr.m.call_stats()
    mgr.shutdown()
    mgr.m.close()
    assert len(ref_calls) == 92 and len(ref_probes) == 9
    assert calls <= 32 and probes <= 8


def test_null_manager():
    m = B.Manager(B.BACKEND_NULL)
    m.init()
    assert m.get_devices() == []
    with pytest.raises(B.GfdError):
        m.get_driver_version()
    with pytest.raises(B.GfdError):
        m.get_cuda_driver_version()
    m.shutdown()
    m.close()


def test_getters_outside_bracket_fail(tmp_path):
    load_fake()
    use_node(NODES["c2_1xB200"], tmp_path)
    m = B.Manager()
    with pytest.raises(B.GfdError):
        m.get_devices()
    m.init()
    d = m.get_devices()[0]
    assert d.get_name() == "NVIDIA B200"
    with pytest.raises(B.GfdError):
        d.get_attributes()  # "GetAttributes is not supported for non-MIG devices"
    with pytest.raises(B.GfdError):
        d.get_device_handle_from_mig_device_handle()
    m.shutdown()
    with pytest.raises(B.GfdError):
        d.get_name()
    m.close()


def test_mig_device_kind_errors(tmp_path):
    load_fake()
    use_node(NODES["c4_single_1x2x3g90"], tmp_path)
    m = B.Manager()
    m.init()
    mig = m.get_devices()[0].get_mig_devices()[0]
    assert mig.get_name() == "3g.90gb" and mig.get_total_memory_mb() == 91648
    assert mig.get_device_handle_from_mig_device_handle().get_name() == "NVIDIA B200"
    for fn in (mig.is_mig_capable, mig.is_mig_enabled, mig.get_mig_devices, mig.get_cuda_compute_capability):
        with pytest.raises(B.GfdError):
            fn()
    m.shutdown()
    m.close()
