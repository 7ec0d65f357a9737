"""Named simulated nodes (fixtures) shared by the CPU tests.  Each entry builds an oracle.nvml_model.NodeFix;
`NodeFix.dumps()` is the `.nvmlfix` text both the Python RawNvml and the C fake libnvidia-ml.so.1 interpret.
Values follow SURVEY.md 8(d): B200 = "NVIDIA B200", 183359 MiB, cc 10.0, driver 575.57.08, CUDA 12090."""
from oracle.nvml_model import NodeFix, GpuFix, MigFix, b200_gpu

MIB = 1024 * 1024


def a100_gpu(mig=None):
    """A100-SXM4-40GB (profile ids 0..4 + 7): the part the reference's own MIG numbers come from."""
    total = 40536 * MIB
    prof = {0: (1, 4864), 1: (2, 9856), 2: (3, 19968), 3: (4, 19968), 4: (7, 40192), 7: (1, 4864)}
    g = GpuFix(name="NVIDIA A100-SXM4-40GB", memory_total=total, cc=(8, 0), max_mig_count=7,
               mig_mode=(0, 0) if mig is None else (1, 1))
    g.gi_profiles = {i: (i, sl, mb) for i, (sl, mb) in prof.items()}
    ci = {0: 1, 1: 2, 2: 3, 3: 4, 4: 7}
    for gp, (gsl, _) in prof.items():
        for cidx, csl in ci.items():
            if csl <= gsl:
                g.ci_profiles[(gp, cidx)] = (cidx, csl)
    for slot, (gp, cp) in enumerate(mig or []):
        gsl, mb = prof[gp]
        g.migs.append(MigFix(slot, gi_id=slot + 1, ci_id=0, gi_profile_id=gp, ci_profile_id=cp,
                             attrs=(mb, 14 * ci[cp], gsl, ci[cp], gsl, gsl if gsl < 7 else 5, 0, 1 if gsl == 7 else 0, 0)))
    return g


def _node(*gpus, **kw):
    return NodeFix(gpus=list(gpus), **kw)


def nodes():
    n = {}
    n["c2_1xB200"] = _node(b200_gpu())
    n["c3_8xB200"] = _node(*[b200_gpu() for _ in range(8)])
    n["c4_single_8x7x1g23"] = _node(*[b200_gpu(mig=[(0, 0)] * 7) for _ in range(8)])
    n["c4_single_1x2x3g90"] = _node(b200_gpu(mig=[(2, 2), (2, 2)]))
    n["c4_invalid_one_empty"] = _node(b200_gpu(mig=[(0, 0)] * 7), b200_gpu(mig=[]))
    n["c4_invalid_enabled_and_disabled"] = _node(b200_gpu(mig=[(0, 0)] * 7), b200_gpu())
    n["c4_invalid_two_profiles"] = _node(b200_gpu(mig=[(2, 2), (3, 3)]))
    n["c5_mixed_hetero"] = _node(
        b200_gpu(mig=[(2, 2), (3, 3)]), b200_gpu(mig=[(0, 0)] * 7), b200_gpu(mig=[(7, 0)]), b200_gpu(mig=[(4, 4)]),
        b200_gpu(), b200_gpu(), b200_gpu(), b200_gpu())
    n["c5_mixed_partial_ci"] = _node(b200_gpu(mig=[(2, 0), (3, 1)]))  # 1c.3g.90gb, 2c.4g.90gb
    n["gi_profile_id_9_fails"] = _node(b200_gpu(mig=[(9, 0)]))        # 1g.45gb: beyond the reference's id range
    n["gi_id9_but_other_gpu_empty"] = _node(b200_gpu(mig=[(9, 0)]), b200_gpu(mig=[]))
    n["dgx_display_skipped"] = _node(GpuFix(name="NVIDIA DGX Display", memory_total=4096 * MIB, cc=(7, 5)), b200_gpu(), b200_gpu())
    n["only_dgx_display"] = _node(GpuFix(name="NVIDIA DGX Display", memory_total=4096 * MIB, cc=(7, 5)))
    n["no_mig_support"] = _node(GpuFix(name="Tesla T4", memory_total=15360 * MIB, cc=(7, 5), mig_mode=None))
    n["hetero_models"] = _node(a100_gpu(), b200_gpu(), a100_gpu())
    n["hetero_mig_and_full"] = _node(a100_gpu(mig=[(0, 0), (0, 0)]), b200_gpu())
    n["a100_single_7x1g5"] = _node(a100_gpu(mig=[(0, 0)] * 7))
    n["a100_mixed"] = _node(a100_gpu(mig=[(2, 2), (1, 1), (0, 0), (7, 0)]))
    n["zero_gpus"] = _node()
    n["driver_two_part"] = _node(b200_gpu(), driver_version="470.57")
    n["driver_four_part_error"] = _node(b200_gpu(), driver_version="1.2.3.4")
    n["driver_one_part_error"] = _node(b200_gpu(), driver_version="575")
    n["cuda_version_11080"] = _node(b200_gpu(), cuda_driver_version=11080)
    n["zero_memory_zero_cc"] = _node(GpuFix(name="Weird GPU", memory_total=0, cc=(0, 0), mig_mode=(0, 0), max_mig_count=0))
    n["name_with_spaces"] = _node(GpuFix(name="NVIDIA RTX 6000 Ada Generation", memory_total=49140 * MIB, cc=(8, 9)))
    # error injection
    n["err_init"] = _node(b200_gpu(), rc={"nvmlInit_v2": 999})
    n["err_init_driver_not_loaded"] = _node(b200_gpu(), rc={"nvmlInit_v2": 9})
    n["err_count"] = _node(b200_gpu(), rc={"nvmlDeviceGetCount_v2": 999})
    n["err_driver_version"] = _node(b200_gpu(), rc={"nvmlSystemGetDriverVersion": 999})
    n["err_cuda_version"] = _node(b200_gpu(), rc={"nvmlSystemGetCudaDriverVersion": 999})
    g = b200_gpu(); g.rc["nvmlDeviceGetName"] = 999
    n["err_name_gpu1"] = _node(b200_gpu(), g)
    g = b200_gpu(); g.rc["nvmlDeviceGetMemoryInfo"] = 999
    n["err_memory_on_non_representative_gpu"] = _node(g, b200_gpu())   # only the LAST gpu per model is queried
    n["err_memory_on_representative_gpu"] = _node(b200_gpu(), g)
    g = b200_gpu(); g.rc["nvmlDeviceGetMigMode"] = 999
    n["err_mig_mode_gpu1"] = _node(b200_gpu(), g)
    n["err_mig_mode_gpu0"] = _node(g, b200_gpu())
    g = b200_gpu(); g.rc["nvmlDeviceGetCudaComputeCapability"] = 999
    n["err_cc"] = _node(g)
    g = b200_gpu(mig=[(0, 0)] * 2); g.migs[1].rc["nvmlDeviceGetAttributes_v2"] = 999
    n["err_mig_attrs_second"] = _node(g)
    g = b200_gpu(mig=[(0, 0)] * 2); g.rc["nvmlDeviceGetMaxMigDeviceCount"] = 999
    n["err_max_mig_count"] = _node(g)
    g = b200_gpu(mig=[(0, 0)] * 2); g.migs[0].rc["nvmlDeviceGetMigDeviceHandleByIndex"] = 999
    n["err_mig_handle"] = _node(g)
    g = b200_gpu(mig=[(0, 0)] * 2); g.migs[0].rc["nvmlDeviceGetMigDeviceHandleByIndex"] = 6
    n["mig_handle_not_found_is_skipped"] = _node(g)
    g = b200_gpu(mig=[(0, 0)] * 2); g.rc["nvmlDeviceGetGpuInstanceProfileInfo"] = 999
    n["err_gi_profile_info"] = _node(g)
    g = b200_gpu(mig=[(1, 1)]); g.rc["nvmlGpuInstanceGetComputeInstanceProfileInfo"] = 999
    n["err_ci_profile_info"] = _node(g)
    g = b200_gpu(mig=[(0, 0)]); g.migs[0].rc["nvmlDeviceGetDeviceHandleFromMigDeviceHandle"] = 999
    n["err_parent_handle"] = _node(g)
    g = b200_gpu(mig=[(0, 0)]); g.migs[0].rc["nvmlDeviceIsMigDeviceHandle"] = 999
    n["err_is_mig_handle"] = _node(g)
    n["err_shutdown_is_ignored"] = _node(b200_gpu(), rc={"nvmlShutdown": 999})
    n["mig_symbol_hidden"] = _node(b200_gpu(mig=[(0, 0)]), hidden={"nvmlDeviceGetMigMode"})
    return n


STRATEGIES = ("none", "single", "mixed", "bogus")
