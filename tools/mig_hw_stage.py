"""One stage of the MIG-on-real-hardware check (tools/mig_hw_check.sh drives it): with the node in whatever MIG state the
caller set up, run the product CLI, the Python oracle over pynvml and the C replay for every strategy, store all three label
files plus a raw NVML dump under OUT/<tag>/, and print one JSON summary line.  Checker side only: nothing here is product code.

usage: python tools/mig_hw_stage.py <out_dir> <tag>"""
import importlib
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def raw_dump():
    """Every value the MIG path consumes, straight from pynvml (so a disagreement can be traced to its input)."""
    import pynvml as nv
    out = {"gpus": []}
    nv.nvmlInit()
    try:
        s = lambda v: v.decode() if isinstance(v, bytes) else v
        out["driver"] = s(nv.nvmlSystemGetDriverVersion())
        out["cuda"] = nv.nvmlSystemGetCudaDriverVersion()
        for i in range(nv.nvmlDeviceGetCount()):
            h = nv.nvmlDeviceGetHandleByIndex(i)
            g = {"index": i, "name": s(nv.nvmlDeviceGetName(h)), "total_bytes_v1": int(nv.nvmlDeviceGetMemoryInfo(h).total)}
            try:
                g["mig_mode"] = [int(x) for x in nv.nvmlDeviceGetMigMode(h)]
            except nv.NVMLError as e:
                g["mig_mode_err"] = int(e.value)
            try:
                g["max_mig"] = int(nv.nvmlDeviceGetMaxMigDeviceCount(h))
            except nv.NVMLError as e:
                g["max_mig_err"] = int(e.value)
            g["gi_profiles"] = []
            for idx in range(0, 20):
                try:
                    p = nv.nvmlDeviceGetGpuInstanceProfileInfo(h, idx, version=1)
                    g["gi_profiles"].append({"idx": idx, "id": int(p.id), "slices": int(p.sliceCount), "mem_mb": int(p.memorySizeMB),
                                             "instances": int(p.instanceCount), "sms": int(p.multiprocessorCount)})
                except nv.NVMLError as e:
                    g["gi_profiles"].append({"idx": idx, "err": int(e.value)})
            g["mig_devices"] = []
            for j in range(g.get("max_mig", 0)):
                try:
                    m = nv.nvmlDeviceGetMigDeviceHandleByIndex(h, j)
                except nv.NVMLError as e:
                    g["mig_devices"].append({"slot": j, "err": int(e.value)})
                    continue
                a = nv.nvmlDeviceGetAttributes_v2(m)
                gi_id, ci_id = int(nv.nvmlDeviceGetGpuInstanceId(m)), int(nv.nvmlDeviceGetComputeInstanceId(m))
                gi = nv.nvmlDeviceGetGpuInstanceById(h, gi_id)
                gi_info = nv.nvmlGpuInstanceGetInfo(gi)
                ci = nv.nvmlGpuInstanceGetComputeInstanceById(gi, ci_id)
                ci_fn = getattr(nv, "nvmlComputeInstanceGetInfo_v2", None) or nv.nvmlComputeInstanceGetInfo
                ci_info = ci_fn(ci)
                d = {"slot": j, "name": s(nv.nvmlDeviceGetName(m)), "uuid": s(nv.nvmlDeviceGetUUID(m)), "gi_id": gi_id, "ci_id": ci_id,
                     "gi_profile_id": int(gi_info.profileId), "ci_profile_id": int(ci_info.profileId),
                     "attrs": {k: int(getattr(a, k)) for k in ("memorySizeMB", "multiprocessorCount", "gpuInstanceSliceCount",
                                                                "computeInstanceSliceCount", "sharedCopyEngineCount", "sharedDecoderCount",
                                                                "sharedEncoderCount", "sharedJpegCount", "sharedOfaCount")}}
                try:
                    d["mig_total_bytes_v1"] = int(nv.nvmlDeviceGetMemoryInfo(m).total)
                except nv.NVMLError as e:
                    d["mig_total_err"] = int(e.value)
                d["ci_profiles"] = []
                for idx in range(0, 10):
                    try:
                        p = nv.nvmlGpuInstanceGetComputeInstanceProfileInfo(gi, idx, 0, version=1)
                        d["ci_profiles"].append({"idx": idx, "id": int(p.id), "slices": int(p.sliceCount)})
                    except nv.NVMLError as e:
                        d["ci_profiles"].append({"idx": idx, "err": int(e.value)})
                g["mig_devices"].append(d)
            out["gpus"].append(g)
    finally:
        nv.nvmlShutdown()
    return out


def main():
    out_dir, tag = sys.argv[1], sys.argv[2]
    d = os.path.join(out_dir, tag)
    os.makedirs(d, exist_ok=True)
    B = importlib.import_module("gpu-feature-discovery_b200.binding")
    from oracle import gfd_oracle as O, nvml_model as M

    summary = {"tag": tag, "cases": {}}
    try:
        raw = raw_dump()
        json.dump(raw, open(os.path.join(d, "nvml_raw.json"), "w"), indent=1)
        summary["mig_devices"] = sum(len([m for m in g["mig_devices"] if "err" not in m]) for g in raw["gpus"])
        summary["mig_modes"] = [g.get("mig_mode") for g in raw["gpus"]]
    except Exception as e:  # the dump is evidence, not the check: keep going
        summary["raw_dump_error"] = repr(e)

    for strategy in ("none", "single", "mixed"):
        case = {}
        # product CLI, default (reference-parity) tables and the opt-in current tables
        for variant, extra in (("cli", []), ("cli_current_tables", ["--current-tables"])):
            f = os.path.join(d, "%s_%s.labels" % (variant, strategy))
            if os.path.exists(f):
                os.unlink(f)
            p = subprocess.run([B.BIN_PATH, "--oneshot", "--no-timestamp", "--mig-strategy=" + strategy, "-o", f] + extra,
                               capture_output=True, text=True, timeout=300)
            open(os.path.join(d, "%s_%s.stderr" % (variant, strategy)), "w").write(p.stderr)
            case[variant + "_rc"] = p.returncode
            case[variant] = open(f).read() if os.path.exists(f) else None
        # oracle over pynvml
        cfg = O.Config(mig_strategy=strategy, no_timestamp=True)
        try:
            labels, _ = O.run_pass(O.NvmlManager(M.PynvmlRaw()), lambda: O.nvidia_pci_devices(), cfg)
            case["oracle"] = O.render(labels)
        except O.GfdError as e:
            case["oracle"] = None
            case["oracle_error"] = str(e)
        if case["oracle"] is not None:
            open(os.path.join(d, "oracle_%s.labels" % strategy), "w").write(case["oracle"])
        # C replay of the reference's call sequence
        f = os.path.join(d, "replay_%s.labels" % strategy)
        if os.path.exists(f):
            os.unlink(f)
        p = subprocess.run([os.path.join(ROOT, "oracle", "ref_replay"), "--mig-strategy", strategy, "--no-timestamp", "-o", f],
                           capture_output=True, text=True, timeout=300)
        open(os.path.join(d, "replay_%s.stderr" % strategy), "w").write(p.stderr)
        case["replay_rc"] = p.returncode
        case["replay"] = open(f).read() if (p.returncode == 0 and os.path.exists(f)) else None

        ok_oracle = (case["cli"] == case["oracle"]) and ((case["cli_rc"] == 0) == (case["oracle"] is not None))
        ok_replay = (case["cli"] == case["replay"]) and ((case["cli_rc"] == 0) == (case["replay_rc"] == 0))
        summary["cases"][strategy] = {
            "cli_rc": case["cli_rc"], "oracle_fails": case["oracle"] is None, "replay_rc": case["replay_rc"],
            "diff_vs_oracle": 0 if ok_oracle else 1, "diff_vs_replay": 0 if ok_replay else 1,
            "lines": None if case["cli"] is None else len(case["cli"].splitlines()),
            "cli_current_tables_rc": case["cli_current_tables_rc"],
            "current_tables_changed_lines": None if (case["cli"] is None or case["cli_current_tables"] is None) else
            sorted(set(case["cli_current_tables"].splitlines()) ^ set(case["cli"].splitlines())),
            "current_tables_lines": None if case["cli_current_tables"] is None else len(case["cli_current_tables"].splitlines()),
            "oracle_error": case.get("oracle_error"),
        }
    summary["all_diffs_zero"] = all(c["diff_vs_oracle"] == 0 and c["diff_vs_replay"] == 0 for c in summary["cases"].values())
    json.dump(summary, open(os.path.join(d, "summary.json"), "w"), indent=1)
    print(json.dumps(summary))


if __name__ == "__main__":
    main()
