#!/usr/bin/env python3
"""bench.py -- the labeling-pass benchmark (BASELINE.json metric) + the opt-in HBM probe's roofline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

A "step" is ONE LABELING PASS on the configuration BASELINE.json quotes the metric on (configs[1]: a B200
node, MIG strategy none, --no-timestamp, full label set from the real NVML, vGPU scan of the real sysfs, atomic
file write).  The reference has no GPU compute on this path, so:

  metric / value   labeling_pass_ms: steady-state wall-clock of one pass inside the running labeler
                   (gpu-feature-discovery --passes K+W --sleep-interval 0s --timing; mean of the last K).
  e2e              the same work as a user gets it: exec -> exit of `gpu-feature-discovery --oneshot
                   --no-timestamp -o FILE` plus reading FILE back (median over the K runs; the mean is kept beside
                   it: process spawns on a shared host have 40-270 ms outliers).  The path moves no tensor data, so
                   h2d/d2h bytes are 0 and stated as such; `e2e_probe_host` carries the host-buffer form of the
                   probe kernel (pinned host -> HBM -> checksum -> 16 bytes back) for completeness.
  label_diff       lines that differ between the product's file and (a) the CPU baseline's file, (b) the Python
                   oracle fed with pynvml-harvested values.  Must be 0.
  roofline         the dominant (only) GPU kernel, hbm_read_fold, 4 GiB per launch (>> 126 MB L2: no flush
                   needed), timed live with CUDA events on the launching stream; algorithmic bytes = buffer size.
  cpu_baseline     oracle/ref_replay: a call-sequence-faithful C replay of the reference's pass (the Go reference
                   cannot be built on this image), single-threaded like the reference.  kind = "port".

`--impl reference` times that replay alone (rank 0 only under torchrun) and prints the same line shape.
N > 1: the pass is per node and does not shard across ranks, so rank 0 runs it once over all N GPUs (the labeler's own
per-GPU host threads are the multi-GPU path) while every rank probes its own GPU (replicas); probe GB/s is the sum.
Barriers between the regions go over a gloo group: a rank waiting in an NCCL barrier spins a kernel on its GPU, which
is one of the GPUs rank 0's pass is attaching to.
"""
import argparse
import ctypes
import importlib
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

PROBE_BYTES = 4 << 30
REPLAY = os.path.join(ROOT, "oracle", "ref_replay")


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, torch copy_ read+write)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region."""

    def __init__(self, index):
        threading.Thread.__init__(self, daemon=True)
        self.index, self.rows, self.proc = index, [], None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([c.strip() for c in line.split(",")])
        except OSError:
            pass

    def stop(self):
        if self.proc:
            self.proc.terminate()
        self.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


PHASES = {}  # filled by parse_timing: mean ms of the NVML bracket's phases over the steady-state passes of the last run


def parse_timing(stderr):
    out, ph = [], []
    for line in stderr.splitlines():
        if "timing: pass " in line:
            out.append(float(line.split("timing: pass ")[1].split(" ms")[0]))
            if "[init " in line:
                body = line.split("[init ")[1].rstrip("]")
                vals = [float(x.split()[-1]) for x in ("init " + body).split(",")]
                ph.append(vals)
    if ph:
        PHASES.clear()
        PHASES.update({k: round(sum(v[i] for v in ph) / len(ph), 3) for i, k in enumerate(("init", "attach", "labels", "shutdown"))})
    return out


def product_passes(binary, passes, out_file, extra=()):
    p = subprocess.run([binary, "--passes", str(passes), "--sleep-interval", "0s", "--timing", "--no-timestamp", "--mig-strategy=none",
                        "-o", out_file] + list(extra), capture_output=True, text=True, timeout=600)
    if p.returncode != 0:
        raise RuntimeError("gpu-feature-discovery failed:\n" + p.stderr[-2000:])
    ms = parse_timing(p.stderr)
    if len(ms) != passes:
        raise RuntimeError("expected %d timing lines, got %d" % (passes, len(ms)))
    return ms


def product_oneshot(binary, out_file):
    t0 = time.perf_counter()
    p = subprocess.run([binary, "--oneshot", "--no-timestamp", "--mig-strategy=none", "-o", out_file], capture_output=True, timeout=120)
    with open(out_file, "rb") as f:
        data = f.read()
    dt = (time.perf_counter() - t0) * 1e3
    if p.returncode != 0:
        raise RuntimeError("gpu-feature-discovery --oneshot failed:\n" + p.stderr.decode()[-2000:])
    return dt, data


def replay_passes(passes, out_file):
    p = subprocess.run([REPLAY, "--mig-strategy", "none", "--no-timestamp", "-o", out_file, "--passes", str(passes)],
                       capture_output=True, text=True, timeout=600)
    if p.returncode != 0:
        raise RuntimeError("ref_replay failed:\n" + p.stderr[-2000:])
    return json.loads(p.stdout.strip().splitlines()[-1])


def diff_lines(a, b):
    sa, sb = set(a.splitlines()), set(b.splitlines())
    return len(sa ^ sb)


def reference_arm(args, rank, world):
    if rank != 0:
        return 0
    tmp = tempfile.mkdtemp(prefix="gfd-bench-")
    out = os.path.join(tmp, "gfd")
    total = args.warmup + args.steps
    r = replay_passes(total, out)
    ms = r["ms"][args.warmup:]
    value = sum(ms) / len(ms)
    spread = {"median": round(statistics.median(ms), 4), "min": round(min(ms), 4), "max": round(max(ms), 4)}
    e2e = []
    for i in range(total):  # same definition as the product arm's e2e: exec -> exit of a one-shot run + reading the file
        t0 = time.perf_counter()
        p = subprocess.run([REPLAY, "--mig-strategy", "none", "--no-timestamp", "-o", out], capture_output=True, timeout=120)
        with open(out, "rb") as f:
            f.read()
        if p.returncode != 0:
            raise RuntimeError("ref_replay failed: " + p.stderr.decode()[-1000:])
        if i >= args.warmup:
            e2e.append((time.perf_counter() - t0) * 1e3)
    e2e_ms = statistics.median(e2e)
    line = {
        "impl": "reference", "metric": "labeling_pass_ms", "value": round(value, 4), "unit": "ms", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(value, 4), "higher_is_better": False, "scaling": "weak",
        "vs_baseline": None, "dtype": "u64", "data": "real NVML of this host (no synthetic input: the pass reads the driver)",
        "config": workload_config(),
        "cpu_baseline": {"value": round(value, 4), "unit": "ms", "cores": 1, "kind": "port",
                         "sample": "%d back-to-back passes of oracle/ref_replay (C replay of the reference's NVML call sequence: "
                                   "%d NVML calls + %d dlopen cycles per pass, nvmlInit/Shutdown every pass); the Go reference "
                                   "cannot be built on this image" % (len(ms), r["nvml_calls"], r["dl_cycles"]),
                         "init_ms": round(sum(r["init_ms"][args.warmup:]) / len(ms), 4)},
        "e2e": {"value": round(e2e_ms, 4), "unit": "ms", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0, "mean": round(sum(e2e) / len(e2e), 4),
                "note": "MEDIAN over the K steps of exec->exit of a one-shot oracle/ref_replay run + reading the label file (process spawns on "
                        "this box have 40-270 ms outliers for either arm: profiles/r1c_e2e_context_1gpu.txt)"},
        "gpu_launches": 0, "labeling_pass_ms_spread": spread,
    }
    print(json.dumps(line), flush=True)
    return 0


def reduce_over_ranks(pass_ms, e2e_ms, kernel_ms, probe_gbps, host_gbps, world, device):
    """Times: MAX over ranks (the node pass runs on rank 0 only, the others contribute 0).  Probe bandwidths: SUM over
    ranks (one GPU each).  No data-path collective exists; this is result plumbing (NCCL on GPUs, gloo in the CPU test)."""
    if world <= 1:
        return pass_ms, e2e_ms, kernel_ms, probe_gbps, host_gbps
    import torch
    import torch.distributed as dist
    t = torch.tensor([pass_ms, e2e_ms, kernel_ms], device=device, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    g = torch.tensor([probe_gbps, host_gbps], device=device, dtype=torch.float64)
    dist.all_reduce(g, op=dist.ReduceOp.SUM)
    a, b = [float(x) for x in t.cpu()], [float(x) for x in g.cpu()]
    return a[0], a[1], a[2], b[0], b[1]


def persistence_mode():
    """nvidia-smi's persistence_mode of GPU 0 (decides whether nvmlInit / the first handle pay the driver attach); asked
    ONCE, before any timed region."""
    try:
        p = subprocess.run(["nvidia-smi", "-i", "0", "--query-gpu=persistence_mode", "--format=csv,noheader"], capture_output=True, text=True, timeout=30)
        return p.stdout.strip().splitlines()[0].strip() if p.returncode == 0 and p.stdout.strip() else "unknown"
    except Exception:
        return "unknown"


def workload_config():
    return {"persistence_mode": persistence_mode(), "workload": "BASELINE configs[1]: B200 node, MIG strategy none, --no-timestamp, full label set from the real NVML + "
                        "vGPU sysfs scan + atomic file write; one step = one labeling pass",
            "probe": "opt-in hbm_read_fold, %d MiB per launch (input >> 126 MB L2, so no L2 flush between iterations)" % (PROBE_BYTES >> 20),
            "parallelism": "the pass is per node: rank 0 labels all N GPUs (one host thread per GPU), no collective; probes are per-GPU replicas",
            "multi_rank_note": "with N > 1 the waiting ranks hold CUDA contexts on their GPUs (parked in a host-side gloo barrier), and a GPU that is "
                               "already held open makes nvmlInit_v2 cheaper (profiles/r1c_pass_spread_8xB200.txt); the reference arm's other ranks "
                               "exit at once, so its node is idle: the like-for-like comparison is the N = 1 line"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        return reference_arm(args, rank, world)

    import numpy as np
    import torch
    import torch.distributed as dist
    B = importlib.import_module("gpu-feature-discovery_b200.binding")  # raises if libgfd_b200.so is missing: no fallback
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU: the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    host_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        # Barriers go over a gloo (TCP) group: an NCCL barrier is a kernel that SPINS on every waiting rank's GPU (and a
        # polling host thread), and rank 0's labeling passes attach to exactly those GPUs through NVML meanwhile --
        # measured on 2 GPUs: pass median 50 ms with the other rank parked in an NCCL barrier, 21-30 ms without.
        os.environ.setdefault("GLOO_SOCKET_IFNAME", "lo")  # one node: loopback always exists, the hostname may not resolve
        try:
            if os.environ.get("GFD_BENCH_FILE_BARRIER"):  # test knob: exercise the fallback
                raise RuntimeError("forced by GFD_BENCH_FILE_BARRIER")
            host_group = dist.new_group(backend="gloo")
        except Exception as e:  # no gloo transport: fall back to a barrier through the node's file system
            print("bench: gloo group unavailable (%s); using a file barrier" % e, file=sys.stderr)
    token = [tempfile.mkdtemp(prefix="gfd-bench-barrier-") if rank == 0 else None]
    if world > 1 and host_group is None:
        dist.broadcast_object_list(token, src=0)
    barrier_no = [0]

    def barrier():
        torch.cuda.synchronize()
        if world > 1 and host_group is not None:
            dist.barrier(group=host_group)
        elif world > 1:
            barrier_no[0] += 1
            open(os.path.join(token[0], "%d.%d" % (barrier_no[0], rank)), "w").close()
            while not all(os.path.exists(os.path.join(token[0], "%d.%d" % (barrier_no[0], r))) for r in range(world)):
                time.sleep(0.001)
        torch.cuda.synchronize()

    tmp = tempfile.mkdtemp(prefix="gfd-bench-r%d-" % rank)
    out_file = os.path.join(tmp, "gfd")

    # ---- region A: K labeling passes inside the running labeler (W warm-up passes first) ----------------------
    # The pass is per NODE (one labeler per node is the deployment: a DaemonSet pod), so with N ranks it is rank 0
    # that runs it -- once, over all N GPUs, with the labeler's own per-GPU host threads -- while the other ranks
    # wait at the barrier.  N copies of a node-level daemon hammering one driver would measure contention, not GFD.
    barrier()
    pass_ms, product_text, phases, pass_stats = 0.0, "", {}, {}
    if rank == 0:
        ms = product_passes(B.BIN_PATH, args.warmup + args.steps, out_file)[args.warmup:]
        pass_ms = sum(ms) / len(ms)
        pass_stats = {"median": round(statistics.median(ms), 4), "min": round(min(ms), 4), "max": round(max(ms), 4)}
        phases = dict(PHASES)
        product_text = open(out_file).read()
    barrier()

    # ---- region B: K launches of the dominant kernel, CUDA events on the launching stream ---------------------
    buf = torch.empty(PROBE_BYTES, dtype=torch.uint8, device=dev)
    scratch = torch.empty(B.hbm_scratch_bytes(), dtype=torch.uint8, device=dev)
    B.hbm_scratch_init(scratch.data_ptr(), scratch.numel(), torch.cuda.current_stream().cuda_stream)  # once; no per-call memset
    torch.cuda.synchronize()
    d_out = torch.zeros(2, dtype=torch.int64, device=dev)
    stream = torch.cuda.current_stream()
    B.hbm_fill_pattern(buf.data_ptr(), PROBE_BYTES, 0, stream.cuda_stream)
    launches = 1
    # clocks are sampled around the GPU region only: nvidia-smi is itself an NVML client, and polling it while the
    # labeling passes run would lengthen THEIR driver attach (measured: 17 ms -> 38 ms per pass)
    sampler = ClockSampler(local_rank)
    sampler.start()
    time.sleep(0.3)
    for _ in range(args.warmup):
        B.hbm_read_checksum(buf.data_ptr(), PROBE_BYTES, d_out.data_ptr(), scratch.data_ptr(), scratch.numel(), stream.cuda_stream)
        launches += 1
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        B.hbm_read_checksum(buf.data_ptr(), PROBE_BYTES, d_out.data_ptr(), scratch.data_ptr(), scratch.numel(), stream.cuda_stream)
        launches += 1
    e1.record(stream)
    barrier()
    # keep the GPU under the same load a little longer so the 100 ms sampler sees clocks under load
    t_hold = time.time()
    while time.time() - t_hold < 0.5:
        B.hbm_read_checksum(buf.data_ptr(), PROBE_BYTES, d_out.data_ptr(), scratch.data_ptr(), scratch.numel(), stream.cuda_stream)
        launches += 1
        torch.cuda.synchronize()
    kernel_ms = e0.elapsed_time(e1) / args.steps
    probe_gbps = PROBE_BYTES / (kernel_ms * 1e-3) / 1e9
    checksum = tuple(int(v) for v in d_out.cpu().numpy().view(np.uint64))

    sampler.stop()
    clocks = sampler.summary()

    # ---- e2e: the user-facing call, exec -> exit + read the file back ----------------------------------------
    e2e_ms = e2e_mean = 0.0
    if rank == 0:
        e2e = []
        for i in range(args.warmup + args.steps):
            dt, data = product_oneshot(B.BIN_PATH, out_file)
            if i >= args.warmup:
                e2e.append(dt)
        e2e_ms, e2e_mean = statistics.median(e2e), sum(e2e) / len(e2e)
    barrier()

    # host-buffer form of the probe kernel through the C ABI (pinned host memory; copies inside the timed region)
    host_bytes = 1 << 30
    pinned = torch.empty(host_bytes, dtype=torch.uint8).pin_memory()
    pinned.random_(0, 256)
    reps = 5
    for _ in range(2):
        B.checksum_host(local_rank, pinned.data_ptr(), host_bytes)
    launches += 1 * (2 + reps)  # gfd_checksum_host = hbm_read_fold + fold_finish per call
    t0 = time.perf_counter()
    for _ in range(reps):
        host_sum = B.checksum_host(local_rank, pinned.data_ptr(), host_bytes)
    host_gbps = host_bytes * reps / (time.perf_counter() - t0) / 1e9

    # ---- reductions over ranks --------------------------------------------------------------------------------
    pass_ms, e2e_ms, kernel_ms, probe_total, host_total = reduce_over_ranks(pass_ms, e2e_ms, kernel_ms, probe_gbps, host_gbps,
                                                                             world, dev)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    # ---- rank 0 only: checker + CPU baseline (the oracle is the checker / the baseline, never the thing measured) ----
    from oracle import gfd_oracle as O, nvml_model as M
    cfg = O.Config(mig_strategy="none", no_timestamp=True)
    try:
        want, _ = O.run_pass(O.NvmlManager(M.PynvmlRaw()), lambda: O.nvidia_pci_devices(), cfg)
        oracle_text = O.render(want)
    except O.GfdError as e:
        oracle_text = "ORACLE-ERROR %s\n" % e
    replay_out = os.path.join(tmp, "replay")
    r = replay_passes(args.warmup + min(args.steps, 30), replay_out)
    rms = r["ms"][args.warmup:]
    replay_text = open(replay_out).read()
    orc = ctypes.CDLL(os.path.join(ROOT, "oracle", "libgfd_oracle_c.so"))
    orc.oracle_pattern_checksum.argtypes = [ctypes.c_size_t, ctypes.c_uint64, ctypes.POINTER(ctypes.c_uint64)]
    o = (ctypes.c_uint64 * 2)()
    orc.oracle_pattern_checksum(PROBE_BYTES // 8, 0, o)
    checksum_ok = checksum == (int(o[0]), int(o[1]))

    keep = product_passes(B.BIN_PATH, args.warmup + args.steps, out_file, extra=["--keep-nvml-attached"])[args.warmup:]

    peak, peak_src = load_peaks()
    traffic = None  # dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full capture
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "r1c_traffic.json")))["hbm_read_fold_kernel"]
        if t["bytes_per_launch"] == PROBE_BYTES:
            traffic = t["dram_bytes_read"] + t["dram_bytes_write"]
    except Exception:
        pass
    line = {
        "metric": "labeling_pass_ms", "value": round(pass_ms, 4), "unit": "ms", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(pass_ms, 4), "higher_is_better": False, "scaling": "weak", "vs_baseline": None,
        "dtype": "u64", "data": "real NVML of this host (no synthetic input: the pass reads the driver); probe input = deterministic pattern",
        "config": workload_config(),
        "e2e": {"value": round(e2e_ms, 4), "unit": "ms", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0, "mean": round(e2e_mean, 4),
                "note": "MEDIAN over the K steps of exec->exit of gpu-feature-discovery --oneshot + reading the label file (process spawns on this "
                        "box have 40-270 ms outliers for either arm: profiles/r1c_e2e_context_1gpu.txt); the labeling pass moves no tensor data"},
        "e2e_probe_host": {"value": round(host_total, 2), "unit": "GB/s", "h2d_bytes_per_step": host_bytes, "d2h_bytes_per_step": 16,
                           "note": "gfd_checksum_host: pinned host buffer -> HBM -> hbm_read_fold -> 16-byte result"},
        "gpu_launches": launches,
        "label_diff": {"vs_cpu_baseline": diff_lines(product_text, replay_text), "vs_oracle_pynvml": diff_lines(product_text, oracle_text),
                       "labels": len(product_text.splitlines())},
        "labeling_pass_phases_ms": phases,
        "labeling_pass_ms_spread": pass_stats,  # `value` is the mean over the K steps (the contract); a shared host makes single passes jump
        "labeling_pass_ms_keep_attached": round(sum(keep) / len(keep), 4),
        "probe_hbm_read_gbps": round(probe_total, 1),
        "probe_checksum_ok": checksum_ok,
        "roofline": {"bound": "hbm", "achieved": round(probe_gbps, 1), "peak": peak, "unit": "GB/s", "frac": round(probe_gbps / peak, 4),
                     "traffic": traffic, "traffic_unit": "bytes per launch (ncu, profiles/r1c_ncu_full_summary.md)", "peak_source": peak_src, "kernel": "hbm_read_fold_kernel (+ 1-CTA fold_finish_kernel)",
                     "algorithmic_bytes_per_launch": PROBE_BYTES, "ms_per_launch": round(kernel_ms, 5),
                     "frac_of_nominal_7700": round(probe_gbps / 7700.0, 4)},
        "cpu_baseline": {"value": round(sum(rms) / len(rms), 4), "unit": "ms", "cores": 1, "kind": "port",
                         "sample": "%d back-to-back passes of oracle/ref_replay (C replay of the reference's NVML call sequence: %d NVML "
                                   "calls + %d dlopen cycles per pass)" % (len(rms), r["nvml_calls"], r["dl_cycles"]),
                         "init_ms": round(sum(r["init_ms"][args.warmup:]) / len(rms), 4)},
        "clocks": clocks,
        "host_cores": os.cpu_count(),
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
