"""The N>1 plumbing of bench.py on CPU: world_size-2 over gloo (the data path itself has no collective -- the pass is
per node, probes are per-GPU replicas -- so what needs covering is the result reduction and the rank-0-only arms)."""
import json
import os
import subprocess
import sys

import torch.distributed as dist
import torch.multiprocessing as mp

import nodes as N
from shimutil import B, ROOT, use_node


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    import bench
    dist.init_process_group("gloo", rank=rank, world_size=world)
    # rank 0 ran the node pass (103 ms, e2e 70 ms); rank 1 did not (0); each rank probed its own GPU
    vals = (103.0, 70.0, 0.58, 7400.0, 55.0) if rank == 0 else (0.0, 0.0, 0.61, 7300.0, 54.0)
    out = bench.reduce_over_ranks(*vals, world, "cpu")
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, out))


def test_reduce_over_ranks_gloo_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29533
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank in (0, 1):
        pass_ms, e2e_ms, kernel_ms, probe, host = res[rank]
        assert (pass_ms, e2e_ms, kernel_ms) == (103.0, 70.0, 0.61)   # MAX over ranks
        assert (probe, host) == (14700.0, 109.0)                      # SUM over ranks


def test_reference_arm_runs_on_rank0_only(tmp_path):
    """`bench.py --impl reference` under torchrun: rank 0 prints the line, every other rank exits 0 without work."""
    fix = use_node(N.nodes()["c3_8xB200"], tmp_path)
    env = dict(os.environ, GFD_FAKE_NVML=fix, LD_LIBRARY_PATH=B.FAKE_NVML_DIR + ":" + os.environ.get("LD_LIBRARY_PATH", ""), WORLD_SIZE="2",
               GFD_BENCH_SETTLE_S="0")  # no GPU here whose power state could settle
    r1 = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "3", "--warmup", "3"],
                        env=dict(env, RANK="1"), capture_output=True, text=True, timeout=120)
    assert r1.returncode == 0 and r1.stdout.strip() == ""
    r0 = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "3", "--warmup", "3"],
                        env=dict(env, RANK="0"), capture_output=True, text=True, timeout=120)
    assert r0.returncode == 0, r0.stderr
    line = json.loads(r0.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["metric"] == "labeling_pass_ms" and line["unit"] == "ms" and line["n_gpus"] == 2
    assert line["higher_is_better"] is False and line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] == 1
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["value"] > 0
    # --gpus 2 on the 8-GPU node: the device cap makes the baseline label 2 GPUs (and the product arm runs under the same cap)
    assert line["config"]["gpus_labelled"] == 2 and "32 NVML calls + 3 dlopen cycles" in line["cpu_baseline"]["sample"]
    assert line["value"] == line["labeling_pass_ms_spread"]["median"] and line["labeling_pass_ms_spread"]["n"] == 3
    r8 = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "8", "--steps", "6", "--warmup", "3"],
                        env=dict(env, RANK="0"), capture_output=True, text=True, timeout=120)
    line = json.loads(r8.stdout.strip().splitlines()[-1])
    assert line["config"]["gpus_labelled"] == 8 and "92 NVML calls + 9 dlopen cycles" in line["cpu_baseline"]["sample"]
    assert line["labeling_pass_ms_spread"]["n"] == 6 and set(line["cpu_baseline"]["phases_ms"]) == {"init", "shutdown"}


def test_series_times_exactly_k_steps_and_alternates():
    """W warm-up passes untimed, then exactly K timed passes in chunks, A and B alternating chunk by chunk; the first pass of
    every process is dropped (process start-up is e2e's business)."""
    sys.path.insert(0, ROOT)
    import bench
    calls = []

    def mk(tag):
        def run(n):
            calls.append((tag, n))
            return [(100.0 if i == 0 else 10.0 + i, {"init": 1.0}) for i in range(n)]
        return run
    a, b = bench.series(3, 12, mk("A"), mk("B"))
    assert len(a) == 12 and len(b) == 12
    assert calls == [("A", 3), ("B", 3), ("A", 6), ("B", 6), ("A", 6), ("B", 6), ("A", 3), ("B", 3)]
    assert all(ms != 100.0 for ms, _ in a + b)
    st, ph = bench.stats(a)
    assert st["n"] == 12 and st["min"] == 11.0 and ph == {"init": 1.0}
    rows = bench.parse_timing("x\n2026/11/07 00:00:00 timing: pass 17.5 ms (nvml 17.0, vgpu 0.3, write 0.2) [init 7.1, attach 8.8, labels 0.4, shutdown 2.2]\n")
    assert rows == [(17.5, {"init": 7.1, "attach": 8.8, "labels": 0.4, "shutdown": 2.2})]


def test_p2p_section_shape_and_launch_count():
    sys.path.insert(0, ROOT)
    import bench

    class FakeB:
        PROBE_READ, PROBE_PUSH, PROBE_DMA = 0, 2, 4

        @staticmethod
        def probe_p2p_matrix(n, nbytes, warmup, iters, mode):
            base = {0: 670.0, 2: 705.0, 4: 770.0}[mode]
            return [[0.0 if a == b else base + a + b for b in range(n)] for a in range(n)], base * n
    r, launches = bench.p2p_section(FakeB, 4, warmup=2, iters=10)
    assert r["bound"] == "nvlink" and r["peak"] == 900.0 and r["pairs"] == 12 and r["verified"] is True
    assert r["achieved_per_direction_min"] == 671.0 and r["push"]["achieved_per_direction_min"] == 706.0
    assert r["mode"] == "read" and r["push"]["mode"] == "push" and r["achieved_ring_aggregate"] == 2680.0
    # per mode: fill + pattern per GPU (8), 3 rounds x 4 GPUs x 12 launches (144), ring 4 x 12 (48); PUSH re-reads every destination
    # the copy-engine pass launches only the one-off fill / pattern kernels and the verifying re-reads
    assert launches == (8 + 144 + 48) + (8 + 144 + 48 + 3 * 4 + 4) + (8 + 3 * 4 + 4)
    assert r["dma"]["achieved_per_direction_min"] == 771.0 and "copy engine" in r["dma"]["note"]
