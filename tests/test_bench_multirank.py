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
    env = dict(os.environ, GFD_FAKE_NVML=fix, LD_LIBRARY_PATH=B.FAKE_NVML_DIR + ":" + os.environ.get("LD_LIBRARY_PATH", ""), WORLD_SIZE="2")
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
    assert "92 NVML calls + 9 dlopen cycles" in line["cpu_baseline"]["sample"]
