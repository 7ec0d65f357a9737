#!/bin/bash
# Multi-GPU measurements (run under gpurun --gpus N): NVML attach scaling, labeling pass serial vs threaded,
# P2P probe per directed pair, torchrun bench.  Writes into gpurun_out/mgpu_N/.
N=${1:-2}
MODE=${2:-all}
OUT=gpurun_out/mgpu_$N
mkdir -p $OUT
rm -f $OUT/pass_timing.txt $OUT/replay.jsonl $OUT/nvml_phases.txt
nvidia-smi -L > $OUT/smi_L.txt 2>&1
nproc > $OUT/nproc.txt
for f in 0 2; do for t in 1 $N; do ./tools/nvml_init_timing 5 $f $t >> $OUT/nvml_phases.txt 2>&1; done; done
BIN=./gpu-feature-discovery_b200/gpu-feature-discovery
for rep in 1 2 3 4; do
  for t in 1 $N; do
    $BIN --passes 8 --sleep-interval 0s --timing --no-timestamp --prefetch-threads=$t -o /tmp/o_$t 2>&1 | grep "timing:" | sed "s/^/threads=$t /" >> $OUT/pass_timing.txt
  done
  ./oracle/ref_replay --passes 8 --no-timestamp -o /tmp/o_r >> $OUT/replay.jsonl 2>&1
done
$BIN --passes 8 --sleep-interval 0s --timing --no-timestamp --keep-nvml-attached -o /tmp/o_k 2>&1 | grep "timing:" | sed "s/^/keep /" >> $OUT/pass_timing.txt
diff /tmp/o_1 /tmp/o_r > $OUT/label_diff.txt && diff /tmp/o_$N /tmp/o_r >> $OUT/label_diff.txt && diff /tmp/o_k /tmp/o_r >> $OUT/label_diff.txt && echo "label diff 0 (serial, threaded, keep-attached vs replay)" >> $OUT/label_diff.txt
cp /tmp/o_1 $OUT/labels.txt
python - "$OUT" <<'PY'
import json, statistics, sys, collections
out = sys.argv[1]
d = collections.defaultdict(list)
for l in open(out + "/pass_timing.txt"):
    k = l.split()[0]
    d[k].append(float(l.split("timing: pass ")[1].split(" ms")[0]))
r = []
for l in open(out + "/replay.jsonl"):
    try: r += json.loads(l)["ms"]
    except Exception: pass
d["replay"] = r
summ = {k: {"n": len(v), "median": round(statistics.median(v), 3), "min": round(min(v), 3), "p90": round(sorted(v)[int(len(v) * 0.9)], 3)} for k, v in d.items() if v}
open(out + "/pass_summary.json", "w").write(json.dumps(summ, indent=1) + "\n")
PY
[ "$MODE" = "nvml" ] && exit 0
python -m pytest tests/test_gpu_probes.py -m gpu -q -k "p2p" > $OUT/pytest_p2p.log 2>&1
python tools/p2p_matrix.py > $OUT/p2p_matrix.json 2> $OUT/p2p_matrix.err
python bench.py --impl reference --gpus $N --steps 20 --warmup 3 > $OUT/bench_ref.json 2> $OUT/bench_ref.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err
echo "bench rc=$?" >> $OUT/bench.err
