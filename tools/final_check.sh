#!/bin/bash
# Round-end verification on an N-GPU box (gpurun --gpus N): full GPU suite, smoke, labeling-pass phases
# (serial / per-GPU threads), the label file with every opt-in probe, and bench.py both arms at every power of two up to N
# (launched exactly as the driver launches them).  Writes gpurun_out/final_N/.
N=${1:-8}
OUT=gpurun_out/final_$N
mkdir -p $OUT
nvidia-smi -L > $OUT/smi_L.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q > $OUT/pytest_gpu.txt 2>&1; echo "pytest rc=$?" >> $OUT/pytest_gpu.txt
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.txt 2>&1; echo "smoke rc=$?" >> $OUT/smoke.txt
BIN=./gpu-feature-discovery_b200/gpu-feature-discovery
for rep in 1 2; do
  for t in 1 0; do
    $BIN --passes 8 --sleep-interval 0s --timing --no-timestamp --prefetch-threads=$t -o /tmp/o_$t 2>&1 | grep "timing:" | sed "s/^/threads=$t /" >> $OUT/pass_timing.txt
  done
  ./oracle/ref_replay --passes 8 --no-timestamp -o /tmp/o_r >> $OUT/replay.jsonl 2>&1
done
diff /tmp/o_1 /tmp/o_r > $OUT/label_diff.txt && diff /tmp/o_0 /tmp/o_r >> $OUT/label_diff.txt && echo "label diff 0 (serial, threaded vs replay)" >> $OUT/label_diff.txt
timeout 600 $BIN --oneshot --no-timestamp --probe-hbm --probe-nvlink -o $OUT/labels_with_probes.txt 2> $OUT/labels_with_probes.err; echo "probe labels rc=$?" >> $OUT/bench_rc.txt
g=1
while [ $g -le $N ]; do
  if [ $g -eq 1 ]; then
    timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 3 > $OUT/bench_ref_1.json 2> $OUT/bench_ref_1.err
    timeout 600 python bench.py --gpus 1 --steps 20 --warmup 3 > $OUT/bench_1.json 2> $OUT/bench_1.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $g --master-addr 127.0.0.1 --master-port $((29400+g)) bench.py --impl reference --gpus $g --steps 20 --warmup 3 > $OUT/bench_ref_$g.json 2> $OUT/bench_ref_$g.err
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $g --master-addr 127.0.0.1 --master-port $((29500+g)) bench.py --gpus $g --steps 20 --warmup 3 > $OUT/bench_$g.json 2> $OUT/bench_$g.err
  fi
  echo "bench $g rc=$?" >> $OUT/bench_rc.txt
  g=$((g*2))
done
cat $OUT/bench_rc.txt; tail -3 $OUT/pytest_gpu.txt
