#!/bin/bash
# Round-2 check on a 2-GPU lease: GPU suite with peers, both bench arms at N=1 and N=2, ncu NVLink counters for the pair kernels.
OUT=gpurun_out/r2_2gpu
mkdir -p $OUT
nvidia-smi -L > $OUT/smi_L.txt 2>&1
nvidia-smi topo -m > $OUT/topo.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q -x > $OUT/pytest_gpu.txt 2>&1; echo "pytest rc=$?" >> $OUT/pytest_gpu.txt
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $OUT/pytest_zz_alone.txt 2>&1; echo "rc=$?" >> $OUT/pytest_zz_alone.txt
for g in 1 2; do
  timeout 400 python bench.py --impl reference --gpus $g --steps 20 --warmup 3 > $OUT/bench_ref_$g.json 2> $OUT/bench_ref_$g.err; echo "ref $g rc=$?" >> $OUT/rc.txt
  if [ $g -eq 1 ]; then
    timeout 400 python bench.py --gpus 1 --steps 20 --warmup 3 > $OUT/bench_1.json 2> $OUT/bench_1.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $g --master-addr 127.0.0.1 --master-port $((29500+g)) bench.py --gpus $g --steps 20 --warmup 3 > $OUT/bench_$g.json 2> $OUT/bench_$g.err
  fi
  echo "bench $g rc=$?" >> $OUT/rc.txt
done
ncu --query-metrics 2>/dev/null | grep -i -E "^nvl|nvlrx|nvltx" > $OUT/ncu_nvlink_metrics.txt
for mode in read copy push; do
  timeout 300 python tools/p2p_once.py $mode > $OUT/p2p_once_$mode.json 2> $OUT/p2p_once_$mode.err
  timeout 600 ncu --metrics nvlrx__bytes.sum,nvltx__bytes.sum,gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
      -k regex:"hbm_read_fold|hbm_copy" --csv --log-file $OUT/ncu_p2p_$mode.csv python tools/p2p_once.py $mode > $OUT/ncu_p2p_$mode.log 2>&1
  echo "ncu $mode rc=$?" >> $OUT/rc.txt
done
cat $OUT/rc.txt; tail -3 $OUT/pytest_gpu.txt; tail -2 $OUT/pytest_zz_alone.txt
