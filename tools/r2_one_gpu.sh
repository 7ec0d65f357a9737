#!/bin/bash
# Round-2 check on a 1-GPU lease: GPU suite, the no-alloc file alone in a fresh process, smoke, both bench arms, then (each only
# after its command has exited 0 without ncu) the launch lists of smoke() and bench.py and one `ncu --set full` capture of the kernels.
OUT=gpurun_out/r2_1gpu
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q -x > $OUT/pytest_gpu.txt 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $OUT/pytest_zz_alone.txt 2>&1; echo "zz alone rc=$?" | tee -a $OUT/rc.txt
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.txt 2>&1; echo "smoke rc=$?" | tee -a $OUT/rc.txt
timeout 400 python bench.py --impl reference --steps 20 --warmup 3 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "bench ref rc=$?" | tee -a $OUT/rc.txt
timeout 400 python bench.py --steps 20 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" | tee -a $OUT/rc.txt
timeout 300 ./gpu-feature-discovery_b200/gpu-feature-discovery --oneshot --no-timestamp --probe-hbm -o $OUT/labels_probe_hbm.txt 2> $OUT/labels_probe_hbm.err; echo "probe labels rc=$?" | tee -a $OUT/rc.txt
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_smoke.csv python -c "import __graft_entry__ as g; g.smoke()" > $OUT/ncu_smoke.log 2>&1; echo "ncu smoke rc=$?" | tee -a $OUT/rc.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_bench.csv python bench.py --steps 5 --warmup 3 > $OUT/ncu_bench.log 2>&1; echo "ncu bench rc=$?" | tee -a $OUT/rc.txt
timeout 200 python tools/probe_once.py > $OUT/probe_once.log 2>&1; echo "probe_once rc=$?" | tee -a $OUT/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"hbm_read_fold|hbm_copy" -c 6 -o $OUT/prof_r2 -f python tools/probe_once.py > $OUT/ncu_full.log 2>&1; echo "ncu full rc=$?" | tee -a $OUT/rc.txt
ncu -i $OUT/prof_r2.ncu-rep --page raw --csv > $OUT/prof_r2_raw.csv 2>/dev/null
tail -3 $OUT/pytest_gpu.txt; tail -2 $OUT/pytest_zz_alone.txt; cat $OUT/bench.json
