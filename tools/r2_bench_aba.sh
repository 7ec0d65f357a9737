#!/bin/bash
# reference arm, product arm, reference arm again on one lease (how much two runs of the SAME arm differ is the yardstick for
# any difference between the arms), then the label / (f)-row / integration GPU tests.
OUT=gpurun_out/${1:-r2b}
mkdir -p $OUT
python bench.py --impl reference --steps 30 --warmup 5 > $OUT/bench_ref.json 2> $OUT/bench_ref.err
python bench.py --steps 30 --warmup 5 > $OUT/bench.json 2> $OUT/bench.err
python bench.py --impl reference --steps 30 --warmup 5 > $OUT/bench_ref2.json 2> $OUT/bench_ref2.err
python -m pytest tests/test_gpu_labels.py tests/test_gpu_f_rows.py tests/test_integration_shape.py -m gpu -q 2>&1 | tail -2 | tee $OUT/pytest_subset.txt
python - $OUT <<'PY'
import json, sys
for f in ("bench_ref", "bench", "bench_ref2"):
    d = json.loads(open("%s/%s.json" % (sys.argv[1], f)).read().strip().splitlines()[-1])
    print(f, d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"].get("cpu_baseline_median"), d.get("labeling_pass_phases_ms"),
          d["cpu_baseline"]["value"], d["cpu_baseline"].get("phases_ms"), d.get("label_diff"))
PY
