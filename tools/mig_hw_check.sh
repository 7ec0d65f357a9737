#!/bin/bash
# MIG single/mixed on a REAL B200 (BASELINE configs[3]/[4]): enable MIG on GPU 0, walk through a set of partition layouts,
# and at each one diff the product's label file against oracle(pynvml) and the C replay (tools/mig_hw_stage.py).
# An EXIT trap ALWAYS destroys every instance and disables MIG again: a pool GPU is never left partitioned.  No GPU reset is
# requested; if the mode change reports "pending" (or is refused) the script records the exact output and stops there.
# Everything lands under gpurun_out/r2_mig/.
OUT=gpurun_out/r2_mig
mkdir -p $OUT
SMI="timeout 120 nvidia-smi"
BIN=./gpu-feature-discovery_b200/gpu-feature-discovery

restore() {
  {
    echo "== restore $(date -u +%T)"
    $SMI mig -i 0 -dci
    $SMI mig -i 0 -dgi
    $SMI -i 0 -mig 0
    $SMI -i 0 --query-gpu=mig.mode.current,mig.mode.pending --format=csv
    $SMI -L
  } >> $OUT/restore.log 2>&1
}
trap restore EXIT

mode() { $SMI -i 0 --query-gpu=mig.mode.current --format=csv,noheader 2>/dev/null | tr -d ' \r'; }
pending() { $SMI -i 0 --query-gpu=mig.mode.pending --format=csv,noheader 2>/dev/null | tr -d ' \r'; }
stage() { timeout 600 python tools/mig_hw_stage.py $OUT "$1" >> $OUT/stages.jsonl 2>> $OUT/stages.err; echo "stage $1 rc=$?" >> $OUT/log.txt; }
wipe() { { echo "== wipe before $1"; $SMI mig -i 0 -dci; $SMI mig -i 0 -dgi; } >> $OUT/partition.log 2>&1; }
layout() {  # layout <tag> <nvidia-smi mig args...>
  local tag=$1; shift
  wipe $tag
  { echo "== $tag: nvidia-smi mig -i 0 $*"; $SMI mig -i 0 "$@"; echo "rc=$?"; $SMI -L; } >> $OUT/partition.log 2>&1
}

{ id; $SMI -L; $SMI -i 0 --query-gpu=name,driver_version,mig.mode.current,mig.mode.pending,persistence_mode --format=csv; } > $OUT/before.txt 2>&1
ls -la /proc/driver/nvidia/capabilities/ /proc/driver/nvidia/capabilities/mig/ >> $OUT/before.txt 2>&1
env | grep -i -E "^NVIDIA_|^CUDA_VIS" >> $OUT/before.txt 2>&1

stage s0_mig_disabled

$SMI -i 0 -mig 1 > $OUT/enable.txt 2>&1
echo "rc=$?" >> $OUT/enable.txt
echo "current=$(mode) pending=$(pending)" >> $OUT/enable.txt
if [ "$(mode)" != "Enabled" ]; then
  echo "MIG mode could not be enabled on this pool (see enable.txt): no MIG stage runs" >> $OUT/log.txt
else

$SMI mig -i 0 -lgip > $OUT/lgip.txt 2>&1
$SMI mig -i 0 -lcip >> $OUT/lgip.txt 2>&1

# (1) MIG enabled, no instance yet: single must give MIG-INVALID (mig-strategy.go:193-196), mixed only the GPU labels
stage s1_enabled_empty

# (2) config[3]: uniform 7 x 1g.23gb (nvidia-smi profile 19 = NVML GI profile id 0)
layout s2_single_7x1g -cgi 19,19,19,19,19,19,19 -C
stage s2_single_7x1g
# the generic -m gpu label tests (product vs oracle(pynvml), vs replay, C-ABI getters vs pynvml) in this state
timeout 600 python -m pytest tests/test_gpu_labels.py -q -x -k "matches or getters or mig" > $OUT/pytest_s2.txt 2>&1; echo "rc=$?" >> $OUT/pytest_s2.txt
# pass latency with 7 MIG devices: product (default threads / serial) vs the reference's call sequence
for rep in 1 2; do
  $BIN --passes 12 --sleep-interval 0s --timing --no-timestamp --mig-strategy=single -o /tmp/mig_p 2>&1 | grep "timing:" | sed "s/^/product /" >> $OUT/pass_timing_s2.txt
  ./oracle/ref_replay --passes 12 --no-timestamp --mig-strategy single -o /tmp/mig_r >> $OUT/pass_timing_s2.txt 2>&1
done

# (3) config[4]: heterogeneous 3g.90gb + 4g.90gb -> mixed has two resources, single is MIG-INVALID
layout s3_hetero_3g_4g -cgi 9,5 -C
stage s3_hetero_3g_4g

# (4) a GI profile id >= 8 (1g.45gb = nvidia-smi 15 = NVML id 9): the reference fails the pass; --current-tables resolves it
layout s4_1g45gb -cgi 15 -C
stage s4_1g45gb

# (5) +me, 1g and 2g side by side
layout s5_me_1g_2g -cgi 20,19,14 -C
stage s5_me_1g_2g

# (6) the whole GPU as one instance
layout s6_7g -cgi 0 -C
stage s6_7g

# (7) compute-instance split: one 3g.90gb GI carrying 1c compute instances -> "1c.3g.90gb"
wipe s7_ci_split
{ echo "== s7_ci_split"; $SMI mig -i 0 -cgi 9; echo "rc=$?"; $SMI mig -i 0 -lcip; $SMI mig -i 0 -cci 0,0; echo "rc=$?"; $SMI -L; } >> $OUT/partition.log 2>&1
stage s7_ci_split

# (8) 2 x 3g.90gb uniform: single with count 2
layout s8_single_2x3g -cgi 9,9 -C
stage s8_single_2x3g
fi

restore
trap - EXIT
stage s9_restored
echo "done" >> $OUT/log.txt
# the box must be exactly as found: the ordinary GPU suite runs on the restored GPU
timeout 900 python -m pytest tests -m gpu -q -x > $OUT/pytest_gpu_after_restore.txt 2>&1; echo "rc=$?" >> $OUT/pytest_gpu_after_restore.txt
