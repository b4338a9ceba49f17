#!/bin/bash
# GPU box: launch list of the bench step and a full-metric capture of blend_backward alone (after a kernel change).
# Usage: gpurun --timeout 900 -- 'bash tools/capture_bwd.sh r2g'
set -u
tag=${1:-r2g}
out=gpurun_out
mkdir -p $out
export BENCH_NO_SPINUP=1 BENCH_NO_SAMPLER=1
A="--no-cpu-baseline --sub none"
python bench.py $A --steps 3 --warmup 3 > $out/${tag}_plain_short.log 2>&1 &&
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/${tag}_launches_ours.csv \
    python bench.py $A --steps 3 --warmup 3 > $out/${tag}_ncu_ours.log 2>&1
BENCH_E2E_EAGER=1 python bench.py $A --steps 8 --warmup 8 > $out/${tag}_plain_short_eager.log 2>&1 &&
BENCH_E2E_EAGER=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:blend_backward_kernel \
    -s 8 -c 8 -o /tmp/${tag}_full_bwd -f python bench.py $A --steps 8 --warmup 8 > $out/${tag}_ncu_full.log 2>&1
ncu -i /tmp/${tag}_full_bwd.ncu-rep --page raw --csv > $out/${tag}_full_bwd_raw.csv 2>/dev/null
ncu -i /tmp/${tag}_full_bwd.ncu-rep --page source --csv --launch-skip 0 --launch-count 1 2>/dev/null \
    | python tools/ncu_regions.py > $out/${tag}_regions_blend_backward.txt
tail -c 200 $out/${tag}_ncu_full.log
