#!/bin/bash
# GPU box: plain runs first (must exit 0), then the ncu launch lists and one full-metric capture of a step.
# Usage (from the repo root): gpurun --timeout 1200 -- 'bash tools/capture_profiles.sh r1'
set -u
tag=${1:-r1}
out=gpurun_out
mkdir -p $out
timeout 300 python bench.py --impl reference --steps 50 --warmup 5 > $out/${tag}_plain_ref.log 2>&1 || exit 1
timeout 300 python bench.py > $out/${tag}_plain_ours.log 2>&1 || exit 1
export BENCH_NO_SPINUP=1 BENCH_NO_SAMPLER=1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches_ours.csv \
    python bench.py --no-cpu-baseline --steps 3 --warmup 3 > $out/${tag}_ncu_ours.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches_ref.csv \
    python bench.py --impl reference --steps 3 --warmup 3 > $out/${tag}_ncu_ref.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'^(preprocess|tile_scan|scatter_pairs|tile_sort|blend_forward|blend_backward|preprocess_backward)' \
    -c 9 -o $out/${tag}_full_ours -f python bench.py --no-cpu-baseline --steps 1 --warmup 1 > $out/${tag}_ncu_full.log 2>&1
tail -1 $out/${tag}_plain_ours.log | cut -c1-400
tail -1 $out/${tag}_plain_ref.log | cut -c1-300
