#!/bin/bash
# GPU box: plain runs first (must exit 0), then the ncu launch lists and one full-metric capture of a step.
# Usage (from the repo root): gpurun --timeout 1500 -- 'bash tools/capture_profiles.sh r2'
set -u
tag=${1:-r2}
out=gpurun_out
mkdir -p $out
timeout 300 python bench.py --impl reference --steps 50 --warmup 5 > $out/${tag}_plain_ref.json 2> $out/${tag}_plain_ref.err || exit 1
timeout 600 python bench.py > $out/${tag}_plain_ours.json 2> $out/${tag}_plain_ours.err || exit 1
export BENCH_NO_SPINUP=1 BENCH_NO_SAMPLER=1
A="--no-cpu-baseline --sub none"
python bench.py $A --steps 3 --warmup 3 > $out/${tag}_plain_short.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/${tag}_launches_ours.csv \
    python bench.py $A --steps 3 --warmup 3 > $out/${tag}_ncu_ours.log 2>&1
python bench.py --impl reference $A --steps 3 --warmup 3 > $out/${tag}_plain_short_ref.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/${tag}_launches_ref.csv \
    python bench.py --impl reference $A --steps 3 --warmup 3 > $out/${tag}_ncu_ref.log 2>&1
BENCH_E2E_EAGER=1 python bench.py $A --steps 3 --warmup 3 > $out/${tag}_plain_short_eager.log 2>&1 &&
BENCH_E2E_EAGER=1 timeout 900 ncu --set full --clock-control none --import-source on \
    -k regex:'(preprocess_kernel|tile_scan_kernel|scatter_pairs_kernel|tile_sort_kernel|blend_forward_kernel|blend_backward_kernel|preprocess_backward_kernel)' \
    -s 21 -c 7 -o $out/${tag}_full_ours -f python bench.py $A --steps 3 --warmup 3 > $out/${tag}_ncu_full.log 2>&1
tail -c 300 $out/${tag}_ncu_full.log
head -c 400 $out/${tag}_plain_ours.json
head -c 300 $out/${tag}_plain_ref.json
