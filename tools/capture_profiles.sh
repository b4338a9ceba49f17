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
# one full cycle of the eight bench views (eager launches; the first 8 steps are warm-up): 8 x 7 kernels
BENCH_E2E_EAGER=1 python bench.py $A --steps 8 --warmup 8 > $out/${tag}_plain_short_eager.log 2>&1 &&
BENCH_E2E_EAGER=1 timeout 1200 ncu --set full --clock-control none --import-source on \
    -k regex:'(preprocess_kernel|tile_scan_kernel|scatter_pairs_kernel|tile_sort_kernel<4, 256|blend_forward_kernel|blend_backward_kernel|preprocess_backward_kernel)' \
    -s 56 -c 56 -o /tmp/${tag}_full_ours -f python bench.py $A --steps 8 --warmup 8 > $out/${tag}_ncu_full.log 2>&1
# gpurun brings back at most 64 MiB: export the raw metric page and the per-kernel source pages here, drop the report
ncu -i /tmp/${tag}_full_ours.ncu-rep --page raw --csv > $out/${tag}_full_ours_raw.csv 2>/dev/null
for k in blend_backward blend_forward tile_sort scatter_pairs preprocess_kernel preprocess_backward; do
    ncu -i /tmp/${tag}_full_ours.ncu-rep --page source --csv --kernel-name regex:$k --launch-skip 0 --launch-count 1 2>/dev/null \
        | python tools/ncu_regions.py > $out/${tag}_regions_$k.txt
done
tail -c 300 $out/${tag}_ncu_full.log
head -c 400 $out/${tag}_plain_ours.json
head -c 300 $out/${tag}_plain_ref.json
# the fused L1 + SSIM loss at the bench image size (forward, reduce, backward)
cat > /tmp/ssim_run.py <<'PY'
import sys; sys.path.insert(0, ".")
import torch
from splatloc_b200.ssim_loss import l1_ssim_loss
g = torch.Generator().manual_seed(0)
gt = torch.rand((3, 480, 640), generator=g).cuda(); img = (gt + 0.05 * torch.randn((3, 480, 640), generator=g).cuda()).clamp(0, 1)
for _ in range(3):
    x = img.clone().requires_grad_(True); l1_ssim_loss(x, gt, 0.2).backward()
torch.cuda.synchronize(); print("ok")
PY
python /tmp/ssim_run.py > $out/${tag}_ssim_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:ssim_ -s 6 -c 3 -o /tmp/${tag}_full_ssim -f \
    python /tmp/ssim_run.py > $out/${tag}_ncu_ssim.log 2>&1
ncu -i /tmp/${tag}_full_ssim.ncu-rep --page raw --csv > $out/${tag}_full_ssim_raw.csv 2>/dev/null
