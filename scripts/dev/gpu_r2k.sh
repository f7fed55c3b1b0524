#!/bin/bash
# round 2, call K (1 GPU): --set full of nav_kernel and finalize_kernel (the two largest non-K3 pieces of a step)
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-dropin --gt-queries 100"
timeout 600 $B > gpurun_out/r2k_plain.log 2>&1; rc=$?; echo "plain rc=$rc"; tail -c 600 gpurun_out/r2k_plain.log
if [ $rc -eq 0 ]; then
  timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:^(nav_kernel|finalize_kernel)' -s 6 -c 2 -o gpurun_out/prof_r2k $B > gpurun_out/r2k_ncu.log 2>&1; echo "ncu rc=$?"
  tail -n 4 gpurun_out/r2k_ncu.log
fi
