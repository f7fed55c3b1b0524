#!/bin/bash
# round 2, call X (1 GPU): DRAM bytes of K3 with lock-step followers on (does the follower's read hit L2?)
mkdir -p gpurun_out
export LMI_OPTIONS="lockstep=1"
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-dropin --gt-queries 100"
timeout 300 $B > gpurun_out/r2x_plain.log 2>&1; rc=$?; echo "plain rc=$rc"
if [ $rc -eq 0 ]; then
  timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__cycles_elapsed.max --clock-control none -k regex:rescore_t -s 3 -c 2 --csv --log-file gpurun_out/r2x_lockstep_k3.csv $B > gpurun_out/r2x_ncu.log 2>&1; echo "ncu rc=$?"
  cat gpurun_out/r2x_lockstep_k3.csv | tail -14
fi
