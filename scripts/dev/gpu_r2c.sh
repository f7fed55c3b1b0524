#!/bin/bash
# round 2, call C: full GPU suite, per-rank shard costs on one GPU (chunk rule), recall-vs-QPS sweep, ncu evidence
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r2c_gpu_tests.log 2>&1; echo "gpu tests rc=$?"
tail -6 gpurun_out/r2c_gpu_tests.log
timeout 420 python scripts/dev/shard_emulate.py --worlds 8 --ranks 0,5 --chunks 0,8192,16384,32768 --out gpurun_out/r2c_shard8.json > gpurun_out/r2c_shard8.log 2>&1; echo "shard8 rc=$?"
grep -o '"world".*' gpurun_out/r2c_shard8.log | cut -c1-330
timeout 300 python scripts/dev/shard_emulate.py --worlds 2,4 --ranks 0 --chunks 0 --out gpurun_out/r2c_shard24.json > gpurun_out/r2c_shard24.log 2>&1; echo "shard24 rc=$?"
grep -o '"world".*' gpurun_out/r2c_shard24.log | cut -c1-330
timeout 300 python scripts/dev/shard_emulate.py --worlds 8 --ranks 0 --policy stripes --chunks 0,16384 --out gpurun_out/r2c_shard8s.json > gpurun_out/r2c_shard8s.log 2>&1; echo "shard8 stripes rc=$?"
grep -o '"world".*' gpurun_out/r2c_shard8s.log | cut -c1-330
timeout 600 python scripts/sweep.py --buckets 1,2,3,4,8,16 --ks 10,100 --steps 5 --out gpurun_out/r2c_sweep > gpurun_out/r2c_sweep.log 2>&1; echo "sweep rc=$?"
cat gpurun_out/r2c_sweep.md 2>/dev/null | head -30
SKIP_LIST= EXTRA="--no-dropin" timeout 700 bash scripts/ncu_bench.sh r2c; echo "ncu rc=$?"
ls -la gpurun_out | grep r2c
