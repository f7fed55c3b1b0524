#!/bin/bash
# round 2, call A: the reworked CTA-pair kernel (two epilogue warp sets, 32-column half-tile epilogue, relaxed
# remote arrive): diagnose, pair tests, A/B against its own variants, full GPU suite, bench line.
mkdir -p gpurun_out
timeout 240 python scripts/dev/pair_ab.py --diag-only --out gpurun_out/r2a_diag.json > gpurun_out/r2a_diag.log 2>&1; echo "diag rc=$?"
tail -3 gpurun_out/r2a_diag.log
timeout 400 python -m pytest tests/test_gpu_pair_kernel.py -q -x -p no:cacheprovider > gpurun_out/r2a_pair_tests.log 2>&1; echo "pair tests rc=$?"
tail -6 gpurun_out/r2a_pair_tests.log
timeout 600 python scripts/dev/pair_ab.py --exp --skip-diag --steps 10 --out gpurun_out/r2a_ab.json \
  --seq "3:0:0:2,3:0:0:1,1:0:0,3:0:64:1,3:0:0:2,3:0:0:1,3:0:64:2,3:0:2:2,3:0:2:1,3:0:6:2,3:0:10:2,3:0:0:2,3:16128:0:2,3:64000:0:2" \
  > gpurun_out/r2a_ab.log 2>&1; echo "ab rc=$?"
grep -o '"rescore": [0-9], "dbg_flags": [0-9]*, "epi_sets": [0-9].*"k3_sm_mhz": [0-9.]*' gpurun_out/r2a_ab.log | cut -c1-200
timeout 900 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/r2a_gpu_tests.log 2>&1; echo "gpu tests rc=$?"
tail -5 gpurun_out/r2a_gpu_tests.log
timeout 500 python bench.py --steps 20 --warmup 3 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench rc=$?"
tail -c 2500 gpurun_out/r2a_bench.json
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
