#!/bin/bash
# One gpurun call: diagnose + test + A/B the CTA-pair rescoring kernel, then the regular GPU suite.
# Everything is wrapped in `timeout` (a wrong mbarrier protocol hangs, it does not fail).
mkdir -p gpurun_out
export LMI_TEST_PAIR=1
timeout 180 python scripts/dev/pair_ab.py --diag-only --out gpurun_out/pair_diag.json > gpurun_out/pair_diag.log 2>&1; echo "diag rc=$?"
tail -4 gpurun_out/pair_diag.log
timeout 300 python -m pytest tests/test_gpu_pair_kernel.py -q -p no:cacheprovider > gpurun_out/pair_tests.log 2>&1; echo "pair tests rc=$?"
tail -15 gpurun_out/pair_tests.log
timeout 420 python scripts/dev/pair_ab.py --out gpurun_out/pair_ab.json > gpurun_out/pair_ab.log 2>&1; echo "ab rc=$?"
tail -8 gpurun_out/pair_ab.log
unset LMI_TEST_PAIR
timeout 420 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/gpu_tests.log 2>&1; echo "gpu tests rc=$?"
tail -5 gpurun_out/gpu_tests.log
if python -c "import json,sys; d=json.load(open('gpurun_out/pair_diag.json')); sys.exit(0 if d['diag']['rows_mismatch']==0 and d['diag_odd_tile']['rows_mismatch']==0 else 1)"; then
  timeout 400 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --rescore tc_pair > gpurun_out/bench_pair.json 2> gpurun_out/bench_pair.err; echo "bench pair rc=$?"
  tail -c 1500 gpurun_out/bench_pair.json
fi
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
