#!/bin/bash
# round 2, call W (1 GPU): lock-step followers (half pair follows its chunk's full pair) -- per-rank cost on 1/8 and 1/4 shards, unsharded A/B
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_pair_kernel.py -q -x -p no:cacheprovider -k "equals_default or work_list" > gpurun_out/r2w_tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/r2w_tests.log
timeout 400 python scripts/dev/shard_emulate.py --worlds 8,4 --ranks 0,3 --chunks 0 --item-orders 0 --locksteps 0,1,0,1 --out gpurun_out/r2w_shard.json > gpurun_out/r2w_shard.log 2>&1; echo "shard rc=$?"
python - <<'PY'
import json
try:
    for r in json.load(open("gpurun_out/r2w_shard.json")):
        print({k: (round(v, 4) if isinstance(v, float) else v) for k, v in r.items() if k in ("world", "rank", "lockstep", "partial_equal_to_first", "ms_partial_search", "ms_rescore_avg", "ms_finalize", "work_items")})
except Exception as e:
    print("no shard report", e)
PY
tail -3 gpurun_out/r2w_shard.log
timeout 400 python scripts/dev/pair_ab.py --skip-diag --steps 10 --out gpurun_out/r2w_ab.json --seq "3:0:0:0:0:0:0,3:0:0:0:0:0:1,3:0:0:0:0:0:0,3:0:0:0:0:0:1" > gpurun_out/r2w_ab.log 2>&1; echo "ab rc=$?"
python - <<'PY'
import json
try:
    for r in json.load(open("gpurun_out/r2w_ab.json"))["runs"]:
        print({k: (round(v, 3) if isinstance(v, float) else v) for k, v in r.items() if k in ("lockstep", "ms_per_search", "k3_ms_avg", "k3_sm_mhz", "useful_tflops", "ids_equal_to_first")})
except Exception as e:
    print("no ab report", e)
PY
