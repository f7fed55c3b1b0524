#!/bin/bash
# round 2, call I (1 GPU): work-list order A/B (full pairs first vs bucket-major), unsharded and on a 1/8 shard
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_pair_kernel.py tests/test_gpu_seeded.py -q -x -p no:cacheprovider > gpurun_out/r2i_tests.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/r2i_tests.log
timeout 400 python scripts/dev/pair_ab.py --skip-diag --steps 10 --out gpurun_out/r2i_ab.json --seq "3:0:0:0:0,3:0:0:0:1,3:0:0:0:0,3:0:0:0:1,3:0:0:0:0,3:0:0:0:1" > gpurun_out/r2i_ab.log 2>&1; echo "ab rc=$?"
python - <<'PY'
import json
try:
    for r in json.load(open("gpurun_out/r2i_ab.json"))["runs"]:
        print({k: (round(v, 3) if isinstance(v, float) else v) for k, v in r.items() if k in ("item_order", "ms_per_search", "k3_ms_avg", "k3_sm_mhz", "useful_tflops", "ids_equal_to_first")})
except Exception as e:
    print("no ab report", e)
PY
timeout 400 python scripts/dev/shard_emulate.py --worlds 8,4 --ranks 0,3 --chunks 0 --item-orders 0,1,0,1 --out gpurun_out/r2i_shard.json > gpurun_out/r2i_shard.log 2>&1; echo "shard rc=$?"
python - <<'PY'
import json
try:
    for r in json.load(open("gpurun_out/r2i_shard.json")):
        print({k: (round(v, 4) if isinstance(v, float) else v) for k, v in r.items() if k in ("world", "rank", "item_order", "ms_partial_search", "ms_rescore_avg", "ms_finalize", "work_items")})
except Exception as e:
    print("no shard report", e)
PY
