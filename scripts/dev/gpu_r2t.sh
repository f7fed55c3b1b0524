#!/bin/bash
# round 2, call T (1 GPU): L2 eviction hints on the chunk loads (last reader = evict_first), nb = 4 and nb = 8, interleaved
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_pair_kernel.py -q -x -p no:cacheprovider > gpurun_out/r2t_tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/r2t_tests.log
for NB in 4 8; do
timeout 500 python scripts/dev/pair_ab.py --skip-diag --steps 10 --buckets $NB --out gpurun_out/r2t_ab_nb$NB.json --seq "3:0:0:0:0:0,3:0:0:0:0:1,3:0:0:0:0:0,3:0:0:0:0:1,3:0:0:0:1:0,3:0:0:0:1:1" > gpurun_out/r2t_ab_nb$NB.log 2>&1; echo "ab nb=$NB rc=$?"
python - <<PY
import json
try:
    for r in json.load(open("gpurun_out/r2t_ab_nb$NB.json"))["runs"]:
        print({k: (round(v, 3) if isinstance(v, float) else v) for k, v in r.items() if k in ("item_order", "l2_hint", "ms_per_search", "k3_ms_avg", "k3_sm_mhz", "useful_tflops", "ids_equal_to_first")})
except Exception as e:
    print("no ab report", e)
PY
done
