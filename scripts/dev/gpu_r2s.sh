#!/bin/bash
# round 2, call S (1 GPU): readers-per-chunk histogram; planner A/B item_order 0 vs 2 (three-tile buckets as three half pairs)
mkdir -p gpurun_out
timeout 300 python scripts/dev/qt_hist.py --out gpurun_out/r2s_qt_hist.json > gpurun_out/r2s_qt_hist.log 2>&1; echo "hist rc=$?"; tail -22 gpurun_out/r2s_qt_hist.log
timeout 500 python scripts/dev/pair_ab.py --skip-diag --steps 10 --out gpurun_out/r2s_ab.json --seq "3:0:0:0:0,3:0:0:0:2,3:0:0:0:0,3:0:0:0:2,3:0:0:0:0,3:0:0:0:2,3:0:0:0:1" > gpurun_out/r2s_ab.log 2>&1; echo "ab rc=$?"
python - <<'PY'
import json
try:
    for r in json.load(open("gpurun_out/r2s_ab.json"))["runs"]:
        print({k: (round(v, 3) if isinstance(v, float) else v) for k, v in r.items() if k in ("item_order", "ms_per_search", "k3_ms_avg", "k3_sm_mhz", "useful_tflops", "ids_equal_to_first", "work_items")})
except Exception as e:
    print("no ab report", e)
PY
tail -3 gpurun_out/r2s_ab.log
