#!/bin/bash
# round 2, call G (1 GPU): bisection merge (tests + timing), DRAM re-read isolation experiment
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_seeded.py tests/test_gpu_exact_fallback.py -q -p no:cacheprovider > gpurun_out/r2g_tests.log 2>&1; echo "tests rc=$?"
tail -4 gpurun_out/r2g_tests.log
timeout 400 python scripts/dev/pair_ab.py --exp --skip-diag --steps 10 --out gpurun_out/r2g_ab.json --seq "3:0:0:2,3:0:256:2,3:0:0:2,3:0:256:2,3:0:0:2" > gpurun_out/r2g_ab.log 2>&1; echo "ab rc=$?"
python - <<'PY'
import json
try:
    for r in json.load(open("gpurun_out/r2g_ab.json"))["runs"]:
        print({k: (round(v, 3) if isinstance(v, float) else v) for k, v in r.items() if k in ("dbg_flags", "epi_sets", "ms_per_search", "k3_ms_avg", "k3_sm_mhz", "useful_tflops", "ids_equal_to_first")})
except Exception as e:
    print("no ab report", e)
PY
timeout 400 python scripts/dev/shard_emulate.py --worlds 8 --ranks 0 --chunks 0 --k 100 --merge --out gpurun_out/r2g_shard8_k100.json > gpurun_out/r2g_shard8_k100.log 2>&1; echo "shard8 k100 rc=$?"
grep -o '"world".*' gpurun_out/r2g_shard8_k100.log | cut -c1-330
