#!/bin/bash
# round 2, call F (1 GPU): suite after the K4 fix, k = 100 per-rank costs + merge kernel timing, fresh N = 1 bench line
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r2f_gpu_tests.log 2>&1; echo "gpu tests rc=$?"
tail -4 gpurun_out/r2f_gpu_tests.log
timeout 400 python scripts/dev/shard_emulate.py --worlds 8 --ranks 0 --chunks 0 --k 100 --merge --out gpurun_out/r2f_shard8_k100.json > gpurun_out/r2f_shard8_k100.log 2>&1; echo "shard8 k100 rc=$?"
grep -o '"world".*' gpurun_out/r2f_shard8_k100.log | cut -c1-400
timeout 400 python scripts/dev/shard_emulate.py --worlds 8 --ranks 0 --chunks 0 --k 10 --merge --out gpurun_out/r2f_shard8_k10.json > gpurun_out/r2f_shard8_k10.log 2>&1; echo "shard8 k10 rc=$?"
grep -o '"world".*' gpurun_out/r2f_shard8_k10.log | cut -c1-400
timeout 700 python bench.py --steps 20 --warmup 3 > gpurun_out/r2f_bench.json 2> gpurun_out/r2f_bench.err; echo "bench rc=$?"
tail -c 300 gpurun_out/r2f_bench.err
python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/r2f_bench.json").read().strip().splitlines()[-1])
    keep = {k: j[k] for k in ("value", "ms_per_step", "recall_at_10", "e2e", "e2e_dropin", "stages_ms", "exactness", "store_build_seconds")}
    keep["roofline"] = {k: j["roofline"][k] for k in ("achieved", "frac", "frac_vs_sustained_peak", "kernel_ms", "tile_efficiency", "traffic")}
    cb = j.get("cpu_baseline") or {}
    keep["cpu"] = {k: (cb.get(k) if k in ("value", "cores") else {kk: vv for kk, vv in (cb.get(k) or {}).items() if kk != "checker"}) for k in ("value", "cores", "nav_parity", "category_parity", "oracle_parity")}
    keep["clocks"] = j["clocks"]
    print(json.dumps(keep, indent=1)[:5000])
except Exception as e:
    print("no bench line", e)
PY
