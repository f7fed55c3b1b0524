#!/bin/bash
# round 2, call Z (1 GPU): K5 re-scores only the candidates that can reach the top k -- suite, k = 10 / k = 100 lines, 1/8 shard
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/r2z_gpu_suite.log 2>&1; echo "gpu tests rc=$?"; tail -4 gpurun_out/r2z_gpu_suite.log
timeout 600 python bench.py --steps 20 --no-cpu-baseline --no-dropin --no-e2e > gpurun_out/r2z_bench_n1.json 2> gpurun_out/r2z_bench_n1.err; echo "bench rc=$?"
timeout 400 python bench.py --k 100 --steps 5 --no-cpu-baseline --no-dropin --no-e2e > gpurun_out/r2z_bench_k100.json 2> gpurun_out/r2z_bench_k100.err; echo "k100 rc=$?"
python - <<'PY'
import json
for f in ("gpurun_out/r2z_bench_n1.json", "gpurun_out/r2z_bench_k100.json"):
    try:
        j = json.loads(open(f).read().strip().splitlines()[-1])
        print({k: j[k] for k in ("value", "ms_per_step", "recall_at_10", "stages_ms", "parity_sample", "exactness")})
    except Exception as e:
        print("no line", f, e)
PY
timeout 300 python scripts/dev/shard_emulate.py --worlds 8 --ranks 0 --chunks 0 --out gpurun_out/r2z_shard.json > gpurun_out/r2z_shard.log 2>&1; echo "shard rc=$?"
python - <<'PY'
import json
try:
    for r in json.load(open("gpurun_out/r2z_shard.json")):
        print({k: (round(v, 4) if isinstance(v, float) else v) for k, v in r.items() if k in ("world", "rank", "ms_partial_search", "ms_rescore_avg", "ms_finalize")})
except Exception as e:
    print("no shard report", e)
PY
