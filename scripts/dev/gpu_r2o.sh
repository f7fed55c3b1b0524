#!/bin/bash
# round 2, call O (1 GPU): certificate bound derived from d (1.3e-7 * d): suite + bench exactness counters
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/r2o_gpu_suite.log 2>&1; echo "suite rc=$?"
tail -15 gpurun_out/r2o_gpu_suite.log
timeout 600 python bench.py --steps 10 --no-cpu-baseline --no-dropin --no-e2e > gpurun_out/r2o_bench_n1.json 2> gpurun_out/r2o_bench_n1.err; echo "bench rc=$?"
timeout 400 python bench.py --k 100 --steps 5 --no-cpu-baseline --no-dropin --no-e2e > gpurun_out/r2o_bench_k100.json 2> gpurun_out/r2o_bench_k100.err; echo "k100 rc=$?"
python - <<'PY'
import json
for f in ("gpurun_out/r2o_bench_n1.json", "gpurun_out/r2o_bench_k100.json"):
    try:
        j = json.loads(open(f).read().strip().splitlines()[-1])
        print({k: j[k] for k in ("value", "ms_per_step", "recall_at_10", "stages_ms", "parity_sample", "exactness")})
    except Exception as e:
        print("no line", f, e)
PY
