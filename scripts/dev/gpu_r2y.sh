#!/bin/bash
# round 2, call Y (1 GPU): final state -- smoke, whole GPU suite, one short bench line
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('__SMOKE_OK__')" > gpurun_out/r2y_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2y_smoke.log
timeout 900 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/r2y_gpu_suite.log 2>&1; echo "gpu tests rc=$?"; tail -4 gpurun_out/r2y_gpu_suite.log
timeout 600 python bench.py --steps 20 --no-cpu-baseline --no-dropin > gpurun_out/r2y_bench_n1.json 2> gpurun_out/r2y_bench_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/r2y_bench_n1.json").read().strip().splitlines()[-1])
    print({k: j[k] for k in ("value", "ms_per_step", "recall_at_10", "stages_ms", "exactness", "parity_sample")})
    print(j["e2e"], j["roofline"]["frac"], j["roofline"]["kernel_ms"], j["clocks"]["sm_mhz_in_kernel"])
except Exception as e:
    print("no bench line", e)
PY
