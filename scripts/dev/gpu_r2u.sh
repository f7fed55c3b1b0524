#!/bin/bash
# round 2, call U (1 GPU): round-end rehearsal on the final code -- smoke(), reference arm, GPU suite, the default bench line, 1-GPU sweep
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('__SMOKE_OK__')" > gpurun_out/r2u_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/r2u_smoke.log
timeout 900 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/r2u_gpu_suite.log 2>&1; echo "gpu tests rc=$?"; tail -4 gpurun_out/r2u_gpu_suite.log
( time timeout 900 python bench.py > gpurun_out/r2u_bench_n1.json 2> gpurun_out/r2u_bench_n1.err ) 2>&1 | grep real; echo "bench rc=$?"
tail -c 300 gpurun_out/r2u_bench_n1.err
python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/r2u_bench_n1.json").read().strip().splitlines()[-1])
    print({k: j[k] for k in ("value", "ms_per_step", "recall_at_10", "stages_ms", "exactness", "gpu_launches")})
    print("e2e", j["e2e"]); print("dropin", j["e2e_dropin"]); print("roof", j["roofline"]); print("clocks", j["clocks"])
    c = j["cpu_baseline"]; print("cpu", {k: c[k] for k in c if k not in ("note",)})
except Exception as e:
    print("no bench line", e)
PY
( time timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2u_reference_arm.json 2> gpurun_out/r2u_reference_arm.err ) 2>&1 | grep real; echo "reference arm rc=$?"
cut -c1-600 gpurun_out/r2u_reference_arm.json
timeout 600 python scripts/sweep.py --buckets 1,2,3,4,8,16 --ks 10,100 --steps 5 --out gpurun_out/r2u_sweep_10M > gpurun_out/r2u_sweep.log 2>&1; echo "sweep rc=$?"
cat gpurun_out/r2u_sweep_10M.md 2>/dev/null
