#!/bin/bash
# round 2, call B: the reworked CTA-pair kernel, with bail-outs (a wrong barrier / register protocol HANGS, it does not
# fail: every step has a tight timeout and the script stops at the first hang instead of burning the GPU budget).
mkdir -p gpurun_out
free -g | head -2; nproc
check_gpu() { nvidia-smi --query-gpu=utilization.gpu,power.draw --format=csv,noheader; }
timeout 150 python scripts/dev/pair_ab.py --diag-only --diag-sets 1 --out gpurun_out/r2b_diag1.json > gpurun_out/r2b_diag1.log 2>&1; rc1=$?
echo "diag sets=1 rc=$rc1"; tail -2 gpurun_out/r2b_diag1.log | cut -c1-600
if [ $rc1 -ne 0 ]; then echo "BAIL: one-set pair kernel failed"; check_gpu; exit 0; fi
timeout 150 python scripts/dev/pair_ab.py --diag-only --diag-sets 2 --out gpurun_out/r2b_diag2.json > gpurun_out/r2b_diag2.log 2>&1; rc2=$?
echo "diag sets=2 rc=$rc2"; tail -2 gpurun_out/r2b_diag2.log | cut -c1-600
if [ $rc2 -ne 0 ]; then echo "two-set kernel failed: the rest runs with LMI_EPI_SETS=1"; export LMI_EPI_SETS=1; check_gpu; sleep 5; fi
timeout 240 python -m pytest tests/test_gpu_pair_kernel.py -q -x -p no:cacheprovider > gpurun_out/r2b_pair_tests.log 2>&1; rc=$?; echo "pair tests rc=$rc"
tail -4 gpurun_out/r2b_pair_tests.log
if [ $rc -eq 124 ]; then echo "BAIL: pair tests hang"; check_gpu; exit 0; fi
if [ $rc2 -eq 0 ]; then SEQ="3:0:0:2,3:0:0:1,1:0:0,3:0:64:1,3:0:0:2,3:0:0:1,3:0:64:2,3:0:2:2,3:0:2:1,3:0:6:2,3:0:10:2,3:0:0:2,3:16128:0:2,3:64000:0:2"
else SEQ="3:0:0:1,1:0:0,3:0:64:1,3:0:0:1,3:0:2:1,3:0:6:1,3:0:10:1,3:16128:0:1,3:64000:0:1"; fi
timeout 420 python scripts/dev/pair_ab.py --exp --skip-diag --steps 10 --out gpurun_out/r2b_ab.json --seq "$SEQ" > gpurun_out/r2b_ab.log 2>&1; echo "ab rc=$?"
python - <<'PY'
import json
try:
    for r in json.load(open("gpurun_out/r2b_ab.json"))["runs"]:
        print({k: (round(v, 3) if isinstance(v, float) else v) for k, v in r.items() if k in ("rescore", "dbg_flags", "epi_sets", "chunk_objects", "ms_per_search", "k3_ms_avg", "k3_sm_mhz", "useful_tflops", "ms_finalize", "ids_equal_to_first", "n_uncertified")})
except Exception as e:
    print("no ab report", e)
PY
timeout 420 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/r2b_gpu_tests.log 2>&1; echo "gpu tests rc=$?"
tail -8 gpurun_out/r2b_gpu_tests.log
timeout 700 python bench.py --steps 20 --warmup 3 > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err; echo "bench rc=$?"
tail -c 600 gpurun_out/r2b_bench.err
python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/r2b_bench.json").read().strip().splitlines()[-1])
    keep = {k: j[k] for k in ("value", "ms_per_step", "recall_at_10", "e2e", "e2e_dropin", "stages_ms", "parity_sample", "exactness", "store_build_seconds")}
    keep["roofline"] = {k: j["roofline"][k] for k in ("achieved", "frac", "frac_vs_burst_peak", "frac_vs_sustained_peak", "kernel_ms", "tile_efficiency", "kernel")}
    cb = j.get("cpu_baseline") or {}
    keep["cpu"] = {k: cb.get(k) for k in ("value", "cores", "nav_parity", "category_parity", "oracle_parity")}
    keep["clocks"] = j["clocks"]
    print(json.dumps(keep, indent=1)[:6000])
except Exception as e:
    print("no bench line", e)
PY
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
