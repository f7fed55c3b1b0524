#!/bin/bash
# round 2, call P (2 GPUs): GPU suite incl. the two-device tests after the merge-kernel split; N = 2 bench with replicated navigation
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/r2p_gpu_suite.log 2>&1; echo "suite rc=$?"
tail -5 gpurun_out/r2p_gpu_suite.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29541 bench.py --gpus 2 --steps 40 --warmup 5 --no-cpu-baseline --no-e2e --nav-mode replicated > gpurun_out/r2p_bench_n2_replicated.json 2> gpurun_out/r2p_bench_n2_replicated.err; echo "replicated rc=$?"
timeout 600 $TR --master-port 29542 bench.py --gpus 2 --steps 40 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/r2p_bench_n2_sliced.json 2> gpurun_out/r2p_bench_n2_sliced.err; echo "sliced rc=$?"
python - <<'PY'
import json
for f in ("gpurun_out/r2p_bench_n2_replicated.json", "gpurun_out/r2p_bench_n2_sliced.json"):
    try:
        j = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, {k: j[k] for k in ("value", "ms_per_step", "recall_at_10", "parity_sample")}, j["shard_breakdown"])
    except Exception as e:
        print("no line", f, e)
PY
