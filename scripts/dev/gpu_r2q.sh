#!/bin/bash
# round 2, call Q (8 GPUs): navigation sliced (+ list all-gather) vs replicated (no collective on the search stream), both with the pipelined exchange
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 400 $TR --nproc-per-node 8 --master-port 29551 bench.py --gpus 8 --steps 40 --warmup 5 --no-cpu-baseline --no-e2e --nav-mode replicated > gpurun_out/r2q_bench_n8_replicated.json 2> gpurun_out/r2q_bench_n8_replicated.err; echo "replicated rc=$?"
timeout 400 $TR --nproc-per-node 8 --master-port 29552 bench.py --gpus 8 --steps 40 --warmup 5 --no-cpu-baseline --no-e2e --nav-mode sliced > gpurun_out/r2q_bench_n8_sliced.json 2> gpurun_out/r2q_bench_n8_sliced.err; echo "sliced rc=$?"
timeout 400 $TR --nproc-per-node 8 --master-port 29553 bench.py --gpus 8 --steps 40 --warmup 5 --no-cpu-baseline --no-e2e --nav-mode replicated > gpurun_out/r2q_bench_n8_replicated2.json 2> gpurun_out/r2q_bench_n8_replicated2.err; echo "replicated2 rc=$?"
python - <<'PY'
import json
for f in ("r2q_bench_n8_replicated", "r2q_bench_n8_sliced", "r2q_bench_n8_replicated2"):
    try:
        j = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, json.dumps({k: j[k] for k in ("value", "ms_per_step", "recall_at_10", "shard_breakdown")}), j["clocks"]["sm_mhz_in_kernel"], j["parity_sample"]["mismatches"])
    except Exception as e:
        print(f, "no line", e)
PY
