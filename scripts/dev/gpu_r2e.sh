#!/bin/bash
# round 2, call E (8 GPUs): strong scaling of the 10M workload at N = 8 and 4, configs[3] (100M row ranges), sweep on 8 GPUs
mkdir -p gpurun_out
nvidia-smi -L | wc -l
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 500 $TR --nproc-per-node 8 --master-port 29521 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/r2e_bench_n8.json 2> gpurun_out/r2e_bench_n8.err; echo "bench n8 rc=$?"
timeout 500 $TR --nproc-per-node 4 --master-port 29522 bench.py --gpus 4 --steps 20 --warmup 3 > gpurun_out/r2e_bench_n4.json 2> gpurun_out/r2e_bench_n4.err; echo "bench n4 rc=$?"
timeout 900 $TR --nproc-per-node 8 --master-port 29523 bench.py --gpus 8 --rows 100000000 --steps 10 --warmup 3 > gpurun_out/r2e_bench_n8_100M.json 2> gpurun_out/r2e_bench_n8_100M.err; echo "bench 100M rc=$?"
timeout 600 $TR --nproc-per-node 8 --master-port 29524 scripts/sweep.py --buckets 1,2,4,8,16 --ks 10,100 --steps 5 --out gpurun_out/r2e_sweep_n8 > gpurun_out/r2e_sweep_n8.log 2>&1; echo "sweep n8 rc=$?"
python - <<'PY'
import json
for f in ("r2e_bench_n8", "r2e_bench_n4", "r2e_bench_n8_100M"):
    try:
        j = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, json.dumps({k: j[k] for k in ("value", "ms_per_step", "recall_at_10", "parity_sample", "exactness", "shard_breakdown")}))
        print("   e2e", j["e2e"] and j["e2e"]["value"], "roof", {k: j["roofline"][k] for k in ("achieved", "frac", "kernel_ms")}, j["clocks"])
    except Exception as e:
        print(f, "no line", e)
PY
cat gpurun_out/r2e_sweep_n8.md 2>/dev/null
tail -c 300 gpurun_out/r2e_bench_n8.err
