#!/bin/bash
# round 2, call N (8 GPUs): 10M at N = 8 with the pipelined exchange and without, N = 4, configs[3] (100M row ranges)
mkdir -p gpurun_out
nvidia-smi -L | wc -l
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 400 $TR --nproc-per-node 8 --master-port 29531 bench.py --gpus 8 --steps 40 --warmup 5 > gpurun_out/r2n_bench_n8.json 2> gpurun_out/r2n_bench_n8.err; echo "bench n8 rc=$?"
timeout 400 $TR --nproc-per-node 8 --master-port 29532 bench.py --gpus 8 --steps 40 --warmup 5 --no-pipeline --no-e2e > gpurun_out/r2n_bench_n8_onestream.json 2> gpurun_out/r2n_bench_n8_onestream.err; echo "bench n8 one-stream rc=$?"
timeout 400 $TR --nproc-per-node 4 --master-port 29533 bench.py --gpus 4 --steps 40 --warmup 5 > gpurun_out/r2n_bench_n4.json 2> gpurun_out/r2n_bench_n4.err; echo "bench n4 rc=$?"
timeout 700 $TR --nproc-per-node 8 --master-port 29534 bench.py --gpus 8 --rows 100000000 --steps 10 --warmup 3 > gpurun_out/r2n_bench_n8_100M.json 2> gpurun_out/r2n_bench_n8_100M.err; echo "bench 100M rc=$?"
python - <<'PY'
import json
for f in ("r2n_bench_n8", "r2n_bench_n8_onestream", "r2n_bench_n4", "r2n_bench_n8_100M"):
    try:
        j = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, json.dumps({k: j[k] for k in ("value", "ms_per_step", "recall_at_10", "parity_sample", "exactness", "shard_breakdown")}))
        print("   e2e", j["e2e"] and j["e2e"]["value"], "roof", {k: j["roofline"][k] for k in ("achieved", "frac", "kernel_ms")}, j["clocks"])
    except Exception as e:
        print(f, "no line", e)
PY
tail -c 300 gpurun_out/r2n_bench_n8.err
