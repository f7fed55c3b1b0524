#!/bin/bash
# round 2, call V (N GPUs): the driver's scaling command on the final code
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 500 $TR --nproc-per-node $N --master-port 29561 bench.py --gpus $N --steps 40 --warmup 5 > gpurun_out/r2v_bench_n$N.json 2> gpurun_out/r2v_bench_n$N.err; echo "bench n$N rc=$?"
python - <<PY
import json
try:
    j = json.loads(open("gpurun_out/r2v_bench_n$N.json").read().strip().splitlines()[-1])
    print(json.dumps({k: j[k] for k in ("value", "ms_per_step", "recall_at_10", "parity_sample", "exactness", "shard_breakdown")}))
    print("e2e", j["e2e"] and j["e2e"]["value"], "roof", {k: j["roofline"][k] for k in ("achieved", "frac", "kernel_ms")}, j["clocks"])
except Exception as e:
    print("no line", e)
PY
tail -c 300 gpurun_out/r2v_bench_n$N.err
