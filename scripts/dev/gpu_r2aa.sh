#!/bin/bash
# round 2, call AA (N GPUs): navigation look-ahead on / off, same box
N=${1:-4}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29571 bench.py --gpus $N --steps 40 --warmup 5 --no-cpu-baseline --no-e2e --nav-lookahead 1 > gpurun_out/r2aa_bench_n${N}_lookahead.json 2> gpurun_out/r2aa_bench_n${N}_lookahead.err; echo "lookahead rc=$?"
tail -c 400 gpurun_out/r2aa_bench_n${N}_lookahead.err
timeout 300 $TR --master-port 29572 bench.py --gpus $N --steps 40 --warmup 5 --no-cpu-baseline --no-e2e --nav-lookahead 0 > gpurun_out/r2aa_bench_n${N}_plain.json 2> gpurun_out/r2aa_bench_n${N}_plain.err; echo "plain rc=$?"
python - <<PY
import json
for f in ("gpurun_out/r2aa_bench_n${N}_lookahead.json", "gpurun_out/r2aa_bench_n${N}_plain.json"):
    try:
        j = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, {k: j[k] for k in ("value", "ms_per_step", "recall_at_10")}, j["parity_sample"]["mismatches"], j["shard_breakdown"]["rescore_ms_per_rank"])
    except Exception as e:
        print("no line", f, e)
PY
