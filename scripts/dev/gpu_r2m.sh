#!/bin/bash
# round 2, call M (N GPUs, default 2): pipelined exchange vs one-stream exchange, same box
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29521 bench.py --gpus $N --steps 40 --warmup 5 --no-cpu-baseline > gpurun_out/r2m_bench_n${N}_pipelined.json 2> gpurun_out/r2m_bench_n${N}_pipelined.err; echo "pipelined rc=$?"
tail -c 300 gpurun_out/r2m_bench_n${N}_pipelined.err
timeout 600 $TR --master-port 29522 bench.py --gpus $N --steps 40 --warmup 5 --no-cpu-baseline --no-pipeline --no-e2e > gpurun_out/r2m_bench_n${N}_onestream.json 2> gpurun_out/r2m_bench_n${N}_onestream.err; echo "one-stream rc=$?"
tail -c 300 gpurun_out/r2m_bench_n${N}_onestream.err
python - <<PY
import json
for f in ("gpurun_out/r2m_bench_n${N}_pipelined.json", "gpurun_out/r2m_bench_n${N}_onestream.json"):
    try:
        j = json.loads(open(f).read().strip().splitlines()[-1])
        print(f)
        print({k: j[k] for k in ("value", "ms_per_step", "recall_at_10", "parity_sample")})
        print(j.get("e2e")); print(j.get("shard_breakdown")); print(j["config"].get("pipeline"))
    except Exception as e:
        print("no line", f, e)
PY
