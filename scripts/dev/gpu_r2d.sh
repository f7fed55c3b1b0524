#!/bin/bash
# round 2, call D (2 GPUs): two-device tests, the sharded step on real GPUs (pair kernel, sliced navigation), k = 100 check
mkdir -p gpurun_out
nvidia-smi -L
timeout 600 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r2d_gpu_tests.log 2>&1; echo "gpu tests rc=$?"
tail -5 gpurun_out/r2d_gpu_tests.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r2d_bench_n2.json 2> gpurun_out/r2d_bench_n2.err; echo "bench n2 rc=$?"
tail -c 400 gpurun_out/r2d_bench_n2.err
python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/r2d_bench_n2.json").read().strip().splitlines()[-1])
    print(json.dumps({k: j[k] for k in ("value", "ms_per_step", "recall_at_10", "e2e", "stages_ms", "parity_sample", "exactness", "shard_breakdown")}, indent=1))
    print({k: j["roofline"][k] for k in ("achieved", "frac", "kernel_ms", "kernel", "traffic")})
except Exception as e:
    print("no bench line", e)
PY
timeout 400 python bench.py --k 100 --steps 5 --no-cpu-baseline --no-dropin --no-e2e > gpurun_out/r2d_bench_k100.json 2> gpurun_out/r2d_bench_k100.err; echo "k100 rc=$?"
python - <<'PY'
import json
try:
    j = json.loads(open("gpurun_out/r2d_bench_k100.json").read().strip().splitlines()[-1])
    print({k: j[k] for k in ("value", "ms_per_step", "recall_at_10", "stages_ms", "parity_sample", "chunk_objects")})
except Exception as e:
    print("no k100 line", e)
PY
