#!/bin/bash
# round 2, call H (1 GPU): what the driver runs at round end -- smoke(), the reference arm, the GPU suite
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('__SMOKE_OK__')" > gpurun_out/r2h_smoke.log 2>&1; echo "smoke rc=$?"; tail -5 gpurun_out/r2h_smoke.log
( time timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2h_reference_arm.json 2> gpurun_out/r2h_reference_arm.err ) 2>&1 | grep real; echo "reference arm rc=$?"
tail -c 300 gpurun_out/r2h_reference_arm.err; cut -c1-1500 gpurun_out/r2h_reference_arm.json
timeout 600 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r2h_gpu_tests.log 2>&1; echo "gpu tests rc=$?"
tail -3 gpurun_out/r2h_gpu_tests.log
