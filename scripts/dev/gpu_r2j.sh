#!/bin/bash
# round 2, call J (1 GPU): allocation-free navigation entry point -- alone in a fresh process, then the whole GPU suite
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -x -m gpu -p no:cacheprovider > gpurun_out/r2j_noalloc_alone.log 2>&1; rc1=$?; echo "alone rc=$rc1"
tail -5 gpurun_out/r2j_noalloc_alone.log
timeout 1500 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/r2j_gpu_suite.log 2>&1; echo "suite rc=$?"
tail -5 gpurun_out/r2j_gpu_suite.log
