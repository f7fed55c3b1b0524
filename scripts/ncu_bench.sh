# ncu evidence for one bench.py configuration (run under gpurun, 1 GPU):
#   1. launch list with per-launch device time of every kernel of this library
#   2. one `--set full` capture of the rescoring kernel
TAG=${1:-r1a}
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-dropin --gt-queries 100 $EXTRA"
K='regex:^(nav_|plan_|gather_|rescore_|finalize_|exact_|merge_|cast_)'
if [ -z "$SKIP_LIST" ]; then $B > gpurun_out/plain_$TAG.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 200 --csv --log-file gpurun_out/launches_$TAG.csv $B > gpurun_out/ncu1_$TAG.log 2>&1; fi
$B > gpurun_out/plain2_$TAG.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:rescore_t -s 3 -c 1 -o gpurun_out/prof_$TAG $B > gpurun_out/ncu2_$TAG.log 2>&1
tail -n 3 gpurun_out/ncu2_$TAG.log
