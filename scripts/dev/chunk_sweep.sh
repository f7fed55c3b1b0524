#!/bin/bash
# developer aid: effect of the object-chunk size on the rescoring kernel (k = 10 and k = 100)
for k in 10 100; do
  for c in 4096 8192 16384 32768 65536; do
    python bench.py --k $k --chunk-objects $c --steps 10 --no-cpu-baseline --no-e2e --gt-queries 1000 --parity-queries 0 2>/dev/null \
      | python -c "import sys,json; j=json.loads(sys.stdin.read()); print('k=$k chunk=$c', 'ms/step=%.2f'%j['ms_per_step'], 'k3=%.2f'%j['roofline']['kernel_ms'], 'fin=%.2f'%j['stages_ms']['ms_finalize'], 'items=%d'%j['work_items'], 'recall=%.4f'%j['recall_at_10'])"
  done
done
