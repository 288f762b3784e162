#!/bin/bash
# s4_traverse_kernel: resident-CTA instantiation x stack depth at the per-GPU shares of strong scaling (cfg2 / 8, / 4, / 1)
for n in 125000 250000 1000000; do
  for cfg in "10 2" "10 1" "12 1" "14 1" "16 1"; do
    set -- $cfg
    B200LK_S4_MINB=$1 B200LK_STACK_DEPTH=$2 python bench.py --patterns $n --steps 40 --warmup 10 --no-cpu 2>/dev/null | \
      python -c "import json,sys; d=json.loads(sys.stdin.read()); print('patterns $n minb $1 depth $2: ms/step %.4f kernel %.4f frac %.3f' % (d['ms_per_step'], d['roofline']['kernel_ms_per_step'], d['roofline']['frac']))"
  done
done
