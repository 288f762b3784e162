mkdir -p gpurun_out
for cfg in "4 3" "5 2" "5 1" "6 1" "6 2" "4 2"; do
  set -- $cfg
  B200LK_S4_MINB=$1 B200LK_STACK_DEPTH=$2 python bench.py --steps 10 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('minb=$1 depth=$2 ms_per_step=%.3f kernel_ms=%.3f lnl=%.6f'%(d['ms_per_step'],d['roofline']['kernel_ms_per_step'],d['lnl_df_ddf'][0]))"
done
