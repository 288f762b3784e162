mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 10 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('cfg2 ms_per_step=%.3f kernel_ms=%.3f e2e_ms=%.3f value=%.4g lnl=%r'%(d['ms_per_step'],d['roofline']['kernel_ms_per_step'],d['e2e']['ms_per_step'],d['value'],d['lnl_df_ddf']))"
