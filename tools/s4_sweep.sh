mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for w in cfg4 cfg5 cfg3; do
python bench.py --workload $w --steps 5 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$w ms_per_step=%.3f kernel_ms=%s value=%.4g'%(d['ms_per_step'],d.get('roofline',{}).get('kernel_ms_per_step'),d['value']))"
done
