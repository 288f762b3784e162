mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for w in cfg5; do
python bench.py --workload $w --steps 5 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$w ms_per_step=%.3f value=%.4g'%(d['ms_per_step'],d['value']))"
done
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_cfg5_r1t.csv -k regex:'dense20|s4_traverse|tables_kernel|lnl_kernel|val_kernel|final_reduce|batch_gather' python bench.py --workload cfg5 --steps 1 --warmup 3 > gpurun_out/ncu_l5_r1t.log 2>&1
grep s4_traverse gpurun_out/launches_cfg5_r1t.csv | tail -2 | cut -d, -f5,12-
