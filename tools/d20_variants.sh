#!/bin/bash
# dense20_rp_kernel: row tiles per warp x residency target, cfg3 (LG+G4, 500 x 200k) and cfg4 (LG+C60+G4, 12.5k patterns)
for wl in cfg3 cfg4; do
for v in "" mt1 mt1b5 mt2b4; do
  lib=""; [ -n "$v" ] && lib=$PWD/iqtree2_b200/libb200lk_$v.so
  B200LK_LIB=$lib python bench.py --workload $wl --steps 5 --warmup 3 --cpu-sample 4000 --cpu-steps 1 --shapes none 2>/dev/null | \
    python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$wl variant ${v:-base}: ms/step %.3f kernel %.3f frac %.3f parity %s clocks %s' % (d['ms_per_step'], d['roofline']['kernel_ms_per_step'], d['roofline']['frac'], d['parity']['ok'], d['clocks']['sm_mhz']))"
done; done
