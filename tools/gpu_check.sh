#!/bin/bash
# One GPU-box pass.  Usage (repo root, on the GPU box): bash tools/gpu_check.sh <tag> [stages...]
# stages: test bench bench45 sweep launches launches5 traffic prof_s4 prof_d20   (default: test bench)
# gpurun copies back at most 64 MiB of gpurun_out/: keep full ncu captures to a few launches.
TAG=${1:-rX}; shift
STAGES=${@:-test bench}
OUT=gpurun_out
mkdir -p $OUT
has() { [[ " $STAGES " == *" $1 "* ]]; }
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi_$TAG.txt
if has test; then
python -m pytest tests -m gpu -x -q > $OUT/pytest_$TAG.log 2>&1; echo "pytest rc=$?" >> $OUT/pytest_$TAG.log
python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke_$TAG.log 2>&1; echo "smoke rc=$?" >> $OUT/smoke_$TAG.log
tail -3 $OUT/pytest_$TAG.log; tail -2 $OUT/smoke_$TAG.log
fi
if has bench; then
# the driver's command: the headline cfg2 line with parity, cpu_baseline and the cfg3 / cfg4-full / cfg5 / sweep sub-results
python bench.py --steps 20 --warmup 5 > $OUT/bench_cfg2_$TAG.json 2> $OUT/bench_cfg2_$TAG.err; echo "bench rc=$?"
python bench.py --impl reference --steps 20 --warmup 5 > $OUT/bench_ref_$TAG.json 2> $OUT/bench_ref_$TAG.err; echo "reference rc=$?"
cut -c1-300 $OUT/bench_cfg2_$TAG.json; cut -c1-400 $OUT/bench_ref_$TAG.json
fi
if has launches; then
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_cfg2_$TAG.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --shapes none > $OUT/ncu_l2_$TAG.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/launches_cfg3_$TAG.csv \
    python bench.py --workload cfg3 --patterns 50000 --steps 2 --warmup 3 --no-cpu --shapes none > $OUT/ncu_l3_$TAG.log 2>&1
fi
if has bench45; then
python bench.py --workload cfg4 --steps 5 > $OUT/bench_cfg4_$TAG.json 2> $OUT/bench_cfg4_$TAG.err
python bench.py --workload cfg5 --steps 5 > $OUT/bench_cfg5_$TAG.json 2> $OUT/bench_cfg5_$TAG.err
cut -c1-300 $OUT/bench_cfg4_$TAG.json; cut -c1-300 $OUT/bench_cfg5_$TAG.json
fi
if has sweep; then
# one optimizeAllBranches sweep of cfg2 through the host mirror against the reference's own sweep on the host cores
python tools/sweep_bench.py > $OUT/sweep_cfg2_$TAG.json 2> $OUT/sweep_cfg2_$TAG.err
cut -c1-400 $OUT/sweep_cfg2_$TAG.json
fi
if has launches5; then
# cfg5 through the batched partition dispatch: 4 evaluations (3 warm-up + 1 timed) of ~30 launches each
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/launches_cfg5_$TAG.csv \
    -k regex:'dense20|s4_traverse|tables_kernel|lnl_kernel|val_kernel|final_reduce|batch_gather' \
    python bench.py --workload cfg5 --steps 1 --warmup 3 > $OUT/ncu_l5_$TAG.log 2>&1
fi
if has traffic; then
# DRAM bytes of every partial-update launch of cfg2 / cfg3 at the bench's full size (roofline.traffic)
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:'s4_traverse' \
    --csv --log-file $OUT/traffic_cfg2_$TAG.csv python bench.py --steps 1 --warmup 3 --no-cpu --shapes none > $OUT/ncu_t2_$TAG.log 2>&1
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:'dense20' \
    --csv --log-file $OUT/traffic_cfg3_$TAG.csv python bench.py --workload cfg3 --steps 1 --warmup 3 --no-cpu --shapes none > $OUT/ncu_t3_$TAG.log 2>&1
fi
if has prof_s4; then
ncu --set full --clock-control none --import-source on -k regex:'s4_traverse|eval_tail' --launch-skip 9 -c 3 -f -o $OUT/prof_s4_$TAG \
    python bench.py --steps 1 --warmup 3 --no-cpu --shapes none > $OUT/ncu_s4_$TAG.log 2>&1
fi
if has prof_d20; then
# one step of cfg3 at 50k patterns has 30 dense20_rp launches (levels); four of them from the fourth evaluation
ncu --set full --clock-control none --import-source on -k regex:dense20_rp --launch-skip 92 -c 4 -f -o $OUT/prof_dense20_$TAG \
    python bench.py --workload cfg3 --patterns 50000 --steps 1 --warmup 3 --no-cpu --shapes none > $OUT/ncu_d20_$TAG.log 2>&1
fi
du -sh $OUT
