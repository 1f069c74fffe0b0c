#!/bin/bash
# Round-2 evidence on ONE B200 (run under gpurun from the repo root): bench lines of every workload, the
# reference arm, the ncu launch list of the bench command and one ncu --set full capture of every kernel of a
# cfg3 step (read here with profiles/ncu_traffic.py).  A number printed under ncu is never a bench value.
OUT=gpurun_out
TAG=${1:-r02}
python bench.py --steps 20 --warmup 5 > $OUT/${TAG}_bench_cfg3.json 2> $OUT/${TAG}_bench_cfg3.err; echo "cfg3 rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > $OUT/${TAG}_bench_reference_arm.json 2>> $OUT/${TAG}_bench_cfg3.err; echo "ref rc=$?"
for W in cfg2 cfg2s cfg4 cfg5; do
  python bench.py --workload $W --steps 10 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_bench_$W.json 2>> $OUT/${TAG}_bench_cfg3.err; echo "$W rc=$?"
done
CTCG_SERIAL=1 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $OUT/${TAG}_bench_cfg3_serial.json 2>> $OUT/${TAG}_bench_cfg3.err
CTCG_GROUPS=1 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $OUT/${TAG}_bench_cfg3_groups1.json 2>> $OUT/${TAG}_bench_cfg3.err
# (the ncu passes run with the captured-step cache off: the same kernels in the same order, launched one by one,
#  so that the launch counts skipped below are those of the warm-up alone)
export CTCG_STEP_GRAPHS=0
# launch list of the bench command (every launch with its device time: compare SHARES, not absolutes)
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $OUT/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches_cfg3.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $OUT/ncu_list.log 2>&1; echo "ncu list rc=$?"
# every kernel of one step, full sections (kernels back to back: ncu serialises them anyway)
export CTCG_SERIAL=1
KERN="regex:bisweep|bicombine|gather_kernel|num_sweep|num_combine|final_kernel|scale_grad"
python bench.py --no-cpu-baseline --steps 3 --warmup 3 > $OUT/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "$KERN" -s 21 -c 7 -o $OUT/${TAG}_full_cfg3 \
    python bench.py --no-cpu-baseline --steps 3 --warmup 3 > $OUT/ncu_full.log 2>&1; echo "ncu full rc=$?"
python bench.py --workload cfg5 --no-cpu-baseline --steps 2 --warmup 2 > $OUT/plain.log 2>&1 &&
ncu --set full --clock-control none -k "$KERN" -s 14 -c 7 -o $OUT/${TAG}_full_cfg5 \
    python bench.py --workload cfg5 --no-cpu-baseline --steps 2 --warmup 2 > $OUT/ncu_full5.log 2>&1; echo "ncu full cfg5 rc=$?"
