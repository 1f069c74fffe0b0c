#!/bin/bash
# Round-1 evidence run (one GPU): tests, bench lines, ncu launch list and one --set full capture.
# Everything lands in gpurun_out/; the summaries are copied to profiles/ by hand afterwards.
set -x
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/e_tests.log 2>&1; tail -2 $O/e_tests.log
timeout 300 python bench.py > $O/e_bench_cfg3.json 2> $O/e_bench_cfg3.err; tail -c 600 $O/e_bench_cfg3.json
timeout 200 python bench.py --workload cfg2 --steps 100 --no-cpu-baseline > $O/e_bench_cfg2.json 2>/dev/null
timeout 300 python bench.py --workload cfg4 --steps 30 --no-cpu-baseline > $O/e_bench_cfg4.json 2>/dev/null
timeout 300 python bench.py --workload cfg5 --steps 30 --no-cpu-baseline > $O/e_bench_cfg5.json 2>/dev/null
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/e_bench_ref.json 2>/dev/null
# launch list of the bench command (after it exited 0 without ncu above)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/e_launches.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $O/e_ncu_list.log 2>&1
# one full capture of each of the main kernels
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"bisweep|bicombine|num_sweep|num_combine" -s 8 -c 4 \
    -o $O/e_full python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $O/e_ncu_full.log 2>&1
ls -la $O
