#!/bin/bash
# Experiment: two bisweep CTAs per SM (6 + 5 ring rows, <= 72 registers) against the product build, large-B workloads.
# Build the variant first (here, not on the box):
#   CTCG_EXTRA_NVCC="-DCTCG_NRAW=6 -DCTCG_NE=5 -DCTCG_SWEEP_MINB=2" python pytorch-asr_b200/build.py
# then: gpurun -- 'bash profiles/micro/sweep2.sh two'   (and 'base' with the product build)
set -u
tag=${1:-two}
out=gpurun_out/sweep2; mkdir -p $out
python -m pytest tests/test_gpu_parity.py -q -x -k "bench_configurations or workspace_and" > $out/tests_$tag.log 2>&1; tail -3 $out/tests_$tag.log
for w in cfg5 cfg4 cfg3; do
  python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline > $out/${tag}_$w.json 2> $out/${tag}_$w.err
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/sweep2/*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['ms_per_step'],4), round(d['roofline']['frac'],4), {k:round(v,4) for k,v in d['roofline']['kernel_ms'].items()})
    except Exception as e: print(f,'ERR',e)
PY
