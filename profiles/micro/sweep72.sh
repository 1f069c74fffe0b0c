#!/bin/bash
# Experiment: bisweep capped at 72 registers (launch bounds min 2 blocks; shared memory still allows one CTA per SM),
# so that a full bicombine CTA fits beside it.  Build first:  CTCG_EXTRA_NVCC="-DCTCG_SWEEP_MINB=2" python pytorch-asr_b200/build.py
set -u
tag=${1:-r72}
out=gpurun_out/sweep72; mkdir -p $out
python -m pytest tests/test_gpu_parity.py -q -x -k "bench_configurations" > $out/tests_$tag.log 2>&1; tail -2 $out/tests_$tag.log
for w in cfg3 cfg5; do
  python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline > $out/${tag}_$w.json 2> $out/${tag}_$w.err
  CTCG_SERIAL=1 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline > $out/${tag}_serial_$w.json 2> $out/${tag}_serial_$w.err
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/sweep72/*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['ms_per_step'],4), round(d['roofline']['frac'],4), {k:round(v,4) for k,v in d['roofline']['kernel_ms'].items()})
    except Exception as e: print(f,'ERR',e)
PY
