#!/bin/bash
# bisweep alone (CTCG_SERIAL=1: every kernel on the caller's stream) with one or two CTAs per SM; see sweep2.sh
set -u
tag=${1:-two}
out=gpurun_out/sweep2; mkdir -p $out
for w in cfg5 cfg4; do
  CTCG_SERIAL=1 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline > $out/${tag}_serial_$w.json 2> $out/${tag}_serial_$w.err
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/sweep2/*serial*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['ms_per_step'],4), round(d['roofline']['frac'],4), {k:round(v,4) for k,v in d['roofline']['kernel_ms'].items()})
    except Exception as e: print(f,'ERR',e)
PY
