#!/bin/bash
# A/B of two prebuilt libraries on the same box: profiles/micro/_<tag>_lib.so for each tag given
set -u
out=gpurun_out/ab; mkdir -p $out
for rep in 1 2; do
for tag in "$@"; do
  cp profiles/micro/_${tag}_lib.so pytorch-asr_b200/libctcg_b200.so
  python bench.py --workload cfg3 --steps 20 --warmup 5 --no-cpu-baseline > $out/${tag}_${rep}.json 2> $out/${tag}_${rep}.err
done; done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/ab/*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['ms_per_step'],4), round(d['c_abi_direct']['ms_per_step'],4), {k:round(v,4) for k,v in d['roofline']['kernel_ms'].items()})
    except Exception as e: print(f,'ERR',e)
PY
