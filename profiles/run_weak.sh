#!/bin/bash
# Weak scaling of the default workload on one box: N=1 and N=$1 back to back (same box, same commit).
N=$1
OUT=gpurun_out
for s in "20 5" "200 10"; do
  set -- $s
  timeout 600 python bench.py --gpus 1 --steps $1 --warmup $2 --no-cpu-baseline > $OUT/r3w_s$1_n1of$N.json 2> $OUT/r3w_s$1_n1of$N.err
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus $N --steps $1 --warmup $2 --no-cpu-baseline > $OUT/r3w_s$1_n$N.json 2> $OUT/r3w_s$1_n$N.err
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r3w_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d['n_gpus'], d['steps'], round(d['ms_per_step'],4), round(d['value']/1e6/d['n_gpus'],2), 'M/GPU  e2e', round(d['e2e']['value']/1e6/d['n_gpus'],2))
    except Exception as e: print(f,'ERR',e)
PY
