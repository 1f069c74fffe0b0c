#!/bin/bash
# Multi-GPU bench lines of round 2 (run under `gpurun --gpus N`): weak scaling of the default workload and the
# strong-scaling sweep of BASELINE config 5 (B=2048 in total, T in {400, 800, 1600}).   usage: run_scaling.sh N
N=$1
OUT=gpurun_out
run() {  # name, args...
  local name=$1; shift
  if [ "$N" = 1 ]; then
    timeout 900 python bench.py --gpus 1 "$@" > $OUT/${name}_n$N.json 2> $OUT/${name}_n$N.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
      bench.py --gpus $N "$@" > $OUT/${name}_n$N.json 2> $OUT/${name}_n$N.err
  fi
  echo "$name n=$N rc=$?"
}
run r2s_weak_cfg3 --steps 20 --warmup 5 --no-cpu-baseline
run r2s_weak_cfg3_200 --steps 200 --warmup 10 --no-cpu-baseline
for T in 400 800 1600; do
  run r2s_strong_T$T --workload cfg5 --scaling strong --T $T --steps 5 --warmup 3 --no-cpu-baseline
done
