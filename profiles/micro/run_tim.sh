for m in 256 259 319; do echo "== mask $m"; CTCG_SWEEP_DEBUG=$m CTCG_SERIAL=1 timeout 100 python bench.py --steps 2 --warmup 1 --no-cpu-baseline 2>&1 | grep "^sweep" | tail -6; done
