for m in ${MASKS:-256}; do echo "== mask $m"; CTCG_SWEEP_DEBUG=$m CTCG_SERIAL=1 timeout 100 python bench.py --steps 12 --warmup 3 --no-cpu-baseline 2>&1 | grep "^sweep" | tail -4; done
