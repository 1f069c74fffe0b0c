"""Timing experiments on the split-sweep denominator: CTCG_SWEEP_DEBUG masks out one role's work at a
time (results are wrong by construction; only the kernel times are read)."""
import os, sys, subprocess, json
here = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
names = {1: "main warps' recursion", 2: "producers' exp + store", 4: "backward-sweep CTAs", 8: "scale warp", 16: "producers' loads + max", 32: "TMA copies"}
masks = [int(a) for a in sys.argv[1:]] or [0, 1, 2, 3, 4]
for mask in masks:
    env = dict(os.environ, CTCG_SWEEP_DEBUG=str(mask), CTCG_SERIAL="1")
    r = subprocess.run([sys.executable, os.path.join(here, "bench.py"), "--workload", "cfg3", "--steps", "30",
                        "--warmup", "3", "--no-cpu-baseline"], capture_output=True, text=True, env=env)
    try:
        d = json.loads(r.stdout.strip().splitlines()[-1])
        what = " + ".join(v for k, v in names.items() if mask & k) or "nothing"
        km = d["roofline"]["kernel_ms"]
        print("mask %2d  bisweep %.4f ms  bicombine %.4f ms   skipped: %s" % (mask, km["bisweep"], km["bicombine"], what), flush=True)
    except Exception as e:
        print("mask", mask, "failed", r.stderr[-300:])
