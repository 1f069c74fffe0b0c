"""Timing experiments on the bi-char denominator kernel: CTCG_BICHAR_DEBUG masks out one role's work
at a time (results are wrong by construction; only the kernel time is read)."""
import os, sys, subprocess, json
here = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
names = {1: "aux bwd (exp + alpha recompute)", 2: "main bwd", 4: "main fwd", 8: "aux fwd (row max)", 16: "io bwd stores"}
for mask in (0, 1, 2, 3, 4, 8, 12, 16, 31):
    env = dict(os.environ, CTCG_BICHAR_DEBUG=str(mask), CTCG_SERIAL="1")
    r = subprocess.run([sys.executable, os.path.join(here, "bench.py"), "--workload", "cfg3", "--steps", "30",
                        "--warmup", "3", "--no-cpu-baseline"], capture_output=True, text=True, env=env)
    try:
        d = json.loads(r.stdout.strip().splitlines()[-1])
        what = " + ".join(v for k, v in names.items() if mask & k) or "nothing"
        print("mask %2d  bichar_den %.4f ms   skipped: %s" % (mask, d["roofline"]["kernel_ms"]["bichar_den"], what), flush=True)
    except Exception as e:
        print("mask", mask, "failed", r.stderr[-300:])
