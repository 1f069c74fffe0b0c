"""Timing experiments on the numerator kernel: CTCG_NUM_DEBUG masks out one role's work at a time."""
import os, sys, subprocess, json
here = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for mask in (0, 31, 31 | 32, 31 | 64, 31 | 32 | 64, 32, 64, 32 | 64):
    env = dict(os.environ, CTCG_NUM_DEBUG=str(mask), CTCG_SERIAL="1")
    r = subprocess.run([sys.executable, os.path.join(here, "bench.py"), "--workload", "cfg2", "--steps", "30",
                        "--warmup", "3", "--no-cpu-baseline"], capture_output=True, text=True, env=env)
    try:
        d = json.loads(r.stdout.strip().splitlines()[-1])
        print("mask %2d  num_log %.4f ms" % (mask, d["roofline"]["kernel_ms"]["num_log"]), flush=True)
    except Exception as e:
        print("mask", mask, "failed", r.stderr[-300:])
