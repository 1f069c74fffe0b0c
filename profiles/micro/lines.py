"""Print the essentials of bench JSON lines (skips non-JSON lines such as NCCL's version banner)."""
import glob
import json
import sys

for pat in sys.argv[1:]:
    for f in sorted(glob.glob(pat)):
        d = None
        for ln in open(f):
            ln = ln.strip()
            if ln.startswith("{"):
                try:
                    d = json.loads(ln)
                except ValueError:
                    pass
        if d is None:
            print(f.split("/")[-1], "no JSON line")
            continue
        r = d.get("roofline", {})
        print("%-34s n=%d steps=%-3d %9.4f ms/step %8.1f M f/s  step-frac %.3f  c_abi %.4f ms" % (
            f.split("/")[-1], d["n_gpus"], d["steps"], d["ms_per_step"], d["value"] / 1e6, r.get("frac", 0) or 0,
            d.get("c_abi_direct", {}).get("ms_per_step", 0)))
