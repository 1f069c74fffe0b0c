"""ncu --set full capture of one bench step -> profiles/r02_ncu_traffic.json (what bench.py's
``roofline.traffic`` reads) and a markdown summary.

    python profiles/ncu_traffic.py <workload> <report.ncu-rep> [<commit>]      # here, no GPU needed

Per kernel of the step: dram__bytes_read.sum + dram__bytes_write.sum per launch (mean over the captured
launches), duration, registers, achieved occupancy, DRAM throughput %.  The JSON keeps one entry per
workload; re-running for a workload replaces its entry and the commit hash.
"""
import csv
import io
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SHORT = [("bisweep_kernel", "bisweep"), ("bicombine_kernel", "bicombine"), ("gather_kernel", "gather"),
         ("num_sweep_kernel", "num_sweep"), ("num_combine_kernel", "num_combine"), ("final_kernel", "final"),
         ("scale_grad_kernel", "scale_grad"), ("rows_ctx_kernel", "rows_ctx"), ("rows_kernel", "rows"),
         ("scatter_sub_kernel", "scatter_sub"), ("path_fwd_bwd_kernel", "path_fwd_bwd"), ("viterbi_kernel", "viterbi")]
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}


def main():
    workload, rep = sys.argv[1], sys.argv[2]
    commit = sys.argv[3] if len(sys.argv) > 3 else subprocess.run(
        ["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True, cwd=HERE).stdout.strip()
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}

    def val(r, name):
        i = col[name]
        return float(r[i].replace(",", "")) * UNIT.get(units[i], 1.0)

    per = {}
    for r in data:
        fn = r[col["Kernel Name"]]
        short = next((s for k, s in SHORT if k in fn), None)
        if short is None:
            continue
        e = per.setdefault(short, {"n": 0, "bytes": 0.0, "us": 0.0, "regs": 0, "dram_pct": 0.0, "warps_pct": 0.0})
        e["n"] += 1
        e["bytes"] += val(r, "dram__bytes_read.sum") + val(r, "dram__bytes_write.sum")
        e["us"] += val(r, "gpu__time_duration.sum")
        e["regs"] = int(val(r, "launch__registers_per_thread"))
        e["dram_pct"] += val(r, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed")
        e["warps_pct"] += val(r, "sm__warps_active.avg.pct_of_peak_sustained_active")
    out_path = os.path.join(HERE, "r02_ncu_traffic.json")
    doc = json.load(open(out_path)) if os.path.exists(out_path) else {"workloads": {}, "detail": {}}
    doc["commit"] = commit
    doc["how"] = ("ncu --set full --clock-control none on `python bench.py --workload <w> --steps 3 --warmup 3` with "
                  "CTCG_SERIAL=1 (kernels back to back: ncu serialises them anyway); bytes = dram__bytes_read.sum + "
                  "dram__bytes_write.sum per launch, mean over the captured launches")
    doc["workloads"][workload] = {k: int(round(e["bytes"] / e["n"])) for k, e in per.items()}
    doc["detail"][workload] = {k: {"launches": e["n"], "us": round(e["us"] / e["n"], 2), "registers": e["regs"],
                                   "dram_throughput_pct": round(e["dram_pct"] / e["n"], 1),
                                   "warps_active_pct": round(e["warps_pct"] / e["n"], 1)} for k, e in per.items()}
    json.dump(doc, open(out_path, "w"), indent=1, sort_keys=True)
    tot = sum(doc["workloads"][workload].values())
    print("%s @ %s: %d kernels, %.1f MB DRAM per step" % (workload, commit, len(per), tot / 1e6))
    for k, e in sorted(per.items(), key=lambda kv: -kv[1]["bytes"]):
        print("  %-12s %8.1f MB  %8.1f us  %3d regs  dram %5.1f%%  warps %5.1f%%" % (
            k, e["bytes"] / e["n"] / 1e6, e["us"] / e["n"], e["regs"], e["dram_pct"] / e["n"], e["warps_pct"] / e["n"]))


if __name__ == "__main__":
    main()
