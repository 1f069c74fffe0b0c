"""Benchmark of the CTC / CTC-G loss+grad hot path (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg3]
    python bench.py --impl reference ...       # the CPU implementation of the same path

One "step" = one fused loss+grad call (``ctcg_loss_fwd_bwd`` through the C ABI)
over one synthetic WSJ-shaped batch.  ``value`` = utterance-frames/s with the
inputs resident in HBM; ``e2e`` = the same metric through the public Python API
with pinned HOST logits (H2D copy of logits, graph compile + upload from
labels, D2H read of the loss inside the timed region).  Prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# name -> (variant, S, B per GPU, T, graph_build_args, normalize_by_dim, use_den)
WORKLOADS = {
    "cfg1": ("mono-char CTC", 1, 49, 16, 400, {}, 0, False),
    "cfg2": ("bi-char CTC", 2, 49, 64, 400, {}, 0, False),
    "cfg2s": ("bi-char CTC V=1296", 2, 36, 64, 400, {}, 0, False),
    "cfg3": ("CTC-G bi-char", 2, 49, 64, 400, {}, None, True),
    "cfg4": ("CTC-GB contextual blanks", 2, 49, 256, 800, {"use_contextual_blanks": True}, None, True),
    "cfg5": ("CTC-G sweep shard", 2, 49, 256, 800, {}, None, True),
}
# dram__bytes_read.sum + dram__bytes_write.sum per launch (ncu --set full, profiles/r01_e_ncu_full_summary.md)
NCU_TRAFFIC = {("cfg3", "bisweep"): 402519296, ("cfg3", "bicombine"): 448670208, ("cfg3", "num_sweep"): 9504512,
               ("cfg3", "num_combine"): 203017216}
CFG_INDEX = {"cfg1": 1, "cfg2": 2, "cfg2s": 2, "cfg3": 3, "cfg4": 4, "cfg5": 5}


def make_batch(workload, rank=0):
    """Seeded synthetic batch of SURVEY 8(d): logits ~ 3*N(0,1), lens[0]=T, others
    uniform in [0.6T, T] sorted descending, U_b = len_b // 4 labels in 1..S-1."""
    name, order, S, B, T, args, nbd, use_den = WORKLOADS[workload]
    seed = 1000 + CFG_INDEX[workload] + 7919 * rank
    rng = np.random.default_rng(seed)
    lens = sorted([T] + rng.integers(int(0.6 * T), T + 1, size=B - 1).tolist(), reverse=True)
    seqs = [rng.integers(1, S, size=max(1, ln // 4)).tolist() for ln in lens]
    g = torch.Generator().manual_seed(seed)
    V = S ** order
    return dict(name=name, order=order, S=S, B=B, T=T, V=V, args=args, nbd=nbd, use_den=use_den,
                lens=lens, seqs=seqs, gen=g, seed=seed)


class ClockSampler(object):
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self._stop = threading.Event()
        self._thr = None

    def _nvml(self):
        """Fast path: NVML in-process (sub-millisecond per sample)."""
        try:
            import pynvml as N
            N.nvmlInit()
            h = N.nvmlDeviceGetHandleByIndex(self.index)
            mx = N.nvmlDeviceGetMaxClockInfo(h, N.NVML_CLOCK_SM)
            bits = {"hw_slowdown": N.nvmlClocksThrottleReasonHwSlowdown,
                    "hw_thermal_slowdown": N.nvmlClocksThrottleReasonHwThermalSlowdown,
                    "sw_thermal_slowdown": N.nvmlClocksThrottleReasonSwThermalSlowdown,
                    "sw_power_cap": N.nvmlClocksThrottleReasonSwPowerCap}

            def sample():
                sm = N.nvmlDeviceGetClockInfo(h, N.NVML_CLOCK_SM)
                r = N.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                pw = N.nvmlDeviceGetPowerUsage(h) / 1000.0
                return [str(sm), str(mx), str(pw)] + ["Active" if (r & bits[k]) else "Not Active" for k in
                                                       ("hw_slowdown", "hw_thermal_slowdown",
                                                        "sw_thermal_slowdown", "sw_power_cap")]
            sample()
            return sample
        except Exception:
            return None

    def _run(self):
        sample = self._nvml()
        while not self._stop.is_set():
            try:
                if sample is not None:
                    self.rows.append(sample())
                    self._stop.wait(0.002)
                    continue
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True,
                                     timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.1)

    def start(self):
        self._thr = threading.Thread(target=self._run, daemon=True)
        self._thr.start()

    def stop(self):
        self._stop.set()
        if self._thr:
            self._thr.join(timeout=10)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for nm, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons), "samples": len(sm)}


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md, 6.65 TB/s)"


def cpu_path_time(batch, sample_b, threads, repeats=1):
    """Time the oracle's torch-CPU restatement of the reference path (fp32) on the
    first ``sample_b`` utterances of the batch; returns (seconds, frames)."""
    from oracle import engine, graphs
    torch.set_num_threads(threads)
    S, order = batch["S"], batch["order"]
    ogen = graphs.GraphGen(order, S, batch["args"])
    seqs = batch["seqs"][:sample_b]
    lens = batch["lens"][:sample_b]
    U = max(len(s) for s in seqs)
    lab = np.zeros((sample_b, U), dtype=np.int64)
    for i, s in enumerate(seqs):
        lab[i, :len(s)] = s
    nm = ogen.get_training_matrices_batch(lab, [len(s) for s in seqs])
    dm = ogen.get_decoding_matrices() if batch["use_den"] else None
    g = torch.Generator().manual_seed(batch["seed"])
    logits = torch.randn(lens[0], sample_b, batch["V"], generator=g) * 3
    best = None
    for _ in range(repeats):
        t0 = time.perf_counter()
        engine.fst_loss(logits, lens, nm, dm, batch["nbd"], S, order, dtype=torch.float32)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return best, int(sum(lens))


def run_reference(args):
    """--impl reference: the CPU implementation of the same path on the host cores.

    The reference is pure Python on torch ops (no compiled sources to build into
    oracle/_ref) and cannot travel to the GPU box, so the timed code is the
    oracle's restatement of it (same tensor-op sequence, fp32, all host threads).
    Each step is a bounded sample (the first ``sample_b`` utterances of the batch)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    batch = make_batch(args.workload)
    threads = os.cpu_count() or 1
    sample_b = 2 if batch["use_den"] else min(batch["B"], 16)
    if batch["use_den"] and batch["T"] > 400:
        sample_b = 1
    for _ in range(max(args.warmup, 0) and 1):
        cpu_path_time(batch, sample_b, threads)
    times, frames = [], 0
    budget_s = float(os.environ.get("CTCG_REFERENCE_BUDGET_S", "120"))   # keep the arm within minutes
    for _ in range(args.steps):
        dt, frames = cpu_path_time(batch, sample_b, threads)
        times.append(dt)
        if sum(times) > budget_s:
            break
    total = sum(times)
    value = frames * len(times) / total
    sample = "first %d of %d utterances of %s, T=%d, V=%d, fp32" % (
        sample_b, batch["B"], args.workload, batch["T"], batch["V"])
    line = {
        "impl": "reference", "metric": "CTC(-G) loss+grad utterance-frames/sec", "value": value,
        "unit": "utterance-frames/s", "n_gpus": args.gpus, "steps": len(times), "steps_requested": args.steps,
        "warmup": min(args.warmup, 1),
        "ms_per_step": 1e3 * total / len(times), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.workload, batch, args.gpus),
        "cpu_baseline": {"value": value, "unit": "utterance-frames/s", "cores": threads, "kind": "port",
                         "sample": sample},
        "e2e": {"value": value, "unit": "utterance-frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def workload_config(workload, batch, n_gpus):
    return {"workload": "%s: %s, B=%d per GPU, T=%d, V=%d (S=%d), U=len/4, lens in [0.6T,T]" % (
                workload, batch["name"], batch["B"], batch["T"], batch["V"], batch["S"]),
            "global_batch": batch["B"] * n_gpus, "frames_per_step_per_gpu": int(sum(batch["lens"])),
            "parallelism": "utterance-sharded dp%d" % n_gpus,
            "l2": ("logits+grad per step (%.0f MB) exceed the 126 MB L2; 3 input buffers rotated"
                   if 2 * batch["T"] * batch["B"] * batch["V"] * 4 > 126e6 else
                   "logits+grad per step (%.0f MB) fit the 126 MB L2 (parity-size workload, not the bench "
                   "configuration); 3 input buffers rotated") % (2 * batch["T"] * batch["B"] * batch["V"] * 4 / 1e6)}


def run_ours(args):
    import pytorch_asr_b200 as P
    from pytorch_asr_b200 import parallel

    rank, world, local = parallel.init_from_env()
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    batch = make_batch(args.workload, rank)
    T, B, V, S = batch["T"], batch["B"], batch["V"], batch["S"]
    gen = P.CTCGraphGen(context_order=batch["order"], graph_build_args=batch["args"], num_symbols=S)
    num = gen.get_training_graphs(batch["seqs"])
    den = gen.get_decoding_graph() if batch["use_den"] else None
    mode = {None: P.NORM_NONE, 0: P.NORM_SOFTMAX}[batch["nbd"]]
    lens = batch["lens"]
    frames = int(sum(lens))
    NBUF = 3
    host = [(torch.randn(T, B, V, generator=batch["gen"]) * 3).pin_memory() for _ in range(NBUF)]
    xs = [h.to(dev) for h in host]
    ones = None

    meter = parallel.LossMeter(dev)

    def step(i):
        # utterances are independent: no exchange per step; the logged scalar [sum loss, frames] is
        # accumulated on the device and all-reduced once, at the end of the timed region
        loss, grad = P._lib.loss_fwd_bwd(xs[i % NBUF], lens, num, den, mode, S, grad_scale=ones)
        if world > 1:
            meter.add(loss, frames)
        return loss

    for i in range(args.warmup):
        step(i)
    if world > 1:
        meter.total()                        # warm the collective (communicator set-up, first-call costs)
    parallel.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = P._lib.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    meter = parallel.LossMeter(dev)          # restart the running scalar for the timed steps
    for i in range(args.steps):
        out = step(i)
    if world > 1:
        total = meter.total()                # the one NCCL all-reduce (16 bytes), inside the timed region
    ev1.record()
    torch.cuda.synchronize()
    parallel.barrier()
    launches = P._lib.launch_count() - l0
    if world == 1:
        out = parallel.reduce_loss(out, frames)
    else:
        out = total / args.steps
    ms = ev0.elapsed_time(ev1)
    ms = parallel.max_over_ranks(ms)
    clocks = sampler.stop() if rank == 0 else None
    total_frames = parallel.sum_over_ranks(frames)

    # ---- end to end through the public API with host buffers -------------------------------
    def e2e_step(i):
        x = host[i % NBUF].to(dev, non_blocking=True).requires_grad_(True)
        g_num = gen.get_training_graphs(batch["seqs"])          # graph compile + upload from labels
        loss = P.fst_loss(x, lens, g_num, den, normalize_by_dim=batch["nbd"], num_symbols=S)
        tot = loss.sum()
        tot.backward()
        return float(tot.item()), x.grad                         # D2H read of the step's loss

    for i in range(min(args.warmup, 3)):
        e2e_step(i)
    parallel.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e2e_steps = max(3, min(args.steps, 10))
    for i in range(e2e_steps):
        e2e_step(i)
    torch.cuda.synchronize()
    e2e_s = parallel.max_over_ranks(time.perf_counter() - t0)
    e2e_value = total_frames * e2e_steps / e2e_s

    # ---- per-kernel attribution (outside the timed regions) --------------------------------
    kern = {}
    nprof_steps = 10
    if rank == 0:
        P._lib.profile(True)
        nprof = nprof_steps
        for i in range(nprof):
            P._lib.loss_fwd_bwd(xs[i % NBUF], lens, num, den, mode, S)
        for name, t in P._lib.profile_collect():
            kern[name] = kern.get(name, 0.0) + t / nprof
        P._lib.profile(False)

    if rank != 0:
        parallel.finalize()
        return
    ms_per_step = ms / args.steps
    value = total_frames * args.steps / (ms * 1e-3)
    peak, peak_src = measured_peak()
    alg_bytes = 8.0 * V * frames                     # read V logits + write V gradients per valid frame
    dom = max(kern, key=kern.get) if kern else None
    step_gbs = alg_bytes / (ms_per_step * 1e-3) / 1e9
    # dominant kernel: one launch processes the whole batch (all `frames` utterance-frames)
    dom_gbs = alg_bytes / (kern[dom] * 1e-3) / 1e9 if dom else step_gbs
    # DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture
    # (profiles/r01_e_ncu_full_summary.md); only captured for the default workload
    traffic = NCU_TRAFFIC.get((args.workload, dom))
    # every kernel against the bytes it must move itself (fp32): the sweeps read the logits once per
    # direction, the combine reads them once and writes the gradient once, the row kernels of plain CTC
    # read + write once; the compact-column kernels of the numerator move far less than 8*V per frame
    own_bytes = {"bisweep": 8.0 * V * frames, "bicombine": 8.0 * V * frames,
                 "rows": 8.0 * V * frames, "rows_ctx": 8.0 * V * frames}
    per_kernel = {}
    for k, v in kern.items():
        ent = {"ms": round(v, 4)}
        if k in own_bytes and v > 0:
            ent["own_algorithmic_bytes"] = own_bytes[k]
            ent["own_gbs"] = own_bytes[k] / (v * 1e-3) / 1e9
            ent["own_frac"] = ent["own_gbs"] / peak
        tr = NCU_TRAFFIC.get((args.workload, k))
        if tr:
            ent["ncu_dram_bytes"] = tr
        per_kernel[k] = ent
    if "bisweep" in kern:
        conc = ("denominator (bisweep -> bicombine) on a side stream, concurrent with the numerator "
                "(gather -> num_sweep -> num_combine) on the caller's stream; joined before the numerator subtracts its posteriors (num_combine) and final")
    else:
        conc = "single stream"
    roof = {"bound": "hbm", "achieved": dom_gbs, "peak": peak, "unit": "GB/s", "frac": dom_gbs / peak,
            "traffic": traffic, "peak_source": peak_src,
            "kernel": dom, "algorithmic_bytes_per_launch": alg_bytes,
            "scope": "dominant kernel: algorithmic bytes of one launch (8*V*sum(len)) / its mean CUDA-event "
                     "duration over %d steps run right after the timed region" % nprof_steps,
            "kernel_ms": {k: round(v, 4) for k, v in kern.items()},
            "per_kernel": per_kernel,
            "dominant_kernel_share": (kern[dom] / sum(kern.values())) if kern else None,
            "step_achieved": step_gbs, "step_frac": step_gbs / peak,
            "concurrency": conc}
    line = {
        "metric": "CTC(-G) loss+grad utterance-frames/sec", "value": value, "unit": "utterance-frames/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": workload_config(args.workload, batch, world),
        "clocks": clocks, "gpu_launches": int(launches),
        "e2e": {"value": e2e_value, "unit": "utterance-frames/s",
                "h2d_bytes_per_step": int(T * B * V * 4 + 4 * B), "d2h_bytes_per_step": 4,
                "steps": e2e_steps, "includes": "H2D logits from pinned host, numerator graph compile+upload "
                "from labels, fused loss+grad, D2H loss"},
        "roofline": roof,
        "loss_per_frame": float(out[0]) / max(float(out[1]), 1.0),
    }
    if not args.no_cpu_baseline and world == 1:       # reported on rank 0 at N=1 only
        threads = os.cpu_count() or 1
        sample_b = 2 if batch["use_den"] else min(B, 16)
        if batch["use_den"] and T > 400:
            sample_b = 1
        dt, fr = cpu_path_time(batch, sample_b, threads)
        line["cpu_baseline"] = {"value": fr / dt, "unit": "utterance-frames/s", "cores": threads, "kind": "port",
                                "sample": "first %d of %d utterances, one pass, fp32 torch-CPU restatement of "
                                          "path_reduction (%.1f s)" % (sample_b, B, dt)}
    print(json.dumps(line))
    parallel.finalize()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg3", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
