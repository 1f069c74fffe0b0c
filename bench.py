"""Benchmark of the CTC / CTC-G loss+grad hot path (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg3]
    python bench.py --workload cfg5 --scaling strong [--T 800]     # B=2048 in total, striped over the ranks
    python bench.py --impl reference ...                           # the CPU implementation of the same path

One "step" = one loss+grad over one synthetic WSJ-shaped batch through the public API, exactly what a
trainer runs: ``P.fst_loss(logits, ...).sum().backward()`` (one ``ctcg_loss_fwd_bwd`` through the C ABI
plus ``ctcg_scale_grad`` in backward).  ``value`` = utterance-frames/s with the inputs resident in HBM;
``e2e`` = the same metric with pinned HOST logits (H2D copy of the logits, graph compile + upload from
the labels, D2H read of the loss inside the timed region); ``e2e_host`` = the C entry point for host
buffers (``ctcg_loss_host``), gradient copied back too.  ``roofline.frac`` is the WHOLE STEP:
8*V*sum(len) bytes / ms_per_step / measured HBM peak.  Prints ONE JSON line.
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

# name -> (variant, order, S, B per GPU, T, graph_build_args, normalize_by_dim, use_den)
WORKLOADS = {
    "cfg1": ("mono-char CTC", 1, 49, 16, 400, {}, 0, False),
    "cfg2": ("bi-char CTC", 2, 49, 64, 400, {}, 0, False),
    "cfg2s": ("bi-char CTC V=1296", 2, 36, 64, 400, {}, 0, False),
    "cfg3": ("CTC-G bi-char", 2, 49, 64, 400, {}, None, True),
    "cfg4": ("CTC-GB contextual blanks", 2, 49, 256, 800, {"use_contextual_blanks": True}, None, True),
    "cfg5": ("CTC-G sweep shard", 2, 49, 256, 800, {}, None, True),
}
CFG_INDEX = {"cfg1": 1, "cfg2": 2, "cfg2s": 2, "cfg3": 3, "cfg4": 4, "cfg5": 5}
STRONG_TOTAL_B = 2048          # BASELINE.json configs[4]: B=2048 in total, T in {400, 800, 1600}
TRAFFIC_JSON = os.path.join(ROOT, "profiles", "r02_ncu_traffic.json")   # written by profiles/ncu_traffic.py


def make_batch(workload, rank=0, world=1, scaling="weak", T_override=None):
    """Seeded synthetic batch of SURVEY 8(d): logits ~ 3*N(0,1), lens[0]=T, others uniform in
    [0.6T, T] sorted descending, U_b = len_b // 4 labels in 1..S-1.  ``strong``: one batch of
    STRONG_TOTAL_B utterances for the whole job, utterance i -> rank i mod world."""
    name, order, S, B, T, args, nbd, use_den = WORKLOADS[workload]
    if T_override:
        T = int(T_override)
    V = S ** order
    if scaling == "strong":
        seed = 1000 + CFG_INDEX[workload]
        rng = np.random.default_rng(seed)
        total = STRONG_TOTAL_B
        lens_all = sorted([T] + rng.integers(int(0.6 * T), T + 1, size=total - 1).tolist(), reverse=True)
        seqs_all = [rng.integers(1, S, size=max(1, ln // 4)).tolist() for ln in lens_all]
        lens, seqs = lens_all[rank::world], seqs_all[rank::world]
        B = len(lens)
        T = lens[0]                      # this rank's longest utterance
    else:
        # weak scaling = the SAME work on every GPU: every rank's batch has the lengths and transcripts of the
        # N=1 batch (the seed does not depend on the rank), its own logits (their seed does, below).  With a
        # per-rank draw the ranks' frame counts differ by a few per cent and the whole-job figure measures the luck
        # of the draw next to the overheads.
        seed = 1000 + CFG_INDEX[workload]
        rng = np.random.default_rng(seed)
        lens = sorted([T] + rng.integers(int(0.6 * T), T + 1, size=B - 1).tolist(), reverse=True)
        seqs = [rng.integers(1, S, size=max(1, ln // 4)).tolist() for ln in lens]
    return dict(name=name, order=order, S=S, B=B, T=T, V=V, args=args, nbd=nbd, use_den=use_den,
                lens=lens, seqs=seqs, seed=seed + 7919 * rank, scaling=scaling)


class ClockSampler(object):
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []          # (host time, fields)
        self._stop = threading.Event()
        self._thr = None
        self._sample = self._nvml()      # NVML set up on the caller's thread, before any barrier: the sampling
        self.window = None               # thread must not hold the GIL (imports) while the timed loop starts

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
        sample = self._sample
        while not self._stop.is_set():
            try:
                if sample is not None:
                    self.rows.append((time.perf_counter(), sample()))
                    self._stop.wait(0.002)
                    continue
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True,
                                     timeout=5).stdout.strip()
                if out:
                    self.rows.append((time.perf_counter(), [c.strip() for c in out.split(",")]))
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
        rows = self.rows
        if self.window is not None:      # the thread starts before the barrier: keep the samples of the timed region
            inside = [r for r in rows if self.window[0] <= r[0] <= self.window[1]]
            rows = inside if inside else rows[-1:]
        for _, r in rows:
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


def ncu_traffic(workload):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of every kernel of the step, from the ncu
    --set full capture committed under profiles/ (profiles/ncu_traffic.py writes the file, with the commit
    it was taken at); None when the workload has no capture."""
    try:
        with open(TRAFFIC_JSON) as f:
            d = json.load(f)
        ent = d["workloads"].get(workload)
        if not ent:
            return None, None
        return {k: int(v) for k, v in ent.items()}, d.get("commit")
    except Exception:
        return None, None


def cpu_sample_size(batch):
    """the CTC-G / CTC-GB CPU path is timed on a slice of the batch (B=8 at T=400;
    fewer utterances for the long ones so that one pass stays near ten seconds) and scaled linearly."""
    if not batch["use_den"]:
        return min(batch["B"], 16)
    return max(1, min(batch["B"], 8 * 400 // max(batch["T"], 400)))


def cpu_path_time(batch, sample_b, threads):
    """One pass of the oracle's torch-CPU restatement of the reference path (fp32: get_normalized_acts,
    the max-shift / numerator - denominator glue, path_reduction forward and backward) on the first
    ``sample_b`` utterances of the batch; returns (seconds, frames)."""
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
    t0 = time.perf_counter()
    engine.fst_loss(logits, lens, nm, dm, batch["nbd"], S, order, dtype=torch.float32)
    return time.perf_counter() - t0, int(sum(lens))


def cpu_baseline(batch, passes=3, budget_s=150.0):
    """The ONE protocol both arms report: 1 warm-up pass + best of ``passes``."""
    threads = os.cpu_count() or 1
    sample_b = cpu_sample_size(batch)
    t_start = time.perf_counter()
    cpu_path_time(batch, sample_b, threads)                 # warm-up (allocator, thread pool, graph build)
    times, frames = [], 0
    for _ in range(max(passes, 1)):
        dt, frames = cpu_path_time(batch, sample_b, threads)
        times.append(dt)
        if time.perf_counter() - t_start > budget_s:
            break
    best = min(times)
    return {"value": frames / best, "unit": "utterance-frames/s", "cores": threads, "kind": "port",
            "sample": "first %d of %d utterances (T=%d, V=%d), fp32 torch-CPU restatement of the reference path, "
                      "1 warm-up + best of %d passes (%.2f s), per-utterance cost scales linearly to the batch"
                      % (sample_b, batch["B"], batch["T"], batch["V"], len(times), best),
            "pass_seconds": [round(t, 3) for t in times]}


def workload_config(workload, batch, n_gpus):
    tb = 2 * batch["T"] * batch["B"] * batch["V"] * 4
    cfg = {"workload": "%s: %s, B=%d per GPU, T=%d, V=%d (S=%d), U=len/4, lens in [0.6T,T]" % (
               workload, batch["name"], batch["B"], batch["T"], batch["V"], batch["S"]),
           "global_batch": batch["B"] * n_gpus if batch["scaling"] == "weak" else STRONG_TOTAL_B,
           "frames_per_step_per_gpu": int(sum(batch["lens"])),
           "parallelism": "utterance-sharded dp%d" % n_gpus + (
               "; every rank's batch has the lengths and transcripts of the N=1 batch and its own logits"
               if n_gpus > 1 and batch["scaling"] == "weak" else ""),
           "l2": ("logits+grad per step (%.0f MB) exceed the 126 MB L2; input buffers rotated" if tb > 126e6 else
                  "logits+grad per step (%.0f MB) fit the 126 MB L2 (parity-size workload, not the bench "
                  "configuration); input buffers rotated") % (tb / 1e6)}
    if batch["scaling"] == "strong":
        cfg["workload"] = "%s strong: %s, B=%d in total striped over %d GPU(s) (B=%d on rank 0), T=%d, V=%d" % (
            workload, batch["name"], STRONG_TOTAL_B, n_gpus, batch["B"], batch["T"], batch["V"])
    return cfg


def run_reference(args):
    """--impl reference: the CPU implementation of the same path on the host cores.

    The reference is pure Python on torch ops (no compiled sources to build into oracle/_ref) and cannot
    travel to the GPU box, so the timed code is the oracle's restatement of it (same tensor-op sequence,
    fp32, all host threads); each step is one pass over a bounded sample of the workload.  Same protocol
    and the same function as the ``cpu_baseline`` of the main line (1 warm-up + best pass)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    batch = make_batch(args.workload, 0, 1, "weak", args.T)
    budget_s = float(os.environ.get("CTCG_REFERENCE_BUDGET_S", "150"))     # keep the arm within minutes
    cb = cpu_baseline(batch, passes=max(1, min(args.steps, 20)), budget_s=budget_s)
    n = len(cb["pass_seconds"])
    best = min(cb["pass_seconds"])
    line = {
        "impl": "reference", "metric": "CTC(-G) loss+grad utterance-frames/sec", "value": cb["value"],
        "unit": "utterance-frames/s", "n_gpus": args.gpus, "steps": n, "steps_requested": args.steps, "warmup": 1,
        "ms_per_step": 1e3 * best, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.workload, batch, args.gpus),
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": "utterance-frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def run_ours(args):
    import pytorch_asr_b200 as P
    from pytorch_asr_b200 import parallel

    rank, world, local = parallel.init_from_env()
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    batch = make_batch(args.workload, rank, world, args.scaling, args.T)
    T, B, V, S = batch["T"], batch["B"], batch["V"], batch["S"]
    gen = P.CTCGraphGen(context_order=batch["order"], graph_build_args=batch["args"], num_symbols=S)
    num = gen.get_training_graphs(batch["seqs"])
    den = gen.get_decoding_graph() if batch["use_den"] else None
    mode = {None: P.NORM_NONE, 0: P.NORM_SOFTMAX}[batch["nbd"]]
    lens = batch["lens"]
    frames = int(sum(lens))
    buf_bytes = T * B * V * 4
    NBUF = 3 if buf_bytes <= (4 << 30) else (2 if buf_bytes <= (12 << 30) else 1)   # one 31 GB buffer is >> L2 too
    host_ok = buf_bytes <= (2 << 30)          # pinned host copies (for the end-to-end legs) only at sane sizes
    if host_ok:
        g = torch.Generator().manual_seed(batch["seed"])
        host = [(torch.randn(T, B, V, generator=g) * 3).pin_memory() for _ in range(NBUF)]
        xs = [h.to(dev).requires_grad_(True) for h in host]
    else:
        g = torch.Generator(device=dev).manual_seed(batch["seed"])
        host = None
        xs = [(torch.randn(T, B, V, generator=g, device=dev) * 3).requires_grad_(True) for _ in range(NBUF)]
    meter = parallel.LossMeter(dev)

    def step(i):
        # What a trainer runs.  Utterances are independent: no exchange per step; the logged scalar
        # [sum loss, frames] is accumulated on the device and all-reduced once per logging interval.
        x = xs[i % NBUF]
        x.grad = None
        loss = P.fst_loss(x, lens, num, den, normalize_by_dim=batch["nbd"], num_symbols=S)
        loss.sum().backward()
        if world > 1:
            meter.add(loss, frames)
        return loss

    # One-time initialisation, before the warm-up: ctcg_loss_fwd_bwd replays a captured CUDA graph for calls whose
    # arguments repeat (third sighting; a capture costs milliseconds) -- with NBUF rotating buffers that is
    # 3 * NBUF steps until every signature is captured.  Like any first-call cost it belongs in front of the warm-up.
    graphs_on = os.environ.get("CTCG_STEP_GRAPHS", "1") != "0"
    init_steps = 4 * NBUF if graphs_on else 0
    for i in range(init_steps):
        step(i)
    for i in range(args.warmup):
        step(i)
    if world > 1:
        meter.total()                        # warm the collective (communicator set-up, first-call costs)
    # everything that costs host time on one rank only (sampler thread, NVML) happens before the barrier, so the
    # ranks leave it together: a rank that starts late is waited for in the window's all-reduce by all the others
    sampler = ClockSampler(local) if rank == 0 else None
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    meter = parallel.LossMeter(dev)          # restart the running scalar for the timed steps
    l0 = P._lib.launch_count()
    sg0 = P._lib.step_graph_info()
    parallel.barrier()
    torch.cuda.synchronize()
    w0 = time.perf_counter()
    trace = [] if os.environ.get("BENCH_TRACE") else None   # debugging aid: an event after every step (stderr)
    ev0.record()
    for i in range(args.steps):
        out = step(i)
        if trace is not None:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            trace.append((e, time.perf_counter() - w0))
    if world > 1:
        total = meter.total()                # the one NCCL all-reduce (16 bytes), inside the timed region
    ev1.record()
    torch.cuda.synchronize()
    if trace is not None:
        sys.stderr.write("rank %d: step end on the GPU (ms since start) / enqueued by the host (ms): %s | window %.3f\n" % (
            rank, " ".join("%.2f/%.2f" % (ev0.elapsed_time(e), 1e3 * h) for e, h in trace), ev0.elapsed_time(ev1)))
    if sampler is not None:
        sampler.window = (w0, time.perf_counter())
    parallel.barrier()
    launches = P._lib.launch_count() - l0
    sg1 = P._lib.step_graph_info()
    step_graphs = {"replays": sg1[0] - sg0[0], "captures": sg1[1] - sg0[1], "ordinary_path": sg1[2] - sg0[2],
                   "init_steps_before_warmup": init_steps}
    if world == 1:
        out = parallel.reduce_loss(out.detach(), frames)
    else:
        out = total / args.steps
    ms = parallel.max_over_ranks(ev0.elapsed_time(ev1))
    clocks = sampler.stop() if rank == 0 else None
    total_frames = parallel.sum_over_ranks(frames)

    # ---- the same steps straight through the C ABI (no autograd around the call) ----------------
    xd = [x.detach() for x in xs]
    for x in xs:
        x.grad = None                        # the gradient of the last public-API step (memory at B=2048, T=1600)
    for i in range(init_steps + min(args.warmup, 3)):
        P._lib.loss_fwd_bwd(xd[i % NBUF], lens, num, den, mode, S)
    torch.cuda.synchronize()
    ev0.record()
    for i in range(args.steps):
        P._lib.loss_fwd_bwd(xd[i % NBUF], lens, num, den, mode, S)
    ev1.record()
    torch.cuda.synchronize()
    ms_cabi = parallel.max_over_ranks(ev0.elapsed_time(ev1))

    # ---- end to end through the public API with host buffers -------------------------------
    e2e = e2e_host = None
    if host_ok:
        # What a trainer with a host-side producer does: the upload of batch i+1 runs on a copy stream while the
        # kernels of batch i run (every step's H2D copy and D2H loss read are inside the timed region).
        copy_stream = torch.cuda.Stream(device=dev)
        main_stream = torch.cuda.current_stream(dev)

        def e2e_upload(i):
            with torch.cuda.stream(copy_stream):
                x = P.upload_logits(host[i % NBUF], lens)                # live frames only (sum(len)*V floats)
                ev = torch.cuda.Event()
                ev.record(copy_stream)
            return x, ev

        def e2e_step(i, cur, last=False):
            x, ev = cur
            main_stream.wait_event(ev)
            x.record_stream(main_stream)
            x.requires_grad_(True)
            g_num = gen.get_training_graphs(batch["seqs"])          # graph compile + upload from labels
            loss = P.fst_loss(x, lens, g_num, den, normalize_by_dim=batch["nbd"], num_symbols=S)
            tot = loss.sum()
            tot.backward()
            nxt = None if last else e2e_upload(i + 1)                # queued behind nothing: overlaps this step's kernels
            return float(tot.item()), x.grad, nxt                    # D2H read of the step's loss

        cur = e2e_upload(0)
        for i in range(min(args.warmup, 3)):
            cur = e2e_step(i, cur)[2]
        parallel.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e2e_steps = max(3, min(args.steps, 10))
        cur = e2e_upload(0)
        for i in range(e2e_steps):
            cur = e2e_step(i, cur, last=(i == e2e_steps - 1))[2]
        torch.cuda.synchronize()
        e2e_s = parallel.max_over_ranks(time.perf_counter() - t0)
        e2e = {"value": total_frames * e2e_steps / e2e_s, "unit": "utterance-frames/s",
               "h2d_bytes_per_step": int(frames * V * 4 + 4 * B), "d2h_bytes_per_step": 4, "steps": e2e_steps,
               "includes": "H2D of the live frames of the logits from pinned host (ctcg_upload_logits, on a copy stream: the upload of step i+1 overlaps the kernels of step i), numerator graph compile+upload from labels, fused "
                           "loss+grad (gradient stays on the device for the trainer's backprop), D2H loss"}
        if rank == 0 and world == 1:
            # the C entry point for host buffers: logits H2D, loss AND gradient D2H inside the call
            gh, lh = torch.empty_like(host[0]).pin_memory(), torch.empty(B).pin_memory()
            P._lib.loss_host(host[0], lens, num, den, mode, S, loss_out=lh, grad_out=gh)
            t0 = time.perf_counter()
            hsteps = 3
            for i in range(hsteps):
                P._lib.loss_host(host[i % NBUF], lens, num, den, mode, S, loss_out=lh, grad_out=gh)
            dt = time.perf_counter() - t0
            e2e_host = {"value": frames * hsteps / dt, "unit": "utterance-frames/s", "entry": "ctcg_loss_host",
                        "h2d_bytes_per_step": int(frames * V * 4 + 4 * B), "d2h_bytes_per_step": int(buf_bytes + 4 * B),
                        "steps": hsteps}

    # ---- per-kernel attribution (outside the timed regions) --------------------------------
    kern = {}
    nprof_steps = 10
    if rank == 0:
        P._lib.profile(True)
        for i in range(nprof_steps):
            P._lib.loss_fwd_bwd(xd[i % NBUF], lens, num, den, mode, S)
        for name, t in P._lib.profile_collect():
            kern[name] = kern.get(name, 0.0) + t / nprof_steps
        P._lib.profile(False)

    if rank != 0:
        parallel.finalize()
        return
    ms_per_step = ms / args.steps
    value = total_frames * args.steps / (ms * 1e-3)
    peak, peak_src = measured_peak()
    alg_bytes = 8.0 * V * frames                     # read V logits + write V gradients per valid frame
    step_gbs = alg_bytes / (ms_per_step * 1e-3) / 1e9
    dom = max(kern, key=kern.get) if kern else None
    traffic_k, traffic_commit = ncu_traffic(args.workload if args.scaling == "weak" and not args.T else "-")
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
        if traffic_k and k in traffic_k:
            ent["ncu_dram_bytes"] = traffic_k[k]
        per_kernel[k] = ent
    if "bisweep" in kern:
        conc = ("gather on the caller's stream; then the denominator sweeps (bisweep) on a side stream next to the "
                "numerator (num_sweep -> num_combine); the denominator's combine (bicombine) waits for both and "
                "writes gamma_den - gamma_num in one store; final")
    else:
        conc = "single stream"
    traffic = sum(traffic_k.values()) if traffic_k else None
    roof = {"bound": "hbm", "achieved": step_gbs, "peak": peak, "unit": "GB/s", "frac": step_gbs / peak,
            "traffic": traffic, "peak_source": peak_src,
            "algorithmic_bytes_per_step": alg_bytes,
            "scope": "whole step: 8*V*sum(len) algorithmic bytes / ms_per_step (CUDA events around the timed steps); "
                     "no single kernel does the 8*V work, so the step is the unit",
            "traffic_scope": ("sum over the step's kernels of dram__bytes_read.sum + dram__bytes_write.sum per launch, "
                              "ncu --set full at commit %s (profiles/r02_ncu_traffic.json)" % traffic_commit)
            if traffic_k else "no ncu capture for this workload",
            "traffic_over_algorithmic": (traffic / alg_bytes) if traffic else None,
            "dominant_kernel": dom, "kernel_ms": {k: round(v, 4) for k, v in kern.items()},
            "kernel_ms_scope": "mean CUDA-event duration per kernel on its own stream over %d steps run after the "
                               "timed region (kernels of different streams overlap: they do not add up to the step)"
                               % nprof_steps,
            "per_kernel": per_kernel, "concurrency": conc}
    line = {
        "metric": "CTC(-G) loss+grad utterance-frames/sec", "value": value, "unit": "utterance-frames/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": workload_config(args.workload, batch, world),
        "clocks": clocks, "gpu_launches": int(launches), "step_graphs": step_graphs,
        "timed_call": "P.fst_loss(x, lens, num, den).sum().backward() (public API; device-resident inputs)",
        "c_abi_direct": {"ms_per_step": ms_cabi / args.steps,
                         "value": total_frames * args.steps / (ms_cabi * 1e-3),
                         "call": "ctcg_loss_fwd_bwd through ctypes, no autograd"},
        "e2e": e2e if e2e else {"value": None, "unit": "utterance-frames/s", "h2d_bytes_per_step": 0,
                                "d2h_bytes_per_step": 0,
                                "skipped": "logits of this workload (%.1f GB per buffer) are generated on the "
                                           "device; no pinned host copy" % (buf_bytes / 1e9)},
        "roofline": roof,
        "loss_per_frame": float(out[0]) / max(float(out[1]), 1.0),
    }
    if e2e_host:
        line["e2e_host"] = e2e_host
    if not args.no_cpu_baseline and world == 1:       # reported on rank 0 at N=1 only
        line["cpu_baseline"] = cpu_baseline(make_batch(args.workload, 0, 1, "weak", args.T))
    print(json.dumps(line))
    parallel.finalize()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg3", choices=sorted(WORKLOADS))
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="strong: one batch of %d utterances striped over the ranks (BASELINE config 5)" % STRONG_TOTAL_B)
    ap.add_argument("--T", type=int, default=None, help="override the workload's T (config 5: 400 / 800 / 1600)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
