"""Utterance-sharded multi-GPU execution: one process per GPU, no data-path collective.

The reference is single-process / single-GPU (no DataParallel, no
torch.distributed).  Utterances are independent, so the batch is striped over
ranks -- utterance ``i`` of the length-sorted batch goes to rank ``i mod G`` so
every rank sees the same length mix and its shard stays sorted -- each rank runs
the fused loss on its shard, the gradients stay on the GPU that owns the
utterance, and the only exchange is one all-reduce (sum) of the 2-element fp64
vector ``[sum of losses, number of frames]`` per step (NCCL over NVLink on the
box; gloo in the CPU tests).
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def init_from_env(backend=None):
    """Join the process group described by RANK / WORLD_SIZE / MASTER_* (torchrun).
    Returns ``(rank, world_size, local_rank)``; a lone process needs no group."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        kwargs = {}
        if backend == "nccl":
            torch.cuda.set_device(local)
            kwargs["device_id"] = torch.device("cuda", local)
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kwargs)
    return rank, world, local


def finalize():
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()


def stripe(n_items, rank, world):
    """Indices of the length-sorted batch owned by ``rank`` (i mod G striping)."""
    return list(range(rank, n_items, world))


def shard_batch(lens, seqs, rank, world):
    """Stripe a length-sorted batch; the shard stays sorted descending."""
    idx = stripe(len(lens), rank, world)
    return idx, [lens[i] for i in idx], [seqs[i] for i in idx]


def _group_active():
    return dist.is_initialized() and dist.get_world_size() > 1


_side_streams = {}


def reduce_loss(loss_per_utt, n_frames, blocking=True):
    """All-reduce ``[sum_b loss_b, frames]`` over ranks; returns the 2-element fp64
    tensor (on the loss's device).  Stream-ordered after the loss kernels; the
    gradient tensors are untouched, so backprop on each rank can proceed without
    waiting for it.  ``blocking=False`` (CUDA only) runs the collective on a side
    stream: the caller's stream is not held up by the NCCL launch latency, and the
    result must be read after a device (or ``torch.cuda.current_stream().wait_stream``)
    synchronisation."""
    buf = torch.empty(2, dtype=torch.float64, device=loss_per_utt.device)
    buf[0].copy_(loss_per_utt.sum(dtype=torch.float64))
    buf[1].fill_(float(n_frames))       # kernels only: no host <-> device synchronisation
    if _group_active():
        if not blocking and buf.is_cuda:
            dev = buf.device
            side = _side_streams.get(dev.index)
            if side is None:
                side = _side_streams[dev.index] = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                dist.all_reduce(buf, op=dist.ReduceOp.SUM)
            buf.record_stream(side)
        else:
            dist.all_reduce(buf, op=dist.ReduceOp.SUM)
    return buf


class LossMeter(object):
    """Running ``[sum of losses, frames]`` of this rank, kept on the device; ``total()`` does the
    one all-reduce when the trainer actually wants the number (a logging interval, an epoch).
    Utterances are independent, so nothing has to be exchanged per step."""

    def __init__(self, device, fold_every=256):
        self.acc = torch.zeros((), dtype=torch.float64, device=device)
        self.frames = 0
        self.pending = []              # per-step (B,) loss tensors, summed lazily: no kernel per step
        self.fold_every = fold_every

    def add(self, loss_per_utt, n_frames):
        self.pending.append(loss_per_utt.detach())
        self.frames += int(n_frames)
        if len(self.pending) >= self.fold_every:
            self._fold()

    def _fold(self):
        if self.pending:
            self.acc += torch.cat([p.reshape(-1) for p in self.pending]).sum(dtype=torch.float64)
            self.pending = []

    def total(self):
        """``[sum of losses, frames]`` over all ranks, on the device.  Nothing here waits for the GPU: the
        buffer is filled by kernels (a ``torch.tensor(x, device=...)`` would be a synchronous copy that
        stalls the host until every queued step has run, and the fold, the NCCL launch and the next step's
        launches would then sit exposed behind it), so the collective is enqueued while the last steps are
        still running."""
        self._fold()
        buf = torch.empty(2, dtype=torch.float64, device=self.acc.device)
        buf[0].copy_(self.acc)
        buf[1].fill_(float(self.frames))
        if _group_active():
            dist.all_reduce(buf, op=dist.ReduceOp.SUM)
        return buf


def barrier():
    if _group_active():
        dist.barrier()


def max_over_ranks(value):
    if not _group_active():
        return float(value)
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    t = torch.tensor([float(value)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(value):
    if not _group_active():
        return value
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    t = torch.tensor([float(value)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return type(value)(t.item())


def sharded_fst_loss(logits_shard, lens_shard, numerator, denominator, normalize_by_dim=None,
                     num_symbols=0):
    """Fused loss on this rank's shard + the scalar reduction.

    Returns ``(loss_per_utt (local, differentiable), [global loss sum, global frames])``.
    """
    from .ctc_losses import fst_loss
    loss = fst_loss(logits_shard, lens_shard, numerator, denominator, normalize_by_dim, num_symbols)
    totals = reduce_loss(loss.detach(), int(sum(int(l) for l in lens_shard)))
    return loss, totals
