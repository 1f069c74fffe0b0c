"""World-size-2 checks of the utterance-sharded driver over gloo on CPU.

The data path has no collective (utterances are independent); what is covered here
is the host logic of ``pytorch-asr_b200/parallel.py``: i-mod-G striping of the
length-sorted batch, the one scalar all-reduce ``[sum of losses, frames]`` and the
max/sum helpers ``bench.py`` uses for its timing.  The per-shard loss is the CPU
oracle (test infrastructure) standing in for the CUDA call, so that the sharded total
can be compared with the unsharded one.
"""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import pytorch_asr_b200 as P
from pytorch_asr_b200 import parallel


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _batch():
    g = torch.Generator().manual_seed(11)
    S, T = 5, 14
    lens = [14, 13, 11, 10, 9, 7]
    seqs = [[1, 2, 2], [3, 1], [4, 4, 1], [2], [1, 3], [2, 2]]
    logits = torch.randn(T, len(lens), S * S, generator=g) * 2
    return S, T, lens, seqs, logits


def _oracle_loss(S, logits, lens, seqs):
    from oracle import engine, graphs
    ogen = graphs.GraphGen(2, S, {})
    U = max(len(s) for s in seqs)
    lab = np.zeros((len(seqs), U), dtype=np.int64)
    for i, s in enumerate(seqs):
        lab[i, :len(s)] = s
    nm = ogen.get_training_matrices_batch(lab, [len(s) for s in seqs])
    dm = ogen.get_decoding_matrices()
    loss, _ = engine.fst_loss(logits[:lens[0]], lens, nm, dm, None, S, 2, dtype=torch.float64)
    return loss


def _worker(rank, world, port, out):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    r, w, _ = parallel.init_from_env(backend="gloo")
    assert (r, w) == (rank, world)
    S, T, lens, seqs, logits = _batch()
    idx, lens_s, seqs_s = parallel.shard_batch(lens, seqs, rank, world)
    assert lens_s == sorted(lens_s, reverse=True)             # a shard stays sorted
    loss = _oracle_loss(S, logits[:, idx], lens_s, seqs_s)
    tot = parallel.reduce_loss(loss, sum(lens_s))
    meter = parallel.LossMeter(torch.device("cpu"))       # running scalar, one all-reduce on demand
    meter.add(loss, sum(lens_s))
    meter.add(loss, sum(lens_s))
    mtot = meter.total()
    assert abs(float(mtot[0]) - 2 * float(tot[0])) < 1e-9 * abs(float(tot[0])) and float(mtot[1]) == 2 * float(tot[1])
    mx = parallel.max_over_ranks(float(rank + 1))
    sm = parallel.sum_over_ranks(len(idx))
    parallel.barrier()
    out[rank] = (idx, loss.tolist(), tot.tolist(), mx, sm)
    parallel.finalize()


def test_striping_covers_the_batch_once():
    for n in (1, 5, 8, 64):
        for world in (1, 2, 4, 8):
            seen = sorted(i for r in range(world) for i in parallel.stripe(n, r, world))
            assert seen == list(range(n))
            sizes = [len(parallel.stripe(n, r, world)) for r in range(world)]
            assert max(sizes) - min(sizes) <= 1


def test_world_size_two_sharded_loss_equals_unsharded():
    world, port = 2, _free_port()
    with mp.Manager() as m:
        out = m.dict()
        mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
        res = {k: out[k] for k in range(world)}
    S, T, lens, seqs, logits = _batch()
    full = _oracle_loss(S, logits, lens, seqs)
    for rank in range(world):
        idx, loss, tot, mx, sm = res[rank]
        assert idx == list(range(rank, len(lens), world))
        np.testing.assert_allclose(loss, full[idx].numpy(), rtol=1e-12)      # utterances are independent
        np.testing.assert_allclose(tot[0], float(full.sum()), rtol=1e-12)    # the one all-reduce
        assert tot[1] == float(sum(lens)) and mx == float(world) and sm == len(lens)


def test_single_process_needs_no_group():
    assert not dist.is_initialized()
    t = parallel.reduce_loss(torch.tensor([1.0, 2.5]), 7)
    assert t.tolist() == [3.5, 7.0]
    assert parallel.max_over_ranks(3.0) == 3.0 and parallel.sum_over_ranks(4) == 4
    assert hasattr(P, "fst_loss")
