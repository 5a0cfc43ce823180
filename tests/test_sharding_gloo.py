"""World-size-2 gloo tests of the image-sharded host logic.  The per-rank partial
that tss_proto_update_partial computes on the GPU is formed here with the oracle's closed-form weights, so the
exchange (all_gather of counts, all_reduce of the weighted partial) and the combination are what is tested:
two ranks must reproduce the reference's SEQUENTIAL update over the concatenated batch."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import label_path as O


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _weights(num0, M, mode, momentum, start_mean, cap=3000.0):
    """closed-form coefficients of M sequential updates from count num0 (fp64): (w0, w[M], num_after)."""
    w0, w, n = 1.0, np.zeros(M), float(num0)
    for j in range(M):
        mean = mode == "mean" or (start_mean and n < 100)
        a, b = (n / (n + 1.0), 1.0 / (n + 1.0)) if mean else (1.0 - momentum, momentum)
        w0 *= a
        w[:j] *= a
        w[j] = b
        n = min(n + 1.0, cap)
    return w0, w, n


def _worker(rank, ws, port, mode, start_mean, steps, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    from b200seg import sharding

    C, K, n_per_rank, momentum = 5, 16, 3, 0.1
    protos, num = torch.zeros(C, K), torch.zeros(C)
    hist = torch.zeros(C, C, dtype=torch.int64)
    bad = torch.zeros(1, dtype=torch.int64)
    for step in range(steps):
        r = np.random.RandomState(100 + step)                          # the same global batch on every rank
        vec = r.standard_normal((ws * n_per_rank, C, K)).astype(np.float32)
        ok = r.rand(ws * n_per_rank, C) < 0.7
        lo, hi = sharding.shard_range(ws * n_per_rank, rank, ws)
        counts = torch.from_numpy(ok[lo:hi].sum(0).astype(np.int32))
        first, total = sharding.prototype_positions(counts)
        partial = torch.zeros(C, K)
        scale0, num_after = torch.zeros(C), torch.zeros(C)
        for c in range(C):
            w0, w, na = _weights(num[c].item(), int(total[c]), mode, momentum, start_mean)
            mine = vec[lo:hi][ok[lo:hi, c], c]                         # this rank's contributing vectors, in order
            wj = w[int(first[c]): int(first[c]) + len(mine)]
            partial[c] = torch.from_numpy((wj[:, None] * mine.astype(np.float64)).sum(0).astype(np.float32))
            scale0[c], num_after[c] = w0, na
        sharding.apply_prototype_partial(protos, num, partial, scale0, num_after)
        hist += torch.full((C, C), rank + 1, dtype=torch.int64)
        bad += rank
    sharding.all_reduce_confusion(hist, bad)
    loss = sharding.all_reduce_loss(torch.tensor(2.0 * (rank + 1)), torch.tensor(4.0), eps=1e-6)
    if rank == 0:
        out.put((protos.numpy(), num.numpy(), hist.numpy(), int(bad), float(loss)))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("mode,start_mean", [("mean", True), ("moving_average", False), ("moving_average", True)])
def test_two_ranks_reproduce_the_sequential_update(mode, start_mean):
    ws, steps = 2, 3
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, ws, port, mode, start_mean, steps, out)) for r in range(ws)]
    for p in procs:
        p.start()
    protos, num, hist, bad, loss = out.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # sequential oracle over the concatenated batches (the reference's update_objective_SingleVector)
    C, K, n_per_rank = 5, 16, 3
    bank = O.PrototypeBank(C, K, momentum=0.1)
    for step in range(steps):
        r = np.random.RandomState(100 + step)
        vec = r.standard_normal((ws * n_per_rank, C, K)).astype(np.float32)
        ok = r.rand(ws * n_per_rank, C) < 0.7
        for n in range(ws * n_per_rank):
            for c in range(C):
                if ok[n, c]:
                    bank.update(c, torch.from_numpy(vec[n, c]), mode, start_mean)
    assert np.array_equal(num, bank.num.numpy())
    assert np.allclose(protos, bank.vectors.numpy(), rtol=1e-5, atol=1e-6)
    assert (hist == steps * 3).all() and bad == steps * 1
    assert abs(loss - 6.0 / (8.0 + 1e-6)) < 1e-6


def test_single_process_fallbacks():
    from b200seg import sharding
    assert sharding.world() == (0, 1)
    counts = torch.tensor([3, 0, 2], dtype=torch.int32)
    first, total = sharding.prototype_positions(counts)
    assert first.tolist() == [0, 0, 0] and total.tolist() == [3, 0, 2]
    h = torch.ones(3, 3, dtype=torch.int64)
    assert sharding.all_reduce_confusion(h) is h
