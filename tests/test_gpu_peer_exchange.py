"""Two-GPU test of the peer-memory prototype exchange (sharding.PeerGather + tss_proto_finalize_scatter):
two ranks, each with its own images, must both end with the bank that ONE process gets from the sequential
update over the concatenated batch (the reference's global image order), over several steps: bit-identical
between the ranks, and within fp32 summation-order noise of a separate run (the segmented sums are accumulated
with floating-point atomics, so two runs of tss_proto_accumulate differ in the last bits).
Skipped on boxes with fewer than two GPUs."""
import os
import socket

import pytest
import torch
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _inputs(step, rank, n, c, k, h, w, dev):
    g = torch.Generator(device="cpu").manual_seed(9000 + 10 * step + rank)
    feat = torch.randn(n, k, h, w, generator=g)
    logits = 3.0 * torch.randn(n, c, h, w, generator=g)
    return feat.to(dev), logits.to(dev)


def _worker_graphed(rank, ws, port, steps, out_dir):
    """The form bench.py times: per input slot ONE captured graph for labels + accumulate and ONE for the tail
    (finalize + NVLink scatter, barrier, sequential update) replayed on a tail stream beside the next step."""
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(ws))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=ws, device_id=dev)
    from b200seg import ops, sharding
    from b200seg.graphs import GraphedCall

    n, c, k, h, w = 3, 13, 64, 17, 29
    bank = sharding.ShardedPrototypes(c, k, dev, ops.UPDATE_MOVING_AVERAGE, True, 0.9)
    gather = sharding.PeerGather(n, c, k, dev)
    slots = 2
    bufs = [tuple(torch.empty_like(t) for t in _inputs(0, rank, n, c, k, h, w, dev)) for _ in range(slots)]
    for s in range(slots):                      # the capture warm-ups run real steps: feed them step 0 / 1 and start from there
        for d, t in zip(bufs[s], _inputs(s, rank, n, c, k, h, w, dev)):
            d.copy_(t)
    mains = [GraphedCall(lambda s=s: ops.class_sums(bufs[s][0], bufs[s][1], c), warmup=1) for s in range(slots)]

    def tail_fn(s):
        means = ops.class_finalize(mains[s].out, scatter=gather.targets(s))
        bank.update_from_peers(gather, s)
        return means

    tails = [GraphedCall(lambda s=s: tail_fn(s), warmup=1) for s in range(slots)]
    bank.protos.zero_()
    bank.num.zero_()
    tail = torch.cuda.Stream(dev)
    ev_main = [torch.cuda.Event() for _ in range(slots)]
    ev_tail = [torch.cuda.Event() for _ in range(slots)]
    cur = torch.cuda.current_stream(dev)
    torch.cuda.synchronize()
    dist.barrier()
    for step in range(steps):
        s = step % slots
        if step >= slots:
            cur.wait_event(ev_tail[s])
        for d, t in zip(bufs[s], _inputs(step, rank, n, c, k, h, w, dev)):
            d.copy_(t)
        mains[s].replay()
        ev_main[s].record(cur)
        with torch.cuda.stream(tail):
            tail.wait_event(ev_main[s])
            tails[s].replay()
            ev_tail[s].record(tail)
    cur.wait_stream(tail)
    torch.cuda.synchronize()
    torch.save({"protos": bank.protos.cpu(), "num": bank.num.cpu()}, os.path.join(out_dir, "rank%d.pt" % rank))
    dist.barrier()
    dist.destroy_process_group()


def _worker(rank, ws, port, steps, out_dir):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(ws))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=ws, device_id=dev)
    from b200seg import ops, sharding

    n, c, k, h, w = 3, 13, 64, 17, 29
    bank = sharding.ShardedPrototypes(c, k, dev, ops.UPDATE_MOVING_AVERAGE, True, 0.9)
    gather = sharding.PeerGather(n, c, k, dev)
    for step in range(steps):
        feat, logits = _inputs(step, rank, n, c, k, h, w, dev)
        ops.class_means(feat, logits, c, scatter=gather.targets(step % 2))
        bank.update_from_peers(gather, step % 2)
    torch.cuda.synchronize()
    torch.save({"protos": bank.protos.cpu(), "num": bank.num.cpu()}, os.path.join(out_dir, "rank%d.pt" % rank))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("worker", [_worker, _worker_graphed], ids=["eager", "graphs_tail_stream"])
def test_peer_exchange_equals_sequential_update(tmp_path, worker):
    from b200seg import ops

    ws, steps = 2, 5
    mp.spawn(worker, args=(ws, _free_port(), steps, str(tmp_path)), nprocs=ws, join=True)
    got = [torch.load(os.path.join(str(tmp_path), "rank%d.pt" % r)) for r in range(ws)]
    # single-process reference: the same kernels over the concatenated batch, rank order = image order
    n, c, k, h, w = 3, 13, 64, 17, 29
    dev = torch.device("cuda", 0)
    protos, num = torch.zeros(c, k, device=dev), torch.zeros(c, device=dev)
    for step in range(steps):
        parts = [_inputs(step, r, n, c, k, h, w, dev) for r in range(ws)]
        feat, logits = torch.cat([p[0] for p in parts]), torch.cat([p[1] for p in parts])
        means = ops.class_means(feat, logits, c)
        ops.prototype_update(protos, num, means, ops.UPDATE_MOVING_AVERAGE, True, 0.9)
    assert torch.equal(got[0]["protos"], got[1]["protos"]) and torch.equal(got[0]["num"], got[1]["num"]), "ranks disagree"
    for r in range(ws):
        torch.testing.assert_close(got[r]["protos"], protos.cpu(), rtol=1e-5, atol=1e-7)
        assert torch.equal(got[r]["num"], num.cpu())
