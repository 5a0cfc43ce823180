"""Image-sharded form of the label path: one process per GPU, rank r owns a
contiguous range of the global image order, and only three tiny tensors ever cross NVLink --

  * the [C,C] int64 confusion matrix          all_reduce(SUM), exact and order-free;
  * (sum loss, valid pixels)                  all_reduce(SUM), divided afterwards;
  * the prototype update                      all_gather of per-class contribution counts [W,C] int32, then
                                              all_reduce(SUM) of the closed-form weighted partial [C,K+2].

The reference's update_objective_SingleVector (cal_prototype.py:153-169, self_training.py:211-227) is
sequential in the global (image, class) order but linear in the per-(image, class) means, so each rank can
weight its own means with the closed-form coefficient of their global position (computed in fp64 inside
tss_proto_update_partial) and one all-reduce reproduces the sequential result to fp32 rounding.

Pseudo-labels, distances, weights, rectified labels and per-pixel losses are per-image independent: no
data-path collective exists for them.  torch.distributed is plumbing (NCCL on the GPUs, gloo in CPU tests).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.distributed as dist

__all__ = ["shard_range", "world", "all_reduce_confusion", "all_reduce_loss", "prototype_positions",
           "apply_prototype_partial", "ShardedPrototypes", "PeerGather"]


def world(group=None) -> Tuple[int, int]:
    """(rank, world_size); (0, 1) when torch.distributed is not initialised."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def shard_range(total: int, rank: int, world_size: int) -> Tuple[int, int]:
    """Contiguous [begin, end) of rank's images; the first total % W ranks hold one more."""
    if not 0 <= rank < world_size:
        raise ValueError("rank %d outside [0, %d)" % (rank, world_size))
    base, extra = divmod(int(total), int(world_size))
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def all_reduce_confusion(hist: torch.Tensor, bad: Optional[torch.Tensor] = None, group=None) -> torch.Tensor:
    """Sum the per-rank [C,C] int64 matrices (and the out-of-range counter) in place."""
    if world(group)[1] > 1:
        if bad is not None:
            packed = torch.cat([hist.reshape(-1), bad.reshape(-1)])
            dist.all_reduce(packed, op=dist.ReduceOp.SUM, group=group)
            hist.copy_(packed[:-bad.numel()].reshape(hist.shape))
            bad.copy_(packed[-bad.numel():].reshape(bad.shape))
        else:
            dist.all_reduce(hist, op=dist.ReduceOp.SUM, group=group)
    return hist


def all_reduce_loss(loss_sum: torch.Tensor, valid: torch.Tensor, eps: float = 0.0, group=None) -> torch.Tensor:
    """Global mean of a per-pixel loss from per-rank (sum, valid-count): CE uses eps=0, RCE eps=1e-6
    (self_training.py:173)."""
    packed = torch.stack([loss_sum.double().reshape(()), valid.double().reshape(())])
    if world(group)[1] > 1:
        dist.all_reduce(packed, op=dist.ReduceOp.SUM, group=group)
    return (packed[0] / (packed[1] + eps)).to(torch.float32)


def prototype_positions(local_counts: torch.Tensor, group=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """local_counts [C] int32 = how many (image, class) vectors of THIS rank's batch contribute per class.
    Returns (first_index [C], total [C]) int32: where this rank's contributions start in the global order of
    this step (ranks hold consecutive image ranges, so rank order = image order) and the global count."""
    rank, ws = world(group)
    counts = local_counts.to(torch.int32).contiguous()
    if ws == 1:
        return torch.zeros_like(counts), counts.clone()
    gathered = [torch.empty_like(counts) for _ in range(ws)]
    dist.all_gather(gathered, counts, group=group)
    table = torch.stack(gathered)                       # [W, C]
    first = table[:rank].sum(dim=0, dtype=torch.int32) if rank > 0 else torch.zeros_like(counts)
    return first.contiguous(), table.sum(dim=0, dtype=torch.int32).contiguous()


def apply_prototype_partial(protos: torch.Tensor, num: torch.Tensor, partial: torch.Tensor, scale0: torch.Tensor,
                            num_after: torch.Tensor, group=None) -> None:
    """protos = scale0 * protos + sum_ranks(partial); num = num_after (identical on every rank).  In place."""
    if world(group)[1] > 1:
        dist.all_reduce(partial, op=dist.ReduceOp.SUM, group=group)
    protos.mul_(scale0.reshape(-1, 1)).add_(partial)
    num.copy_(num_after)


class ShardedPrototypes:
    """A prototype bank replicated on every rank and updated from image-sharded batches.  Rank r's batch
    must hold the images that precede rank r+1's in the reference's order (shard_range)."""

    def __init__(self, num_classes: int, dim: int, device, mode: int, start_mean: bool, momentum: float,
                 cap: float = 3000.0, group=None):
        self.protos = torch.zeros(num_classes, dim, device=device)
        self.num = torch.zeros(num_classes, device=device)
        self.mode, self.start_mean, self.momentum, self.cap, self.group = mode, start_mean, momentum, cap, group
        self._comm = None          # side stream of update_overlapped

    def update(self, means) -> None:
        """means: ops.ClassMeans of this rank's batch.  Two kernels + (all_gather, all_reduce) per step."""
        from . import ops

        counts = ops.prototype_update_partial(means, self.num, self.mode, self.start_mean, self.momentum, None, None,
                                              self.cap)
        first, total = prototype_positions(counts, self.group)
        partial, scale0, num_after = ops.prototype_update_partial(means, self.num, self.mode, self.start_mean,
                                                                  self.momentum, first, total, self.cap)
        apply_prototype_partial(self.protos, self.num, partial, scale0, num_after, self.group)

    def update_gathered(self, means) -> None:
        """Exact form for equal per-rank batches: ONE all_gather of the per-(image, class) mean vectors in rank
        order = global image order, then the SAME sequential update kernel as the single-GPU path over the
        gathered list.  Vectors that fail the validity rules are all-zero, and the kernel re-evaluates the
        reference's `vector.sum() == 0` skip itself, so no flags travel.  Every rank ends with the
        bit-identical bank the reference's loop would produce; one small collective (n*C*K*4 bytes per rank:
        213 KB at 16 x 13 x 256) and one launch per step."""
        from . import ops

        rank, ws = world(self.group)
        if ws == 1:
            ops.prototype_update(self.protos, self.num, means, self.mode, self.start_mean, self.momentum, self.cap)
            return
        n, c, k = means.vectors.shape
        if getattr(self, "_gkey", None) != (n, c, k):
            self._gv = torch.empty((ws * n, c, k), dtype=torch.float32, device=self.protos.device)
            self._gkey = (n, c, k)
        dist.all_gather_into_tensor(self._gv, means.vectors, group=self.group)
        gathered = ops.ClassMeans(self._gv, None, None, None, None, None, None)
        ops.prototype_update(self.protos, self.num, gathered, self.mode, self.start_mean, self.momentum, self.cap)

    def update_from_peers(self, gather: "PeerGather", slot: int) -> None:
        """The exchange already happened inside the finalize kernel (ops.class_means(scatter=gather.targets(slot))):
        wait for every rank's stores, then the sequential update over the gathered list -- the same kernel and the
        same bit-identical bank as update_gathered."""
        from . import ops

        gather.barrier()
        vec, vsum = gather.gathered(slot)
        gathered = ops.ClassMeans(vec, None, vsum, None, None, None, None)
        ops.prototype_update(self.protos, self.num, gathered, self.mode, self.start_mean, self.momentum, self.cap)

    def update_overlapped(self, means, gathered: bool = True):
        """Same as update(), on a side stream: the exchange is two tiny kernels and two latency-bound
        collectives, and nothing on the compute stream reads the bank until the NEXT exchange, so it overlaps
        the following step's label kernels.  ``means`` must stay untouched until the returned event has
        completed (wait on it before overwriting those tensors); call join() before reading protos / num."""
        cur = torch.cuda.current_stream(self.protos.device)
        if self._comm is None:
            self._comm = torch.cuda.Stream(self.protos.device)
        ready = torch.cuda.Event()
        ready.record(cur)
        with torch.cuda.stream(self._comm):
            self._comm.wait_event(ready)
            if gathered:
                self.update_gathered(means)
            else:
                self.update(means)
            done = torch.cuda.Event()
            done.record(self._comm)
        return done

    def join(self) -> None:
        """Make the current stream wait for every overlapped exchange issued so far."""
        if self._comm is not None:
            torch.cuda.current_stream(self.protos.device).wait_stream(self._comm)


class PeerGather:
    """Gather buffer for the per-(image, class) mean vectors in peer-mapped (symmetric) memory: every rank's
    finalize kernel stores its vectors straight into ALL ranks' buffers over NVLink (ops.class_means(scatter=...)),
    one barrier follows, and each rank runs the sequential update kernel over its own complete copy
    [ranks*N, C, K] -- rank order = the reference's global image order.  No NCCL kernel, no staging copy, and the
    stores, the barrier and the update are ordinary stream work, so the whole step (exchange included) replays
    as ONE CUDA graph.  Two slots alternate: a rank can only start writing slot s again after every rank has
    passed the barrier of the step in between, i.e. has finished reading slot s.

    Needs torch.distributed._symmetric_memory (CUDA peer mapping inside one node); ShardedPrototypes.update_gathered
    is the NCCL form of the same exchange."""

    SLOTS = 2

    def __init__(self, n_img: int, num_classes: int, dim: int, device, group=None):
        import torch.distributed._symmetric_memory as symm

        self.group = group if group is not None else dist.group.WORLD
        self.rank, self.ws = world(group)
        if self.ws > 8:
            raise ValueError("PeerGather: at most 8 ranks (one NVLink domain), got %d" % self.ws)
        self.shape = (int(n_img), int(num_classes), int(dim))
        self.vec_elems = self.ws * n_img * num_classes * dim       # per slot: [ranks,N,C,K] vectors, then [ranks,N,C] sums
        self.slot_elems = self.vec_elems + self.ws * n_img * num_classes
        self.buf = symm.empty(self.SLOTS * self.slot_elems, dtype=torch.float32, device=device)
        self.hdl = symm.rendezvous(self.buf, self.group)
        self.buf.zero_()
        self._ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        self._step = 0
        torch.cuda.synchronize(device)
        self.hdl.barrier(channel=0)

    def targets(self, slot: int):
        """scatter= argument of ops.class_means: this rank's [N,C,K] / [N,C] slices of the local buffer and the
        device addresses of the same slices on every peer."""
        n, c, k = self.shape
        base = slot * self.slot_elems
        ov = base + self.rank * n * c * k
        os_ = base + self.vec_elems + self.rank * n * c
        local_v = self.buf[ov:ov + n * c * k].view(n, c, k)
        local_s = self.buf[os_:os_ + n * c].view(n, c)
        peers = [p + 4 * ov for r, p in enumerate(self._ptrs) if r != self.rank]
        peer_sums = [p + 4 * os_ for r, p in enumerate(self._ptrs) if r != self.rank]
        return local_v, local_s, peers, peer_sums

    def gathered(self, slot: int):
        """([ranks*N, C, K] mean vectors, [ranks*N, C] sums) of every rank, in rank = global image order."""
        n, c, k = self.shape
        base = slot * self.slot_elems
        vec = self.buf[base:base + self.vec_elems].view(self.ws * n, c, k)
        sm = self.buf[base + self.vec_elems:base + self.slot_elems].view(self.ws * n, c)
        return vec, sm

    def barrier(self) -> None:
        """All ranks' stores of this step have landed (stream-ordered on every rank)."""
        self.hdl.barrier(channel=0)
