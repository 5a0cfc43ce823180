"""Host-fed label pipeline: the stage loop of generate_pseudo_label.py:60-96 + cal_prototype.py:63-82 +
segmentation_evaluate.py:38-44,77-79 for batches that live in HOST memory.

The reference moves every batch H2D, runs the label math, and copies every prediction back before the next
batch starts.  Here six streams overlap the legs of consecutive batches: batch i+1 is uploaded (copy engine,
pinned buffers) while batch i runs through the kernels and batch i-1's labels drain to pinned host memory, with
two device-side input slots and two host-side output slots; the label kernel of a batch (it needs only the logits)
starts as soon as they are up, while the features still upload and the previous batch's kernels run.  Inside a batch the XU-bound pseudo-label kernel and
the HBM-bound accumulate kernel share the SMs (two streams, co-resident CTAs), and the latency-bound end of the
prototype stage -- means, the exchange between GPUs when the batch is image-sharded, the sequential update -- runs
on a tail stream beside the next batch.  Per batch the PCIe transfer (4*(C+K)*h*w bytes in per frame) is the
bound, so the kernels hide completely behind it.
"""
from __future__ import annotations

from typing import Optional, Sequence

import torch

from . import dropin, ops


class HostLabelPipeline:
    """submit(logits, feat, gt) per host batch (pinned fp32 / uint8 tensors); results of batch i are in
    ``out_label[i % 2]`` / ``out_conf[i % 2]`` (pinned) once ``wait(i)`` returns.  Prototypes accumulate in
    ``features`` (dropin.Class_Features) -- or, image-sharded over several GPUs, in ``bank``
    (sharding.ShardedPrototypes) through ``gather`` (sharding.PeerGather: the finalize kernel's own NVLink stores,
    one barrier, the sequential update over every rank's vectors) -- and the confusion matrix in ``meter``."""

    SLOTS = 2

    def __init__(self, device, num_classes: int, dim: int, size: Sequence[int], thr=0.3, keep=(5, 11),
                 ignore_index: int = 12, update_prototypes: bool = True, overlap: bool = True, bank=None, gather=None):
        self.dev = torch.device(device)
        self.C, self.K, self.size = int(num_classes), int(dim), (int(size[0]), int(size[1]))
        self.thr, self.keep, self.ignore = thr, keep, int(ignore_index)
        self.update_prototypes = update_prototypes
        self.overlap = overlap
        self.bank, self.gather = bank, gather
        if (bank is None) != (gather is None):
            raise ValueError("bank and gather go together (image-sharded prototype exchange)")
        self.features = dropin.Class_Features(self.dev, self.C, self.K)
        self.meter = dropin.ConfusionMeter(self.C, self.dev)
        self.s_in, self.s_out = torch.cuda.Stream(self.dev), torch.cuda.Stream(self.dev)
        self.s_side, self.s_tail = torch.cuda.Stream(self.dev, priority=-1), torch.cuda.Stream(self.dev)
        self.s_pre = torch.cuda.Stream(self.dev)    # label kernel of a batch: as soon as its logits are up, beside the previous batch
        self._in = [None] * self.SLOTS          # device input slots (allocated at first submit)
        self._ready = [torch.cuda.Event() for _ in range(self.SLOTS)]
        self._free = [torch.cuda.Event() for _ in range(self.SLOTS)]
        self._drained = [torch.cuda.Event() for _ in range(self.SLOTS)]
        self.out_label = [None] * self.SLOTS
        self.out_conf = [None] * self.SLOTS
        self._count = 0
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    def _slot(self, i, logits, feat, gt):
        s = i % self.SLOTS
        if self._in[s] is None:
            # allocated under the upload stream: the caching allocator then never hands these blocks to work that is
            # still pending on the compute stream
            with torch.cuda.stream(self.s_in):
                self._in[s] = tuple(None if t is None else torch.empty_like(t, device=self.dev) for t in (logits, feat, gt))
            n = logits.shape[0]
            self.out_label[s] = torch.empty((n,) + self.size, dtype=torch.uint8).pin_memory()
            self.out_conf[s] = torch.empty((n,) + self.size, dtype=torch.float16).pin_memory()
        return s

    def submit(self, logits: torch.Tensor, feat: torch.Tensor, gt: Optional[torch.Tensor] = None) -> int:
        i = self._count
        s = self._slot(i, logits, feat, gt)
        cur = torch.cuda.current_stream(self.dev)
        d_logits, d_feat, d_gt = self._in[s]
        with torch.cuda.stream(self.s_in):
            if i >= self.SLOTS:
                self.s_in.wait_event(self._free[s])       # batch i-2 no longer reads this slot
            d_logits.copy_(logits, non_blocking=True)
            logits_up = torch.cuda.Event()
            logits_up.record(self.s_in)
            d_feat.copy_(feat, non_blocking=True)
            if gt is not None:
                d_gt.copy_(gt, non_blocking=True)
            self._ready[s].record(self.s_in)
        cur.wait_event(self._ready[s])
        if self.overlap:
            # the label kernel needs only the logits (16 MB of the 356 MB of a cfg2 batch): it runs on its own stream while the
            # features are still uploading and the previous batch's kernels run
            with torch.cuda.stream(self.s_pre):
                self.s_pre.wait_event(logits_up)
                cs = ops.class_labels(d_logits, self.C, self.K, overflow_channel=self.C)
                labelled = torch.cuda.Event()
                labelled.record(self.s_pre)
            for t in (cs.sums, cs.records, cs.seg_label):
                t.record_stream(cur)
                t.record_stream(self.s_side)
            # pseudo-labels (XU / issue bound) and the segmented sums (HBM bound) on the same SMs at the same time
            self.s_side.wait_event(labelled)
            self.s_side.wait_event(self._ready[s])
            with torch.cuda.stream(self.s_side):
                ops.class_accumulate(cs, d_feat, ops.CORESIDENT_CONSUMERS)
            pl = ops.pseudo_label(d_logits, size=self.size, thr=self.thr, keep=self.keep, ignore_index=self.ignore,
                                  flags=ops.CORESIDENT_PL_FLAGS)
            if gt is not None:
                self.meter.add(pl.label, d_gt)
            cur.wait_stream(self.s_side)
        else:
            pl = dropin.pseudo_label_batch(d_logits, size=self.size, thr=self.thr, keep=self.keep, ignore_index=self.ignore)
            cs = ops.class_sums(d_feat, d_logits, self.C, overflow_channel=self.C)
            if gt is not None:
                self.meter.add(pl.label, d_gt)
        self._free[s].record(cur)
        # the latency-bound end of the prototype stage: beside the next batch's kernels
        self.s_tail.wait_stream(cur)
        with torch.cuda.stream(self.s_tail):
            if self.gather is not None:
                slot = i % self.gather.SLOTS
                ops.class_finalize(cs, scatter=self.gather.targets(slot))
                self.bank.update_from_peers(self.gather, slot)
            elif self.update_prototypes:
                f = self.features
                ops.class_finalize(cs, update=(f.objective_vectors, f.objective_vectors_num, ops.UPDATE_MEAN, True,
                                               f.proto_momentum))
            else:
                ops.class_finalize(cs)
        for t in (cs.sums, cs.records):
            t.record_stream(self.s_tail)
        with torch.cuda.stream(self.s_out):
            self.s_out.wait_event(self._free[s])
            self.out_label[s].copy_(pl.label_thr, non_blocking=True)
            self.out_conf[s].copy_(pl.conf16, non_blocking=True)
            pl.label_thr.record_stream(self.s_out)
            pl.conf16.record_stream(self.s_out)
            self._drained[s].record(self.s_out)
        self.h2d_bytes = sum(t.numel() * t.element_size() for t in (logits, feat, gt) if t is not None)
        self.d2h_bytes = self.out_label[s].numel() + 2 * self.out_conf[s].numel()
        self._count += 1
        return i

    def wait(self, i: int) -> None:
        """Block the host until batch i's labels are in out_label / out_conf."""
        self._drained[i % self.SLOTS].synchronize()

    def join(self) -> None:
        """Make the current stream wait for everything submitted so far (no host block)."""
        cur = torch.cuda.current_stream(self.dev)
        for s in range(min(self._count, self.SLOTS)):
            cur.wait_event(self._drained[s])
        cur.wait_stream(self.s_tail)
