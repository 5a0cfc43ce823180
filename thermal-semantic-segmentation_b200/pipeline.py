"""Host-fed label pipeline: the stage loop of generate_pseudo_label.py:60-96 + cal_prototype.py:63-82 +
segmentation_evaluate.py:38-44,77-79 for batches that live in HOST memory.

The reference moves every batch H2D, runs the label math, and copies every prediction back before the next
batch starts.  Here three streams overlap the three legs: batch i+1 is uploaded (copy engine, pinned
buffers) while batch i runs through the kernels and batch i-1's labels drain to pinned host memory, with
two device-side input slots and two host-side output slots.  Per batch the PCIe transfer (4*(C+K)*h*w
bytes in per frame) is the bound, so the kernels hide completely behind it.
"""
from __future__ import annotations

from typing import Optional, Sequence

import torch

from . import dropin, ops


class HostLabelPipeline:
    """submit(logits, feat, gt) per host batch (pinned fp32 / uint8 tensors); results of batch i are in
    ``out_label[i % 2]`` / ``out_conf[i % 2]`` (pinned) once ``wait(i)`` returns.  Prototypes accumulate in
    ``features`` (dropin.Class_Features), the confusion matrix in ``meter`` (dropin.ConfusionMeter)."""

    SLOTS = 2

    def __init__(self, device, num_classes: int, dim: int, size: Sequence[int], thr=0.3, keep=(5, 11),
                 ignore_index: int = 12, update_prototypes: bool = True):
        self.dev = torch.device(device)
        self.C, self.K, self.size = int(num_classes), int(dim), (int(size[0]), int(size[1]))
        self.thr, self.keep, self.ignore = thr, keep, int(ignore_index)
        self.update_prototypes = update_prototypes
        self.features = dropin.Class_Features(self.dev, self.C, self.K)
        self.meter = dropin.ConfusionMeter(self.C, self.dev)
        self.s_in, self.s_out = torch.cuda.Stream(self.dev), torch.cuda.Stream(self.dev)
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
            self._in[s] = tuple(torch.empty_like(t, device=self.dev) for t in (logits, feat, gt))
            n = logits.shape[0]
            self.out_label[s] = torch.empty((n,) + self.size, dtype=torch.uint8).pin_memory()
            self.out_conf[s] = torch.empty((n,) + self.size, dtype=torch.float16).pin_memory()
        return s

    def submit(self, logits: torch.Tensor, feat: torch.Tensor, gt: Optional[torch.Tensor]) -> int:
        i = self._count
        s = self._slot(i, logits, feat, gt)
        cur = torch.cuda.current_stream(self.dev)
        d_logits, d_feat, d_gt = self._in[s]
        with torch.cuda.stream(self.s_in):
            if i >= self.SLOTS:
                self.s_in.wait_event(self._free[s])       # batch i-2 no longer reads this slot
            d_logits.copy_(logits, non_blocking=True)
            d_feat.copy_(feat, non_blocking=True)
            d_gt.copy_(gt, non_blocking=True)
            self._ready[s].record(self.s_in)
        cur.wait_event(self._ready[s])
        pl = dropin.pseudo_label_batch(d_logits, size=self.size, thr=self.thr, keep=self.keep, ignore_index=self.ignore)
        if self.update_prototypes:
            self.features.update_from_batch(d_feat, d_logits)
        else:
            self.features.class_means(d_feat, d_logits)
        self.meter.add(pl.label, d_gt)
        self._free[s].record(cur)
        with torch.cuda.stream(self.s_out):
            self.s_out.wait_event(self._free[s])
            if i >= self.SLOTS:
                pass  # the host slot is reused: the caller has consumed batch i-2 after wait(i-2)
            self.out_label[s].copy_(pl.label_thr, non_blocking=True)
            self.out_conf[s].copy_(pl.conf16, non_blocking=True)
            pl.label_thr.record_stream(self.s_out)
            pl.conf16.record_stream(self.s_out)
            self._drained[s].record(self.s_out)
        self.h2d_bytes = sum(t.numel() * t.element_size() for t in (logits, feat, gt))
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
