"""Host-side mirror of the reference's Python entry points for the label path.

Same names, argument meaning and error behaviour as the reference callables, so that the reference
scripts run unmodified once these are bound in their place (``install()``).  Every piece
of arithmetic goes through the C ABI (``ops``) to the sm_100a kernels; tensors must live on a CUDA
device and there is no CPU path.

    cal_prototype.Class_Features              -> Class_Features
    self_training.SelfTrain (label methods)   -> SelfTrain
    generate_pseudo_label.generate_pl :60-96  -> generate_pl (pseudo_label_batch / soft_label_batch + formats)
    data/freiburg_dataset.py:115-125          -> threshold_block, prepare_freiburg_labels (Freiburg ctor hook)
    utils.eval_tools._fast_hist / evaluate    -> _fast_hist / evaluate (+ ConfusionMeter, on-device)
    utils.loss.SemanticConsistency / FocalLoss-> SemanticConsistency / FocalLoss
    segmentation_evaluate.seg_validate :24-80 -> seg_validate (validate_batch per batch)

Label range checks: torch raises (CPU) or device-asserts (CUDA) on targets outside [0, C); the kernels count such
pixels instead, which needs no host sync.  With ``STRICT_LABELS`` (env ``B200SEG_STRICT=1``) every loss call reads
the counter back and raises like the reference; otherwise only ``SelfTrain.rce`` (whose reference form raises on
the host, in ``one_hot``) does.  INTEGRATION.md section 1 lists the remaining deliberate differences.
"""
from __future__ import annotations

import copy
import os
from typing import Iterable, Optional, Sequence

import numpy as np
import torch

from . import ops

__all__ = ["Class_Features", "SelfTrain", "pseudo_label_batch", "soft_label_batch", "threshold_block",
           "_fast_hist", "evaluate", "metrics_from_hist", "ConfusionMeter", "validate_batch", "seg_validate",
           "generate_pl", "prepare_freiburg_labels", "SemanticConsistency", "FocalLoss", "install", "STRICT_LABELS"]

STRICT_LABELS = os.environ.get("B200SEG_STRICT", "0") not in ("", "0")


def _raise_if_bad_labels(bad: torch.Tensor, what: str) -> None:
    """One device->host read of the kernel's out-of-range label counter (strict mode only)."""
    n = int(bad.item())
    if n:
        raise IndexError("%s: %d target values are out of bounds (torch's CrossEntropyLoss raises here)" % (what, n))


# ------------------------------------------------------------------------------------------------
# cal_prototype.py:85-169
class Class_Features:
    """Per-class prototype bank (cal_prototype.py:85-169).  State lives on ``device`` (the reference
    keeps it on the CPU and copies every vector over, :75; the values are the same)."""

    def __init__(self, device, numbers: int = 13, dim: int = 256):
        self.class_numbers = numbers
        self.device = torch.device(device)
        self.num = np.zeros(numbers)
        self.objective_vectors = torch.zeros([numbers, dim], device=self.device)
        self.objective_vectors_num = torch.zeros([numbers], device=self.device)
        self.proto_momentum = 0.9999

    # -- batched fast path: no host round trip at all
    def class_means(self, feat_cls, outputs, labels_val=None) -> ops.ClassMeans:
        return ops.class_means(feat_cls, outputs, self.class_numbers, labels_val=labels_val,
                               overflow_channel=self.class_numbers)

    def update_from_batch(self, feat_cls, outputs, labels_val=None, name: str = "mean", start_mean: bool = True):
        """calculate_mean_vector + the update loop of cal_prototype.py:72-75 for one batch, on device."""
        return ops.class_means(feat_cls, outputs, self.class_numbers, labels_val=labels_val,
                               overflow_channel=self.class_numbers,
                               update=(self.objective_vectors, self.objective_vectors_num, _mode(name), start_mean,
                                       self.proto_momentum))

    # -- reference signatures
    def calculate_mean_vector(self, feat_cls, outputs, labels_val=None):
        return _means_to_lists(self.class_means(feat_cls, outputs, labels_val))

    def calculate_mean_vector_by_output(self, feat_cls, outputs, model):
        return _means_to_lists(ops.class_means(feat_cls, outputs, self.class_numbers,
                                               overflow_channel=_overflow_of(model, self.class_numbers)))

    def process_label(self, label):
        return ops.one_hot_overflow(label, self.class_numbers, self.class_numbers)

    def update_objective_SingleVector(self, id, vector, name: str = "moving_average", start_mean: bool = True):
        ops.prototype_update_single(self.objective_vectors, self.objective_vectors_num, id,
                                    vector.to(self.device), _mode(name), start_mean, self.proto_momentum)


def _mode(name: str) -> int:
    if name == "moving_average":
        return ops.UPDATE_MOVING_AVERAGE
    if name == "mean":
        return ops.UPDATE_MEAN
    raise NotImplementedError("no such updating way of objective vectors {}".format(name))


def _overflow_of(model, default: int) -> int:
    args = getattr(model, "args", None)
    return int(getattr(args, "ignore_index", default)) if args is not None else default


def _means_to_lists(m: ops.ClassMeans):
    """(vectors, ids) python lists in the reference's order: image-major, then class; one D2H of N*C bytes."""
    valid = m.valid.cpu().numpy().astype(bool)
    vectors, ids = [], []
    for n, t in zip(*np.nonzero(valid)):
        vectors.append(m.vectors[n, t].reshape(-1, 1, 1))
        ids.append(int(t))
    return vectors, ids


# ------------------------------------------------------------------------------------------------
# generate_pseudo_label.py:74-85, :91-94
def soft_label_batch(out: torch.Tensor) -> torch.Tensor:
    """args.soft branch (:75): low-resolution class probabilities [N,C,h,w]."""
    return ops.softmax_nchw(out)


def pseudo_label_batch(out: torch.Tensor, flip_out: Optional[torch.Tensor] = None, size: Optional[Sequence[int]] = None,
                       flip_dim: int = 1, thr=None, keep=None, ignore_index: int = 255, want_conf32: bool = False):
    """Hard branch of generate_pl.  Without ``flip_out``: Softmax(dim=1)(out).max(1) (:85) at the logits'
    resolution (or, with ``size``, after bilinear upsampling of the logits -- the fused headline path).
    With ``flip_out``: the flip branch as committed (:81-85) -- softmax, upsample, average with the flipped
    second pass (``flip_dim=1`` is what ``torch.fliplr`` does to NCHW, 3 is the width mirror), a second
    softmax, then max.  Returns ops.PseudoLabels (uint8 label, fp16 confidence as stored at :91,:94, and the
    thresholded label of data/freiburg_dataset.py:115-123 when ``thr``/``keep`` are given; the fp32 confidence
    before the storage cast only with ``want_conf32``)."""
    if flip_out is None:
        return ops.pseudo_label(out, size=size, mode=ops.PL_UP_SOFTMAX, thr=thr, keep=keep, ignore_index=ignore_index,
                                want_conf32=want_conf32)
    if size is None:
        raise ValueError("the flip branch upsamples to the image size: pass size=(H, W)")
    if flip_dim not in (1, 3):
        raise ValueError("flip_dim must be 1 (torch.fliplr on NCHW) or 3 (width)")
    return ops.pseudo_label(ops.softmax_nchw(out), size=size, mode=ops.PL_PROB_UP_SOFTMAX,
                            src_flip=ops.softmax_nchw(flip_out),
                            flip=ops.FLIP_CLASS if flip_dim == 1 else ops.FLIP_WIDTH, thr=thr, keep=keep,
                            ignore_index=ignore_index, want_conf32=want_conf32)


def threshold_block(label: torch.Tensor, conf16: Optional[torch.Tensor], num_classes: int, threshold=0.3,
                    ignore_index: int = 12, keep: Optional[Iterable[int]] = (5, 11),
                    size_wh: Optional[Sequence[int]] = None) -> torch.Tensor:
    """data/freiburg_dataset.py:115-125 on device: label[conf <= thr] = ignore (fp16 compare), class
    whitelist, then PIL NEAREST resize to ``size_wh`` = (W, H) when given."""
    out = ops.threshold_labels(label, conf16, num_classes, thr=threshold if conf16 is not None else None, keep=keep,
                               ignore_index=ignore_index)
    if size_wh is not None:
        out = ops.resize_nearest_u8(out, (int(size_wh[1]), int(size_wh[0])), rule="pil")
    return out


# ------------------------------------------------------------------------------------------------
# utils/eval_tools.py
FREIBURG_NAMES = ["IoU road,parking", "IoU ground,sidewalk", "IoU building,", "IoU curb", "IoU fence",
                  "IoU pole,traffic light,traffic sign", "IoU vegetation", "IoU terrain", "IoU sky",
                  "IoU person,rider", "IoU car,truck,bus,train", "IoU motorcycle,bicycle"]
CITYSCAPES_NAMES = ["IoU road", "IoU sidewalk", "IoU building,", "IoU wall", "IoU fence", "IoU pole",
                    "IoU traffic light", "IoU traffic sign", "IoU vegetation", "IoU terrain", "IoU sky",
                    "IoU person", "IoU rider", "IoU car", "IoU truck", "IoU bus", "IoU train",
                    "IoU motorcycle", "IoU bicycle"]


def _as_device_labels(a, device) -> torch.Tensor:
    """Label maps arrive as numpy arrays (seg_validate hands over .cpu().numpy() int64) or tensors."""
    if isinstance(a, torch.Tensor):
        t = a
    else:
        arr = np.ascontiguousarray(a)
        if arr.dtype not in (np.uint8, np.int32, np.int64):
            arr = arr.astype(np.int64)
        t = torch.from_numpy(arr)
    return t.to(device, non_blocking=True) if t.device != device else t


class ConfusionMeter:
    """Running [C,C] confusion matrix on device (rows = ground truth), the accumulation of
    evaluate() :49-51 without the D2H of every prediction."""

    def __init__(self, num_classes: int, device="cuda"):
        self.num_classes = num_classes
        self.device = torch.device(device)
        self.hist = torch.zeros(num_classes, num_classes, dtype=torch.int64, device=self.device)
        self.bad = torch.zeros(1, dtype=torch.int64, device=self.device)

    def add(self, pred, gt) -> None:
        ops.confusion_matrix(_as_device_labels(pred, self.device).reshape(-1),
                             _as_device_labels(gt, self.device).reshape(-1), self.num_classes, self.hist, self.bad)

    def result(self) -> np.ndarray:
        """Host int64 matrix.  Raises where np.bincount/reshape would (a counted pixel with pred >= C or < 0)."""
        both = torch.cat([self.hist.reshape(-1), self.bad]).cpu().numpy()
        if both[-1] != 0:
            raise ValueError("cannot reshape array of size > %d into shape (%d,%d): %d predictions outside [0, %d)"
                             % (self.num_classes ** 2, self.num_classes, self.num_classes, both[-1], self.num_classes))
        return both[:-1].reshape(self.num_classes, self.num_classes)


def _fast_hist(label_pred, label_true, num_classes, device="cuda"):
    """utils/eval_tools.py:4-9 -> ndarray [C,C] int64."""
    m = ConfusionMeter(num_classes, device)
    m.add(label_pred, label_true)
    return m.result()


def metrics_from_hist(hist, num_classes: int):
    """utils/eval_tools.py:53-67 -- float64 arithmetic on <= 361 numbers stays on the host in numpy, the
    same expressions in the same order, so the metrics are bit-identical given an identical matrix."""
    hist = np.asarray(hist, dtype=np.float64)
    acc = np.diag(hist).sum() / hist.sum()
    acc_cls = np.diag(hist) / hist.sum(axis=1)
    acc_cls = np.nanmean(acc_cls)
    iu = np.diag(hist) / (hist.sum(axis=1) + hist.sum(axis=0) - np.diag(hist))
    if num_classes == 13:
        mean_iu = np.nanmean(iu[:11])
        cls_iu = dict(zip(FREIBURG_NAMES, iu))
    elif num_classes == 19:
        mean_iu = np.nanmean(iu)
        cls_iu = dict(zip(CITYSCAPES_NAMES, iu))
    else:
        raise ValueError("invalid dataset for evaluation.")
    freq = hist.sum(axis=1) / hist.sum()
    fwavacc = (freq[freq > 0] * iu[freq > 0]).sum()
    return acc, acc_cls, mean_iu, fwavacc, cls_iu


def evaluate(predictions, gts, num_classes, device="cuda"):
    """utils/eval_tools.py:12-67.  Counts are exact integers (< 2**53), so the float64 running sum of the
    reference and one int64 matrix converted at the end hold the same values."""
    if num_classes not in (13, 19):
        raise ValueError("invalid dataset for evaluation.")
    m = ConfusionMeter(num_classes, device)
    if isinstance(predictions, (np.ndarray, torch.Tensor)) and isinstance(gts, (np.ndarray, torch.Tensor)) \
            and len(predictions) == len(gts) and _numel(predictions) == _numel(gts):
        # one stacked array each -- what seg_validate passes: predictions [M,H,W], labels [M,H,W] or (after ToTensor)
        # [M,1,H,W] (segmentation_evaluate.py:42-44,77-78); image-wise flatten order is the same: a single launch
        m.add(predictions.reshape(-1), gts.reshape(-1))
    else:
        for lp, lt in zip(predictions, gts):
            m.add(lp, lt)
    return metrics_from_hist(m.result(), num_classes)


def _numel(a) -> int:
    return int(a.numel()) if isinstance(a, torch.Tensor) else int(a.size)


def validate_batch(outputs: torch.Tensor, label: torch.Tensor, meter: ConfusionMeter, size=(256, 512),
                   ignore_index: Optional[int] = None):
    """seg_validate's per-batch work (segmentation_evaluate.py:38-44) without leaving the device: bilinear
    upsample + argmax fused, histogram accumulated.  Returns (predictions uint8 [N,H,W], CE loss or None)."""
    n = outputs.shape[0]
    if isinstance(label, torch.Tensor) and label.is_cuda and label.dtype == torch.uint8 and label.numel() == n * size[0] * size[1]:
        # uint8 ground truth already on the device: the matrix is accumulated inside the upsample + argmax kernel
        pl = ops.pseudo_label(outputs, size=size, mode=ops.PL_UP_ARGMAX, want_conf16=False, gt=label.contiguous(), hist=meter.hist)
    else:
        pl = ops.pseudo_label(outputs, size=size, mode=ops.PL_UP_ARGMAX, want_conf16=False)
        meter.add(pl.label, label)
    loss = None
    if ignore_index is not None:
        loss = ops.cross_entropy(ops.upsample_bilinear(outputs, size), label, ignore_index)
    return pl.label, loss


def seg_validate(args, sem_net, val_data, loss_func, device, vis=None, num_classes=13):
    """segmentation_evaluate.seg_validate (:24-80) with the per-batch arithmetic on the device: upsample + argmax in
    one kernel, the confusion matrix accumulated where the predictions are (no int64 D2H of every prediction, no
    np.concatenate of the whole validation set), the loss accumulated on the device and read once at the end.
    Returns (mean_iu, average loss, cls_iu) like the reference.  ``loss_func``: a mean-reduced unweighted
    nn.CrossEntropyLoss / dropin._SegLoss runs in the fused CE kernel, anything else is called on the upsampled
    logits as the reference does.  The prediction-saving branch (:55-75) is presentation and is not reproduced."""
    print('validating...')
    sem_net.eval()
    dev = torch.device(device)
    meter = ConfusionMeter(num_classes, dev)
    loss_sum = torch.zeros((), dtype=torch.float64, device=dev)
    count = 0
    ignore = _ce_ignore_index(loss_func)
    for item in val_data:
        image = item['image'].to(dev)
        label = item['label'].to(dev)
        outputs = sem_net(image)['out']
        n = image.size(0)
        if ignore is not None:
            _, loss = validate_batch(outputs, label, meter, size=(256, 512), ignore_index=ignore)
        else:
            validate_batch(outputs, label, meter, size=(256, 512))
            loss = loss_func(ops.upsample_bilinear(outputs, (256, 512)), label)
        loss_sum += loss.detach().double() * n          # AverageMeter.update(loss.item(), n) without the sync (:46)
        count += n
    hist = meter.result()
    acc, acc_cls, mean_iu, fwavacc, cls_iu = metrics_from_hist(hist, num_classes)
    return mean_iu, float(loss_sum.item()) / max(count, 1), cls_iu


def _ce_ignore_index(loss_func) -> Optional[int]:
    """ignore_index when ``loss_func`` is a plain mean cross-entropy the fused kernel reproduces, else None."""
    if isinstance(loss_func, _SegLoss):
        return loss_func.ignore_index
    if isinstance(loss_func, torch.nn.CrossEntropyLoss) and loss_func.reduction == "mean" and loss_func.weight is None \
            and getattr(loss_func, "label_smoothing", 0.0) == 0.0:
        return int(loss_func.ignore_index)
    return None


# ------------------------------------------------------------------------------------------------
# generate_pseudo_label.py:60-96 and data/freiburg_dataset.py:108-125
def generate_pl(net, dataloader, device, args):
    """generate_pseudo_label.generate_pl with the reference's signature and output directory
    (``<root>/pseudo_labels/<pseudo_type>/<checkpoint>``); the arithmetic runs in the fused kernels and the files are
    the reference's formats (formats.generate_pl)."""
    from . import formats

    save_dir = os.path.join(args.root, 'pseudo_labels', args.pseudo_type, args.checkpoint_name.replace('.pth', ''))
    return formats.generate_pl(net, dataloader, device, save_dir, soft=bool(args.soft), flip=bool(getattr(args, "flip", False)))


def prepare_freiburg_labels(args, image_names: Iterable[str], device="cuda", size_wh=(700, 320), batch: int = 64):
    """The threshold block of Freiburg.__getitem__ (data/freiburg_dataset.py:112-125) for a whole pseudo-label
    directory, on the device: every ``<name>.png`` + ``<name>_conf.npy`` under ``args.path_lp`` goes through
    label[conf <= thr] = ignore (fp16 compare), the {5, 11} whitelist and the PIL-NEAREST resize to the 700x320 crop
    the dataset uses, and is written as ``<name>.png`` into ``<path_lp>_b200thr``.  Returns a copy of ``args`` with
    ``path_lp`` pointing there and ``threshold`` cleared, so that the unmodified ``__getitem__`` (which runs in
    DataLoader workers, where CUDA is not available) loads the finished labels and skips its numpy block."""
    from . import formats

    out_dir = os.path.normpath(args.path_lp) + "_b200thr"
    os.makedirs(out_dir, exist_ok=True)
    names = [os.path.basename(n) for n in image_names]
    by_shape = {}
    for name in names:     # label files of one size go through the kernels together
        lab, _ = formats.read_hard_files(args.path_lp, name, with_conf=False)
        by_shape.setdefault(lab.shape, []).append(name)
    for group in by_shape.values():
        for i in range(0, len(group), batch):
            part = group[i:i + batch]
            out = formats.load_hard_batch(args.path_lp, part, args.num_classes if hasattr(args, "num_classes") else 13, device,
                                          threshold=args.threshold, ignore_index=args.ignore_index, keep=(5, 11), size_wh=size_wh)
            host = out.cpu().numpy()
            for k, name in enumerate(part):
                formats.Image.fromarray(host[k]).save(os.path.join(out_dir, name))
    new_args = copy.copy(args)
    new_args.path_lp = out_dir
    new_args.threshold = None
    return new_args


# ------------------------------------------------------------------------------------------------
# utils/loss.py:142-184
class SemanticConsistency(torch.nn.Module):
    """utils/loss.py:142-150: classes listed in ``ignore_index`` are rewritten to 255 IN PLACE in the
    target, then CrossEntropyLoss(ignore_index=255, reduction)."""

    def __init__(self, ignore_index=(), reduction="mean"):
        super().__init__()
        if reduction not in ("mean", "sum"):
            raise ValueError("reduction must be 'mean' or 'sum'")
        self.ignore_index = ignore_index
        self.reduction = reduction

    def forward(self, input: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        for class_idx in self.ignore_index:
            target[target == class_idx] = 255
        return ops.cross_entropy(input, target, 255, reduction=self.reduction)


class FocalLoss(torch.nn.Module):
    """utils/loss.py:153-184: -(1 - pt)^gamma * alpha[y] * log pt with pt detached, no ignore handling."""

    def __init__(self, gamma=0, alpha=None, size_average=True):
        super().__init__()
        self.gamma = gamma
        self.alpha = alpha
        if isinstance(alpha, (float, int)):
            self.alpha = torch.Tensor([alpha, 1 - alpha])
        if isinstance(alpha, list):
            self.alpha = torch.Tensor(alpha)
        self.size_average = size_average

    def forward(self, input, target):
        alpha = self.alpha
        if alpha is not None and (alpha.device != input.device or alpha.dtype != torch.float32):
            alpha = self.alpha = alpha.to(device=input.device, dtype=torch.float32)
        return ops.focal_loss(input, target, float(self.gamma), alpha, self.size_average)


# ------------------------------------------------------------------------------------------------
# self_training.py:17-227
class _SegLoss:
    """nn.CrossEntropyLoss(ignore_index=args.ignore_index, reduction='mean') (self_training.py:27)."""

    def __init__(self, ignore_index: int):
        self.ignore_index = ignore_index

    def __call__(self, input, target):
        loss, bad = ops.cross_entropy_checked(input, target, self.ignore_index)
        if STRICT_LABELS:
            _raise_if_bad_labels(bad, "seg_loss")
        return loss


class SelfTrain:
    """The label mathematics of self_training.SelfTrain with the reference's method names.  ``step`` is the
    reference's step() (:49-134) with the post-network arithmetic replaced by the fused kernels; the two
    network forwards, backward, optimiser and EMA-of-weights lines are the caller's models, untouched."""

    def __init__(self, args, seg_net, ema_net, optimizer_seg, device, logger=None, dim: int = 256):
        self.args = args
        self.seg_net = seg_net
        self.ema_net = ema_net
        self.num_classes = self.args.num_classes
        self.optimizer_seg = optimizer_seg
        self.device = torch.device(device)
        self.objective_vectors = torch.zeros([self.num_classes, dim], device=self.device)
        self.objective_vectors_num = torch.zeros([self.num_classes], device=self.device)
        self.logger = logger
        self.seg_loss = _SegLoss(args.ignore_index)
        self.scale_rate = 4
        self.counter = 0

    # -- A7/A8/A9
    def full2weak(self, feat, target_weak_params):
        return ops.full2weak(feat, target_weak_params)

    def feat_prototype_distance(self, feat):
        return ops.prototype_distance(feat, self._protos(), want_dist=True)[0]

    def get_prototype_weight(self, feat, target_weak_params):
        feat = self.full2weak(feat, target_weak_params)
        return ops.prototype_distance(feat, self._protos(), temperature=self.args.proto_temperature, want_dist=False,
                                      want_weight=True)[1]

    def _protos(self) -> torch.Tensor:
        # self_train_main.py:72 rebinds objective_vectors to a loaded tensor; accept whatever is there
        p = self.objective_vectors
        if p.device != self.device or p.dtype != torch.float32 or not p.is_contiguous():
            p = self.objective_vectors = p.to(device=self.device, dtype=torch.float32).contiguous()
        return p

    # -- A10 (self_training.py:103-112); distance + weights + rectification in ONE pass over feat
    def rectify_labels(self, feat_weak, soft, hard_lp, want_weight: bool = False):
        _, wt, lab = ops.prototype_distance(feat_weak, self._protos(), temperature=self.args.proto_temperature,
                                            want_dist=False, want_weight=want_weight, soft=soft, hard=hard_lp,
                                            train_thred=self.args.train_thred, ignore_index=self.args.ignore_index)
        return (lab, wt) if want_weight else lab

    # -- A12
    def rce(self, prediction, label):
        """SelfTrain.rce (:167-174).  one_hot(label, C+1) raises on the host for labels > C or < 0, and so does this
        (one read of the kernel's counter); a label equal to C is accepted like there."""
        lab = label.to(self.device)
        _, _, rce, bad, _ = ops.ce_rce(prediction, lab, self.args.ignore_index, 0.0, 1.0)
        if int(bad.item()) != 0:
            raise RuntimeError("Class values must be smaller than num_classes.")
        return rce

    def target_loss(self, logits, label):
        """self_training.py:114-118 in one forward kernel: alpha*CE + beta*RCE (or CE alone)."""
        if getattr(self.args, "rce", False):
            total, _, _, bad, over = ops.ce_rce(logits, label, self.args.ignore_index, self.args.rce_alpha, self.args.rce_beta)
            if STRICT_LABELS:
                _raise_if_bad_labels(bad + over, "SelfTrain.step target loss")
            return total
        return self.seg_loss(logits, label)

    # -- A5/A6
    def class_means(self, feat_cls, outputs, labels=None, thresh=None) -> ops.ClassMeans:
        return ops.class_means(feat_cls, outputs, self.num_classes, labels_val=labels,
                               overflow_channel=self.args.ignore_index, use_conf_gate=True,
                               conf_thresh=-1 if thresh is None else thresh)

    def calculate_mean_vector(self, feat_cls, outputs, labels=None, thresh=None):
        return _means_to_lists(self.class_means(feat_cls, outputs, labels, thresh))

    def process_label(self, label):
        return ops.one_hot_overflow(label, self.num_classes, self.args.ignore_index)

    def update_objective_SingleVector(self, id, vector, name="moving_average", start_mean=True):
        self._protos()
        ops.prototype_update_single(self.objective_vectors, self._num(), id, vector.to(self.device), _mode(name),
                                    start_mean, self.args.proto_momentum)

    def _num(self) -> torch.Tensor:
        n = self.objective_vectors_num
        if n.device != self.device or n.dtype != torch.float32 or not n.is_contiguous():
            n = self.objective_vectors_num = n.to(device=self.device, dtype=torch.float32).contiguous()
        return n

    def update_prototypes(self, feat_cls, outputs, name="moving_average", start_mean=False):
        """self_training.py:123-126 for one batch on device: segmented means + the sequential update."""
        return ops.class_means(feat_cls, outputs, self.num_classes, overflow_channel=self.args.ignore_index,
                               use_conf_gate=True, conf_thresh=-1,
                               update=(self._protos(), self._num(), _mode(name), start_mean, self.args.proto_momentum))

    # -- the step
    def step(self, source_image, source_label, target_image, target_image_full, target_lp_soft, target_lp_hard,
             target_weak_params, target_label=None):
        a = self.args
        source_out = self.seg_net(source_image)
        scaled_size = (int(source_label.size()[1] / self.scale_rate), int(source_label.size()[2] / self.scale_rate))
        source_up = _Upsample.apply(source_out["out"], (source_label.size()[1], source_label.size()[2]))
        loss_source = self.seg_loss(source_up, source_label)
        loss_source.backward()

        if a.proto_rectify:
            threshold_arg = ops.upsample_bilinear(target_lp_soft, scaled_size)
        else:
            threshold_arg = ops.upsample_bilinear(target_lp_hard.unsqueeze(1).float(), scaled_size).long()

        ema_out = ema_feat8 = quarter = None
        if a.ema:
            with torch.no_grad():
                ema_out = self.ema_net(target_image_full)
            quarter = (int(target_image_full.shape[2] / 4), int(target_image_full.shape[3] / 4))
            ema_feat8 = ema_out["feat"].detach()        # the network's own 1/8-stride features (:66-67)

        target_out = self.seg_net(target_image)
        target_out["out"] = _Upsample.apply(target_out["out"], tuple(threshold_arg.shape[2:]))
        batch, _, w, h = threshold_arg.shape
        if a.proto_rectify:
            # :69 (interpolation to the quarter map) + :82,:136-165 (full2weak, distances, weights) + :103-112
            # (rectify): at the reference's batch sizes ONE kernel reading the 1/8-stride features, no [N,K,H/4,W/4] /
            # [N,K,64,128] tensors (ops.rectify_from_stride8 switches to the materialising form for large batches)
            hard_lp = ops.resize_nearest_u8(target_lp_hard.to(torch.uint8), tuple(threshold_arg.shape[2:]), rule="torch")
            threshold_arg = ops.rectify_from_stride8(
                ema_feat8, quarter, target_weak_params, self._protos(), a.proto_temperature, threshold_arg, hard_lp,
                a.train_thred, a.ignore_index)[1]

        loss_target_pseudo = self.target_loss(target_out["out"], threshold_arg.reshape([batch, w, h]))
        loss_target_pseudo.backward()
        self.optimizer_seg.step()

        if a.moving_prototype:
            # :124-126 takes the class means of the INTERPOLATED maps (:69-72): those two are materialised here
            self.update_prototypes(ops.upsample_bilinear(ema_feat8, quarter),
                                   ops.upsample_bilinear(ema_out["out"].detach(), quarter), start_mean=False)

        if a.ema:
            for param_q, param_k in zip(self.seg_net.parameters(), self.ema_net.parameters()):
                param_k.data = param_k.data.clone() * 0.999 + param_q.data.clone() * (1. - 0.999)
            for buffer_q, buffer_k in zip(self.seg_net.buffers(), self.ema_net.buffers()):
                buffer_k.data = buffer_q.data.clone()
        self.counter += 1
        return loss_target_pseudo, loss_source


class _Upsample(torch.autograd.Function):
    """Bilinear align_corners=True upsample with its adjoint, so that losses taken on the upsampled logits
    (self_training.py:53-57,75,114) back-propagate into the network."""

    @staticmethod
    def forward(ctx, x, size):
        ctx.in_hw = (x.shape[-2], x.shape[-1])
        return ops.upsample_bilinear(x, size)

    @staticmethod
    def backward(ctx, g):
        return ops.upsample_bilinear_adjoint(g, ctx.in_hw), None


# ------------------------------------------------------------------------------------------------
def _hook_freiburg(cls, device="cuda") -> None:
    """Wrap data.freiburg_dataset.Freiburg.__init__: a self-training dataset with a confidence threshold gets its
    stored pseudo-labels thresholded / whitelisted / resized on the device once (prepare_freiburg_labels), in the
    process that builds the dataset; __getitem__ itself stays the reference's."""
    if getattr(cls.__init__, "_b200seg_hooked", False):
        return
    original = cls.__init__

    def __init__(self, args, *a, **kw):
        original(self, args, *a, **kw)
        if getattr(self, "self_train", False) and getattr(self, "with_label", False) \
                and getattr(args, "proto_rectify", False) and getattr(args, "threshold", None):
            self.args = prepare_freiburg_labels(args, self.data_list, device=device)

    __init__._b200seg_hooked = True
    cls.__init__ = __init__


def install(reference_modules: dict) -> None:
    """Rebind the reference's names to this package.  ``reference_modules`` maps module
    names ('cal_prototype', 'self_training', 'self_train_main', 'generate_pseudo_label', 'segmentation_evaluate',
    'segmentation_train', 'data', 'data.freiburg_dataset', 'utils', 'utils.eval_tools', 'utils.loss', 'train') to the
    imported reference modules; absent ones are skipped."""
    def bind(mod, name, obj):
        m = reference_modules.get(mod)
        if m is not None and hasattr(m, name):
            setattr(m, name, obj)

    bind("cal_prototype", "Class_Features", Class_Features)
    for mod in ("self_training", "self_train_main"):
        bind(mod, "SelfTrain", SelfTrain)
    for mod in ("utils.eval_tools", "utils", "segmentation_evaluate", "segmentation_train"):
        bind(mod, "evaluate", evaluate)
        bind(mod, "_fast_hist", _fast_hist)
    for mod in ("utils.loss", "utils", "train"):
        bind(mod, "SemanticConsistency", SemanticConsistency)
        bind(mod, "FocalLoss", FocalLoss)
    bind("generate_pseudo_label", "generate_pl", generate_pl)                      # generate_pseudo_label.py:60-96
    for mod in ("segmentation_evaluate", "segmentation_train", "self_train_main"):
        bind(mod, "seg_validate", seg_validate)                                    # segmentation_evaluate.py:24-80
    for mod in ("data.freiburg_dataset", "data"):                                  # data/freiburg_dataset.py:108-125
        m = reference_modules.get(mod)
        if m is not None and hasattr(m, "Freiburg"):
            _hook_freiburg(m.Freiburg)
