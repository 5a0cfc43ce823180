"""CPU oracle for the self-training label path.  TEST INFRASTRUCTURE ONLY.

This module restates, on CPU tensors, the arithmetic of the reference's label
pipeline (MuliangDu-sudo/thermal-semantic-segmentation @ 8a37af1).  It is the
*checker* for the CUDA path and the timed CPU baseline of ``bench.py``; nothing
in the product (``b200seg`` / ``thermal-semantic-segmentation_b200``) may
import it.  Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs do.

The reference is pure Python on stock ``torch`` / ``numpy`` ops, so the
restatement issues the *same library ops in the same order* (torch 2.11 ATen
CPU kernels, numpy 2.3 ``bincount``); each function cites the reference lines
it follows.  Parity is pinned by ``tests/test_oracle_golden.py`` against
``tests/golden/reference_golden.npz``, which was produced by running the
unmodified reference in the dev container (``tests/golden/make_golden.py``).
The reference itself ships no tests or golden vectors.

Explicit-formula numpy restatements (``*_np``) are kept beside the library
versions where the CUDA kernels implement a formula rather than call a library
(bilinear taps, PIL nearest resize, closed-form prototype update).
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F

__all__ = [
    "upsample_bilinear_ac", "upsample_bilinear_ac_np", "pseudo_label_soft", "pseudo_label_hard",
    "pseudo_label_flip", "pseudo_label_upsampled", "eval_prediction", "threshold_hard_label",
    "resize_nearest_pil_np", "one_hot_with_overflow", "class_mean_vectors", "PrototypeBank",
    "closed_form_update", "full2weak", "prototype_distance", "prototype_weight", "rectify_labels",
    "cross_entropy", "reverse_cross_entropy", "semantic_consistency", "focal_loss", "fast_hist",
    "evaluate", "self_train_step",
]


# --------------------------------------------------------------------------- A1 upsample
def upsample_bilinear_ac(x: torch.Tensor, size) -> torch.Tensor:
    """Bilinear, align_corners=True (generate_pseudo_label.py:82-83,
    segmentation_evaluate.py:39, self_training.py:53-55,62,69-77)."""
    return F.interpolate(x, size=tuple(int(s) for s in size), mode="bilinear", align_corners=True)


def bilinear_taps_np(n_in: int, n_out: int):
    """Source index pair and weights of one axis, fp32, as ATen computes them
    (area_pixel_compute_scale / compute_source_index with align_corners=True)."""
    scale = np.float32(0.0) if n_out <= 1 else np.float32(np.float32(n_in - 1) / np.float32(n_out - 1))
    dst = np.arange(n_out, dtype=np.float32)
    src = (scale * dst).astype(np.float32)
    i0 = np.minimum(src.astype(np.int64), n_in - 1)
    i1 = i0 + (i0 < n_in - 1)
    l1 = (src - i0.astype(np.float32)).astype(np.float32)
    l0 = (np.float32(1.0) - l1).astype(np.float32)
    return i0, i1, l0, l1


def upsample_bilinear_ac_np(x: np.ndarray, size) -> np.ndarray:
    """Explicit formula the CUDA kernel implements:
    out = ly0*(lx0*a + lx1*b) + ly1*(lx0*c + lx1*d), all fp32."""
    x = np.asarray(x, np.float32)
    H, W = int(size[0]), int(size[1])
    y0, y1, ly0, ly1 = bilinear_taps_np(x.shape[-2], H)
    x0, x1, lx0, lx1 = bilinear_taps_np(x.shape[-1], W)
    top = x[..., y0, :][..., :, x0] * lx0 + x[..., y0, :][..., :, x1] * lx1
    bot = x[..., y1, :][..., :, x0] * lx0 + x[..., y1, :][..., :, x1] * lx1
    return (ly0[:, None] * top + ly1[:, None] * bot).astype(np.float32)


# --------------------------------------------------------------------------- A2/A3 pseudo labels
def pseudo_label_soft(logits: torch.Tensor) -> torch.Tensor:
    """args.soft branch: low-res class probabilities (generate_pseudo_label.py:74-78)."""
    return F.softmax(logits, dim=1)


def pseudo_label_hard(logits: torch.Tensor):
    """Hard branch without flip: softmax then max at network resolution
    (generate_pseudo_label.py:85); stored as fp16 / uint8 (:91,:94)."""
    conf, lab = torch.nn.Softmax(dim=1)(logits).max(1, keepdim=True)
    return conf[:, 0], lab[:, 0]


def pseudo_label_flip(logits: torch.Tensor, flip_logits: torch.Tensor, size, flip_dim: int = 1):
    """Flip-TTA branch (generate_pseudo_label.py:80-85): softmax, upsample, average with the
    "un-flipped" second pass, then a SECOND softmax over the probabilities, then max.

    Reference quirk kept on purpose: ``fliplr`` there is ``torch.fliplr`` (:10), which reverses
    dim 1 -- on an NCHW tensor that is the CLASS axis, not the width.  So the committed code averages
    p[c] with p_flip[C-1-c] (and its ``fliplr(images)`` is a no-op on 1-channel input).  ``flip_dim=1``
    reproduces that; ``flip_dim=3`` is the evidently intended width mirror."""
    a = upsample_bilinear_ac(F.softmax(logits, dim=1), size)
    b = upsample_bilinear_ac(F.softmax(flip_logits, dim=1), size)
    avg = (a + torch.flip(b, dims=[flip_dim])) / 2
    conf, lab = torch.nn.Softmax(dim=1)(avg).max(1, keepdim=True)
    return conf[:, 0], lab[:, 0]


def pseudo_label_upsampled(logits: torch.Tensor, size):
    """North-star order: upsample logits, softmax, max (the order of self_training.py:53-55
    + :85's softmax/max, used for the fused headline kernel)."""
    conf, lab = F.softmax(upsample_bilinear_ac(logits, size), dim=1).max(1, keepdim=True)
    return conf[:, 0], lab[:, 0]


def eval_prediction(logits: torch.Tensor, size) -> torch.Tensor:
    """seg_validate: Upsample then max(1)[1] (segmentation_evaluate.py:38-42)."""
    return upsample_bilinear_ac(logits, size).max(1)[1]


# --------------------------------------------------------------------------- A4 threshold
def threshold_hard_label(label_u8: np.ndarray, conf_f16: np.ndarray, threshold, ignore_index: int,
                         keep=(5, 11)) -> np.ndarray:
    """data/freiburg_dataset.py:115-123.  ``conf`` is fp16 on disk and numpy compares it with the
    python-float threshold in fp16; ``threshold`` may be a scalar or a per-class
    vector (the build's generalisation, reference = scalar)."""
    lab = np.array(label_u8, dtype=np.uint8, copy=True)
    conf = np.asarray(conf_f16)
    thr = np.asarray(threshold, dtype=np.float64)
    if thr.ndim == 0:
        if conf.dtype == np.float16:
            reject = conf <= np.float16(thr)          # what numpy 2 does with a python float
        else:
            reject = conf <= conf.dtype.type(thr)
    else:
        per_px = thr[np.minimum(lab, thr.size - 1)]
        reject = conf <= per_px.astype(conf.dtype)
    lab[reject] = ignore_index
    kept = np.full(lab.shape, ignore_index, dtype=np.int64)
    classes = range(int(lab.max()) + 1) if keep is None else keep
    for c in classes:
        kept[lab == c] = c
    return kept.astype(np.uint8)


def pil_nearest_index_np(n_in: int, n_out: int) -> np.ndarray:
    """Source index table of Pillow's NEAREST resize along one axis.  Pillow (ImagingScaleAffine,
    Geometry.c) starts at 0.5*scale and ACCUMULATES ``+= scale`` in double, truncating each value, so
    exact-integer positions such as (262+0.5)*88/700 = 33 can land one below the closed form."""
    scale = float(n_in) / float(n_out)
    pos = 0.0 + scale * 0.5
    out = np.empty(n_out, np.int64)
    for i in range(n_out):
        out[i] = int(pos)
        pos += scale
    return np.minimum(out, n_in - 1)


def resize_nearest_pil_np(a: np.ndarray, size_wh) -> np.ndarray:
    """PIL ``Image.resize(size, NEAREST)`` (data/freiburg_dataset.py:124)."""
    W, H = int(size_wh[0]), int(size_wh[1])
    h, w = a.shape
    return a[pil_nearest_index_np(h, H)][:, pil_nearest_index_np(w, W)]


# --------------------------------------------------------------------------- A5 class means
def one_hot_with_overflow(label: torch.Tensor, num_classes: int, overflow_channel: int) -> torch.Tensor:
    """process_label (cal_prototype.py:146-151 with overflow_channel = C;
    self_training.py:204-209 with overflow_channel = args.ignore_index)."""
    n, _, a, b = label.shape
    canvas = torch.zeros(n, num_classes + 1, a, b)
    idx = torch.where(label < num_classes, label, torch.tensor([float(overflow_channel)]))
    return canvas.scatter_(1, idx.long(), 1)


def class_mean_vectors(feat: torch.Tensor, logits: torch.Tensor, num_classes: int, labels_val=None,
                       thresh=None, overflow_channel=None, use_conf_mask=False):
    """Per-(image, class) masked mean of ``feat`` by argmax label.

    ``use_conf_mask=False`` -> Class_Features.calculate_mean_vector (cal_prototype.py:118-144);
    ``use_conf_mask=True``  -> SelfTrain.calculate_mean_vector (self_training.py:176-202), whose
    divisor and product carry ``mask = conf >= thresh`` while the <10 rule uses the unmasked map.
    Returns (list of [K,1,1] tensors, list of class ids), image-major then class."""
    if overflow_channel is None:
        overflow_channel = num_classes
    prob = F.softmax(logits, dim=1)
    hot = one_hot_with_overflow(prob.argmax(dim=1, keepdim=True).float(), num_classes, overflow_channel)
    if labels_val is not None:
        hot = one_hot_with_overflow(labels_val, num_classes, overflow_channel) * hot
    if use_conf_mask:
        gate = prob.max(dim=1, keepdim=True)[0].ge(-1 if thresh is None else thresh)
        frac = F.adaptive_avg_pool2d(hot * gate, 1)
    else:
        gate = None
        frac = F.adaptive_avg_pool2d(hot, 1)
    vectors, ids = [], []
    for n in range(feat.shape[0]):
        for t in range(num_classes):
            if frac[n][t].item() == 0:
                continue
            if (hot[n][t] > 0).sum() < 10:
                continue
            prod = feat[n] * hot[n][t]
            if gate is not None:
                prod = prod * gate[n]
            vectors.append(F.adaptive_avg_pool2d(prod, 1) / frac[n][t])
            ids.append(t)
    return vectors, ids


# --------------------------------------------------------------------------- A6 prototype update
class PrototypeBank:
    """objective_vectors / objective_vectors_num and their sequential update
    (cal_prototype.py:91-93,153-169; self_training.py:24-25,211-227)."""

    def __init__(self, num_classes: int, dim: int = 256, momentum: float = 0.9999):
        self.vectors = torch.zeros(num_classes, dim)
        self.num = torch.zeros(num_classes)
        self.momentum = momentum

    def update(self, cls: int, vector: torch.Tensor, name: str = "moving_average", start_mean: bool = True):
        if vector.sum().item() == 0:
            return
        if start_mean and self.num[cls].item() < 100:
            name = "mean"
        v = vector.squeeze()
        if name == "moving_average":
            self.vectors[cls] = self.vectors[cls] * (1 - self.momentum) + self.momentum * v
            self.num[cls] += 1
            self.num[cls] = min(self.num[cls], 3000)
        elif name == "mean":
            self.vectors[cls] = self.vectors[cls] * self.num[cls] + v
            self.num[cls] += 1
            self.vectors[cls] = self.vectors[cls] / self.num[cls]
            self.num[cls] = min(self.num[cls], 3000)
        else:
            raise NotImplementedError(name)


def closed_form_update(vec0: np.ndarray, num0: float, contribs: np.ndarray, mode: str, momentum: float,
                       cap: int = 3000):
    """Closed form of M sequential updates of ONE class; fp64.
    'mean':           obj_j = (obj_{j-1} * n_{j-1} + v_j) / (n_{j-1} + 1), n_j = min(n_{j-1} + 1, cap)
    'moving_average': obj_j = obj_{j-1} * (1 - m) + m * v_j
    Returns (vector fp64, num).  The CUDA update kernel evaluates exactly these weights."""
    obj = np.asarray(vec0, np.float64).copy()
    n = float(num0)
    w0, w = 1.0, np.zeros(len(contribs), np.float64)
    for j in range(len(contribs)):
        if mode == "mean":
            a, b = n / (n + 1.0), 1.0 / (n + 1.0)
        else:
            a, b = 1.0 - momentum, momentum
        w0 *= a
        w[:j] *= a
        w[j] = b
        n = min(n + 1.0, float(cap))
    if len(contribs):
        obj = w0 * obj + (w[:, None] * np.asarray(contribs, np.float64)).sum(0)
    return obj, n


# --------------------------------------------------------------------------- A9 full2weak
def full2weak(feat: torch.Tensor, weak_params) -> torch.Tensor:
    """self_training.py:145-158: per image resize to (int(h/4), int(w/4)), crop, optional W flip."""
    parts = []
    sized, crop, flip = weak_params["RandomSized"], weak_params["RandomCrop"], weak_params["RandomHorizontallyFlip"]
    for i in range(feat.shape[0]):
        hh, ww = sized[0][i], sized[1][i]
        f = upsample_bilinear_ac(feat[i:i + 1], (int(hh / 4), int(ww / 4)))
        y1, y2, x1, x2 = crop[0][i], crop[1][i], crop[2][i], crop[3][i]
        ys, th, xs, tw = int(y1 / 4), int((y2 - y1) / 4), int(x1 / 4), int((x2 - x1) / 4)
        f = f[:, :, ys:ys + th, xs:xs + tw]
        if flip[i]:
            f = torch.flip(f, dims=[3])
        parts.append(f)
    return torch.cat(parts, 0)


# --------------------------------------------------------------------------- A7/A8 distance, weights
def prototype_distance(feat: torch.Tensor, protos: torch.Tensor) -> torch.Tensor:
    """feat_prototype_distance (self_training.py:160-165): L2 norm of (proto_c - feat) over K."""
    n, _, h, w = feat.shape
    c = protos.shape[0]
    out = -torch.ones((n, c, h, w))
    for i in range(c):
        out[:, i] = torch.norm(protos[i].reshape(-1, 1, 1).expand(-1, h, w) - feat, 2, dim=1)
    return out


def prototype_weight(feat: torch.Tensor, protos: torch.Tensor, temperature: float, weak_params=None) -> torch.Tensor:
    """get_prototype_weight (self_training.py:136-143)."""
    if weak_params is not None:
        feat = full2weak(feat, weak_params)
    d = prototype_distance(feat, protos)
    nearest, _ = d.min(dim=1, keepdim=True)
    return F.softmax(-(d - nearest) * temperature, dim=1)


# --------------------------------------------------------------------------- A10 rectify
def rectify_labels(weights: torch.Tensor, soft: torch.Tensor, hard_lp: torch.Tensor, train_thred, ignore_index: int):
    """self_training.py:106-112.  ``hard_lp`` is already at the weights' resolution ([N,1,h,w] int64).
    Returns ([N,1,h,w] int64 labels, normalised rectified probabilities)."""
    rect = weights * soft
    lab = rect.max(1, keepdim=True)[1]
    rect = rect / rect.sum(1, keepdim=True)
    top = rect.max(1, keepdim=True)[0]
    lab[top < train_thred] = ignore_index
    lab = torch.where(hard_lp != ignore_index, hard_lp, lab)
    return lab, rect


def resize_nearest_torch(label: torch.Tensor, size) -> torch.Tensor:
    """F.interpolate(mode='nearest') of the stored hard label (self_training.py:103)."""
    return F.interpolate(label.float(), size=tuple(size), mode="nearest").long()


# --------------------------------------------------------------------------- A11-A13 losses
def cross_entropy(logits: torch.Tensor, target: torch.Tensor, ignore_index: int, reduction: str = "mean"):
    """nn.CrossEntropyLoss(ignore_index, reduction) (self_training.py:27; utils/loss.py:145)."""
    return torch.nn.CrossEntropyLoss(ignore_index=ignore_index, reduction=reduction)(logits, target)


def reverse_cross_entropy(logits: torch.Tensor, label: torch.Tensor, num_classes: int, ignore_index: int):
    """SelfTrain.rce (self_training.py:167-174)."""
    p = torch.clamp(F.softmax(logits, dim=1), min=1e-7, max=1.0)
    valid = (label != ignore_index).float()
    hot = F.one_hot(label, num_classes + 1).float()
    hot = torch.clamp(hot.permute(0, 3, 1, 2)[:, :-1], min=1e-4, max=1.0)
    return -(torch.sum(p * torch.log(hot), dim=1) * valid).sum() / (valid.sum() + 1e-6)


def semantic_consistency(logits: torch.Tensor, target: torch.Tensor, ignore_classes=(), reduction: str = "mean"):
    """SemanticConsistency.forward (utils/loss.py:142-150): listed classes become 255 IN PLACE."""
    for c in ignore_classes:
        target[target == c] = 255
    return cross_entropy(logits, target, 255, reduction)


def focal_loss(logits: torch.Tensor, target: torch.Tensor, gamma=0, alpha=None, size_average: bool = True):
    """FocalLoss.forward (utils/loss.py:164-184); pt is detached (``.data``); no ignore handling."""
    if isinstance(alpha, (float, int)):
        alpha = torch.tensor([alpha, 1 - alpha], dtype=torch.float32)
    elif isinstance(alpha, (list, tuple)):
        alpha = torch.tensor(list(alpha), dtype=torch.float32)
    x = logits
    if x.dim() > 2:
        x = x.view(x.size(0), x.size(1), -1).transpose(1, 2).contiguous().view(-1, x.size(1))
    t = target.view(-1, 1)
    logpt = F.log_softmax(x, dim=1).gather(1, t).view(-1)
    pt = logpt.detach().exp()
    if alpha is not None:
        logpt = logpt * alpha.type_as(x).gather(0, t.view(-1))
    loss = -1 * (1 - pt) ** gamma * logpt
    return loss.mean() if size_average else loss.sum()


# --------------------------------------------------------------------------- the whole step
def self_train_step(net_src_out, net_ema_feat, net_ema_out, net_tgt_out, source_label, full_hw, lp_soft, lp_hard,
                    weak_params, protos, num, num_classes=13, ignore_index=12, train_thred=0.0, temperature=0.5,
                    rce_alpha=0.1, rce_beta=1.0, momentum=1e-4, scale_rate=4):
    """SelfTrain.step (self_training.py:49-134) on RECORDED network outputs (the three forward passes are the
    caller's): source CE (:51-58), soft-label upsample (:62), EMA maps to the quarter resolution (:66-72), target
    logits upsample (:75), prototype weights (:82), rectify (:103-112), CE + RCE (:114-118), backward (:58,:120),
    moving-average prototype update (:123-126).  Returns a dict with the losses, the rectified labels, the
    gradients left on the raw logits and the bank after the update."""
    src = net_src_out.detach().clone().requires_grad_(True)
    tgt = net_tgt_out.detach().clone().requires_grad_(True)
    H, W = source_label.shape[1], source_label.shape[2]
    scaled = (int(H / scale_rate), int(W / scale_rate))
    loss_source = cross_entropy(upsample_bilinear_ac(src, (H, W)), source_label, ignore_index)
    loss_source.backward()
    soft_up = upsample_bilinear_ac(lp_soft, scaled)
    quarter = (int(full_hw[0] / 4), int(full_hw[1] / 4))
    ema_feat = upsample_bilinear_ac(net_ema_feat, quarter)
    ema_out = upsample_bilinear_ac(net_ema_out, quarter)
    tgt_up = upsample_bilinear_ac(tgt, soft_up.shape[2:])
    weights = prototype_weight(ema_feat, protos, temperature, weak_params)
    hard = resize_nearest_torch(lp_hard, soft_up.shape[2:])
    label, _ = rectify_labels(weights, soft_up, hard, train_thred, ignore_index)
    b, _, hh, ww = soft_up.shape
    lab = label.reshape(b, hh, ww)
    loss_target = rce_alpha * cross_entropy(tgt_up, lab, ignore_index) + \
        rce_beta * reverse_cross_entropy(tgt_up, lab.clone(), num_classes, ignore_index)
    loss_target.backward()
    bank = PrototypeBank(num_classes, protos.shape[1], momentum=momentum)
    bank.vectors, bank.num = protos.clone(), num.clone()
    vecs, ids = class_mean_vectors(ema_feat, ema_out, num_classes, overflow_channel=ignore_index, use_conf_mask=True)
    for t, v in zip(ids, vecs):
        bank.update(t, v, "moving_average", start_mean=False)
    return {"loss_target": loss_target.detach(), "loss_source": loss_source.detach(), "rectified_label": lab,
            "weights": weights, "soft_up": soft_up, "target_logits": tgt_up.detach(), "src_out_grad": src.grad,
            "tgt_out_grad": tgt.grad, "protos_after": bank.vectors, "num_after": bank.num}


# --------------------------------------------------------------------------- A14/A15 evaluation
def fast_hist(pred: np.ndarray, gt: np.ndarray, num_classes: int) -> np.ndarray:
    """_fast_hist (utils/eval_tools.py:4-9): rows = ground truth, columns = prediction."""
    ok = (gt >= 0) & (gt < num_classes)
    flat = np.bincount(num_classes * gt[ok].astype(int) + pred[ok], minlength=num_classes ** 2)
    return flat.reshape(num_classes, num_classes)


FREIBURG_NAMES = ["IoU road,parking", "IoU ground,sidewalk", "IoU building,", "IoU curb", "IoU fence",
                  "IoU pole,traffic light,traffic sign", "IoU vegetation", "IoU terrain", "IoU sky",
                  "IoU person,rider", "IoU car,truck,bus,train", "IoU motorcycle,bicycle"]
CITYSCAPES_NAMES = ["IoU road", "IoU sidewalk", "IoU building,", "IoU wall", "IoU fence", "IoU pole",
                    "IoU traffic light", "IoU traffic sign", "IoU vegetation", "IoU terrain", "IoU sky",
                    "IoU person", "IoU rider", "IoU car", "IoU truck", "IoU bus", "IoU train",
                    "IoU motorcycle", "IoU bicycle"]


def metrics_from_hist(hist: np.ndarray, num_classes: int):
    """utils/eval_tools.py:53-67 on an accumulated float64 matrix."""
    hist = np.asarray(hist, np.float64)
    diag = np.diag(hist)
    with np.errstate(divide="ignore", invalid="ignore"):
        acc = diag.sum() / hist.sum()
        acc_cls = np.nanmean(diag / hist.sum(axis=1))
        iu = diag / (hist.sum(axis=1) + hist.sum(axis=0) - diag)
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


def evaluate(predictions, gts, num_classes: int):
    """evaluate (utils/eval_tools.py:12-67): float64 accumulation image by image."""
    hist = np.zeros((num_classes, num_classes))
    for lp, lt in zip(predictions, gts):
        hist += fast_hist(lp.flatten(), lt.flatten(), num_classes)
    return metrics_from_hist(hist, num_classes)
