"""Pseudo-label files, byte-compatible with what the reference writes and reads.

Writers  -- generate_pseudo_label.py:78 (soft: ``<name>.npy`` = fp32 [C,h,w] probabilities),
            :91 (hard: ``<name>.png`` = 8-bit grayscale class map), :94 (``<name>_conf.npy`` = fp16 [H,W]).
Readers  -- data/freiburg_dataset.py:108-125 (np.load of the soft map; PNG + _conf.npy, then the threshold /
            whitelist block and the NEAREST resize, which run on the device here), and the loader-side bilinear
            resize of the soft map to the image size (utils/self_train_augmentation.py:21-24).
Driver   -- generate_pl (generate_pseudo_label.py:60-96) with the fused kernels; files written by either
            implementation are read by the other.

File IO is host work (PIL / numpy, as in the reference).  A batch leaves the device in ONE copy into pinned
memory instead of one ``.cpu()`` per image and tensor.
"""
from __future__ import annotations

import os
from typing import Iterable, Optional, Sequence

import numpy as np
import torch
from PIL import Image

from . import dropin, ops


def _names(filenames: Iterable[str]):
    return [os.path.basename(f) for f in filenames]


def _to_host(t: torch.Tensor) -> np.ndarray:
    if t.is_cuda:
        host = torch.empty(t.shape, dtype=t.dtype).pin_memory()
        host.copy_(t, non_blocking=True)
        torch.cuda.current_stream(t.device).synchronize()
        return host.numpy()
    return t.detach().numpy()


def save_soft_batch(save_dir: str, filenames: Iterable[str], prob: torch.Tensor) -> None:
    """generate_pseudo_label.py:75-78: one fp32 [C,h,w] ``.npy`` per image, named after the image."""
    os.makedirs(save_dir, exist_ok=True)
    host = _to_host(prob.to(torch.float32))
    for k, name in enumerate(_names(filenames)):
        np.save(os.path.join(save_dir, name.replace(".png", ".npy")), host[k])


def save_hard_batch(save_dir: str, filenames: Iterable[str], label: torch.Tensor, conf16: torch.Tensor) -> None:
    """generate_pseudo_label.py:87-94: ``<name>.png`` (uint8 class map, PIL mode 'L') and
    ``<name>_conf.npy`` (fp16).  label [N,H,W] uint8, conf16 [N,H,W] float16 (ops.PseudoLabels)."""
    if label.dtype != torch.uint8 or conf16.dtype != torch.float16:
        raise TypeError("save_hard_batch expects uint8 labels and float16 confidences (as the reference stores them)")
    os.makedirs(save_dir, exist_ok=True)
    lab, conf = _to_host(label), _to_host(conf16)
    for k, name in enumerate(_names(filenames)):
        Image.fromarray(lab[k]).save(os.path.join(save_dir, name))
        np.save(os.path.join(save_dir, name.replace(".png", "_conf.npy")), conf[k])


def read_soft(path_soft: str, img_name: str) -> np.ndarray:
    """data/freiburg_dataset.py:109-110."""
    return np.load(os.path.join(path_soft, os.path.basename(img_name).replace(".png", ".npy")))


def read_hard_files(path_lp: str, img_name: str, with_conf: bool = True):
    """data/freiburg_dataset.py:112-118: the stored class map (uint8 [H,W]) and, when thresholding, the stored
    fp16 confidence.  Host arrays, exactly as the reference loads them."""
    base = os.path.basename(img_name)
    label = np.array(Image.open(os.path.join(path_lp, base)), dtype=np.uint8)
    conf = np.load(os.path.join(path_lp, base.replace(".png", "_conf.npy"))) if with_conf else None
    return label, conf


def load_hard_batch(path_lp: str, img_names: Sequence[str], num_classes: int, device, threshold: Optional[float] = 0.3,
                    ignore_index: int = 12, keep: Optional[Iterable[int]] = (5, 11),
                    size_wh: Optional[Sequence[int]] = None) -> torch.Tensor:
    """The reader side for a batch of same-sized label files: files -> pinned host -> device, then the
    threshold block + class whitelist + PIL-NEAREST resize of data/freiburg_dataset.py:115-125 in two kernels
    (dropin.threshold_block).  ``threshold=None`` mirrors ``args.threshold`` unset (files used as they are).
    Returns uint8 [N,H',W'] on ``device``."""
    labs, confs = [], []
    for name in img_names:
        lab, conf = read_hard_files(path_lp, name, with_conf=threshold is not None)
        labs.append(lab)
        if conf is not None:
            if conf.dtype != np.float16:
                conf = conf.astype(np.float16)
            confs.append(conf)
    dev = torch.device(device)
    lab_t = torch.from_numpy(np.stack(labs)).pin_memory().to(dev, non_blocking=True)
    if threshold is None:
        out = lab_t
        if size_wh is not None:
            out = ops.resize_nearest_u8(out, (int(size_wh[1]), int(size_wh[0])), rule="pil")
        return out
    conf_t = torch.from_numpy(np.stack(confs)).pin_memory().to(dev, non_blocking=True)
    return dropin.threshold_block(lab_t, conf_t, num_classes, threshold=threshold, ignore_index=ignore_index, keep=keep,
                                  size_wh=size_wh)


def load_soft_batch(path_soft: str, img_names: Sequence[str], size_hw: Sequence[int], device) -> torch.Tensor:
    """The reader side of the soft pseudo-labels: ``<name>.npy`` fp32 [C,h,w] maps (data/freiburg_dataset.py:109-110)
    -> pinned host -> device, then the loader's bilinear align_corners=True resize to the image size
    (utils/self_train_augmentation.py:21-24: ``F.interpolate(lpsoft.unsqueeze(0), size=[H, W], ...)``) in one kernel
    for the whole batch.  Returns fp32 [N,C,H,W] on ``device`` (what ``step()`` upsamples again at self_training.py:62)."""
    maps = [read_soft(path_soft, name) for name in img_names]
    host = torch.from_numpy(np.stack(maps).astype(np.float32, copy=False)).pin_memory()
    return ops.upsample_bilinear(host.to(torch.device(device), non_blocking=True), (int(size_hw[0]), int(size_hw[1])))


@torch.no_grad()
def generate_pl(net, dataloader, device, save_dir: str, soft: bool = False, flip: bool = False,
                log_every: int = 100) -> int:
    """generate_pl (generate_pseudo_label.py:60-96): network forward, then the fused label kernels, then the
    reference's files.  The colour visualisation (:89,:93) is presentation, not part of the label path, and is
    not written.  Returns the number of images processed."""
    net.eval()
    count = 0
    for i, data_i in enumerate(dataloader):
        images = data_i["image"].to(device)
        filenames = data_i["img_path"]
        out = net(images)["out"]
        if soft:
            save_soft_batch(save_dir, filenames, dropin.soft_label_batch(out))
        else:
            if flip:
                flip_out = net(torch.fliplr(images))["out"]      # fliplr on NCHW, as committed (:10,:81)
                pl = dropin.pseudo_label_batch(out, flip_out=flip_out, size=images.shape[2:], flip_dim=1)
            else:
                pl = dropin.pseudo_label_batch(out)
            save_hard_batch(save_dir, filenames, pl.label, pl.conf16)
        count += images.shape[0]
        if log_every and i % log_every == 0:
            print("pseudo label generation: [{}/{}]".format(i, len(dataloader)))
    return count
