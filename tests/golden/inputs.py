"""Seeded synthetic inputs shared by make_golden.py (reference side) and the parity tests.

Only numpy's legacy RandomState is used: its stream is frozen by numpy's compatibility
policy, so regenerating an input from its seed on the GPU box yields the bytes the
golden output was computed from.  Fixtures therefore store outputs + seeds, not inputs.
"""
import numpy as np


def rs(seed):
    return np.random.RandomState(seed)


def randn(seed, *shape, scale=1.0):
    return (rs(seed).standard_normal(shape) * scale).astype(np.float32)


def softmax_np(x, axis):
    x = x - x.max(axis=axis, keepdims=True)
    e = np.exp(x)
    return (e / e.sum(axis=axis, keepdims=True)).astype(np.float32)


# ---------------------------------------------------------------- pseudo-label cases
PL_CASES = {
    # name: (N, C, h, w, H, W, seed)
    "pl_small": (2, 13, 9, 17, 64, 128, 101),
    "pl_odd": (1, 19, 7, 11, 41, 77, 102),
}


def pl_logits(name):
    n, c, h, w, H, W, seed = PL_CASES[name]
    return randn(seed, n, c, h, w, scale=3.0), randn(seed + 5000, n, c, h, w, scale=3.0)


# ---------------------------------------------------------------- threshold case
def threshold_inputs(seed=201, h=40, w=88, num_classes=13):
    r = rs(seed)
    label = r.randint(0, num_classes, size=(h, w)).astype(np.uint8)
    conf = r.uniform(0.05, 0.95, size=(h, w)).astype(np.float32)
    # plant values around the fp16-rounded threshold (0.3 -> 0.300048828125)
    edge = np.array([0.3, 0.30001, 0.30003, 0.30004, 0.300048828125, 0.30006, 0.3001, 0.3002,
                     0.2999, 0.29999, 0.3000488, 0.3000489, 0.30005], dtype=np.float32)
    conf.reshape(-1)[: edge.size] = edge
    label.reshape(-1)[: edge.size] = 5
    return label, conf


# ---------------------------------------------------------------- prototype cases
def proto_inputs(seed=301, n=3, k=256, c=13, h=9, w=17):
    """feat/logits for Class_Features; plants (i) a class with <10 px, (ii) an all-zero class."""
    r = rs(seed)
    feat = r.standard_normal((n, k, h, w)).astype(np.float32)
    logits = (r.standard_normal((n, c, h, w)) * 3.0).astype(np.float32)
    # image 0: class 2 only at 4 pixels (count < 10 -> skipped)
    logits[0, 2] = -50.0
    logits[0, 2, 0, :4] = 50.0
    # image 1: class 4 features all zero (vector.sum()==0 -> update skipped)
    lab1 = logits[1].argmax(0)
    feat[1][:, lab1 == 4] = 0.0
    # image 2: class 7 absent altogether
    logits[2, 7] = -50.0
    labels_val = r.randint(0, c + 2, size=(n, 1, h, w)).astype(np.float32)  # includes >= C values
    # make roughly half agree with the prediction so the intersection is non-trivial
    pred = logits.argmax(1)[:, None]
    agree = r.uniform(size=pred.shape) < 0.6
    labels_val = np.where(agree, pred, labels_val).astype(np.float32)
    return feat, logits, labels_val


def selftrain_feat(seed=401, n=2, k=256, h=16, w=32):
    return randn(seed, n, k, h, w)


def selftrain_protos(seed=402, c=13, k=256):
    return randn(seed, c, k)


def ce_inputs(seed=501, n=2, c=13, h=16, w=32, ignore=12):
    r = rs(seed)
    logits = (r.standard_normal((n, c, h, w)) * 2.0).astype(np.float32)
    label = r.randint(0, c, size=(n, h, w)).astype(np.int64)  # contains ignore (= c-1) naturally
    return logits, label


def hist_inputs(seed=601, n=4, h=32, w=64, c=19):
    r = rs(seed)
    gt = r.randint(0, c + 1, size=(n, h, w)).astype(np.int64)
    gt[r.uniform(size=gt.shape) < 0.02] = 255
    gt[0, 0, :5] = -1
    pred = np.where(r.uniform(size=gt.shape) < 0.7, np.clip(gt, 0, c - 1), r.randint(0, c, size=gt.shape))
    return pred.astype(np.int64), gt


# ---------------------------------------------------------------- SelfTrain.step() case
WEAK_PARAMS = {"RandomSized": [[72, 96], [144, 192]],
               "RandomCrop": [[4, 20], [68, 84], [8, 40], [136, 168]],
               "RandomHorizontallyFlip": [True, False]}


def full2weak_feat(seed=403):
    return randn(seed, 2, 256, 20, 40)


def step_inputs(seed=404):
    """Inputs of the whole-step golden case, drawn in the order make_golden.py draws them."""
    r = rs(seed)
    src_img = r.standard_normal((2, 1, 64, 128)).astype(np.float32)
    src_lab = r.randint(0, 13, size=(2, 64, 128)).astype(np.int64)
    tgt_img = r.standard_normal((2, 1, 64, 128)).astype(np.float32)
    tgt_full = r.standard_normal((2, 1, 80, 160)).astype(np.float32)
    lp_soft = softmax_np(r.standard_normal((2, 13, 64, 128)).astype(np.float32) * 3, 1)
    lp_hard = r.choice([5, 11, 12, 12, 12], size=(2, 1, 64, 128)).astype(np.int64)
    return dict(src_img=src_img, src_lab=src_lab, tgt_img=tgt_img, tgt_full=tgt_full, lp_soft=lp_soft,
                lp_hard=lp_hard)


def cap_vectors(seed=303, m=3100, k=256):
    return rs(seed).standard_normal((m, k)).astype(np.float32)
