#!/usr/bin/env python
"""bench.py -- pseudo-labelled px/s of the self-training label path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--config cfg1|cfg2|cfg3|cfg4|cfg5]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Default workload (BASELINE.json configs[1], "cfg2"): one STEP = one batch of 16 synthetic Freiburg-shape frames
(640x1920, 13 classes; DeepLabv2 1/8-stride(+1) logits [16,13,81,241] and 256-ch features [16,256,81,241]) through

    1. fused bilinear upsample -> softmax -> argmax/confidence -> fp16 cast -> class-wise threshold + whitelist
       (uint8 label, fp16 confidence, thresholded uint8 label at 640x1920)            [XU / issue bound]
    2. argmax label + per-(image, class) segmented feature sums                        [HBM bound]
       -- 1 and 2 run AT THE SAME TIME on every SM (two capture streams, co-resident CTAs); the label kernel of 2 needs
          only the logits, so the one of step i+1 is issued on its own stream beside step i (`coresident_prelabel`; each
          timed step still runs one, and the timed region waits for the one its last step issued) --
    3. confusion-matrix accumulation of the labels against a uint8 ground-truth map (mIoU at the end),
    4. means + sequential prototype update (+ the exchange between GPUs) on a tail stream, beside the next step.

value = frames * 640 * 1920 / device time, inputs resident in HBM.  Images shard across ranks (weak scaling: 16
frames per rank per step); the only collectives are the prototype mean vectors (the finalize kernel's own NVLink
stores + one barrier, once per step) and the [13,13] int64 confusion matrix (once, inside the timed region).
e2e = the same step through b200seg.pipeline.HostLabelPipeline with pinned HOST buffers, H2D + D2H inside the timed
region.  The other BASELINE.json configurations run with --config (cfg1, cfg3, cfg4, cfg5); the default line also
times the distance / rectify kernel and the CE + RCE kernels at their BASELINE shapes (aux_stages).
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time
import types

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

METRIC = "pseudo-labelled px/sec (+proto update, mIoU)"
UNIT = "px/s"
N_SETS = 4  # rotating input sets (cfg2: 4 x (116 MB in + 79 MB out) = 780 MB working set >> 126 MB L2)

CONFIGS = {
    # BASELINE.json configs[0]: 4 frames 640x320 -> DeepLabv2 1/8(+1) maps 41x81; pseudo-label + prototypes + mIoU
    "cfg1": dict(frames=4, H=320, W=640, h=41, w=81, C=13, K=256, thr=0.3, keep=(5, 11), ignore=12, kind="label_step",
                 title="cfg1: 4 thermal frames 640x320, 13 classes (random-init DeepLabv2 shapes): pseudo-labels + "
                       "256-ch prototype update + confusion matrix"),
    # configs[1] -- the headline
    "cfg2": dict(frames=16, H=640, W=1920, h=81, w=241, C=13, K=256, thr=0.3, keep=(5, 11), ignore=12, kind="label_step",
                 title="cfg2: Freiburg 1920x640, 13 classes, batch 16/GPU: fused upsample/softmax/argmax/threshold "
                       "pseudo-labels + 256-ch prototype update + confusion matrix"),
    # configs[2]: Cityscapes shape, 2048-ch 1/8-stride features: prototype accumulation + distance-weighted labels
    "cfg3": dict(frames=4, H=1024, W=2048, h=129, w=257, C=19, K=2048, ignore=255, kind="proto_distance",
                 title="cfg3: Cityscapes 2048x1024, 19 classes, 2048-ch 1/8-stride features, 4 images/GPU/step: prototype "
                       "accumulation + update, distance -> weights -> rectified labels (unit: feature px)"),
    # configs[3]: full self-training round over 10k frames
    "cfg4": dict(frames=16, H=320, W=640, h=41, w=81, C=13, K=256, thr=0.3, keep=(5, 11), ignore=12, kind="self_train_round",
                 total_frames=10000,
                 title="cfg4: self-training round over 10k synthetic 640x320 frames, 16 frames/GPU/step: pseudo-labels, "
                       "prototype update (exchange), rectified labels from the 1/8 features (interpolation to 1/4, full2weak, "
                       "distances, weights), CE+RCE forward+backward at 1/4 with the upsample and its adjoint"),
    # configs[4]: evaluation sweep over 50k label maps
    "cfg5": dict(frames=1024, H=256, W=512, C=19, kind="eval_sweep", total_frames=50000,
                 title="cfg5: confusion-matrix mIoU over 50k synthetic 19-class 256x512 uint8 label maps, 1024 maps/GPU/step"),
}
CFG = CONFIGS["cfg2"]   # tools/ import this


def workload_name(cfg=None):
    return (cfg or CFG)["title"]


def algorithmic_bytes(cfg=None):
    """Compulsory HBM traffic of one step per stage; every input element read once,
    every output written once."""
    c = cfg or CFG
    n = c["frames"]
    if c["kind"] == "eval_sweep":
        return {"confusion": n * c["H"] * c["W"] * 2}
    P, px = c["h"] * c["w"], c["H"] * c["W"]
    proto = n * P * (4 * c["K"] + 4 * c["C"]) + 4 * n * c["C"] * (c["K"] + 1)
    if c["kind"] == "proto_distance":
        dist = n * P * (4 * c["K"] + 4 * c["C"] + 1 + 8)        # feat + soft in, uint8 hard in, int64 label out
        return {"prototype": proto, "distance_rectify": dist}
    pl = 4 * c["C"] * P * n + n * px * (1 + 2 + 1)                  # logits in; label, conf16, label_thr out
    hist = n * px * 2                                               # uint8 pred + uint8 gt
    out = {"pseudo_label": pl, "prototype": proto, "confusion": hist}
    if c["kind"] == "self_train_round":
        q = (c["H"] // 4) * (c["W"] // 4)
        out["rectify_sampled"] = n * (P * 4 * c["K"] + q * (4 * c["C"] + 1 + 8))
        out["ce_rce_fwd_bwd"] = n * q * ((4 * c["C"] + 8) * 2 + 4 * c["C"]) + n * q * 4 * c["C"] * 2   # + the upsample round trip
    return out


# ------------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md clocks line)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag, self.proc = index, [], threading.Event(), None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                if self.stop_flag.is_set():
                    break
                self.samples.append((time.perf_counter(), [f.strip() for f in line.split(",")]))
        except OSError:
            pass

    def finish(self, t0: float, t1: float):
        self.stop_flag.set()
        if self.proc is not None:
            self.proc.kill()
        rows = [f for (t, f) in self.samples if t0 <= t <= t1 and len(f) >= 6] or \
               [f for (_, f) in self.samples if len(f) >= 6][-3:]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        mhz = sorted(int(r[0]) for r in rows if r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in rows for i in range(4) if r[2 + i].lower().startswith("active")})
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else None,
                "sm_max_mhz": int(rows[0][1]) if rows[0][1].isdigit() else None, "reasons": reasons,
                "samples": len(rows)}


def ncu_traffic(stage: str):
    """Steady-state DRAM bytes per launch of a stage's dominant kernel from the committed ncu --set full capture
    (profiles/r02_traffic.json: a LATE launch of a multi-launch run, so dirty output lines of earlier launches are
    evicted inside the measured launch as they are in the timed loop), or None."""
    for name in ("r02_traffic.json", "r01_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as fh:
                return int(json.load(fh)[stage])
        except (OSError, KeyError, ValueError):
            continue
    return None


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, STREAM-style copy)"
    except (OSError, KeyError, ValueError):
        return 6650.0, "fallback (B200_PROFILING.md, 6.65 TB/s)"


# ------------------------------------------------------------------------------------------------
# synthetic inputs: logits 3*randn so that confidences straddle 0.3
def make_inputs(device, seed: int, frames: int, cfg=None):
    c = cfg or CFG
    g = torch.Generator(device=device).manual_seed(seed)
    if c["kind"] == "eval_sweep":
        gt = torch.randint(0, c["C"] + 1, (frames, c["H"], c["W"]), generator=g, device=device, dtype=torch.int16)
        gt = torch.where(torch.rand(gt.shape, generator=g, device=device) < 0.01, torch.full_like(gt, 255), gt)
        same = torch.rand(gt.shape, generator=g, device=device) < 0.7
        rnd = torch.randint(0, c["C"], gt.shape, generator=g, device=device, dtype=torch.int16)
        pred = torch.where(same, gt.clamp(max=c["C"] - 1), rnd)
        return pred.to(torch.uint8), gt.to(torch.uint8)
    logits = 3.0 * torch.randn(frames, c["C"], c["h"], c["w"], generator=g, device=device)
    feat = torch.randn(frames, c["K"], c["h"], c["w"], generator=g, device=device)
    if c["kind"] == "proto_distance":
        soft = torch.softmax(3.0 * torch.randn(frames, c["C"], c["h"], c["w"], generator=g, device=device), 1)
        hard = torch.full((frames, 1, c["h"], c["w"]), 255, dtype=torch.uint8, device=device)
        return logits, feat, soft, hard
    gt = torch.randint(0, c["C"], (frames, c["H"], c["W"]), generator=g, device=device, dtype=torch.int16).to(torch.uint8)
    if c["kind"] == "self_train_round":
        q = (c["H"] // 4, c["W"] // 4)
        soft = torch.softmax(3.0 * torch.randn(frames, c["C"], *q, generator=g, device=device), 1)
        hard = torch.where(torch.rand((frames, 1) + q, generator=g, device=device) < 0.2,
                           torch.full((frames, 1) + q, 5, dtype=torch.int16, device=device),
                           torch.full((frames, 1) + q, c["ignore"], dtype=torch.int16, device=device)).to(torch.uint8)
        return logits, feat, gt, soft, hard
    return logits, feat, gt


def deeplab_inputs(cfg, device):
    """cfg1: out / feat of the reference's random-init DeepLabv2-ResNet101 (baseline/_ref/models, torch.manual_seed(0))
    on 4 synthetic frames, or None when the reference tree is not there (the caller falls back to randn maps)."""
    ref = load_reference()
    if ref is None or not hasattr(ref, "models"):
        return None
    try:
        import contextlib
        with contextlib.redirect_stdout(sys.stderr):      # the reference's model code prints; stdout carries ONE JSON line
            torch.manual_seed(0)
            net = ref.models.Deeplab(torch.nn.BatchNorm2d, num_classes=cfg["C"], num_channels=1, freeze_bn=False, get_feat=True)
            net = net.to(device).eval()
            g = torch.Generator().manual_seed(1234)
            x = torch.rand(cfg["frames"], 1, cfg["H"], cfg["W"], generator=g).to(device)
            with torch.no_grad():
                out = net(x)
        return out["out"].float().contiguous(), out["feat"].float().contiguous()
    except Exception as e:  # noqa: BLE001
        print("bench.py: DeepLab input generation failed (%s: %s); using randn maps" % (type(e).__name__, e), file=sys.stderr)
        return None


# ------------------------------------------------------------------------------------------------
# The reference's own modules for the CPU arm: baseline/_ref/ (staged by __graft_entry__.build() from the unmodified
# reference; pure Python, so it cannot be pip-installed -- no setup.py / pyproject).
_REF = [False, None]


def load_reference():
    if _REF[0]:
        return _REF[1]
    _REF[0] = True
    root = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isfile(os.path.join(root, "self_training.py")):
        return None
    try:
        import importlib
        import scipy.signal
        import scipy.signal.windows
        if not hasattr(scipy.signal, "gaussian"):            # library drift, not arithmetic
            scipy.signal.gaussian = scipy.signal.windows.gaussian
        if "visdom" not in sys.modules:
            vis = types.ModuleType("visdom")
            vis.Visdom = lambda *a, **k: None
            sys.modules["visdom"] = vis
        if "torchvision.models.utils" not in sys.modules:
            import torch.hub
            tvu = types.ModuleType("torchvision.models.utils")
            tvu.load_state_dict_from_url = torch.hub.load_state_dict_from_url
            sys.modules["torchvision.models.utils"] = tvu
        sys.path.insert(0, root)
        ns = types.SimpleNamespace()
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            ns.cal_prototype = importlib.import_module("cal_prototype")
            ns.self_training = importlib.import_module("self_training")
            ns.eval_tools = importlib.import_module("utils.eval_tools")
            ns.models = importlib.import_module("models")
        _REF[1] = ns
    except Exception as e:  # noqa: BLE001
        print("bench.py: baseline/_ref present but not importable (%s: %s); CPU arm = oracle port" % (type(e).__name__, e),
              file=sys.stderr)
    return _REF[1]


class CpuArm:
    """One step of a workload on CPU tensors as the reference computes it.  With baseline/_ref the reference's OWN
    functions run wherever the path is a function there (Class_Features.calculate_mean_vector +
    update_objective_SingleVector cal_prototype.py:72-75, SelfTrain.feat_prototype_distance / get_prototype_weight /
    rce / seg_loss / calculate_mean_vector, utils.eval_tools._fast_hist); the inline blocks of the scripts
    (upsample -> softmax -> max generate_pseudo_label.py:83-85 / self_training.py:53-55, fp16 / uint8 storage casts
    :91,:94, the threshold block data/freiburg_dataset.py:115-123, the rectify block self_training.py:103-112) are
    the same torch / numpy calls in the same order (oracle/label_path.py, pinned to the reference by the goldens)."""

    def __init__(self, cfg):
        from oracle import label_path as O
        self.O, self.cfg = O, cfg
        self.ref = load_reference()
        self.kind = "reference" if self.ref is not None else "port"
        c = cfg
        if c["kind"] == "eval_sweep":
            self.hist = np.zeros((c["C"], c["C"]))
            return
        if self.ref is not None:
            self.cf = self.ref.cal_prototype.Class_Features(torch.device("cpu"), numbers=c["C"])
            if c["K"] != 256:
                self.cf.objective_vectors = torch.zeros([c["C"], c["K"]])
            a = types.SimpleNamespace(num_classes=c["C"], ignore_index=c["ignore"], proto_temperature=0.5, proto_momentum=1e-4,
                                      rce_alpha=0.1, rce_beta=1.0)
            self.st = self.ref.self_training.SelfTrain(a, None, None, None, torch.device("cpu"), None)
            self.st.objective_vectors = torch.randn(c["C"], c["K"], generator=torch.Generator().manual_seed(5))
        else:
            self.bank = O.PrototypeBank(c["C"], c["K"])
            self.protos = torch.randn(c["C"], c["K"], generator=torch.Generator().manual_seed(5))
        self.hist = np.zeros((c["C"], c["C"]))

    def fast_hist(self, pred, gt, C):
        if self.ref is not None:
            return self.ref.eval_tools._fast_hist(pred, gt, C)
        return self.O.fast_hist(pred, gt, C)

    def prototypes(self, feat, logits):
        c = self.cfg
        if self.ref is not None:
            vectors, ids = self.cf.calculate_mean_vector(feat, logits)
            for t in range(len(ids)):
                self.cf.update_objective_SingleVector(ids[t], vectors[t].detach().cpu(), "mean")
        else:
            vectors, ids = self.O.class_mean_vectors(feat, logits, c["C"])
            for t, v in zip(ids, vectors):
                self.bank.update(t, v, "mean")

    def label_step(self, logits, feat, gt):
        O, c = self.O, self.cfg
        conf, lab = O.pseudo_label_upsampled(logits, (c["H"], c["W"]))
        lab_u8 = lab.numpy().astype(np.uint8)
        conf16 = conf.numpy().astype(np.float16)
        for i in range(lab_u8.shape[0]):
            O.threshold_hard_label(lab_u8[i], conf16[i], c["thr"], c["ignore"], keep=c["keep"])
        self.prototypes(feat, logits)
        g = gt.numpy()
        for i in range(lab_u8.shape[0]):
            self.hist += self.fast_hist(lab_u8[i].reshape(-1).astype(np.int64), g[i].reshape(-1).astype(np.int64), c["C"])

    def proto_distance(self, logits, feat, soft, hard):
        O, c = self.O, self.cfg
        self.prototypes(feat, logits)
        if self.ref is not None:
            d = self.st.feat_prototype_distance(feat)
            near, _ = d.min(dim=1, keepdim=True)
            w = torch.nn.functional.softmax(-(d - near) * 0.5, dim=1)       # get_prototype_weight :139-142 (no full2weak here)
        else:
            w = O.prototype_weight(feat, self.protos, 0.5)
        O.rectify_labels(w, soft, hard.long(), 0.0, c["ignore"])

    def self_train_round(self, logits, feat, gt, soft, hard):
        O, c = self.O, self.cfg
        self.label_step(logits, feat, gt)
        q = (c["H"] // 4, c["W"] // 4)
        fq = O.upsample_bilinear_ac(feat, q)                                  # self_training.py:69
        if self.ref is not None:
            d = self.st.feat_prototype_distance(fq)
            near, _ = d.min(dim=1, keepdim=True)
            w = torch.nn.functional.softmax(-(d - near) * 0.5, dim=1)
        else:
            w = O.prototype_weight(fq, self.protos, 0.5)
        lab, _ = O.rectify_labels(w, soft, hard.long(), 0.0, c["ignore"])
        x = O.upsample_bilinear_ac(logits, q).requires_grad_(True)
        y = lab.reshape(lab.shape[0], *q)
        if self.ref is not None:
            loss = 0.1 * self.st.seg_loss(x, y) + 1.0 * self.st.rce(x, y.clone())
        else:
            loss = 0.1 * O.cross_entropy(x, y, c["ignore"]) + 1.0 * O.reverse_cross_entropy(x, y.clone(), c["C"], c["ignore"])
        loss.backward()

    def eval_sweep(self, pred, gt):
        c = self.cfg
        p, g = pred.numpy().astype(np.int64), gt.numpy().astype(np.int64)
        for i in range(p.shape[0]):                                           # evaluate() :49-51
            self.hist += self.fast_hist(p[i].flatten(), g[i].flatten(), c["C"])

    def step(self, data):
        getattr(self, self.cfg["kind"])(*data)

    def miou(self):
        return float(self.O.metrics_from_hist(self.hist, self.cfg["C"])[2]) if self.cfg["C"] in (13, 19) else float("nan")


CPU_SAMPLE = {"label_step": 2, "proto_distance": 1, "self_train_round": 2, "eval_sweep": 64}   # frames per CPU step


def run_reference(args, cfg):
    """--impl reference: rank 0 only, all host threads, each step a bounded sample of the workload."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    frames = min(CPU_SAMPLE[cfg["kind"]], cfg["frames"])
    if cfg is CONFIGS["cfg1"]:
        frames = cfg["frames"]
    sets = [make_inputs("cpu", 1234 + s, frames, cfg) for s in range(2)]
    arm = CpuArm(cfg)
    steps = args.steps if args.steps is not None else 20
    if cfg["kind"] == "proto_distance":
        steps = min(steps, 4)                                                # ~15 s of CPU per Cityscapes image
    for i in range(min(args.warmup, 2)):
        arm.step(sets[i % 2])
    t0 = time.perf_counter()
    for i in range(steps):
        arm.step(sets[i % 2])
    dt = time.perf_counter() - t0
    px = frames * (cfg["h"] * cfg["w"] if cfg["kind"] == "proto_distance" else cfg["H"] * cfg["W"]) * steps
    val = px / dt
    sample = "%d of %d frames per step (same shapes, same stages), %d steps" % (frames, cfg["frames"], steps)
    how = ("reference's own functions from baseline/_ref (Class_Features, SelfTrain.*, _fast_hist) + its inline blocks as the "
           "same torch/numpy calls" if arm.kind == "reference" else "oracle port of the reference's torch/numpy op sequence")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": min(args.warmup, 2), "ms_per_step": 1e3 * dt / steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(cfg), "sample": sample, "miou": arm.miou(), "how": how},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": torch.get_num_threads(), "kind": arm.kind, "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def cpu_baseline(cfg, budget_s: float = 12.0):
    """The reference's CPU path timed on this box's host cores on a bounded sample of the same workload (rank 0, N=1)."""
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    frames = min(CPU_SAMPLE[cfg["kind"]], cfg["frames"])
    data = make_inputs("cpu", 1234, frames, cfg)
    arm = CpuArm(cfg)
    arm.step(data)  # warm-up
    reps, t0 = 0, time.perf_counter()
    while reps < 2 or (time.perf_counter() - t0 < budget_s and reps < 50):
        arm.step(data)
        reps += 1
    dt = time.perf_counter() - t0
    px = frames * (cfg["h"] * cfg["w"] if cfg["kind"] == "proto_distance" else cfg["H"] * cfg["W"])
    return {"value": px * reps / dt, "unit": UNIT, "cores": torch.get_num_threads(), "kind": arm.kind,
            "sample": "%d of %d frames per step x %d steps in %.1f s (%s, all host threads)" % (
                frames, cfg["frames"], reps, dt, "reference's own functions from baseline/_ref + its inline blocks"
                if arm.kind == "reference" else "oracle = the reference's own torch/numpy op sequence")}


# ------------------------------------------------------------------------------------------------
def parity_check(cfg, ops, out, data):
    """One frame of the step the bench is about to time, against the oracle (rank 0, before timing)."""
    from oracle import label_path as O

    res = {}
    if cfg["kind"] == "eval_sweep":
        pred, gt = data
        h, bad = ops.confusion_matrix(pred[:8].reshape(-1), gt[:8].reshape(-1), cfg["C"])
        ref = O.fast_hist(pred[:8].cpu().numpy().reshape(-1).astype(np.int64), gt[:8].cpu().numpy().reshape(-1).astype(np.int64), cfg["C"])
        res["hist_exact"] = bool(np.array_equal(h.cpu().numpy(), ref)) and int(bad) == 0
        return res
    logits, feat = data[0][:1].cpu(), data[1][:1].cpu()
    C = cfg["C"]
    if "pl" in out:
        pl = out["pl"]
        up = O.upsample_bilinear_ac(logits, (cfg["H"], cfg["W"]))
        prob = torch.softmax(up, 1)
        conf_ref, lab_ref = prob.max(1)
        top2 = prob.topk(2, dim=1)[0]
        safe = ((top2[:, 0] - top2[:, 1]) > 1e-6)[0].numpy()
        lab = pl.label[0].cpu().numpy()
        res["label_mismatch_outside_1e-6_margin"] = int((lab[safe] != lab_ref[0].numpy()[safe]).sum())
        same = lab == lab_ref[0].numpy()
        c16 = pl.conf16[0].cpu().numpy().astype(np.float32)
        res["conf_max_abs_err_fp16_stored"] = float(np.abs(c16 - conf_ref[0].numpy().astype(np.float16).astype(np.float32))[same].max())
        gt0 = data[2][0].cpu().numpy().reshape(-1).astype(np.int64)
        h, _ = ops.confusion_matrix(pl.label[0].reshape(-1), data[2][0].reshape(-1), C)
        res["hist_exact"] = bool(np.array_equal(h.cpu().numpy(), O.fast_hist(lab.reshape(-1).astype(np.int64), gt0, C)))
        if data[2].dtype == torch.uint8:   # the same matrix from inside the pseudo-label kernel (tss_pseudo_label_hist)
            hf = torch.zeros(C, C, dtype=torch.int64, device=data[0].device)
            plf = ops.pseudo_label(data[0][:1], size=(cfg["H"], cfg["W"]), thr=cfg["thr"], keep=cfg["keep"], ignore_index=cfg["ignore"],
                                   flags=ops.CORESIDENT_PL_FLAGS, gt=data[2][:1], hist=hf)
            res["hist_exact"] = res["hist_exact"] and bool(torch.equal(hf, h)) and bool(torch.equal(plf.label, pl.label[:1]))
    if "means" in out:
        m = out["means"]
        vr, ir = O.class_mean_vectors(feat, logits, C)
        valid = list(np.nonzero(m.valid[0].cpu().numpy())[0])
        err = 0.0
        for t, v in zip(ir, vr):
            a, b = m.vectors[0, t].cpu().numpy().astype(np.float64), v.reshape(-1).numpy().astype(np.float64)
            err = max(err, float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30)))
        res["class_ids_equal"] = valid == ir
        res["proto_mean_max_rel_err"] = err
    if "dist_label" in out:
        protos = out["protos"].cpu()
        soft, hard = data[2][:1].cpu(), data[3][:1].cpu().long()
        w_ref = O.prototype_weight(feat, protos, 0.5)
        lab_ref, rect = O.rectify_labels(w_ref, soft, hard, 0.0, cfg["ignore"])
        d_ref = O.prototype_distance(feat, protos)
        bad = (out["dist_label"][:1].cpu() != lab_ref).numpy()
        srt = np.sort(rect.numpy(), axis=1)
        eta = 2 * 0.5 * 1e-5 * float(d_ref.max())
        ok = (srt[:, -1:] - srt[:, -2:-1]) < 1e-6 + 2 * eta * srt[:, -1:]
        res["rectified_label_mismatch_outside_margin"] = int((bad & ~ok).sum())
        res["rectified_label_mismatch_total"] = int(bad.sum())
    return res


def parity_ok(res) -> bool:
    return (res.get("label_mismatch_outside_1e-6_margin", 0) == 0 and res.get("hist_exact", True)
            and res.get("class_ids_equal", True) and res.get("proto_mean_max_rel_err", 0.0) <= 1e-5
            and res.get("conf_max_abs_err_fp16_stored", 0.0) <= 2.0 ** -11      # one fp16 ulp below 1.0
            and res.get("rectified_label_mismatch_outside_margin", 0) == 0)


# ------------------------------------------------------------------------------------------------
def bind_near_gpu(local_rank: int):
    """Pin this rank's threads (and with them the first-touch placement of its pinned host buffers) to the host NUMA
    node of its GPU when the box has more than one.  Returns a short description for the JSON line."""
    try:
        import pynvml
        pynvml.nvmlInit()
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(local_rank)).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        node_file = "/sys/bus/pci/devices/%s/numa_node" % bus.lower()[-12:]
        node = int(open(node_file).read().strip())
        nodes = [d for d in os.listdir("/sys/devices/system/node") if d.startswith("node") and d[4:].isdigit()]
        if node < 0 or len(nodes) < 2:
            return "single host NUMA node (%d listed): nothing to bind" % len(nodes)
        cpus = []
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            a, _, b = part.partition("-")
            cpus += list(range(int(a), int(b or a) + 1))
        os.sched_setaffinity(0, cpus)
        return "bound to host NUMA node %d (%d cpus)" % (node, len(cpus))
    except Exception as e:  # noqa: BLE001
        return "not bound (%s)" % type(e).__name__


def run_b200(args, cfg):
    import torch.distributed as dist

    from b200seg import dropin, ops, sharding
    from b200seg.graphs import GraphedCall

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the B200 path has no CPU fallback (use --impl reference "
                         "for the CPU arm)")
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    ws = int(os.environ.get("WORLD_SIZE", "1"))
    numa = bind_near_gpu(local_rank) if ws > 1 else "single rank"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if ws > 1:
        # stdout carries ONE JSON line: NCCL prints a version banner to fd 1 when its communicator is created, so
        # the init and the first collective run with fd 1 pointing at stderr
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    c = cfg
    kind = c["kind"]
    frames, C = c["frames"], c["C"]
    size = (c["H"], c["W"])
    steps = args.steps
    if steps is None:
        steps = 50 if "total_frames" not in c else max(3, math.ceil(c["total_frames"] / (frames * ws)))
    warm = max(args.warmup, 3)
    n_sets = N_SETS if kind != "eval_sweep" else 8
    data_note = "synthetic"
    sets = [make_inputs(dev, 1234 + 1000 * rank + s, frames, c) for s in range(n_sets)]
    if c is CONFIGS["cfg1"]:
        dl = deeplab_inputs(c, dev)
        if dl is not None:   # the network's outputs on 4 synthetic frames (random-init weights, torch.manual_seed(0))
            sets = [(dl[0], dl[1], s[2]) for s in sets]
            data_note = "synthetic frames through the reference's random-init DeepLabv2-ResNet101 (baseline/_ref/models)"
        else:
            data_note = "synthetic maps of the DeepLabv2 output shapes (reference model code not staged)"

    has_proto = kind != "eval_sweep"
    bank = sharding.ShardedPrototypes(C, c.get("K", 32), dev, ops.UPDATE_MEAN, True, 0.9999) if has_proto else None
    if kind in ("proto_distance", "self_train_round"):
        bank.protos.copy_(torch.randn(C, c["K"], generator=torch.Generator().manual_seed(5)).to(dev))
        bank.num.fill_(200.0)
        bank.mode = ops.UPDATE_MOVING_AVERAGE            # self_training.py:125-126 (start_mean=False, m = 1e-4)
        bank.start_mean, bank.momentum = False, 1e-4
    meter = dropin.ConfusionMeter(C, dev)
    # N > 1: the mean vectors are exchanged by the finalize kernel's own stores into every peer's gather buffer
    # (sharding.PeerGather, symmetric memory over NVLink); NCCL all_gather is the fallback when peer mapping is
    # not available.  All ranks must agree, so the outcome is all-reduced.
    gather, exchange = None, "none at N=1"
    if ws > 1 and has_proto:
        ok = torch.ones(1, device=dev)
        if args.exchange == "peer":
            try:
                gather = sharding.PeerGather(frames, C, c["K"], dev)
            except Exception as e:  # noqa: BLE001
                print("bench.py: rank %d: peer-memory exchange unavailable (%s: %s), using NCCL all_gather" % (rank, type(e).__name__, e),
                      file=sys.stderr, flush=True)
                ok.zero_()
        else:
            ok.zero_()
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if ok.item() == 0:
            gather = None
        exchange = ("finalize kernel stores its mean vectors into every rank's symmetric-memory gather buffer over NVLink, one barrier, "
                    "sequential update kernel; on the tail stream, beside the next step" if gather is not None else
                    "one NCCL all_gather of the mean vectors + sequential update kernel on the tail stream")

    # capture fork: the HBM-bound accumulate kernel beside the XU-bound pseudo-label kernel; high priority, so that the block
    # scheduler places its one CTA per SM before it fills the SM with pseudo-label CTAs
    side = torch.cuda.Stream(dev, priority=-1)
    # confusion matrix, finalize / exchange / update of step i beside the kernels of step i+1
    tail = torch.cuda.Stream(dev, priority=-1) if args.tail_priority else torch.cuda.Stream(dev)
    co_pl, co_nc = ops.CORESIDENT_PL_FLAGS, ops.CORESIDENT_CONSUMERS
    if args.coresident:
        v = [int(x) for x in args.coresident.split(",")]
        co_pl, co_nc = ops.pl_flags(sm_warps=v[0], cta_warps=v[1], waves=v[2]), v[3]
    identity_weak = None
    if kind == "self_train_round":
        q = (c["H"] // 4, c["W"] // 4)
        identity_weak = {"RandomSized": [torch.full((frames,), 4 * q[0]), torch.full((frames,), 4 * q[1])],
                         "RandomCrop": [torch.zeros(frames), torch.full((frames,), 4.0 * q[0]), torch.zeros(frames),
                                        torch.full((frames,), 4.0 * q[1])],
                         "RandomHorizontallyFlip": [False] * frames}
        identity_weak = ops.weak_table(identity_weak, frames, dev)      # device table built once, outside the graph capture

    # ---- stages
    def stage_pl(i, flags=0, with_hist=False):
        # with_hist: the confusion matrix of (gt, label) is accumulated inside the pseudo-label kernel (tss_pseudo_label_hist)
        return ops.pseudo_label(sets[i % n_sets][0], size=size, thr=c["thr"], keep=c["keep"], ignore_index=c["ignore"], flags=flags,
                                gt=sets[i % n_sets][2] if with_hist else None, hist=meter.hist if with_hist else None)

    def stage_sums(i, consumer_warps=0, accumulate_stream=None):
        return ops.class_sums(sets[i % n_sets][1], sets[i % n_sets][0], C, consumer_warps=consumer_warps,
                              accumulate_stream=accumulate_stream)

    def stage_hist(i, pl):
        ops.confusion_matrix(pl.label.reshape(-1), sets[i % n_sets][2].reshape(-1), C, meter.hist, meter.bad)

    def stage_tail(i, cs, pl=None):
        """confusion matrix of the step's labels, then means + validity + sequential update; at N > 1 the exchange in
        between (slots alternate with i).  Nothing here is read by the next step's kernels: it all runs beside them."""
        if pl is not None:
            stage_hist(i, pl)
        if ws == 1:
            return ops.class_finalize(cs, update=(bank.protos, bank.num, bank.mode, bank.start_mean, bank.momentum))
        if gather is not None:
            means = ops.class_finalize(cs, scatter=gather.targets(i % 2))
            bank.update_from_peers(gather, i % 2)
            return means
        return ops.class_finalize(cs)          # NCCL form: bank.update_gathered(means) follows outside the graph

    def stage_distance(i):
        logits, feat, soft, hard = sets[i % n_sets]
        return ops.prototype_distance(feat, bank.protos, temperature=0.5, want_dist=False, want_weight=False, soft=soft,
                                      hard=hard, train_thred=0.0, ignore_index=c["ignore"])[2]

    def stage_round_tail(i):
        """cfg4: rectified labels from the 1/8 features (fused sampling), then CE + RCE forward and backward at 1/4"""
        logits, feat, gt, soft, hard = sets[i % n_sets]
        q = (c["H"] // 4, c["W"] // 4)
        lab = ops.rectify_from_stride8(feat, q, identity_weak, bank.protos, 0.5, soft, hard, 0.0, c["ignore"])[1]
        x = ops.upsample_bilinear(logits, q).requires_grad_(True)
        loss = ops.ce_rce(x, lab.reshape(frames, *q), c["ignore"], 0.1, 1.0)[0]
        g, = torch.autograd.grad(loss, x)
        return lab, loss, ops.upsample_bilinear_adjoint(g, (c["h"], c["w"]))

    def pre_step(i):
        """coresident_prelabel: the label kernel of step i (logits -> tile records, zeroed sums).  It needs nothing of step
        i - 1, so it is issued on its own stream BESIDE the kernels of step i - 1 instead of in front of step i's."""
        return ops.class_labels(sets[i % n_sets][0], C, c["K"])

    def main_step(i, mode, pre=None):
        """everything of step i that the NEXT step's kernels must not overtake; returns what the tail needs"""
        cur = torch.cuda.current_stream(dev)
        if kind == "eval_sweep":
            pred, gt = sets[i % n_sets]
            ops.confusion_matrix(pred.reshape(-1), gt.reshape(-1), C, meter.hist, meter.bad)
            return None, None
        if kind == "proto_distance":
            cs = stage_sums(i)
            lab = stage_distance(i)
            return cs, lab
        if mode == "coresident_prelabel":
            side.wait_stream(cur)                # accumulate kernel (records from the pre graph) on the high-priority fork
            with torch.cuda.stream(side):
                cs = ops.class_accumulate(pre.out, sets[i % n_sets][1], co_nc)
            pl = stage_pl(i, co_pl)
            extra = stage_round_tail(i) if kind == "self_train_round" else None
            cur.wait_stream(side)
        elif mode in ("coresident", "coresident_hist_main", "coresident_hist_fused"):
            cs = stage_sums(i, co_nc, side)      # label kernel here, accumulate kernel on the high-priority fork
            pl = stage_pl(i, co_pl, with_hist=(mode == "coresident_hist_fused"))
            if mode == "coresident_hist_main":   # confusion matrix behind the pseudo-label kernel instead of on the tail stream
                stage_hist(i, pl)
            extra = stage_round_tail(i) if kind == "self_train_round" else None
            cur.wait_stream(side)
        else:
            pl = stage_pl(i)
            cs = stage_sums(i)
            extra = stage_round_tail(i) if kind == "self_train_round" else None
        return cs, (pl, extra)

    def build_graphs(mode):
        pres = [GraphedCall(lambda i=i: pre_step(i), warmup=1) for i in range(n_sets)] if mode == "coresident_prelabel" else None
        mains = [GraphedCall(lambda i=i: main_step(i, mode, pres[i] if pres else None), warmup=1) for i in range(n_sets)]
        hist_in_tail = kind != "proto_distance" and mode not in ("coresident_hist_main", "coresident_hist_fused")
        tails = [GraphedCall(lambda i=i: stage_tail(i, mains[i].out[0], mains[i].out[1][0] if hist_in_tail else None),
                             warmup=1) for i in range(n_sets)] if has_proto else None
        return mains, tails, pres

    ev_main = [torch.cuda.Event() for _ in range(n_sets)]
    ev_tail = [torch.cuda.Event() for _ in range(n_sets)]
    ev_pre = [torch.cuda.Event() for _ in range(n_sets)]
    pre_stream = torch.cuda.Stream(dev)
    # --main-streams 2: consecutive steps' kernel graphs alternate between two streams, so that step i+1's kernels fill the
    # SMs while step i's last CTAs drain (steps read different input sets; the prototype bank is updated in order on the
    # tail stream either way)
    main_streams = [torch.cuda.Stream(dev) for _ in range(args.main_streams)] if args.main_streams > 1 else []
    state = {"issued": 0, "pre_for": -1}

    def join_all():
        cur = torch.cuda.current_stream(dev)
        for st_ in [tail, pre_stream] + main_streams:
            cur.wait_stream(st_)

    def make_step(mains, tails, pres=None):
        def issue_pre(j):
            # label kernel of step j on its own stream; its set's records / sums were last read by the tail of step j - n_sets
            with torch.cuda.stream(pre_stream):
                if tails is not None and j >= n_sets:
                    pre_stream.wait_event(ev_tail[j % n_sets])
                pres[j % n_sets].replay()
                ev_pre[j % n_sets].record(pre_stream)
            state["pre_for"] = j

        def step(_i=None):
            # the set (and with it the gather slot, set % 2) follows ONE running counter, whichever loop calls: consecutive
            # steps always alternate the slot
            i = state["issued"]
            if main_streams:
                cur = main_streams[i % len(main_streams)]
                with torch.cuda.stream(cur):
                    step_on(cur, i)
            else:                                        # (no stream context: cfg1's 22 us steps are bound by host launch time)
                step_on(torch.cuda.current_stream(dev), i)
            state["issued"] += 1

        def step_on(cur, i):
            s = i % n_sets
            if tails is not None and state["issued"] >= n_sets:
                cur.wait_event(ev_tail[s])              # the tail of step i - n_sets has read this set's sums / records
            if pres is not None:
                if state["pre_for"] != i:               # first step of this form: nothing was issued ahead
                    issue_pre(i)
                cur.wait_event(ev_pre[s])
            mains[s].replay()
            if pres is not None:
                issue_pre(i + 1)                        # beside this step's kernels
            if tails is not None:
                ev_main[s].record(cur)
                with torch.cuda.stream(tail):
                    tail.wait_event(ev_main[s])
                    means = tails[s].replay()
                    if ws > 1 and gather is None:
                        bank.update_gathered(means)
                    ev_tail[s].record(tail)
        return step

    def fence():
        join_all()
        torch.cuda.synchronize()
        if ws > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def time_steps(step, n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        fence()
        e0.record()
        for i in range(n):
            step(i)
        join_all()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if ws > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) / n

    # ---- choose the step form: the co-resident form (pseudo-label on 12 warps/SM + accumulate on 5) against the serial
    # one, each timed over a few untimed-region replays; every rank takes the same decision (all-reduced times)
    modes = ["serial"] if kind in ("eval_sweep", "proto_distance") else (
        (["coresident", "coresident_prelabel", "coresident_hist_main", "serial"] if ws > 1 else
         ["coresident", "coresident_prelabel", "serial"]) if args.overlap == "auto"
        else [args.overlap])   # N > 1: the longer tail (exchange, update over every rank's images) prefers the confusion matrix on the main stream
    trial = {}
    built = {}
    for m in modes:
        built[m] = build_graphs(m)
        st = make_step(*built[m])
        for i in range(2 * n_sets):      # consecutive steps must alternate the gather slot: always whole rounds of the sets
            st(i)
        # N > 1: the ranks enter a trial with some skew and the exchange barrier turns it into step time, so the trials are
        # longer there (and the better of two counts) -- a 12-step trial could not tell forms 10 us apart
        trial[m] = time_steps(st, 3 * n_sets) if ws == 1 else min(time_steps(st, 12 * n_sets), time_steps(st, 12 * n_sets))
    mode = min(trial, key=trial.get)
    for m in list(built):
        if m != mode:
            del built[m]
    mains, tails, pres = built[mode]
    step = make_step(mains, tails, pres)
    launches_per_step = mains[0].launches + (tails[0].launches if tails else 0) + (pres[0].launches if pres else 0) + (1 if ws > 1 and has_proto and gather is None else 0)

    # ---- G consecutive steps as ONE graph (--steps-per-graph): the same kernels with the same dependencies -- label kernel
    # of step j+1 beside step j, accumulate fork, tail stream -- captured once for G steps, so that the host issues one
    # launch per G steps instead of three graph launches and six event calls per step.  Built to test whether the step loop
    # at N = 8 (0.18 ms/step in every form, 0.154 at N = 1) is bound by host launch work: it is not (see below).
    class _Held:
        def __init__(self, out):
            self.out = out

    def capture_block(G):
        keep = []
        hist_in_tail = mode not in ("coresident_hist_main", "coresident_hist_fused")
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            cur = torch.cuda.current_stream(dev)
            pre_j, ev_pre_j = None, None

            def fork_pre(j):
                pre_stream.wait_stream(cur)
                with torch.cuda.stream(pre_stream):
                    held = _Held(pre_step(j))
                    ev = torch.cuda.Event()
                    ev.record(pre_stream)
                keep.append(held)
                return held, ev
            if mode == "coresident_prelabel":
                pre_j, ev_pre_j = fork_pre(0)
            for j in range(G):
                if mode == "coresident_prelabel":
                    cur.wait_event(ev_pre_j)
                    this_pre = pre_j
                    if j + 1 < G:
                        pre_j, ev_pre_j = fork_pre(j + 1)     # beside step j's kernels
                    cs, rest = main_step(j, mode, this_pre)
                else:
                    cs, rest = main_step(j, mode)
                keep.append((cs, rest))
                ev = torch.cuda.Event()
                ev.record(cur)
                with torch.cuda.stream(tail):
                    tail.wait_event(ev)
                    keep.append(stage_tail(j, cs, rest[0] if hist_in_tail else None))
            for st_ in ((tail, pre_stream) if mode == "coresident_prelabel" else (tail,)):
                cur.wait_stream(st_)          # every stream that joined the capture returns to it
        return g, keep

    block = {"graph": None, "G": 0}

    def run_steps(n):
        """issue exactly n steps: whole blocks where a block graph exists (they start on an even step: the gather slot
        alternates with the step index), single steps otherwise"""
        done = 0
        while done < n:
            G = block["G"]
            if block["graph"] is not None and n - done >= G and state["issued"] % 2 == 0:
                join_all()                       # after whatever single steps left on the tail / label streams
                block["graph"].replay()
                state["issued"] += G
                state["pre_for"] = -1
                done += G
            else:
                step()
                done += 1

    def time_run(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        fence()
        e0.record()
        run_steps(n)
        join_all()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if ws > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) / n

    # ---- in-run parity: one frame of this very step against the oracle, before anything is timed
    parity = None
    if rank == 0 and not args.no_parity:
        out = {}
        if kind == "eval_sweep":
            pass
        elif kind == "proto_distance":
            out["means"] = ops.class_finalize(stage_sums(0))
            out["dist_label"], out["protos"] = stage_distance(0), bank.protos.clone()
        else:
            out["pl"] = stage_pl(0, co_pl if mode != "serial" else 0)
            out["means"] = ops.class_finalize(stage_sums(0, co_nc if mode != "serial" else 0))
        torch.cuda.synchronize()
        parity = parity_check(c, ops, out, sets[0])
        if not parity_ok(parity):
            raise SystemExit("bench.py: the step's outputs do not match the oracle: %s" % json.dumps(parity))
    if ws > 1:
        fence()

    if ws > 1:
        # setup, not warm-up steps: the first multi-GPU replays also pay for peer mappings, NCCL channel set-up and
        # rank skew; keep replaying in chunks of 12 steps until two consecutive chunks agree within 10 % on the
        # slowest rank (the decision is all-reduced: every rank replays the same number of steps), at most 40 chunks
        prev = None
        for _ in range(40):
            cur_t = time_steps(step, 12)
            if prev is not None and abs(cur_t - prev) <= 0.1 * min(cur_t, prev):
                break
            prev = cur_t
        scratch = meter.hist.clone()
        sharding.all_reduce_confusion(scratch, meter.bad.clone())
        fence()
    for i in range(((warm + n_sets - 1) // n_sets) * n_sets):
        step(i)
    fence()
    block_trial = None
    want_G = args.steps_per_graph
    # (measured: 0.163 vs 0.153 ms at N = 1, 0.166 vs 0.166 at N = 2, 0.183 vs 0.180 at N = 8 -- the host keeps up with one graph
    # per step at every N, so the default stays 1; 0 = time both and keep the faster)
    if kind == "label_step" and has_proto and mode != "serial" and not (ws > 1 and gather is None) and want_G != 1:
        G = want_G if want_G > 1 else 8
        G -= G % 2
        n_try = 6 * G
        t_single = min(time_run(n_try), time_run(n_try))
        block["graph"], block["keep"] = capture_block(G)
        block["G"] = G
        run_steps(2 * G)
        t_block = min(time_run(n_try), time_run(n_try))
        block_trial = {"steps_per_graph": G, "single_ms": round(t_single, 5), "block_ms": round(t_block, 5)}
        if want_G <= 1 and t_block >= t_single:          # automatic: keep the blocks only where they are faster
            block["graph"], block["G"] = None, 0
        fence()

    # ---- device-resident timed region
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
        time.sleep(0.35)
    meter.hist.zero_()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    fence()
    t_wall0 = time.perf_counter()
    ev0.record()
    run_steps(steps)
    join_all()    # the timed region ends when the last exchange / update has landed (and the label kernel issued ahead by the last step)
    sharding.all_reduce_confusion(meter.hist, meter.bad)
    ev1.record()
    fence()
    t_wall1 = time.perf_counter()
    launches = launches_per_step * steps
    ms = torch.tensor([ev0.elapsed_time(ev1)], device=dev)
    ms_per_rank = [float(ms.item())]
    if ws > 1:
        every = [torch.empty_like(ms) for _ in range(ws)]
        dist.all_gather(every, ms)
        ms_per_rank = [float(t.item()) for t in every]
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    with np.errstate(all="ignore"):
        miou = float(dropin.metrics_from_hist(meter.result(), C)[2]) if kind != "proto_distance" else None

    # ---- N > 1: the replicated bank must be bit-identical on every rank, and equal to what ONE process gets from the
    # sequential update over all ranks' images in global order (rank 0 gathers two steps' inputs and recomputes)
    bank_spread, bank_vs_seq = 0.0, None
    if ws > 1 and has_proto:
        hi, lo = bank.protos.clone(), bank.protos.clone()
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        bank_spread = float((hi - lo).abs().max().item())
        bank.protos.zero_()
        bank.num.zero_()
        fence()
        vsteps, base = 2, state["issued"]
        for i in range(vsteps):
            step(i)
        fence()
        ref_p, ref_n = torch.zeros_like(bank.protos), torch.zeros_like(bank.num)
        for i in range(vsteps):
            lg, ft = sets[(base + i) % n_sets][0], sets[(base + i) % n_sets][1]
            all_lg = [torch.empty_like(lg) for _ in range(ws)] if rank == 0 else None
            all_ft = [torch.empty_like(ft) for _ in range(ws)] if rank == 0 else None
            dist.gather(lg, all_lg, dst=0)
            dist.gather(ft, all_ft, dst=0)
            if rank == 0:
                seq = ops.class_means(torch.cat(all_ft), torch.cat(all_lg), C)
                ops.prototype_update(ref_p, ref_n, seq, bank.mode, bank.start_mean, bank.momentum, bank.cap)
                del all_lg, all_ft, seq
        if rank == 0:
            bank_vs_seq = float(((bank.protos - ref_p).abs().max() / ref_p.abs().max().clamp_min(1e-30)).item())
            if not torch.equal(bank.num, ref_n) or bank_vs_seq > 1e-5:
                raise SystemExit("bench.py: sharded prototype bank differs from the sequential update: rel %.3e" % bank_vs_seq)
        fence()

    # ---- per-stage device time: one graph per (stage, input set), CUDA events around every replay -> roofline
    stage_ms = {}
    if kind == "eval_sweep":
        stage_ms["confusion"] = ms / steps
    else:
        stage_defs = []
        if kind != "proto_distance":
            plg = [GraphedCall(lambda i=i: stage_pl(i)) for i in range(n_sets)]
            stage_defs.append(("pseudo_label", plg))
        prg = [GraphedCall(lambda i=i: ops.class_finalize(stage_sums(i), update=(bank.protos, bank.num, bank.mode, bank.start_mean,
                                                                                 bank.momentum)), warmup=1) for i in range(n_sets)]
        stage_defs.append(("prototype", prg))
        if kind == "proto_distance":
            stage_defs.append(("distance_rectify", [GraphedCall(lambda i=i: stage_distance(i)) for i in range(n_sets)]))
        else:
            stage_defs.append(("confusion", [GraphedCall(lambda i=i: stage_hist(i, plg[i].out)) for i in range(n_sets)]))
        if kind == "self_train_round":
            stage_defs.append(("rectify_ce_rce", [GraphedCall(lambda i=i: stage_round_tail(i)) for i in range(n_sets)]))
        reps = min(steps, 50)
        fence()
        evs = []
        for i in range(reps):
            e = [torch.cuda.Event(enable_timing=True) for _ in range(len(stage_defs) + 1)]
            e[0].record()
            for j, (_, graphs) in enumerate(stage_defs):
                graphs[i % n_sets].replay()
                e[j + 1].record()
            evs.append(e)
        fence()
        for j, (name, _) in enumerate(stage_defs):
            stage_ms[name] = sum(e[j].elapsed_time(e[j + 1]) for e in evs) / reps
        del stage_defs
    aux = aux_stages(ops, dev, GraphedCall) if (c is CONFIGS["cfg2"] and rank == 0 and not args.no_aux) else None
    if ws > 1:
        fence()
    clocks = sampler.finish(t_wall0, time.perf_counter()) if sampler else None

    # ---- end to end through the drop-in API: pinned host buffers, H2D + D2H inside the timed region
    e2e = run_e2e(args, c, dev, rank, ws, dist, dropin, ops, sharding, steps, gather is not None, numa)

    if rank != 0:
        if ws > 1:
            dist.destroy_process_group()
        return
    unit_px = (c["h"] * c["w"] if kind == "proto_distance" else c["H"] * c["W"])
    px_step = frames * unit_px * ws
    alg = algorithmic_bytes(c)
    peak, peak_src = measured_peak_gbs()
    top = max(stage_ms, key=stage_ms.get)
    achieved = alg.get(top, 0) / (stage_ms[top] * 1e-3) / 1e9 if top in alg else None
    step_bytes = sum(alg.values())
    roof = {"bound": "xu" if top == "pseudo_label" else "hbm", "kernel": top, "achieved": achieved, "peak": peak, "unit": "GB/s",
            "frac": (achieved / peak) if achieved else None, "traffic": ncu_traffic(top), "peak_source": peak_src,
            "algorithmic_bytes_per_launch": alg.get(top),
            "step": {"algorithmic_bytes": step_bytes, "achieved": step_bytes / (ms / steps * 1e-3) / 1e9,
                     "frac": step_bytes / (ms / steps * 1e-3) / 1e9 / peak}}
    if top == "pseudo_label":
        # MUFU floor of the fused low-resolution form: C ex2 + 1 rcp per pixel at 16 lanes/clk/SM on 148 SMs at the sampled clock
        mhz = (clocks or {}).get("sm_mhz") or 1965
        floor_ms = frames * c["H"] * c["W"] * (C + 1) / (148 * 16 * mhz * 1e6) * 1e3
        roof["xu_floor_ms"] = floor_ms
        roof["xu_frac"] = floor_ms / stage_ms[top]
        roof["note"] = ("the dominant stage (fused upsample->softmax->argmax->threshold from LOW-RES logits) is bound by MUFU.EX2 + "
                        "issue slots, not HBM (`bound: xu`): frac = HBM fraction of its 4 B/px algorithmic traffic, xu_frac = time of "
                        "C ex2 + 1 rcp per pixel at 16 lanes/clk/SM over measured time; in the timed step it shares the SMs with the "
                        "HBM-bound accumulate kernel (roofline.step = whole step against HBM)")
    out = {
        "metric": METRIC, "value": px_step * steps / (ms * 1e-3), "unit": UNIT, "n_gpus": ws, "steps": steps,
        "warmup": warm, "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8" if kind == "eval_sweep" else "f32", "data": data_note,
        "config": {"workload": workload_name(c), "frames_per_gpu_per_step": frames,
                   "l2": "inputs rotate over %d resident sets (working set %d MB > 126 MB L2)" % (
                       n_sets, n_sets * sum(t.numel() * t.element_size() for t in sets[0]) // 2 ** 20),
                   "launch": ("ONE CUDA-graph replay per %d consecutive steps (kernels, label-ahead and tail streams captured with their "
                              "dependencies; leftover steps as below); per step: " % block["G"] if block["graph"] is not None else "") +
                             "one CUDA-graph replay per step for the kernels (%s form: %s) + one for the prototype tail on its own "
                             "stream, %d kernels per step" % (mode, {"coresident": "pseudo-label kernel in 4-warp CTAs, two waves, and the "
                             "accumulate kernel (6 consumer warps, high-priority capture fork) resident on the same SMs; confusion "
                             "matrix on the tail stream", "coresident_hist_main": "as coresident, confusion matrix behind the "
                             "pseudo-label kernel", "coresident_hist_fused": "as coresident, confusion matrix accumulated inside "
                             "the pseudo-label kernel (tss_pseudo_label_hist)", "coresident_prelabel": "as coresident, and the label "
                             "kernel of step i+1 (logits -> tile records) issued on its own stream beside the kernels of step i: every "
                             "step still runs one label kernel inside the timed region", "serial": "one stream"}[mode],
                             launches_per_step),
                   "step_form_trials_ms": {k: round(v, 5) for k, v in trial.items()},
                   "steps_per_graph": block["G"] if block["graph"] is not None else 1, "steps_per_graph_trial": block_trial,
                   "parallelism": "image-sharded x%d; collectives: prototype means [%d,%d,%s] fp32 per rank and step (%s), "
                                  "confusion [%d,%d] int64 once" % (ws, frames, C, c.get("K", "-"), exchange, C, C),
                   "miou": miou, "bank_max_spread_over_ranks": bank_spread, "bank_vs_sequential_rel": bank_vs_seq,
                   "ms_per_step_by_rank": [round(t / steps, 5) for t in ms_per_rank],
                   "host_numa": numa},
        "parity_checked": parity is not None, "parity": parity,
        "stages_ms": {k: round(v, 5) for k, v in stage_ms.items()},
        "stages_gbs": {k: round(alg[k] / (stage_ms[k] * 1e-3) / 1e9, 1) for k in stage_ms if k in alg},
        "roofline": roof,
        "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
        "wall_ms_per_step": 1e3 * (t_wall1 - t_wall0) / steps,
    }
    if aux:
        out["aux_stages"] = aux
    if ws == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(c)
    print(json.dumps(out))
    if ws > 1:
        dist.destroy_process_group()


def aux_stages(ops, dev, GraphedCall):
    """Driver-timed numbers for the kernels the cfg2 step does not contain (A7-A12, cfg5's histogram), at their BASELINE
    shapes: device time between CUDA-graph replays over rotating inputs, algorithmic bytes, fraction of the HBM peak."""
    peak, _ = measured_peak_gbs()
    res = {}

    def run(name, fn, sets, alg_bytes, reps=20):
        graphs = [GraphedCall(lambda s=s: fn(*s)) for s in sets]
        for i in range(3):
            graphs[i % len(graphs)].replay()
        torch.cuda.synchronize()
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
        evs[0].record()
        for i in range(reps):
            graphs[i % len(graphs)].replay()
            evs[i + 1].record()
        torch.cuda.synchronize()
        t = sorted(evs[i].elapsed_time(evs[i + 1]) for i in range(reps))[reps // 2]
        res[name] = {"ms": round(t, 5), "algorithmic_MB": round(alg_bytes / 1e6, 1), "GBps": round(alg_bytes / (t * 1e-3) / 1e9, 1),
                     "frac_of_hbm_peak": round(alg_bytes / (t * 1e-3) / 1e9 / peak, 4)}

    g = torch.Generator(device=dev).manual_seed(77)
    # cfg3: distance -> weights -> rectified labels, 2 and 8 Cityscapes images, K = 2048, C = 19
    for n in (2, 8):
        K, C, P = 2048, 19, 129 * 257
        sets = []
        for _ in range(2):
            sets.append((torch.randn(n, K, 129, 257, generator=g, device=dev), torch.randn(C, K, generator=g, device=dev),
                         torch.softmax(3 * torch.randn(n, C, 129, 257, generator=g, device=dev), 1),
                         torch.full((n, 1, 129, 257), 255, dtype=torch.uint8, device=dev)))
        run("distance_rectify_cfg3_k2048_%dimg" % n,
            lambda f, p, s, h: ops.prototype_distance(f, p, 0.5, want_dist=False, soft=s, hard=h, train_thred=0.0, ignore_index=255),
            sets, n * P * (4 * K + 4 * C + 1 + 8))
        del sets
    # CE + RCE forward / backward, 4 x 13 x 640 x 1920, int64 targets (what the reference passes)
    n, C, P = 4, 13, 640 * 1920
    sets = [(torch.randn(n, C, 640, 1920, generator=g, device=dev),
             torch.randint(0, C, (n, 640, 1920), generator=g, device=dev, dtype=torch.int64)) for _ in range(2)]
    run("ce_rce_fwd_4x13x640x1920", lambda x, y: ops.ce_rce(x, y, 255, 0.1, 1.0)[0], sets, n * P * (4 * C + 8))

    def fwd_bwd(x, y):
        xr = x.detach().requires_grad_(True)
        return torch.autograd.grad(ops.ce_rce(xr, y, 255, 0.1, 1.0)[0], xr)[0]
    run("ce_rce_fwd_bwd_4x13x640x1920", fwd_bwd, sets, n * P * (4 * C + 8) * 2 + n * P * 4 * C)
    del sets
    # cfg5: confusion matrix over 1024 uint8 maps 256 x 512, 19 classes
    sets = [make_inputs(dev, 88 + s, 1024, CONFIGS["cfg5"]) for s in range(2)]
    hist = torch.zeros(19, 19, dtype=torch.int64, device=dev)
    bad = torch.zeros(1, dtype=torch.int64, device=dev)
    run("confusion_cfg5_1024x256x512_u8", lambda p, t: ops.confusion_matrix(p.reshape(-1), t.reshape(-1), 19, hist, bad), sets,
        1024 * 256 * 512 * 2)
    return res


def run_e2e(args, c, dev, rank, ws, dist, dropin, ops, sharding, steps, use_peer, numa):
    """The same step through the public host-fed API: every step's inputs come from pinned HOST memory (H2D inside
    the timed region), every step's results go back to pinned host memory (D2H inside the timed region).  At N > 1 the
    step is the FULL step: the prototype exchange runs through sharding.PeerGather on the pipeline's tail stream."""
    kind = c["kind"]
    frames, C = c["frames"], c["C"]
    steps = min(steps, 50)
    if kind == "eval_sweep":
        host = [tuple(t.pin_memory() for t in make_inputs("cpu", 4321 + 1000 * rank + s, frames, c)) for s in range(2)]
        meter = dropin.ConfusionMeter(C, dev)
        s_in = torch.cuda.Stream(dev)
        slots = [tuple(torch.empty_like(t, device=dev) for t in host[0]) for _ in range(2)]
        ready = [torch.cuda.Event() for _ in range(2)]
        free = [torch.cuda.Event() for _ in range(2)]

        def submit(i):
            s = i % 2
            cur = torch.cuda.current_stream(dev)
            with torch.cuda.stream(s_in):
                if i >= 2:
                    s_in.wait_event(free[s])
                for d, h in zip(slots[s], host[s]):
                    d.copy_(h, non_blocking=True)
                ready[s].record(s_in)
            cur.wait_event(ready[s])
            meter.add(slots[s][0], slots[s][1])
            free[s].record(cur)

        for i in range(3):
            submit(i)
        torch.cuda.synchronize()
        if ws > 1:
            dist.barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        for i in range(steps):
            submit(i)
        sharding.all_reduce_confusion(meter.hist, meter.bad)
        hist = meter.result()
        ev1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([ev0.elapsed_time(ev1)], device=dev)
        if ws > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        ms = float(ms.item())
        return {"value": frames * c["H"] * c["W"] * ws * steps / (ms * 1e-3), "unit": UNIT,
                "h2d_bytes_per_step": int(sum(t.numel() * t.element_size() for t in host[0])),
                "d2h_bytes_per_step": int(hist.nbytes // max(steps, 1)), "ms_per_step": ms / steps,
                "api": "b200seg.dropin.ConfusionMeter.add on double-buffered device slots fed from pinned host label maps"}
    if kind != "label_step":
        return {"value": None, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                "note": "end-to-end arm is reported for the label-step configurations (cfg1, cfg2) and cfg5"}
    from b200seg.pipeline import HostLabelPipeline

    size = (c["H"], c["W"])
    host = []
    for s in range(2):
        lg, ft, gt = make_inputs("cpu", 4321 + 1000 * rank + s, frames, c)
        host.append((lg.pin_memory(), ft.pin_memory(), gt.pin_memory()))
    bank = gather = None
    if ws > 1 and use_peer:
        bank = sharding.ShardedPrototypes(C, c["K"], dev, ops.UPDATE_MEAN, True, 0.9999)
        gather = sharding.PeerGather(frames, C, c["K"], dev)
    pipe = HostLabelPipeline(dev, C, c["K"], size, thr=c["thr"], keep=c["keep"], ignore_index=c["ignore"],
                             update_prototypes=True, overlap=True, bank=bank, gather=gather)
    full_step = ws == 1 or gather is not None   # without peer memory the e2e arm keeps the prototype update rank-local

    for i in range(4):
        pipe.submit(*host[i % 2])
    pipe.join()
    torch.cuda.synchronize()
    if ws > 1:
        dist.barrier()

    # bare-copy ceiling: the same bytes per step, no kernels -- what the host / PCIe side of this box can deliver to all
    # ranks at once
    s_in, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    d_in = [torch.empty_like(t, device=dev) for t in host[0]]
    d_lab = torch.empty((frames,) + size, dtype=torch.uint8, device=dev)
    d_conf = torch.empty((frames,) + size, dtype=torch.float16, device=dev)

    def bare(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        if ws > 1:
            dist.barrier()
        e0.record()
        for i in range(n):
            with torch.cuda.stream(s_in):
                for d, h in zip(d_in, host[i % 2]):
                    d.copy_(h, non_blocking=True)
            with torch.cuda.stream(s_out):
                pipe.out_label[i % 2].copy_(d_lab, non_blocking=True)
                pipe.out_conf[i % 2].copy_(d_conf, non_blocking=True)
        cur = torch.cuda.current_stream(dev)
        cur.wait_stream(s_in)
        cur.wait_stream(s_out)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if ws > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) / n

    bare(3)
    bare_ms = bare(10)

    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    if ws > 1:
        dist.barrier()
    ev0.record()
    for i in range(steps):
        pipe.submit(*host[i % 2])
    pipe.join()
    sharding.all_reduce_confusion(pipe.meter.hist, pipe.meter.bad)
    hist = pipe.meter.result()       # D2H of the [C,C] matrix: the metric the user reads
    ev1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([ev0.elapsed_time(ev1)], device=dev)
    if ws > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    h2d = int(pipe.h2d_bytes)
    return {"value": frames * c["H"] * c["W"] * ws * steps / (ms * 1e-3), "unit": UNIT,
            "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": int(pipe.d2h_bytes + hist.nbytes // max(steps, 1)),
            "ms_per_step": ms / steps,
            "bare_copy_ms_per_step": bare_ms, "frac_of_bare_copy_ceiling": bare_ms / (ms / steps),
            "h2d_GBps_aggregate": h2d * ws / (ms / steps * 1e-3) / 1e9, "host_numa": numa,
            "full_step_at_every_n": bool(full_step),
            "api": "b200seg.pipeline.HostLabelPipeline.submit (pseudo-label kernel beside ops.class_sums, ConfusionMeter.add, "
                   "ops.class_finalize + prototype update%s on the tail stream; upload / kernels / tail / download on five "
                   "streams, two slots)" % (" through sharding.PeerGather" if gather is not None else "")}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="cfg2", choices=sorted(CONFIGS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the in-run oracle check of one frame")
    ap.add_argument("--no-aux", action="store_true", help="skip the aux_stages timings of the default line")
    ap.add_argument("--overlap", default="auto", choices=["auto", "coresident", "coresident_prelabel", "coresident_hist_main", "coresident_hist_fused", "serial"],
                    help="step form: pseudo-label and accumulate kernels resident together, one after the other, or the faster "
                         "of the two measured before the timed region (default)")
    ap.add_argument("--steps-per-graph", type=int, default=1,
                    help="G > 1: G consecutive steps are captured as one graph (same kernels and dependencies, one host launch per G "
                         "steps); 1 (default): one graph per step; 0: both are timed before the timed region and the faster is used")
    ap.add_argument("--main-streams", type=int, default=1, help="2 = consecutive steps' kernel graphs alternate between two streams (tuning)")
    ap.add_argument("--tail-priority", type=int, default=0, help="1 = the tail stream gets the high stream priority (tuning)")
    ap.add_argument("--coresident", default="", help="sm_warps,cta_warps,waves,consumer_warps of the co-resident form (tuning)")
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="N > 1 prototype exchange: peer-memory stores from the finalize kernel (default) or NCCL all_gather")
    args = ap.parse_args()
    cfg = CONFIGS[args.config]
    if args.impl == "reference":
        run_reference(args, cfg)
    else:
        run_b200(args, cfg)


if __name__ == "__main__":
    main()
