"""Per-kernel throughput of every stage of the label path at the BASELINE.json shapes.

usage: python tools/bench_stages.py [--reps R] [--only NAME[,NAME]] > stages.jsonl

One JSON line per stage: device time (CUDA events between back-to-back CUDA-graph replays of the stage's launches,
median of R after 3 warm-ups),
algorithmic bytes (each input read once, each output written once), achieved GB/s and the fraction of the
measured HBM peak.  Inputs rotate over enough resident copies that consecutive iterations never re-read the
same bytes out of the 126 MB L2 (or the input itself is larger than L2).
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from b200seg import ops  # noqa: E402
from b200seg.graphs import GraphedCall  # noqa: E402

L2 = 126e6
dev = torch.device("cuda:0")
PEAK, PEAK_SRC = bench.measured_peak_gbs()


def copies_for(nbytes: float) -> int:
    return max(2, int(2.5 * L2 // max(nbytes, 1)) + 1) if nbytes < 2.5 * L2 else 2


def timed(fn, sets, reps):
    """fn(*set) captured once per resident input set into a CUDA graph (b200seg.graphs); `reps` replays are
    enqueued back to back with an event between consecutive replays, so the host's launch cost stays off the
    device timeline.  Returns (median, min) microseconds per replay."""
    graphs = [GraphedCall(lambda st=st: fn(*st)) for st in sets]
    for i in range(3):
        graphs[i % len(graphs)].replay()
    torch.cuda.synchronize()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
    evs[0].record()
    for i in range(reps):
        graphs[i % len(graphs)].replay()
        evs[i + 1].record()
    torch.cuda.synchronize()
    out = sorted(evs[i].elapsed_time(evs[i + 1]) * 1e3 for i in range(reps))
    return out[len(out) // 2], out[0]


def timed_eager(fn, sets, reps):
    for i in range(3):
        fn(*sets[i % len(sets)])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps):
        fn(*sets[i % len(sets)])
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / reps
    return us, us


def report(name, shape, us_med, us_min, alg_bytes, unit_px, launches, note=""):
    gbs = alg_bytes / (us_med * 1e-6) / 1e9
    print(json.dumps({"stage": name, "shape": shape, "us_median": round(us_med, 2), "us_min": round(us_min, 2),
                      "algorithmic_MB": round(alg_bytes / 1e6, 2), "GBps": round(gbs, 1), "frac_of_measured_hbm": round(gbs / PEAK, 4),
                      "peak_GBps": PEAK, "Gpx_per_s": round(unit_px / (us_med * 1e-6) / 1e9, 3), "launches": launches,
                      "note": note}), flush=True)


def g(seed):
    return torch.Generator(device=dev).manual_seed(seed)


def stage_pl_low(reps, N, C, h, w, H, W, tag, sm_warps=0, cta_warps=0, waves=0, rows2=False, regs96=False):
    nb = 4 * N * C * h * w
    sets = [(3.0 * torch.randn(N, C, h, w, generator=g(s), device=dev),) for s in range(copies_for(nb + 4 * N * H * W))]
    thr, keep = 0.3, ((5, 11) if C == 13 else None)
    fn = lambda x: ops.pseudo_label(x, size=(H, W), thr=thr, keep=keep, ignore_index=C - 1 if C == 13 else 255,
                                    flags=ops.pl_flags(sm_warps=sm_warps, cta_warps=cta_warps, waves=waves, rows2=rows2, regs96=regs96))
    med, mn = timed(fn, sets, reps)
    report("pl_low_" + tag + ("_w%d" % sm_warps if sm_warps else "") + ("_rows2" if rows2 else "") + ("_regs96" if regs96 else ""), [N, C, h, w, H, W], med, mn, nb + N * H * W * 4, N * H * W, 1,
           "XU/issue-bound (C ex2 per px), not HBM: see DESIGN.md section 4")


def stage_pl_hist(reps, N, C, h, w, H, W, tag, fused, flags=0):
    """pseudo-labels + confusion matrix: inside the pseudo-label kernel (fused) or as the separate histogram kernel"""
    nb = 4 * N * C * h * w
    sets = [(3.0 * torch.randn(N, C, h, w, generator=g(s), device=dev),
             torch.randint(0, C, (N, H, W), generator=g(70 + s), device=dev, dtype=torch.int64).to(torch.uint8))
            for s in range(copies_for(nb + 6 * N * H * W))]
    hist = torch.zeros(C, C, dtype=torch.int64, device=dev)
    thr, keep = 0.3, ((5, 11) if C == 13 else None)

    def fn(x, gt):
        if fused:
            return ops.pseudo_label(x, size=(H, W), thr=thr, keep=keep, ignore_index=C - 1, flags=flags, gt=gt, hist=hist)
        pl = ops.pseudo_label(x, size=(H, W), thr=thr, keep=keep, ignore_index=C - 1, flags=flags)
        ops.confusion_matrix(pl.label.reshape(-1), gt.reshape(-1), C, hist)
        return pl
    med, mn = timed(fn, sets, reps)
    report("pl_hist_%s_%s" % ("fused" if fused else "separate", tag), [N, C, h, w, H, W], med, mn, nb + N * H * W * 5, N * H * W,
           1 if fused else 2, "pseudo-labels + confusion matrix")


def stage_pl_argmax(reps, N, C, h, w, H, W, tag):
    nb = 4 * N * C * h * w
    sets = [(3.0 * torch.randn(N, C, h, w, generator=g(s), device=dev),) for s in range(copies_for(nb + 4 * N * H * W))]
    fn = lambda x: ops.pseudo_label(x, size=(H, W), mode=ops.PL_UP_ARGMAX, want_conf16=False)
    med, mn = timed(fn, sets, reps)
    report("pl_argmax_" + tag, [N, C, h, w, H, W], med, mn, nb + N * H * W, N * H * W, 1, "upsample + argmax only (no exponentials)")


def stage_pl_full(reps, N, C, H, W, tag):
    nb = 4 * N * C * H * W
    sets = [(3.0 * torch.randn(N, C, H, W, generator=g(s), device=dev),) for s in range(copies_for(nb))]
    fn = lambda x: ops.pseudo_label(x, size=None, thr=0.3, keep=None, ignore_index=255)
    med, mn = timed(fn, sets, reps)
    report("pl_full_" + tag, [N, C, H, W], med, mn, nb + N * H * W * 4, N * H * W, 1, "HBM-bound: (4C+4) B/px")


def stage_proto(reps, N, C, K, h, w, tag, consumer_warps=0, sums_only=False):
    nb = 4 * N * (K + C) * h * w
    sets = []
    for s in range(copies_for(nb)):
        sets.append((torch.randn(N, K, h, w, generator=g(s), device=dev), 3.0 * torch.randn(N, C, h, w, generator=g(100 + s), device=dev)))
    protos = torch.zeros(C, K, device=dev)
    num = torch.zeros(C, device=dev)

    def fn(feat, logits):
        if sums_only:
            return ops.class_sums(feat, logits, C, consumer_warps=consumer_warps)
        return ops.class_means(feat, logits, C, update=(protos, num, ops.UPDATE_MEAN, True, 0.9999), consumer_warps=consumer_warps)
    med, mn = timed(fn, sets, reps)
    report(("sums_" if sums_only else "proto_") + tag + ("_nc%d" % consumer_warps if consumer_warps else ""), [N, C, K, h, w], med, mn,
           nb + 4 * N * C * (K + 1), N * h * w, 2 if sums_only else 3,
           "labels + accumulate" if sums_only else "labels + accumulate + finalize/sequential update")


def stage_overlap(reps, pl_warps, nc, cta_warps=0, waves=0, prio=False, label_first=False, rows2=False, regs96=False):
    """cfg2: the pseudo-label kernel (pl_warps per SM) and labels + accumulate (nc consumer warps) on two forked capture
    streams -- both resident on every SM -- against the same two run one after the other."""
    N, C, K, h, w, H, W = 16, 13, 256, 81, 241, 640, 1920
    sets = []
    for s in range(4):
        sets.append((3.0 * torch.randn(N, C, h, w, generator=g(s), device=dev), torch.randn(N, K, h, w, generator=g(50 + s), device=dev)))
    side = torch.cuda.Stream(dev, priority=-1) if prio else torch.cuda.Stream(dev)

    def fn(logits, feat):
        cur = torch.cuda.current_stream(dev)
        if pl_warps:
            fl = ops.pl_flags(sm_warps=pl_warps, cta_warps=cta_warps, waves=waves, rows2=rows2, regs96=regs96)
            if label_first:   # the label kernel alone first, then accumulate and pseudo-label become ready together
                from b200seg.ops import _ptr, _stream, check, lib
                import ctypes
                n, k, plane = feat.shape[0], feat.shape[1], feat[0, 0].numel()
                L = lib()
                rec = torch.empty(int(L.tss_proto_records_bytes(n, plane)), dtype=torch.uint8, device=dev)
                seg = torch.empty((n, plane), dtype=torch.uint8, device=dev)
                sums = torch.empty((n, C, k), dtype=torch.float32, device=dev)
                check(L.tss_proto_labels(_ptr(logits), None, n, C, plane, C, 0, -1.0, _ptr(seg), _ptr(rec), _ptr(sums), sums.numel(), _stream()), "labels")
                side.wait_stream(cur)
                with torch.cuda.stream(side):
                    check(L.tss_proto_accumulate(_ptr(feat), _ptr(rec), n, k, plane, C, _ptr(sums), 0, nc, _stream()), "acc")
                pl = ops.pseudo_label(logits, size=(H, W), thr=0.3, keep=(5, 11), ignore_index=12, flags=fl)
                cur.wait_stream(side)
                return pl, sums
            side.wait_stream(cur)
            with torch.cuda.stream(side):
                cs = ops.class_sums(feat, logits, C, consumer_warps=nc)
            pl = ops.pseudo_label(logits, size=(H, W), thr=0.3, keep=(5, 11), ignore_index=12, flags=fl)
            cur.wait_stream(side)
        else:
            pl = ops.pseudo_label(logits, size=(H, W), thr=0.3, keep=(5, 11), ignore_index=12)
            cs = ops.class_sums(feat, logits, C, consumer_warps=nc)
        return pl, cs
    med, mn = timed(fn, sets, reps)
    alg = 4 * N * C * h * w + N * H * W * 4 + 4 * N * (K + C) * h * w
    tag = "overlap_pl%d_nc%d" % (pl_warps, nc) + ("_cta%d_wv%d" % (cta_warps, waves) if cta_warps else "") + ("_prio" if prio else "") + \
        ("_lf" if label_first else "") + ("_rows2" if rows2 else "") + ("_regs96" if regs96 else "")
    report(tag if pl_warps else "serial_pl16_nc%d" % (nc or 16), [N, C, K, h, w, H, W], med, mn, alg,
           N * H * W, 3, "pseudo-label + labels + accumulate (no finalize, no confusion)")


def stage_dist(reps, N, C, K, h, w, tag, rectify, form=0):
    nb = 4 * N * K * h * w
    sets = []
    for s in range(copies_for(nb)):
        feat = torch.randn(N, K, h, w, generator=g(s), device=dev)
        soft = torch.softmax(3.0 * torch.randn(N, C, h, w, generator=g(200 + s), device=dev), 1)
        hard = torch.full((N, 1, h, w), 255, dtype=torch.uint8, device=dev)
        sets.append((feat, soft, hard))
    protos = torch.randn(C, K, generator=g(7), device=dev)
    if rectify:
        fn = lambda f, s, hd: ops.prototype_distance(f, protos, 0.5, want_dist=False, want_weight=True, soft=s, hard=hd,
                                                     train_thred=0.0, ignore_index=255, form=form)
        alg = nb + N * h * w * (4 * C + 4 * C + 1 + 8)
    else:
        fn = lambda f, s, hd: ops.prototype_distance(f, protos, 0.5, want_dist=True, want_weight=False, form=form)
        alg = nb + N * h * w * 4 * C
    med, mn = timed(fn, sets, reps)
    report(("dist_rectify_" if rectify else "dist_") + tag + ("_form%x" % form if form else ""), [N, C, K, h, w], med, mn, alg, N * h * w, 1,
           "feat read once; weights+rectified label fused" if rectify else "distance only")


def stage_loss(reps, N, C, H, W, tag):
    nb = 4 * N * C * H * W
    sets = []
    for s in range(copies_for(nb)):
        x = torch.randn(N, C, H, W, generator=g(s), device=dev)
        y = torch.randint(0, C, (N, H, W), generator=g(300 + s), device=dev, dtype=torch.int64)
        sets.append((x, y))

    def fwd(x, y):
        return ops.ce_rce(x, y, 255, 0.1, 1.0)[0]
    med, mn = timed(fwd, sets, reps)
    report("ce_rce_fwd_" + tag, [N, C, H, W], med, mn, nb + 8 * N * H * W, N * H * W, 2, "ops.ce_rce: int64 targets as the reference; includes ~8 scalar torch kernels (mean, alpha/beta mix)")

    out4 = torch.empty(8, dtype=torch.float64, device=dev)
    wsp = torch.zeros(int(ops.lib().tss_loss_workspace_bytes()), dtype=torch.uint8, device=dev)

    def fwd_abi(x, y):   # the two forward launches alone (sums in fp64), without the scalar torch arithmetic of ops.ce_rce
        ops.check(ops.lib().tss_ce_rce_fwd(x.data_ptr(), y.data_ptr(), 2, N, C, H * W, 255, 1, out4.data_ptr(), wsp.data_ptr(),
                                          torch.cuda.current_stream().cuda_stream), "tss_ce_rce_fwd")
    med, mn = timed(fwd_abi, sets, reps)
    report("ce_rce_fwd_abi_" + tag, [N, C, H, W], med, mn, nb + 8 * N * H * W, N * H * W, 2, "tss_ce_rce_fwd alone (kernel + partial reduction)")
    scales = torch.tensor([1.0 / (N * H * W), 0.1 / (N * H * W)], device=dev)

    def bwd(x, y):   # the backward kernel alone, through the C ABI (graph-timed like every other stage)
        grad = torch.empty_like(x)
        ops.check(ops.lib().tss_ce_rce_bwd(x.data_ptr(), y.data_ptr(), 2, N, C, H * W, 255, scales.data_ptr(), grad.data_ptr(),
                                          torch.cuda.current_stream().cuda_stream), "tss_ce_rce_bwd")
        return grad
    med, mn = timed(bwd, sets, reps)
    report("ce_rce_bwd_" + tag, [N, C, H, W], med, mn, 2 * nb + 8 * N * H * W, N * H * W, 1, "logits read once, gradient written once")

    def fwdbwd(x, y):
        xr = x.requires_grad_(True)
        out = ops.ce_rce(xr, y, 255, 0.1, 1.0)
        (out[0] if isinstance(out, (tuple, list)) else out).backward()
        xr.grad = None
    med, mn = timed_eager(fwdbwd, sets, reps)
    report("ce_rce_fwd_bwd_" + tag, [N, C, H, W], med, mn, 3 * nb + 16 * N * H * W, N * H * W, 4,
           "forward + backward through torch.autograd, eager launches: host-bound, not a kernel number")


def stage_hist(reps, n_maps, H, W, C, dtype, tag):
    n = n_maps * H * W
    es = torch.empty(0, dtype=dtype).element_size()
    sets = []
    for s in range(copies_for(2 * n * es)):
        gt = torch.randint(0, C + 1, (n,), generator=g(s), device=dev, dtype=torch.int16).to(dtype)
        gt[gt == C] = 255 if dtype == torch.uint8 else 255
        same = torch.rand(n, generator=g(400 + s), device=dev) < 0.7
        pred = torch.where(same & (gt < C), gt, torch.randint(0, C, (n,), generator=g(500 + s), device=dev, dtype=torch.int16).to(dtype))
        sets.append((pred, gt))
    hist = torch.zeros(C, C, dtype=torch.int64, device=dev)
    bad = torch.zeros(1, dtype=torch.int64, device=dev)
    fn = lambda p, t: ops.confusion_matrix(p, t, C, hist, bad)
    med, mn = timed(fn, sets, reps)
    report("hist_" + tag, [n_maps, H, W, C, str(dtype)], med, mn, 2 * n * es, n, 1, "%d B/px" % (2 * es))


def stage_dist_sweep(reps, N, C, K, h, w, tag):
    for algo in (ops.DIST_EXPANDED,):
        for variant in (2, 4):
            for pf in ((0, 1, 2) if variant == 2 else (0, 1)):
                for split in (1, 2, 4):
                    stage_dist(reps, N, C, K, h, w, tag, True, ops.dist_form(variant, pf, split, algo))


STAGES = {
    "dist_stream_shapes": lambda r: [stage_dist(r, n, 19, 2048, 129, 257, "cfg3_%dimg_shape%d" % (n, sh), True, ops.dist_form(8, sh - 1))
                                     for n in (2, 4, 8) for sh in (1, 2, 3, 4)],
    "dist_stream_shapes_k256": lambda r: [stage_dist(r, 16, 13, 256, 81, 241, "k256_shape%d" % sh, True, ops.dist_form(8, sh - 1))
                                          for sh in (1, 2, 3, 4)],
    "dist_stream_cfg3": lambda r: stage_dist(r, 2, 19, 2048, 129, 257, "cfg3_2img", True, ops.dist_form(8)),
    "dist_stream_cfg3_8": lambda r: stage_dist(r, 8, 19, 2048, 129, 257, "cfg3_8img", True, ops.dist_form(8)),
    "dist_stream_cfg3_4": lambda r: stage_dist(r, 4, 19, 2048, 129, 257, "cfg3_4img", True, ops.dist_form(8)),
    "dist_stream_k256": lambda r: stage_dist(r, 16, 13, 256, 81, 241, "k256", True, ops.dist_form(8)),
    "dist_tlp_k256_best": lambda r: stage_dist(r, 16, 13, 256, 81, 241, "k256", True, ops.dist_form(2, 2, 1, 2)),
    "ov_15_6_c3w2": lambda r: stage_overlap(r, 15, 6, 3, 2, False, False),
    "ov_15_6_c3w2_p": lambda r: stage_overlap(r, 15, 6, 3, 2, True, False),
    "ov_15_6_c3w3_plf": lambda r: stage_overlap(r, 15, 6, 3, 3, True, True),
    "ov_15_6_c3w4_plf": lambda r: stage_overlap(r, 15, 6, 3, 4, True, True),
    "ov_16_6_c4w2_plf": lambda r: stage_overlap(r, 16, 6, 4, 2, True, True),
    "ov_15_6_c5w2_plf": lambda r: stage_overlap(r, 15, 6, 5, 2, True, True),
    "ov_12_6_c3w2_plf": lambda r: stage_overlap(r, 12, 6, 3, 2, True, True),
    "ov_15_6_c3w1_plf": lambda r: stage_overlap(r, 15, 6, 3, 1, True, True),
    "dist_sweep_cfg3": lambda r: stage_dist_sweep(r, 2, 19, 2048, 129, 257, "cfg3_2img"),
    "dist_sweep_cfg3_8": lambda r: stage_dist_sweep(r, 8, 19, 2048, 129, 257, "cfg3_8img"),
    "dist_sweep_k256": lambda r: stage_dist_sweep(r, 16, 13, 256, 81, 241, "k256"),
    "pl_low_cfg2": lambda r: stage_pl_low(r, 16, 13, 81, 241, 640, 1920, "cfg2"),
    "pl_low_city": lambda r: stage_pl_low(r, 4, 19, 129, 257, 1024, 2048, "cityscapes19"),
    "pl_full_cfg2": lambda r: stage_pl_full(r, 16, 13, 640, 1920, "cfg2"),
    "pl_full_city": lambda r: stage_pl_full(r, 4, 19, 1024, 2048, "cityscapes19"),
    "proto_cfg2": lambda r: stage_proto(r, 16, 13, 256, 81, 241, "cfg2_k256"),
    "proto_cfg3": lambda r: stage_proto(r, 2, 19, 2048, 129, 257, "cfg3_k2048"),
    "dist_cfg3": lambda r: stage_dist(r, 2, 19, 2048, 129, 257, "cfg3_k2048", False),
    "dist_rectify_cfg3": lambda r: stage_dist(r, 2, 19, 2048, 129, 257, "cfg3_k2048", True),
    "dist_rectify_k256": lambda r: stage_dist(r, 16, 13, 256, 81, 241, "cfg2_k256", True),
    "loss_cfg2": lambda r: stage_loss(r, 4, 13, 640, 1920, "4x13x640x1920"),
    "dist_rectify_cfg3_direct": lambda r: stage_dist(r, 2, 19, 2048, 129, 257, "cfg3_k2048", True, ops.dist_form(algo=ops.DIST_DIRECT)),
    "dist_rectify_cfg3_8img": lambda r: stage_dist(r, 8, 19, 2048, 129, 257, "cfg3_k2048_8img", True),
    "dist_rectify_cfg3_8img_direct": lambda r: stage_dist(r, 8, 19, 2048, 129, 257, "cfg3_k2048_8img", True, ops.dist_form(algo=ops.DIST_DIRECT)),
    "dist_rectify_k256_direct": lambda r: stage_dist(r, 16, 13, 256, 81, 241, "cfg2_k256", True, ops.dist_form(algo=ops.DIST_DIRECT)),
    "pl_low_cfg2_w12": lambda r: stage_pl_low(r, 16, 13, 81, 241, 640, 1920, "cfg2", 12),
    "pl_low_cfg2_w8": lambda r: stage_pl_low(r, 16, 13, 81, 241, 640, 1920, "cfg2", 8),
    "sums_cfg2_nc16": lambda r: stage_proto(r, 16, 13, 256, 81, 241, "cfg2_k256", 16, True),
    "sums_cfg2_nc8": lambda r: stage_proto(r, 16, 13, 256, 81, 241, "cfg2_k256", 8, True),
    "sums_cfg2_nc4": lambda r: stage_proto(r, 16, 13, 256, 81, 241, "cfg2_k256", 4, True),
    "sums_cfg3_nc4": lambda r: stage_proto(r, 2, 19, 2048, 129, 257, "cfg3_k2048", 4, True),
    "serial_cfg2": lambda r: stage_overlap(r, 0, 0),
    "overlap_cfg2_12_4": lambda r: stage_overlap(r, 12, 4),
    "ov_15_8_c3w2": lambda r: stage_overlap(r, 15, 8, 3, 2),
    "ov_15_8_c3w2_p": lambda r: stage_overlap(r, 15, 8, 3, 2, True),
    "ov_15_8_c3w2_plf": lambda r: stage_overlap(r, 15, 8, 3, 2, True, True),
    "ov_15_8_c3w3_plf": lambda r: stage_overlap(r, 15, 8, 3, 3, True, True),
    "ov_15_16_c3w2_plf": lambda r: stage_overlap(r, 15, 16, 3, 2, True, True),
    "ov_15_6_c3w2_plf": lambda r: stage_overlap(r, 15, 6, 3, 2, True, True),
    "ov_15_4_c3w2_plf": lambda r: stage_overlap(r, 15, 4, 3, 2, True, True),
    "ov_15_8_c5w2_plf": lambda r: stage_overlap(r, 15, 8, 5, 2, True, True),
    "ov_12_8_c6w2_plf": lambda r: stage_overlap(r, 12, 8, 6, 2, True, True),
    "ov_12_6_plf": lambda r: stage_overlap(r, 12, 6, 0, 0, True, True),
    "ov_12_6_p": lambda r: stage_overlap(r, 12, 6, 0, 0, True, False),
    "pl_rows2": lambda r: stage_pl_low(r, 16, 13, 81, 241, 640, 1920, "cfg2", 0, 0, 0, True),
    "pl_rows2_c3w2": lambda r: stage_pl_low(r, 16, 13, 81, 241, 640, 1920, "cfg2_c3w2", 12, 3, 2, True),
    "pl_rows2_c4w2": lambda r: stage_pl_low(r, 16, 13, 81, 241, 640, 1920, "cfg2_c4w2", 12, 4, 2, True),
    "pl_rows2_city": lambda r: stage_pl_low(r, 4, 19, 129, 257, 1024, 2048, "cityscapes19", 0, 0, 0, True),
    "ov_r2_12_6_c3w2": lambda r: stage_overlap(r, 12, 6, 3, 2, True, False, True),
    "ov_r2_12_4_c3w2": lambda r: stage_overlap(r, 12, 4, 3, 2, True, False, True),
    "ov_r2_12_4_c3w3": lambda r: stage_overlap(r, 12, 4, 3, 3, True, False, True),
    "ov_r2_12_4_c4w2": lambda r: stage_overlap(r, 12, 4, 4, 2, True, False, True),
    "ov_r2_12_8_c3w2": lambda r: stage_overlap(r, 12, 8, 3, 2, True, False, True),
    "ov_r2_12_16_c3w2": lambda r: stage_overlap(r, 12, 16, 3, 2, True, False, True),
    "pl96": lambda r: stage_pl_low(r, 16, 13, 81, 241, 640, 1920, "cfg2", 0, 0, 0, False, True),
    "pl96_c4w2": lambda r: stage_pl_low(r, 16, 13, 81, 241, 640, 1920, "cfg2_c4w2", 20, 4, 2, False, True),
    "ov_16_6_c4w2_plf": lambda r: stage_overlap(r, 16, 6, 4, 2, True, True),
    "ov96_12_6_c4w2_plf": lambda r: stage_overlap(r, 12, 6, 4, 2, True, True, False, True),
    "ov96_12_6_c3w2_plf": lambda r: stage_overlap(r, 12, 6, 3, 2, True, True, False, True),
    "ov96_12_6_c4w3_plf": lambda r: stage_overlap(r, 12, 6, 4, 3, True, True, False, True),
    "ov96_13_6_c4w2_plf": lambda r: stage_overlap(r, 13, 6, 4, 2, True, True, False, True),
    "ov96_13_6_c3w2_plf": lambda r: stage_overlap(r, 13, 6, 3, 2, True, True, False, True),
    "ov96_13_6_c4w3_plf": lambda r: stage_overlap(r, 13, 6, 4, 3, True, True, False, True),
    "ov96_16_6_c4w2_plf": lambda r: stage_overlap(r, 16, 6, 4, 2, True, True, False, True),
    "ov96_16_6_c3w2_plf": lambda r: stage_overlap(r, 16, 6, 3, 2, True, True, False, True),
    "ov96_16_6_c4w3_plf": lambda r: stage_overlap(r, 16, 6, 4, 3, True, True, False, True),
    "ov96_20_6_c4w2_plf": lambda r: stage_overlap(r, 20, 6, 4, 2, True, True, False, True),
    "ov96_20_6_c3w2_plf": lambda r: stage_overlap(r, 20, 6, 3, 2, True, True, False, True),
    "ov96_20_6_c4w3_plf": lambda r: stage_overlap(r, 20, 6, 4, 3, True, True, False, True),
    "ov96_20_8_c4w2_plf": lambda r: stage_overlap(r, 20, 8, 4, 2, True, True, False, True),
    "ov96_20_4_c4w2_plf": lambda r: stage_overlap(r, 20, 4, 4, 2, True, True, False, True),
    "ov96_20_6_c5w2_plf": lambda r: stage_overlap(r, 20, 6, 5, 2, True, True, False, True),
    "pl_hist_fused": lambda r: stage_pl_hist(r, 16, 13, 81, 241, 640, 1920, "cfg2", True),
    "pl_hist_separate": lambda r: stage_pl_hist(r, 16, 13, 81, 241, 640, 1920, "cfg2", False),
    "pl_hist_fused_co": lambda r: stage_pl_hist(r, 16, 13, 81, 241, 640, 1920, "cfg2_c4w2", True, ops.CORESIDENT_PL_FLAGS),
    "pl_hist_fused_city": lambda r: stage_pl_hist(r, 4, 19, 129, 257, 1024, 2048, "city19", True),
    "pl_hist_separate_city": lambda r: stage_pl_hist(r, 4, 19, 129, 257, 1024, 2048, "city19", False),
    "pl_argmax_cfg2": lambda r: stage_pl_argmax(r, 16, 13, 81, 241, 640, 1920, "cfg2"),
    "pl_c3w2": lambda r: stage_pl_low(r, 16, 13, 81, 241, 640, 1920, "cfg2_c3w2", 0, 3, 2),
    "overlap_cfg2_10_6": lambda r: stage_overlap(r, 10, 6),
    "overlap_cfg2_12_6": lambda r: stage_overlap(r, 12, 6),
    "sums_cfg2_nc6": lambda r: stage_proto(r, 16, 13, 256, 81, 241, "cfg2_k256", 6, True),
    "pl_low_cfg2_w10": lambda r: stage_pl_low(r, 16, 13, 81, 241, 640, 1920, "cfg2", 10),
    "overlap_cfg2_8_8": lambda r: stage_overlap(r, 8, 8),
    "overlap_cfg2_8_4": lambda r: stage_overlap(r, 8, 4),
    "overlap_cfg2_16_4": lambda r: stage_overlap(r, 16, 4),
    "hist_u8": lambda r: stage_hist(r, 2000, 256, 512, 19, torch.uint8, "cfg5_u8_2000maps"),
    "hist_u8_c13": lambda r: stage_hist(r, 64, 640, 1920, 13, torch.uint8, "cfg2_u8_c13_64maps"),
    "hist_i64": lambda r: stage_hist(r, 500, 256, 512, 19, torch.int64, "cfg5_i64_500maps"),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--only", default="")
    a = ap.parse_args()
    names = [n for n in a.only.split(",") if n] or list(STAGES)
    for n in names:
        STAGES[n](a.reps)
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
