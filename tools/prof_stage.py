"""Run one stage of the bench step a few times (for ncu / quick timing).  usage: prof_stage.py STAGE [reps]
STAGE in pl | proto | hist | all (the bench step) | overlap (pseudo-label beside labels + accumulate) | dist | ce | distce."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from b200seg import ops  # noqa: E402

stage = sys.argv[1] if len(sys.argv) > 1 else "all"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
c = bench.CFG
dev = torch.device("cuda:0")
sets = [bench.make_inputs(dev, 1234 + s, c["frames"]) for s in range(3)]
protos = torch.zeros(c["C"], c["K"], device=dev)
num = torch.zeros(c["C"], device=dev)
hist = torch.zeros(c["C"], c["C"], dtype=torch.int64, device=dev)
bad = torch.zeros(1, dtype=torch.int64, device=dev)
pl0 = ops.pseudo_label(sets[0][0], size=(c["H"], c["W"]), thr=c["thr"], keep=c["keep"], ignore_index=c["ignore"])
times = {}


def timed(name, fn):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    torch.cuda.synchronize()
    times.setdefault(name, []).append(e0.elapsed_time(e1) * 1e3)


for i in range(reps):
    logits, feat, gt = sets[i % 3]
    if stage in ("pl", "all"):
        timed("pl", lambda: ops.pseudo_label(logits, size=(c["H"], c["W"]), thr=c["thr"], keep=c["keep"],
                                             ignore_index=c["ignore"]))
    if stage in ("proto", "all"):
        def f():
            ops.class_means(feat, logits, c["C"], update=(protos, num, ops.UPDATE_MEAN, True, 0.9999))
        timed("proto", f)
    if stage == "overlap":
        side = torch.cuda.Stream(dev)

        def f():
            cur = torch.cuda.current_stream(dev)
            side.wait_stream(cur)
            with torch.cuda.stream(side):
                ops.class_sums(feat, logits, c["C"], consumer_warps=4)
            ops.pseudo_label(logits, size=(c["H"], c["W"]), thr=c["thr"], keep=c["keep"], ignore_index=c["ignore"],
                             flags=ops.pl_flags(sm_warps=12))
            cur.wait_stream(side)
        timed("overlap", f)
    if stage in ("hist", "all"):
        timed("hist", lambda: ops.confusion_matrix(pl0.label.reshape(-1), gt.reshape(-1), c["C"], hist, bad))
if stage in ("dist", "distce"):
    for (N, C, K, h, w) in ((2, 19, 2048, 129, 257), (16, 13, 256, 81, 241)):
        f2 = torch.randn(N, K, h, w, device=dev)
        soft = torch.softmax(3.0 * torch.randn(N, C, h, w, device=dev), 1)
        hard = torch.full((N, 1, h, w), 255, dtype=torch.uint8, device=dev)
        pr = torch.randn(C, K, device=dev)
        for i in range(reps):
            timed("dist_rectify_C%d_K%d" % (C, K), lambda: ops.prototype_distance(
                f2, pr, 0.5, want_dist=False, want_weight=True, soft=soft, hard=hard, train_thred=0.0, ignore_index=255))
        del f2, soft, hard
if stage in ("ce", "distce"):
    x = torch.randn(4, 13, 640, 1920, device=dev)
    y = torch.randint(0, 13, (4, 640, 1920), device=dev, dtype=torch.int64)
    for i in range(reps):
        def g():
            xr = x.requires_grad_(True)
            ops.ce_rce(xr, y, 255, 0.1, 1.0)[0].backward()
            xr.grad = None
        timed("ce_rce_fwd_bwd", g)
for k, v in times.items():
    print(k, "us:", " ".join("%.1f" % t for t in v))
