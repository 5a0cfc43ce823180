"""Fused pseudo-label kernel timing at the three benchmark shapes (CUDA-graph replay, inputs rotated past L2)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from b200seg import ops, lib
from b200seg.graphs import GraphedCall
dev = torch.device("cuda:0")

def run(N, C, h, w, H, W, copies, V):
    xs = [3.0 * torch.randn(N, C, h, w, device=dev) for _ in range(copies)]
    keep = (5, 11) if C == 13 else None
    gs = [GraphedCall(lambda x=x: ops.pseudo_label(x, size=(H, W), thr=0.3, keep=keep, ignore_index=C - 1 if C == 13 else 255)) for x in xs]
    for g in gs: g.replay()
    torch.cuda.synchronize()
    reps = 20
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
    ev[0].record()
    for i in range(reps):
        gs[i % copies].replay(); ev[i + 1].record()
    torch.cuda.synchronize()
    t = sorted(ev[i].elapsed_time(ev[i + 1]) * 1e3 for i in range(reps))[reps // 2]
    out = gs[0].replay()
    torch.cuda.synchronize()
    return t, out

for shape in ((16, 13, 81, 241, 640, 1920), (4, 13, 81, 241, 640, 1920), (4, 19, 129, 257, 1024, 2048)):
    res = {}
    for V in (0,):
        t, out = run(*shape, 4, V)
        res[V] = out
        print("pipeline=%d %s: %.1f us  %.1f Gpx/s" % (V, shape, t, shape[0] * shape[4] * shape[5] / t / 1e3), flush=True)
