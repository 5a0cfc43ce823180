"""Distance kernel A/B: register-pipeline (variant 0) vs thread-level-parallel (variant 1)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from b200seg import ops, lib
from b200seg.graphs import GraphedCall
dev = torch.device("cuda:0")

def run(N, C, K, h, w, copies, V, S):
    lib().tss_tune_distance_variant(V)
    lib().tss_tune_distance_split(S)
    protos = torch.randn(C, K, device=dev)
    feats = [torch.randn(N, K, h, w, device=dev) for _ in range(copies)]
    gs = [GraphedCall(lambda f=f: ops.prototype_distance(f, protos, 0.5, want_dist=True)) for f in feats]
    for g in gs: g.replay()
    torch.cuda.synchronize()
    reps = 12
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
    ev[0].record()
    for i in range(reps):
        gs[i % copies].replay(); ev[i + 1].record()
    torch.cuda.synchronize()
    t = sorted(ev[i].elapsed_time(ev[i + 1]) * 1e3 for i in range(reps))[reps // 2]
    mb = 4 * N * K * h * w / 1e6
    print("V=%d S=%d N=%d C=%d K=%d %dx%d: %.1f us  %.0f GB/s" % (V, S, N, C, K, h, w, t, mb / t * 1e3), flush=True)

for PFD in (2, -1):
    lib().tss_tune_distance_prefetch(PFD)
    for V, Ss in (((2, (1, 2, 4)), (0, (0,))) if PFD == 2 else ((0, (0,)),)):
        for S in Ss:
            print("prefetch", PFD, end=" ")
            run(16, 13, 256, 81, 241, 3, V, S)
            print("prefetch", PFD, end=" ")
            run(2, 19, 2048, 129, 257, 2, V, S)
            print("prefetch", PFD, end=" ")
            run(8, 19, 2048, 129, 257, 2, V, S)
