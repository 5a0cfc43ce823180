"""Distance-kernel experiment: K split and L2-resident vs HBM-streamed inputs."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from b200seg import ops, lib
from b200seg.graphs import GraphedCall
dev = torch.device("cuda:0")

def run(N, C, K, h, w, copies, S):
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
    print("N=%d C=%d K=%d %dx%d copies=%d S=%d: %.1f us  %.0f GB/s  (FP floor %.1f us)" % (
        N, C, K, h, w, copies, S, t, mb / t * 1e3 / 1e3 * 1e3 / 1e3 * 1e0, N * h * w * K * C * 2 / 37.2e12 * 1e6), flush=True)

for S in (0, 1, 2, 8):
    run(4, 13, 256, 81, 241, 1, S)      # 80 MB: L2 resident
    run(16, 13, 256, 81, 241, 3, S)     # streamed
for S in (0, 1, 4):
    run(2, 19, 2048, 129, 257, 2, S)
run(8, 19, 2048, 129, 257, 2, 1)
run(8, 19, 2048, 129, 257, 2, 0)
