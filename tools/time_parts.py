"""Eager per-op device times of the cfg4 round tail (sampled rectify, upsample, CE+RCE forward/backward, upsample adjoint)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from b200seg import ops  # noqa: E402

dev = torch.device("cuda:0")
c = bench.CONFIGS["cfg4"]
frames = c["frames"]
logits, feat, gt, soft, hard = bench.make_inputs(dev, 1, frames, c)
q = (c["H"] // 4, c["W"] // 4)
protos = torch.randn(c["C"], c["K"], device=dev)
weak = {"RandomSized": [torch.full((frames,), 4 * q[0]), torch.full((frames,), 4 * q[1])],
        "RandomCrop": [torch.zeros(frames), torch.full((frames,), 4.0 * q[0]), torch.zeros(frames), torch.full((frames,), 4.0 * q[1])],
        "RandomHorizontallyFlip": [False] * frames}
wt = ops.weak_table(weak, frames, dev)


def timed(name, fn, reps=10):
    for _ in range(3):
        out = fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    print("%-40s %9.1f us" % (name, e0.elapsed_time(e1) * 1e3 / reps), flush=True)
    return out


lab = timed("prototype_distance_sampled", lambda: ops.prototype_distance_sampled(feat, q, wt, protos, temperature=0.5, soft=soft, hard=hard,
                                                                               train_thred=0.0, ignore_index=c["ignore"])[2])
timed("prototype_distance_sampled flags=1", lambda: ops.prototype_distance_sampled(feat, q, wt, protos, temperature=0.5, soft=soft, hard=hard,
                                                                                   train_thred=0.0, ignore_index=c["ignore"], flags=1)[2])
fq = timed("upsample feat to 1/4", lambda: ops.upsample_bilinear(feat, q))
timed("prototype_distance on materialised 1/4", lambda: ops.prototype_distance(fq, protos, 0.5, want_dist=False, soft=soft, hard=hard,
                                                                              train_thred=0.0, ignore_index=c["ignore"])[2])
x = timed("upsample logits to 1/4", lambda: ops.upsample_bilinear(logits, q))
y = lab.reshape(frames, *q)
timed("ce_rce fwd", lambda: ops.ce_rce(x, y, c["ignore"], 0.1, 1.0)[0])
xr = x.detach().requires_grad_(True)


def fb():
    loss = ops.ce_rce(xr, y, c["ignore"], 0.1, 1.0)[0]
    return torch.autograd.grad(loss, xr)[0]


g = timed("ce_rce fwd+bwd (autograd)", fb)
timed("upsample adjoint", lambda: ops.upsample_bilinear_adjoint(g, (c["h"], c["w"])))
