import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from b200seg import ops
dev = torch.device("cuda:0")
protos = torch.randn(13, 256, device=dev)
f = torch.randn(16, 256, 81, 241, device=dev)
for _ in range(3):
    ops.prototype_distance(f, protos, 0.5, want_dist=True)
torch.cuda.synchronize()
print("ok")
