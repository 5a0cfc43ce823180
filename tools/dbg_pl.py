import sys; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests/golden')
import numpy as np, torch
from b200seg import ops
import inputs as gi
n,c,h,w,H,W=2,13,41,81,320,640
logits=torch.from_numpy(gi.randn(900+c+H,n,c,h,w,scale=3.0)).cuda()
a=ops.pseudo_label(logits,(H,W),ops.PL_UP_SOFTMAX,want_conf32=True)
ops.lib().tss_tune_pl_force_tile(1)
b=ops.pseudo_label(logits,(H,W),ops.PL_UP_SOFTMAX,want_conf32=True)
d=(a.conf32-b.conf32).abs()
print('max diff',d.max().item(),'labels differ',(a.label!=b.label).sum().item())
idx=torch.nonzero(d>1e-6)
print(idx[:20].tolist(), len(idx))
for i in idx[:5].tolist():
    print(a.conf32[tuple(i)].item(), b.conf32[tuple(i)].item())
