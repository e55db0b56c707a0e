#!/usr/bin/env python
"""Two launches of the one-pass K2 kernel (rank 16, 65,536 rows) for `ncu -k regex:tb_lowrank_cluster_kernel -s 1 -c 1`."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from taboo_brittleness_b200 import ops, synthetic  # noqa: E402

D = 3584
x = synthetic.make_residuals(512, 128, D, seed=3, device="cuda").to(torch.bfloat16)
w = torch.zeros(D, device="cuda")
U = torch.linalg.qr(torch.randn(D, 16, device="cuda"))[0].to(torch.bfloat16).contiguous()
for _ in range(2):
    ops.lowrank_remove(x, U, w_norm=w, want_x=False, want_h=True, one_pass=True)
torch.cuda.synchronize()
print("ok")
