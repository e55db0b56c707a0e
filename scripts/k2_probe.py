#!/usr/bin/env python
"""K2 (low-rank removal) at one rank, a few calls: run under `ncu --metrics gpu__time_duration.sum` to see
the time of each kernel of the tensor-core path (U prep, coefficient GEMM, apply GEMM, scale)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from taboo_brittleness_b200 import ops, synthetic  # noqa: E402

D = 3584
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
r = int(sys.argv[2]) if len(sys.argv) > 2 else 64
x = torch.empty(rows, D, dtype=torch.bfloat16, device="cuda")
for r0 in range(0, rows, 65536):
    x[r0:r0 + 65536] = synthetic.make_residuals(min(65536, rows - r0) // 128, 128, D, seed=r0 + 1, device="cuda").to(torch.bfloat16)
w_norm = synthetic.make_lens(D, 8, seed=1, device="cuda").w_norm.cuda()
U = torch.linalg.qr(torch.randn(D, r, device="cuda"))[0].to(torch.bfloat16).contiguous()
for _ in range(3):
    ops.lowrank_remove(x, U, w_norm=w_norm, want_x=False, want_h=True)
torch.cuda.synchronize()
