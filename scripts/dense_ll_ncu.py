#!/usr/bin/env python
"""One invocation of the first-condition path (ops.lens_ll_from_logits) on 37,888 rows, for an ncu capture of
tb_umma_gemm_kernel<LensDenseEpi> and tb_logits_ll_colsum_tiled_kernel.   python scripts/dense_ll_ncu.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from taboo_brittleness_b200 import ops, synthetic  # noqa: E402

D, V, P = 3584, 256000, 296
lens = synthetic.make_lens(D, V, seed=42, device="cuda", dtype=torch.bfloat16)
x = synthetic.make_residuals(P, 128, D, seed=1, device="cuda").to(torch.bfloat16)
h = ops.rmsnorm(x, lens.w_norm.cuda())
n = h.shape[0]
cur = torch.randint(0, V, (n,), dtype=torch.int32, device="cuda")
logits = torch.empty(n, V, dtype=torch.float32, device="cuda")
for _ in range(2):
    out, sums, ids = ops.lens_ll_from_logits(h, lens.W_U, cur, None, logits, softcap=30.0)
torch.cuda.synchronize()
print("rows", n, "lse[0]", float(out.lse[0]), "ids[0]", ids[0].tolist())
