#!/usr/bin/env python
"""One rank-update sweep (budgets 64 and 1024) on 296 prompts, for an ncu capture of both variants of
tb_rank_update_lens_kernel.   python scripts/rank_update_ncu.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from taboo_brittleness_b200 import interventions, ops, synthetic  # noqa: E402

D_, T, S, V, P = 3584, 128, 16384, 256000, 296
sae_s = synthetic.make_sae(D_, S, seed=synthetic.SEED_WEIGHTS, device="cuda")
sae = ops.SaeWeights.from_sae_lens_layout(sae_s.W_enc, sae_s.b_enc, sae_s.threshold, sae_s.W_dec, sae_s.b_dec, device="cuda")
del sae_s
lens_s = synthetic.make_lens(D_, V, seed=synthetic.SEED_WEIGHTS, device="cuda", dtype=torch.bfloat16)
eng = interventions.InterventionEngine(sae, ops.LensWeights(w_norm=lens_s.w_norm.cuda(), w_u=lens_s.W_U), softcap=30.0)
del lens_s
x = synthetic.make_residuals(P, T, D_, seed=1234, device="cuda").to(torch.bfloat16)
sid = synthetic.make_secret_ids(P, T, V, seed=1234).cuda()
for _ in range(2):
    out = eng.targeted_budget_sweep_rank_update(x, sid, [64, 1024])
torch.cuda.synchronize()
print("rows", x.shape[0], "lse[0]", float(out[64].lse[0]), float(out[1024].lse[0]))
