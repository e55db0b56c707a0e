#!/usr/bin/env python
"""Nested budgets as rank updates of stored base logits against the dense sweep, at the benchmark's shapes: time per
budget and drift of the readouts.   python scripts/rank_update_probe.py [--prompts 512] [--split]"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from taboo_brittleness_b200 import interventions, ops, synthetic  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--prompts", type=int, default=512)
ap.add_argument("--split", action="store_true")
ap.add_argument("--S", type=int, default=16384)
ap.add_argument("--V", type=int, default=256000)
a = ap.parse_args()
D_, T = 3584, 128
BUDGETS = [1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024]
dev = torch.device("cuda")
sae_s = synthetic.make_sae(D_, a.S, seed=synthetic.SEED_WEIGHTS, device="cuda")
sae = ops.SaeWeights.from_sae_lens_layout(sae_s.W_enc, sae_s.b_enc, sae_s.threshold, sae_s.W_dec, sae_s.b_dec, device="cuda")
del sae_s
lens_s = synthetic.make_lens(D_, a.V, seed=synthetic.SEED_WEIGHTS, device="cuda", dtype=torch.bfloat16)
eng = interventions.InterventionEngine(sae, ops.LensWeights(w_norm=lens_s.w_norm.to(dev), w_u=lens_s.W_U), softcap=30.0)
del lens_s
x = torch.cat([synthetic.make_residuals(min(512, a.prompts - p0), T, D_, seed=1234 + p0, device="cuda").to(torch.bfloat16)
               for p0 in range(0, a.prompts, 512)])
sid = synthetic.make_secret_ids(a.prompts, T, a.V, seed=1234).to(dev)


def timed(fn):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1), out


ms, D = timed(eng.update_matrix)
print(json.dumps({"piece": "D = (W_dec (1+w)) W_U^T, once per weight set", "ms": ms, "GB": D.numel() * 2 / 1e9}), flush=True)
eng.targeted_budget_sweep_rank_update(x[:8 * T], sid[:8 * T], BUDGETS, split=a.split)      # warm-up
eng.targeted_budget_sweep(x[:8 * T], T, sid[:8 * T], BUDGETS)
eng.stats = interventions.SweepStats()
eng.time_lens = True
ms_ru, ru = timed(lambda: eng.targeted_budget_sweep_rank_update(x, sid, BUDGETS, split=a.split))
base = [(s.elapsed_time(e), f) for s, e, f in eng.stats.lens_events]
upd = [(s.elapsed_time(e), f) for s, e, f in eng.stats.ll_events]
n_sub = len(base)
per_budget = [sum(upd[i * len(BUDGETS) + j][0] for i in range(n_sub)) for j in range(len(BUDGETS))]
print(json.dumps({"piece": "rank-update sweep, whole call (encode, target, K1b rstd x 11, base sweeps, updates)", "rows": x.shape[0],
                  "ms": ms_ru, "token_conditions_per_s": x.shape[0] * len(BUDGETS) / ms_ru * 1e3, "split": a.split,
                  "base_sweep_ms": sum(t for t, _ in base), "base_sweep_tflops": sum(f for _, f in base) / sum(t for t, _ in base) / 1e9,
                  "update_ms_per_budget": dict(zip(BUDGETS, [round(t, 2) for t in per_budget]))}), flush=True)
eng.stats = interventions.SweepStats()
ms_d, dn = timed(lambda: eng.targeted_budget_sweep(x, T, sid, BUDGETS))
print(json.dumps({"piece": "dense sweep without LL-Top-k, whole call", "ms": ms_d,
                  "token_conditions_per_s": x.shape[0] * len(BUDGETS) / ms_d * 1e3, "speedup_of_rank_update": ms_d / ms_ru}), flush=True)
for b in BUDGETS:
    r, d = ru[b], dn[b]
    same10 = (r.top_id.sort(1).values == d.top_id.sort(1).values).all(1).float().mean().item()
    print(json.dumps({"budget": b, "max_abs_dlse": (r.lse - d.lse).abs().max().item(),
                      "max_abs_dsecret_prob": (r.secret_prob - d.secret_prob).abs().max().item(),
                      "max_abs_dtop_prob": (r.top_prob[:, 0] - d.top_prob[:, 0]).abs().max().item(),
                      "top1_id_equal": (r.top_id[:, 0] == d.top_id[:, 0]).float().mean().item(), "top10_set_equal": same10}), flush=True)

# Which of the two is closer to the same path WITHOUT any rounding of intermediates (x', h in fp32)? 256 rows, fp32 torch.
n_s = 2 * T
acts = eng.encode(x[:n_s])
ids, _, rank = eng.targeted_rank(acts, T, m=1024)
ranked = ops.rank_lists(acts, rank, T)
cnt, key, val = ranked.count.cpu(), ranked.key.cpu(), ranked.val.cpu()
a_full = torch.zeros(n_s, 1024)
for r in range(n_s):
    base = r * ranked.cap if ranked.row_start is None else int(ranked.row_start[r])
    k = key[base:base + int(cnt[r])]
    a_full[r, (k >> 20).long()] = val[base:base + int(cnt[r])]
a_full = a_full.cuda()
w1 = 1.0 + eng.lens.w_norm.float()
for b in (1, 64, 1024):
    x1 = torch.cat([x[p * T:(p + 1) * T].float() - a_full[p * T:(p + 1) * T, :b] @ eng.sae.w_dec[ids[p, :b].long()].float()
                    for p in range(2)])
    h = x1 * torch.rsqrt(x1.pow(2).mean(1, keepdim=True) + 1e-6) * w1[None]
    lg = 30.0 * torch.tanh(h @ eng.lens.w_u.float().t() / 30.0)
    lse = torch.logsumexp(lg, 1)
    p1 = torch.exp(lg.max(1).values - lse)
    print(json.dumps({"budget": b, "rows": n_s, "against": "fp32 path without rounding of x' and h",
                      "dense_max_abs_dlse": (dn[b].lse[:n_s] - lse).abs().max().item(),
                      "rank_update_max_abs_dlse": (ru[b].lse[:n_s] - lse).abs().max().item(),
                      "dense_max_abs_dtop1_prob": (dn[b].top_prob[:n_s, 0] - p1).abs().max().item(),
                      "rank_update_max_abs_dtop1_prob": (ru[b].top_prob[:n_s, 0] - p1).abs().max().item()}), flush=True)
