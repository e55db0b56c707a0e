#!/usr/bin/env python
"""Where does K1b's time go at small budgets? Variants: no rank table (nothing selected), empty lists."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from taboo_brittleness_b200 import ops, synthetic
D, S, rows = 3584, 16384, 131072
x = torch.cat([synthetic.make_residuals(512, 128, D, seed=i + 1, device="cuda").to(torch.bfloat16) for i in range(2)])
w_norm = synthetic.make_lens(D, 8, seed=1, device="cuda").w_norm.cuda()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
sae_s = synthetic.make_sae(D, S, seed=42, device="cuda")
sae = ops.SaeWeights.from_sae_lens_layout(sae_s.W_enc, sae_s.b_enc, sae_s.threshold, sae_s.W_dec, sae_s.b_dec)
acts = ops.sae_encode(x, sae, cap=128)
mean = ops.latent_mean(acts, 128)
_, _, rank = ops.topm_select(mean, 1024)
def timed(fn, reps=5):
    fn(); torch.cuda.synchronize(); best = None
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1); best = ms if best is None else min(best, ms)
    return best
def run(name, a, rk, b):
    ms = timed(lambda: ops.sae_ablate_decode(x, a, sae, rk, b, rank_rows=128 if rk is not None else 0, w_norm=w_norm, want_x=False, want_h=True))
    print(json.dumps({"variant": name, "ms": ms, "GBps": rows * D * 4 / ms / 1e6}), flush=True)
print("csr" if acts.row_start is not None else "fixed", "mean count", float(acts.count.float().mean()), "max", int(acts.count.max()))
run("budget 1", acts, rank, 1)
run("no rank table", acts, None, 1)
empty = ops.SparseActs(count=torch.zeros_like(acts.count), idx=acts.idx, val=acts.val, S=S, cap=acts.cap, row_start=acts.row_start)
run("empty lists", empty, rank, 1)
cl = ops.SparseActs(count=acts.count.clamp(max=128), idx=acts.idx, val=acts.val, S=S, cap=acts.cap, row_start=acts.row_start)
run("counts clamped to 128", cl, rank, 1)
print(json.dumps({"variant": "rmsnorm", "ms": timed(lambda: ops.rmsnorm(x, w_norm))}))
# bos rows only / normal rows only (re-encoded subsets so the lists match the rows)
xb = x[::128].contiguous(); ab = ops.sae_encode(xb, sae, cap=128)
xn = x.view(-1, 128, D)[:, 1:].reshape(-1, D).contiguous(); an = ops.sae_encode(xn, sae, cap=128)
for name, xx, aa in (("bos rows only (1024)", xb, ab), ("normal rows only", xn, an)):
    ms = timed(lambda: ops.sae_ablate_decode(xx, aa, sae, None, 1, w_norm=w_norm, want_x=False, want_h=True))
    print(json.dumps({"variant": name, "rows": xx.shape[0], "ms": ms, "mean_count": float(aa.count.float().mean()), "csr": aa.row_start is not None}), flush=True)
