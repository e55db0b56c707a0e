#!/usr/bin/env python
"""HBM-bound row kernels against the measured copy bandwidth: K2 (low-rank removal, ranks 1..64),
K1b (ablate + decode, budgets 1..1024), rmsnorm. Prints one JSON line per configuration:
algorithmic bytes = 2 * d * 2 B per row (x in, h out)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from taboo_brittleness_b200 import ops, synthetic  # noqa: E402

D, S = 3584, 16384
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0}
x = torch.empty(rows, D, dtype=torch.bfloat16, device="cuda")
for r0 in range(0, rows, 65536):
    x[r0:r0 + 65536] = synthetic.make_residuals(min(65536, rows - r0) // 128, 128, D, seed=r0 + 1, device="cuda").to(torch.bfloat16)
w_norm = synthetic.make_lens(D, 8, seed=1, device="cuda").w_norm.cuda()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    best = None
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        best = ms if best is None else min(best, ms)
    return best


bytes_alg = rows * D * 2 * 2
ms = timed(lambda: ops.rmsnorm(x, w_norm))
print(json.dumps({"kernel": "rmsnorm", "rows": rows, "ms": ms, "GBps": bytes_alg / ms / 1e6, "frac_of_measured_copy": bytes_alg / ms / 1e6 / peaks["hbm_gbs"]}), flush=True)
for r in (1, 2, 4, 8, 16, 32, 64):
    U = torch.linalg.qr(torch.randn(D, r, device="cuda"))[0].to(torch.bfloat16).contiguous()
    ms = timed(lambda: ops.lowrank_remove(x, U, w_norm=w_norm, want_x=False, want_h=True))
    print(json.dumps({"kernel": "K2 lowrank_remove", "rank": r, "rows": rows, "ms": ms, "GBps": bytes_alg / ms / 1e6,
                      "frac_of_measured_copy": bytes_alg / ms / 1e6 / peaks["hbm_gbs"], "gflops": 4.0 * D * r * rows / ms / 1e6}), flush=True)
sae_s = synthetic.make_sae(D, S, seed=42, device="cuda")
sae = ops.SaeWeights.from_sae_lens_layout(sae_s.W_enc, sae_s.b_enc, sae_s.threshold, sae_s.W_dec, sae_s.b_dec)
n_enc = min(rows, 131072)
acts = ops.sae_encode(x[:n_enc], sae, cap=128)
mean = ops.latent_mean(acts, 128)
_, _, rank = ops.topm_select(mean, 1024)
nnz = float(acts.count.float().mean().item())
ms = timed(lambda: ops.rank_lists(acts, rank, 128))
print(json.dumps({"kernel": "K1b rank_lists (once per ranking)", "rows": n_enc, "ms": ms}), flush=True)
ranked = ops.rank_lists(acts, rank, 128)
print(json.dumps({"mean_ranked_entries_per_row": float(ranked.count.float().mean().item())}), flush=True)
ba = n_enc * D * 2 * 2
tot_old = tot_new = 0.0
for b in (1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024):
    ms = timed(lambda: ops.sae_ablate_decode(x[:n_enc], acts, sae, rank, b, rank_rows=128, w_norm=w_norm, want_x=False, want_h=True))
    ms2 = timed(lambda: ops.sae_ablate_ranked(x[:n_enc], ranked, sae, b, w_norm=w_norm, want_x=False, want_h=True))
    tot_old += ms
    tot_new += ms2
    print(json.dumps({"kernel": "K1b", "budget": b, "rows": n_enc, "mean_active": nnz, "ms_table_lookup": ms, "ms_ranked_prefix": ms2,
                      "GBps_table_lookup": ba / ms / 1e6, "GBps_ranked_prefix": ba / ms2 / 1e6,
                      "frac_of_measured_copy_ranked_prefix": ba / ms2 / 1e6 / peaks["hbm_gbs"]}), flush=True)
print(json.dumps({"kernel": "K1b, 11-budget sweep", "GBps_table_lookup": 11 * ba / tot_old / 1e6, "GBps_ranked_prefix": 11 * ba / tot_new / 1e6,
                  "frac_ranked_prefix": 11 * ba / tot_new / 1e6 / peaks["hbm_gbs"]}), flush=True)
BS = [1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024]
ms = timed(lambda: ops.sae_ablate_ranked_multi(x[:n_enc], ranked, sae, BS, w_norm=w_norm))
print(json.dumps({"kernel": "K1b, all 11 budgets in one pass (tb_sae_ablate_ranked_multi)", "rows": n_enc, "ms": ms,
                  "ms_one_launch_per_budget": tot_new, "GBps_algorithmic": 11 * ba / ms / 1e6,
                  "frac_of_measured_copy": 11 * ba / ms / 1e6 / peaks["hbm_gbs"],
                  "bytes_moved_GB": (n_enc * D * 2 * 12) / 1e9}), flush=True)
for sub in ([1, 2, 4, 8], [128, 256, 512, 1024], [1024]):
    ms = timed(lambda: ops.sae_ablate_ranked_multi(x[:n_enc], ranked, sae, sub, w_norm=w_norm))
    print(json.dumps({"kernel": "K1b multi", "budgets": sub, "ms": ms}), flush=True)
ms = timed(lambda: ops.sae_ablate_ranked_multi(x[:n_enc], ranked, sae, BS, w_norm=w_norm, want_h=False, want_rstd=True))
print(json.dumps({"kernel": "K1b multi, rstd only (no row output)", "ms": ms}), flush=True)
