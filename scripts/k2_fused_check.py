#!/usr/bin/env python
"""K2 in one pass (tb_lowrank_remove_fused) against the three-pass path and a dense fp32 product, then its HBM rate.

    python scripts/k2_fused_check.py [--rows 262144] [--skip-check]
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from taboo_brittleness_b200 import ops, synthetic  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=262144)
ap.add_argument("--skip-check", action="store_true")
a = ap.parse_args()
torch.manual_seed(0)


def problem(n, d, r):
    x = synthetic.make_residuals(max(n // 128, 1), 128, d, seed=3, device="cuda")[:n].to(torch.bfloat16).contiguous()
    if x.shape[0] < n:
        x = torch.cat([x, x[: n - x.shape[0]]]).contiguous()
    U = torch.linalg.qr(torch.randn(d, r, device="cuda", dtype=torch.float64))[0].float().to(torch.bfloat16).contiguous()
    w = (torch.randn(d, device="cuda") * 0.5).to(torch.bfloat16).float()
    return x, U, w


if not a.skip_check:
    for (n, d, r) in [(300, 512, 9), (1000, 512, 40), (128, 256, 8), (700, 3584, 16), (4096, 3584, 64), (5000, 3584, 33)]:
        x, U, w = problem(n, d, r)
        rp = 32 * ((r + 31) // 32)
        dbg = torch.full((n, rp), float("nan"), device="cuda")
        fx, fh = ops.lowrank_remove(x, U, w_norm=w, want_x=True, want_h=True, dbg_c=dbg)
        _, fh2 = ops.lowrank_remove(x, U, w_norm=w, want_x=False, want_h=True)
        os.environ["TB_K2_THREE_PASS"] = "1"
        tx, th = ops.lowrank_remove(x, U, w_norm=w, want_x=True, want_h=True, one_pass=False)
        os.environ.pop("TB_K2_THREE_PASS")
        ax, ah = ops.lowrank_remove(x, U, w_norm=w, want_x=True, want_h=True, one_pass=False)
        _, ah2 = ops.lowrank_remove(x, U, w_norm=w, want_x=False, want_h=True, one_pass=False)
        torch.cuda.synchronize()
        c_ref = x.float() @ U.float()
        c_err = (dbg[:, :r] - c_ref).abs().max().item() / c_ref.abs().max().item()
        pad = dbg[:, r:].abs().max().item() if rp > r else 0.0
        ref = (x.float() - c_ref @ U.float().t()).to(torch.bfloat16)
        print(json.dumps({"n": n, "d": d, "r": r, "coeff_rel_err": c_err, "coeff_pad_max": pad,
                          "x_mismatch_vs_three_pass": (fx != tx).float().mean().item(),
                          "x_mismatch_vs_dense_fp32": (fx != ref).float().mean().item(),
                          "three_pass_mismatch_vs_dense_fp32": (tx != ref).float().mean().item(),
                          "h_mismatch_vs_three_pass": (fh != th).float().mean().item(),
                          "h_alone_equals_h_with_x": bool(torch.equal(fh, fh2)),
                          "apply_cluster_x_mismatch_vs_three_pass": (ax != tx).float().mean().item(),
                          "apply_cluster_x_mismatch_vs_dense_fp32": (ax != ref).float().mean().item(),
                          "apply_cluster_h_mismatch_vs_three_pass": (ah != th).float().mean().item(),
                          "apply_cluster_h_alone_equals_h_with_x": bool(torch.equal(ah, ah2)),
                          "x_max_abs_diff": (fx.float() - tx.float()).abs().max().item(),
                          "h_max_abs_diff": (fh.float() - th.float()).abs().max().item()}), flush=True)

peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json"))) \
    if os.path.exists(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0}
D = 3584
x, _, w = problem(a.rows, D, 8)
big = torch.empty(512 * 1024 * 1024, dtype=torch.uint8, device="cuda")


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    best = None
    for _ in range(reps):
        big.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        best = ms if best is None else min(best, ms)
    return best


bytes_alg = a.rows * D * 4
for r in (8, 16, 32, 64):
    U = torch.linalg.qr(torch.randn(D, r, device="cuda"))[0].to(torch.bfloat16).contiguous()
    for name, one, env in (("one pass (cluster, DSMEM split-K)", True, None), ("coefficients + cluster apply/norm (default)", False, None),
                           ("three passes", False, "1")):
        if env:
            os.environ["TB_K2_THREE_PASS"] = env
        else:
            os.environ.pop("TB_K2_THREE_PASS", None)
        ms = timed(lambda: ops.lowrank_remove(x, U, w_norm=w, want_x=False, want_h=True, one_pass=one))
        os.environ.pop("TB_K2_THREE_PASS", None)
        print(json.dumps({"kernel": "K2 " + name, "rank": r, "rows": a.rows, "ms": ms, "GBps": bytes_alg / ms / 1e6,
                          "frac_of_measured_copy": bytes_alg / ms / 1e6 / peaks["hbm_gbs"]}), flush=True)
