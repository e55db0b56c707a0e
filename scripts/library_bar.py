#!/usr/bin/env python
"""The "library" bar of SURVEY section 8(d): the same lens readout (final RMSNorm -> lm_head -> softcap ->
softmax -> secret probability, per-row top-10, per-prompt LL-Top-10) written the way the reference writes it,
with stock PyTorch ops on the GPU (cuBLAS bf16 GEMM, dense [rows, V] logits and probabilities, torch.topk),
in row batches that fit HBM -- next to the fused kernels on the same rows. Also a secondary parity check over
every row of the batch (ids equal outside near-ties). Prints one JSON line.

    python scripts/library_bar.py [--prompts 512] [--batch-prompts 64]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from taboo_brittleness_b200 import ops, synthetic  # noqa: E402

D, V, T = 3584, 256000, 128


def torch_readout(x, w_norm, W_U, sid, cur, prev, softcap, bp):
    """Unfused: everything the reference materialises is materialised (per batch of bp prompts)."""
    sec, top, ll = [], [], []
    for r0 in range(0, x.shape[0], bp * T):
        xb = x[r0:r0 + bp * T].float()
        h = (xb * torch.rsqrt(xb.pow(2).mean(-1, keepdim=True) + 1e-6) * (1.0 + w_norm)).to(torch.bfloat16)
        try:      # cuBLAS bf16 GEMM with fp32 output: the precision the fused kernels read their accumulators at
            logits = torch.mm(h, W_U.t(), out_dtype=torch.float32)
        except (TypeError, NotImplementedError, RuntimeError):
            logits = (h @ W_U.t()).float()   # bf16-rounded logits: near-ties reorder, fewer equal id sets below
        logits = torch.tanh(logits / softcap) * softcap
        probs = torch.softmax(logits, dim=-1)
        del logits
        rows = torch.arange(xb.shape[0], device=x.device)
        sec.append(probs[rows, sid[r0:r0 + bp * T].long()])
        top.append(torch.topk(probs, 10, dim=-1).indices)
        # aggregate_response_logits (01:35-71): zero the current and the previous token's column per position
        probs[rows, cur[r0:r0 + bp * T].long()] = 0.0
        pv = prev[r0:r0 + bp * T].long()
        ok = pv >= 0
        probs[rows[ok], pv[ok]] = 0.0
        acc = probs.view(-1, T, V).sum(1)
        ll.append(torch.topk(acc, 10, dim=-1).indices)
        del probs, acc
    return torch.cat(sec), torch.cat(top), torch.cat(ll)


def fused_readout(x, w_norm, W_U, sid, cur, prev, softcap):
    h = ops.rmsnorm(x, w_norm)
    out = ops.lens_merge(ops.lens_partials(h, W_U, softcap=softcap, secret_id=sid))
    _, ll = ops.ll_topk(h, W_U, out.lse, cur, prev, softcap=softcap)
    return out.secret_prob, out.top_id, ll


def timed(fn, reps):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--prompts", type=int, default=512)
    ap.add_argument("--batch-prompts", type=int, default=64)
    ap.add_argument("--reps", type=int, default=2)
    a = ap.parse_args()
    ops.device_check()
    lens = synthetic.make_lens(D, V, seed=42, device="cuda", dtype=torch.bfloat16)
    w_norm = lens.w_norm.cuda()
    x = synthetic.make_residuals(a.prompts, T, D, seed=1234, device="cuda").to(torch.bfloat16)
    sid = synthetic.make_secret_ids(a.prompts, T, V).cuda()
    g = torch.Generator().manual_seed(7)
    cur = torch.randint(0, V, (a.prompts * T,), generator=g, dtype=torch.int32).cuda()
    prev = torch.randint(0, V, (a.prompts * T,), generator=g, dtype=torch.int32).cuda()
    prev[::T] = -1
    ms_t, (sec_t, top_t, ll_t) = timed(lambda: torch_readout(x, w_norm, lens.W_U, sid, cur, prev, 30.0, a.batch_prompts), a.reps)
    ms_f, (sec_f, top_f, ll_f) = timed(lambda: fused_readout(x, w_norm, lens.W_U, sid, cur, prev, 30.0), a.reps)
    n = x.shape[0]
    same_top = (top_t.sort(1).values == top_f.long().sort(1).values).all(1).float().mean().item()
    same_ll = (ll_t.sort(1).values == ll_f.long().sort(1).values).all(1).float().mean().item()
    print(json.dumps({
        "what": "lens readout with LL-Top-10 by a second sweep (no nesting): stock PyTorch ops vs the fused kernels",
        "rows": n, "library_ms": ms_t, "fused_ms": ms_f,
        "library_rows_per_s": n / ms_t * 1e3, "fused_rows_per_s": n / ms_f * 1e3, "ratio": ms_t / ms_f,
        "secret_prob_max_abs_diff": (sec_t - sec_f).abs().max().item(),
        "rows_with_equal_top10_sets": same_top, "prompts_with_equal_ll_top10_sets": same_ll,
        "library_peak_mem_gb": torch.cuda.max_memory_allocated() / 1e9}), flush=True)


if __name__ == "__main__":
    main()
