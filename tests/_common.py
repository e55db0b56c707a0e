"""Shared helpers for the parity tests: seeded problems (CPU fp32, bf16-exact) and comparators."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import oracle  # noqa: E402  (tests are allowed to use the oracle as the checker)
from taboo_brittleness_b200 import synthetic  # noqa: E402

BF16 = torch.bfloat16


def cuda_bf16(t):
    return t.to("cuda", BF16).contiguous()


def sae_to_cuda(sae):
    from taboo_brittleness_b200 import ops
    return ops.SaeWeights.from_sae_lens_layout(sae.W_enc, sae.b_enc, sae.threshold, sae.W_dec, sae.b_dec)


def topk_sets_match(kern_ids, ref_probs, k=10, rel_gap=1e-4):
    """Per-row id sets must be equal; a row may differ only where the oracle's own ranking has a
    near-tie at the cut (|p_k - p_{k+1}| <= rel_gap * p_k), or where the differing ids have
    oracle probabilities within rel_gap of the k-th. Returns (n_exact, n_tie, bad_rows)."""
    ref_vals, ref_ids = torch.topk(ref_probs, k + 1, dim=-1)
    n_exact = n_tie = 0
    bad = []
    for r in range(ref_probs.shape[0]):
        ks = set(int(i) for i in kern_ids[r].tolist())
        rs = set(int(i) for i in ref_ids[r, :k].tolist())
        if ks == rs:
            n_exact += 1
            continue
        pk = float(ref_vals[r, k - 1])
        diff = (ks ^ rs)
        ok = all(0 <= i < ref_probs.shape[1] and abs(float(ref_probs[r, i]) - pk) <= rel_gap * max(pk, 1e-30)
                 for i in diff)
        if ok:
            n_tie += 1
        else:
            bad.append(r)
    return n_exact, n_tie, bad
