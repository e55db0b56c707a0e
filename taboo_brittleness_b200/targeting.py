"""Targeting of the interventions as the Execution Plan describes it (Plan:116-124, 144-146) -- prose in
the reference, no code: which token positions ("spike tokens"), which SAE latents ("secret score") and
which residual directions (PCA of spike residuals) an intervention acts on.

* ``spike_positions``: the K response positions with the highest layer-32 secret probability.
* ``secret_scores``: per prompt and latent, (mean activation at the spike positions) x
  max(0, Pearson correlation over the prompt's positions between the latent's activation and the secret
  probability). The four per-latent sums it needs come from ``tb_sae_latent_wsum`` on the sparse
  activations (no dense [rows, S] matrix); ``topm_select`` turns the scores into the rank table the
  ablation kernel reads -- the same path as the mean-activation targeting of 02_run_sae_baseline.py.
* ``pca_directions``: orthonormal top-r principal directions of the mean-centred spike residuals
  (a d x d eigenproblem: torch.linalg.eigh; not a hot path).
* ``spike_row_table``: restricts an ablation to the spike rows (all other rows point at a table that
  ablates nothing), Plan:126 "at each identified spike position".
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import ops
from ._lib import TB_RANK_NONE, TB_TOPM_MAX


def spike_positions(secret_prob: torch.Tensor, rows_per_prompt: int, k: int = 4,
                    exclude: Optional[torch.Tensor] = None) -> torch.Tensor:
    """[P, k] positions (within the prompt) of the k largest secret probabilities, descending.
    ``exclude`` (bool [rows]): positions never chosen (e.g. tokens that are the secret word itself)."""
    p = secret_prob.view(-1, rows_per_prompt).clone()
    if exclude is not None:
        p[exclude.view(-1, rows_per_prompt)] = -1.0
    return torch.topk(p, min(k, rows_per_prompt), dim=1).indices


def spike_mask(spikes: torch.Tensor, rows_per_prompt: int) -> torch.Tensor:
    """bool [P * rows_per_prompt]: True at spike rows."""
    P = spikes.shape[0]
    m = torch.zeros(P, rows_per_prompt, dtype=torch.bool, device=spikes.device)
    m.scatter_(1, spikes, True)
    return m.view(-1)


def secret_scores(acts: ops.SparseActs, secret_prob: torch.Tensor, rows_per_prompt: int,
                  spikes: torch.Tensor) -> torch.Tensor:
    """f32 [P, S] secret score of every latent (Plan:122-124)."""
    T = rows_per_prompt
    p = secret_prob.to(torch.float32).contiguous()
    w_spike = spike_mask(spikes, T).to(torch.float32) / float(spikes.shape[1])
    mean_spike = ops.latent_wsum(acts, T, row_weight=w_spike)            # mean activation at the spike tokens
    s_a = ops.latent_wsum(acts, T)                                       # sum_t a
    s_aa = ops.latent_wsum(acts, T, power=2)                             # sum_t a^2
    s_ap = ops.latent_wsum(acts, T, row_weight=p)                        # sum_t a p
    pp = p.view(-1, T)
    mean_p = pp.mean(1, keepdim=True)
    var_p = (pp - mean_p).pow(2).mean(1, keepdim=True)
    mean_a = s_a / T
    cov = s_ap / T - mean_a * mean_p
    var_a = (s_aa / T - mean_a * mean_a).clamp_min(0.0)
    corr = torch.where((var_a > 0) & (var_p > 0), cov / torch.sqrt(var_a * var_p).clamp_min(1e-30),
                       torch.zeros_like(cov))
    return mean_spike * corr.clamp_min(0.0)


def score_rank_table(scores: torch.Tensor, m: int = TB_TOPM_MAX) -> Tuple[torch.Tensor, torch.Tensor]:
    """(idx i32 [P, m], rank u16 [P, S]) of the m best-scoring latents per prompt, for the ablation kernel."""
    idx, _, rank = ops.topm_select(scores.contiguous(), min(m, scores.shape[1]))
    return idx, rank


def spike_row_table(spikes: torch.Tensor, rows_per_prompt: int) -> torch.Tensor:
    """i32 [rows]: rank-table index per row for ``ops.sae_ablate_decode(row_table=...)`` -- a spike row
    uses its prompt's table p, every other row table P, which ``append_null_table`` makes all-NONE."""
    P = spikes.shape[0]
    table = torch.full((P, rows_per_prompt), P, dtype=torch.int32, device=spikes.device)
    table.scatter_(1, spikes, torch.arange(P, dtype=torch.int32, device=spikes.device)[:, None].expand_as(spikes))
    return table.view(-1).contiguous()


def append_null_table(rank: torch.Tensor) -> torch.Tensor:
    """rank u16 [P, S] -> [P + 1, S] with a last table that selects no latent."""
    null = torch.full((1, rank.shape[1]), TB_RANK_NONE, dtype=torch.int32, device=rank.device).to(torch.uint16)
    return torch.cat([rank, null], 0).contiguous()


def pca_directions(spike_residuals: torch.Tensor, r: int) -> torch.Tensor:
    """Top-r principal directions (orthonormal columns, f32 [d, r]) of the mean-centred rows of
    ``spike_residuals`` [n, d] (Plan:144-146). Sign convention: each direction's largest-magnitude
    component is positive, so the result does not depend on the eigensolver's sign choice."""
    x = spike_residuals.to(torch.float32)
    x = x - x.mean(0, keepdim=True)
    cov = x.t() @ x
    _, vecs = torch.linalg.eigh(cov)
    U = vecs[:, -r:].flip(1).contiguous()
    big = U.abs().argmax(0)
    return U * torch.sign(U[big, torch.arange(r, device=U.device)])[None]
