"""Seeded synthetic inputs of the Gemma-2-9B / Gemma Scope shapes.

There are no weights, tokenizer or datasets offline, so parity tests and the benchmark run on
random-init tensors whose statistics follow the reference's real cached residuals
(src/data/processed/*/prompt_NN.npz): non-<bos> rows ~ N(0, 8.5^2) per element with a few
heavy outlier dimensions (mean |x| about 280 / 125 / 65), first row(s) of each prompt scaled
to the <bos> norm (about 7630). Everything is generated in fp32 and rounded to bf16 ONCE; the
kernels get the bf16 tensors, the oracle gets the same values upcast to fp32.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict

import torch

D_MODEL = 3584
SAE_WIDTH = 16384
VOCAB = 256000
SEED_WEIGHTS = 42     # configs/default.yaml:11 experiment.seed
SEED_DATA = 1234
OUTLIER_DIMS = ((504, 282.0), (1509, 125.0), (1235, 65.0))
BOS_NORM = 7630.0


def _gen(seed: int, device: str = "cpu") -> torch.Generator:
    return torch.Generator(device=device).manual_seed(int(seed))


def bf16_exact(t: torch.Tensor) -> torch.Tensor:
    """Round an fp32 tensor to bf16 and return it as fp32 (bit-exact bf16 values)."""
    return t.to(torch.bfloat16).to(torch.float32)


def make_residuals(n_prompts: int, rows_per_prompt: int, d: int = D_MODEL, seed: int = SEED_DATA,
                   bos_rows: int = 1, device: str = "cpu") -> torch.Tensor:
    """[n_prompts * rows_per_prompt, d] fp32 with bf16-exact values."""
    g = _gen(seed, device)
    n = n_prompts * rows_per_prompt
    x = torch.randn(n, d, generator=g, device=device, dtype=torch.float32) * 8.5
    for dim, mag in OUTLIER_DIMS:
        if dim < d:
            x[:, dim] = mag * (1.0 + 0.1 * torch.randn(n, generator=g, device=device))
    if bos_rows > 0:
        xv = x.view(n_prompts, rows_per_prompt, d)
        head = xv[:, :bos_rows]
        xv[:, :bos_rows] = head * (BOS_NORM / head.norm(dim=-1, keepdim=True))
    return bf16_exact(x)


@dataclass
class SyntheticSae:
    """fp32 (bf16-exact where the kernels use bf16) SAE parameters in sae_lens layout."""
    W_enc: torch.Tensor      # [d, S]
    b_enc: torch.Tensor      # [S]
    threshold: torch.Tensor  # [S]
    W_dec: torch.Tensor      # [S, d]
    b_dec: torch.Tensor      # [d]

    def as_dict(self) -> Dict[str, torch.Tensor]:
        return {"W_enc": self.W_enc, "b_enc": self.b_enc, "threshold": self.threshold,
                "W_dec": self.W_dec, "b_dec": self.b_dec}


def make_sae(d: int = D_MODEL, S: int = SAE_WIDTH, seed: int = SEED_WEIGHTS, target_l0: float = 76.0,
             calib_rows: int = 512, device: str = "cpu") -> SyntheticSae:
    """Random-init JumpReLU SAE: W_enc ~ N(0, 1/d), unit-norm W_dec rows, b_enc ~ N(0, 0.1^2),
    b_dec ~ N(0, 1); each latent's threshold is the (1 - L0/S) quantile of its pre-activation
    on a calibration batch of ordinary (non-<bos>) rows, so the average L0 is about target_l0
    (Gemma Scope id layer_31/width_16k/average_l0_76, 02_run_sae_baseline.py:22)."""
    g = _gen(seed, device)
    W_enc = bf16_exact(torch.randn(d, S, generator=g, device=device) * (d ** -0.5))
    W_dec = torch.randn(S, d, generator=g, device=device)
    W_dec = bf16_exact(W_dec / W_dec.norm(dim=1, keepdim=True))
    b_enc = torch.randn(S, generator=g, device=device) * 0.1
    b_dec = torch.randn(d, generator=g, device=device)
    calib = make_residuals(1, calib_rows, d, seed=seed + 1, bos_rows=0, device=device)
    pre = calib @ W_enc + b_enc
    k = max(1, min(calib_rows, int(round(calib_rows * (1.0 - target_l0 / S)))))
    threshold = torch.kthvalue(pre, k, dim=0).values.clamp_min(1e-3)
    return SyntheticSae(W_enc, b_enc.float(), threshold.float(), W_dec, b_dec.float())


@dataclass
class SyntheticLens:
    w_norm: torch.Tensor   # [d] fp32 (bf16-exact: the checkpoint stores bf16)
    W_U: torch.Tensor      # [V, d] fp32 with bf16-exact values


def make_lens(d: int = D_MODEL, V: int = VOCAB, seed: int = SEED_WEIGHTS, logit_std: float = 7.0,
              device: str = "cpu", dtype: torch.dtype = torch.float32) -> SyntheticLens:
    """w_norm ~ N(0, 0.5^2); W_U ~ N(0, s^2) with s chosen so that logits of a unit-RMS
    normalised row have std about ``logit_std`` (peaked softmax; the 30.0 softcap is exercised)."""
    g = _gen(seed + 7, device)
    w_norm = bf16_exact(torch.randn(d, generator=g, device=device) * 0.5)
    # h = x_hat * (1 + w): E[h^2] ~ 1 + 0.25  ->  logit var = d * 1.25 * s^2
    s = logit_std / (d * 1.25) ** 0.5
    W_U = torch.empty(V, d, device=device, dtype=dtype)
    step = 32768
    for v0 in range(0, V, step):
        v1 = min(V, v0 + step)
        blk = torch.randn(v1 - v0, d, generator=g, device=device) * s
        W_U[v0:v1] = blk.to(torch.bfloat16).to(dtype)
    return SyntheticLens(w_norm, W_U)


def make_secret_ids(n_prompts: int, rows_per_prompt: int, V: int = VOCAB, seed: int = SEED_DATA) -> torch.Tensor:
    """One fixed secret token id per prompt, repeated over its rows -> i32 [rows]."""
    g = _gen(seed + 3)
    ids = torch.randint(0, V, (n_prompts,), generator=g, dtype=torch.int64)
    return ids.repeat_interleave(rows_per_prompt).to(torch.int32)


def plant_secret_direction(lens: SyntheticLens, x: torch.Tensor, secret_id: torch.Tensor,
                           boost: float = 45.0) -> None:
    """Make each secret token's unembedding row point along the normalised residual of the first
    row that carries it, so that row's secret logit rises by ``boost`` and the other rows of the
    prompt (which share the outlier dimensions, cosine ~0.3-0.5) get a partial boost: secret
    probabilities then spread from ~1/V to ~1 instead of all being ~1/V, which gives the 1e-3
    tolerance something to bite on. In place on lens.W_U."""
    xf = x.float()
    h = xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + 1e-6) * (1.0 + lens.w_norm.float())
    seen = set()
    for r, sid in enumerate(secret_id.tolist()):
        if sid in seen:
            continue
        seen.add(sid)
        row = lens.W_U[sid].float() + (boost / float(h[r].norm()) ** 2) * h[r]
        lens.W_U[sid] = row.to(torch.bfloat16).to(lens.W_U.dtype)
