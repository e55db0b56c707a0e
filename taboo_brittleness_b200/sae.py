"""A JumpReLU SAE object with the slice of the ``sae_lens.SAE`` surface the reference uses
(src/02_run_sae_baseline.py:10,30-36,67): ``SAE.from_pretrained(release=, sae_id=, device=) ->
(sae, cfg_dict, sparsity)`` and ``sae.encode(Tensor[T, d]) -> Tensor[T, S]``; the arithmetic runs
in the sm_100a encode kernel (K1).

``sae_lens`` and the Gemma Scope parameters are not available offline, so ``from_pretrained`` reads a
local Gemma Scope ``params.npz`` (keys W_enc [d,S], W_dec [S,d], b_enc, b_dec, threshold) when
``TABOO_SAE_PARAMS`` points at one, and otherwise builds a seeded random-init SAE of the right shapes
(and says so). ``encode`` returns the dense [T, S] activations for drop-in use at the reference's
small T; ``encode_sparse`` is what the sweeps use.
"""
from __future__ import annotations

import os
import re
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import ops
from .models import find_model_response_start

GEMMA_SCOPE_RELEASE = "google/gemma-scope-9b-it-res"            # 02_run_sae_baseline.py:21
GEMMA_SCOPE_ID = "layer_31/width_16k/average_l0_76"             # 02_run_sae_baseline.py:22


class SAE:
    def __init__(self, weights: ops.SaeWeights, cfg: Optional[Dict] = None):
        self.weights = weights
        self.cfg = cfg or {}

    @property
    def d_in(self) -> int:
        return self.weights.d

    @property
    def d_sae(self) -> int:
        return self.weights.S

    @classmethod
    def from_params(cls, W_enc, b_enc, threshold, W_dec, b_dec, device="cuda", cfg=None) -> "SAE":
        return cls(ops.SaeWeights.from_sae_lens_layout(W_enc, b_enc, threshold, W_dec, b_dec, device=device), cfg)

    @classmethod
    def from_pretrained(cls, release: str = GEMMA_SCOPE_RELEASE, sae_id: str = GEMMA_SCOPE_ID,
                        device: str = "cuda") -> Tuple["SAE", Dict, None]:
        """(sae, cfg_dict, sparsity) like sae_lens. Reads ``$TABOO_SAE_PARAMS`` (a Gemma Scope
        params.npz) if set; else a seeded random-init SAE with the width the id names."""
        path = os.environ.get("TABOO_SAE_PARAMS")
        m = re.search(r"width_(\d+)k", sae_id)
        width = int(m.group(1)) * 1024 if m else 16384
        cfg = {"release": release, "sae_id": sae_id, "d_in": 3584, "d_sae": width, "architecture": "jumprelu"}
        if path:
            with np.load(path) as z:
                t = {k: torch.from_numpy(z[k]) for k in ("W_enc", "b_enc", "threshold", "W_dec", "b_dec")}
            cfg.update(d_in=int(t["W_enc"].shape[0]), d_sae=int(t["W_enc"].shape[1]), source=path)
            return cls.from_params(t["W_enc"], t["b_enc"], t["threshold"], t["W_dec"], t["b_dec"], device, cfg), cfg, None
        from . import synthetic
        print(f"[taboo_brittleness_b200] {release}/{sae_id} is not available offline: using a seeded random-init "
              f"JumpReLU SAE of width {width} (set TABOO_SAE_PARAMS to a Gemma Scope params.npz)")
        s = synthetic.make_sae(3584, width, seed=synthetic.SEED_WEIGHTS, device=device)
        cfg["source"] = "random-init"
        return cls.from_params(s.W_enc, s.b_enc, s.threshold, s.W_dec, s.b_dec, device, cfg), cfg, None

    # ---- the reference's call ------------------------------------------------------------------
    def _operand(self, x: torch.Tensor) -> torch.Tensor:
        if x.dim() != 2 or x.shape[1] != self.d_in:
            raise ops.TabooKernelError(f"expected [T, {self.d_in}] residuals, got {tuple(x.shape)}")
        return x.to(self.weights.w_dec.device, torch.bfloat16).contiguous()

    def encode_sparse(self, x: torch.Tensor, cap: int = 128) -> ops.SparseActs:
        return ops.sae_encode(self._operand(x), self.weights, cap=cap)

    def encode(self, x: torch.Tensor) -> torch.Tensor:
        """Dense [T, S] float32 activations (``sae.encode`` of 02:67)."""
        return self.encode_sparse(x).to_dense()


def get_top_sae_features(sae: SAE, residual_stream: torch.Tensor, input_words: Sequence[str], top_k: int) -> List[int]:
    """02_run_sae_baseline.py:53-74 without the dense activations: response slice -> K1 encode ->
    per-latent mean over the response tokens -> top-k latent ids (descending)."""
    start = find_model_response_start(input_words)
    resp = residual_stream[start:]
    if resp.shape[0] == 0:
        return []
    acts = sae.encode_sparse(resp)
    mean = ops.latent_mean(acts, rows_per_prompt=resp.shape[0])
    idx, _, _ = ops.topm_select(mean, top_k)
    return [int(i) for i in idx[0].tolist()]
