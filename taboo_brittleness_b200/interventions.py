"""The intervention layer the reference planned (``interventions.py``, Execution Plan:41) but never
wrote: SAE latent ablation (targeted, Plan:116-126; random control, Plan:128) and low-rank
direction removal (Plan:144-150), each followed by the logit-lens readout.

Everything stays on the GPU between the encode and the merged readout:
    x --K1 encode--> sparse acts --mean_t, top-m--> rank table (all nested budgets at once)
      --K1b ablate+decode (+fused final RMSNorm)--> h --K3 lens--> partials --K4--> readout
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, Iterable, List, Optional, Sequence

import torch

from . import ops
from ._lib import TB_TOPM_MAX

DEFAULT_BUDGETS = (1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024)   # BASELINE.json configs[1]
DEFAULT_RANKS = (1, 2, 4, 8, 16, 32, 64)                           # BASELINE.json configs[3]


def random_rank_table(S: int, seed: int, m: int = TB_TOPM_MAX, device="cuda") -> torch.Tensor:
    """Rank table of the random-latent control: first m entries of a seeded CPU permutation
    (``torch.randperm(S, generator=manual_seed(seed))``, the convention of SURVEY section 5), so the
    host, the oracle and the kernels share the ablated indices bit for bit."""
    g = torch.Generator(device="cpu").manual_seed(int(seed))
    idx = torch.randperm(S, generator=g)[:m]
    return ops.rank_table_from_indices(idx, S).to(device)


@dataclass
class SweepStats:
    launches: int = 0
    lens_events: List = field(default_factory=list)   # (start, end, flops) per lens launch


class InterventionEngine:
    """Holds one SAE + one lens (norm weight and unembedding shard) and runs sweeps over them."""

    def __init__(self, sae: ops.SaeWeights, lens: ops.LensWeights, *, softcap: float = ops.GEMMA2_SOFTCAP,
                 cta_pair: bool = True, n_chunks: Optional[int] = None, encode_cap: int = 128):
        self.sae, self.lens = sae, lens
        self.softcap, self.cta_pair, self.n_chunks, self.encode_cap = softcap, cta_pair, n_chunks, encode_cap
        self.stats = SweepStats()
        self.time_lens = False

    # ---- pieces ---------------------------------------------------------------------------------
    def encode(self, x: torch.Tensor) -> ops.SparseActs:
        acts = ops.sae_encode(x, self.sae, cap=self.encode_cap)
        self.stats.launches += 1 if acts.row_start is None else 2
        return acts

    def targeted_rank(self, acts: ops.SparseActs, rows_per_prompt: int = 0, seg: Optional[torch.Tensor] = None,
                      m: int = TB_TOPM_MAX):
        """(idx [P,m], val [P,m], rank u16 [P,S]) of the top-m mean activations per prompt."""
        mean = ops.latent_mean(acts, rows_per_prompt, seg)
        out = ops.topm_select(mean, min(m, acts.S))
        self.stats.launches += 2
        return out

    def read(self, h: torch.Tensor, secret_id: Optional[torch.Tensor], finalize: bool = True):
        """Lens readout of an already normalised operand."""
        if self.time_lens:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        p = ops.lens_partials(h, self.lens.w_u, softcap=self.softcap, secret_id=secret_id,
                              vocab_offset=self.lens.vocab_offset, n_chunks=self.n_chunks,
                              cta_pair=self.cta_pair)
        if self.time_lens:
            e1.record()
            self.stats.lens_events.append((e0, e1, 2.0 * h.shape[0] * h.shape[1] * self.lens.w_u.shape[0]))
        out = ops.lens_merge(p, finalize=finalize)
        self.stats.launches += 2
        return out

    def ablate_and_read(self, x, acts, rank, budget: int, *, rank_rows: int = 0, row_table=None,
                        secret_id=None, mode: str = "error", finalize: bool = True):
        _, h = ops.sae_ablate_decode(x, acts, self.sae, rank, budget, mode=mode, rank_rows=rank_rows,
                                     row_table=row_table, w_norm=self.lens.w_norm, eps=self.lens.eps,
                                     want_x=False, want_h=True)
        self.stats.launches += 1
        return self.read(h, secret_id, finalize)

    def project_and_read(self, x, U, *, secret_id=None, finalize: bool = True):
        _, h = ops.lowrank_remove(x, U, w_norm=self.lens.w_norm, eps=self.lens.eps, want_x=False, want_h=True)
        self.stats.launches += 1
        return self.read(h, secret_id, finalize)

    # ---- sweeps -----------------------------------------------------------------------------------
    def targeted_budget_sweep(self, x: torch.Tensor, rows_per_prompt: int, secret_id: Optional[torch.Tensor],
                              budgets: Sequence[int] = DEFAULT_BUDGETS, mode: str = "error"
                              ) -> Dict[int, ops.LensReadout]:
        """BASELINE.json configs[1]: encode once, target once (top-1024 of the per-prompt mean),
        then one ablate+decode+lens pass per budget (budgets are prefixes of the sorted list)."""
        acts = self.encode(x)
        _, _, rank = self.targeted_rank(acts, rows_per_prompt, m=min(max(budgets), TB_TOPM_MAX))
        return {int(b): self.ablate_and_read(x, acts, rank, int(b), rank_rows=rows_per_prompt,
                                             secret_id=secret_id, mode=mode) for b in budgets}

    def random_budget_sweep(self, x: torch.Tensor, secret_id, budgets: Sequence[int] = DEFAULT_BUDGETS,
                            seeds: Iterable[int] = range(32), mode: str = "error",
                            acts: Optional[ops.SparseActs] = None) -> Dict[tuple, ops.LensReadout]:
        """BASELINE.json configs[2]: random-latent control, (budget, seed) conditions."""
        acts = acts if acts is not None else self.encode(x)
        out = {}
        for seed in seeds:
            rank = random_rank_table(self.sae.S, seed, min(max(budgets), TB_TOPM_MAX), x.device)
            for b in budgets:
                out[(int(b), int(seed))] = self.ablate_and_read(x, acts, rank, int(b), rank_rows=0,
                                                                secret_id=secret_id, mode=mode)
        return out

    def projection_sweep(self, x: torch.Tensor, U: torch.Tensor, secret_id,
                         ranks: Sequence[int] = DEFAULT_RANKS) -> Dict[int, ops.LensReadout]:
        """BASELINE.json configs[3]: remove the first r columns of U for each r."""
        return {int(r): self.project_and_read(x, U[:, :r].contiguous(), secret_id=secret_id) for r in ranks}
