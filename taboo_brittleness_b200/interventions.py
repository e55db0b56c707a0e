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
    ll_fused_prompts: int = 0        # prompt-conditions that took the one-sweep LL path ...
    ll_fallback_prompts: int = 0     # ... and how many of them had to be redone with the second sweep
    ll_candidates: int = 0           # candidate columns listed by K3F over those prompt-conditions
    lens_events: List = field(default_factory=list)   # (start, end, flops) per lens launch
    ll_events: List = field(default_factory=list)     # (start, end, flops) per LL-Top-k sweep
    row_events: List = field(default_factory=list)    # (start, end, algorithmic bytes) per K1b / K2 call


@dataclass
class LLSpec:
    """What the LL-Top-k readout (01_reproduce_logit_lens.py:35-71, 147-149) needs besides the lens
    operand: the ids zeroed at every response position and how rows map to prompts."""
    cur_id: Optional[torch.Tensor]          # i32 [rows]: id of the token at that position
    prev_id: Optional[torch.Tensor]         # i32 [rows]: id of the previous token (-1: none)
    rows_per_prompt: int = 0                # prompts are consecutive blocks of this many rows, or
    seg: Optional[torch.Tensor] = None      # i64 [P+1] row windows (the response slices)
    k: int = 10


@dataclass
class LLState:
    """What one condition hands to the next, nested one, so that it can take the one-sweep path:
    its rows' exact logsumexp (the next sweep's softmax reference) and, per prompt, the k-th best
    LL sum (the scale of the next candidate threshold)."""
    lse: torch.Tensor       # f32 [rows]
    theta: torch.Tensor     # f32 [P]


@dataclass
class Readout:
    """Per-row lens readout plus, when asked for, the per-prompt LL-Top-k lists."""
    lens: ops.LensReadout
    ll_sum: Optional[torch.Tensor] = None   # f32 [P, k] masked column sums, descending
    ll_id: Optional[torch.Tensor] = None    # i32 [P, k]

    def __getattr__(self, name):            # readout.secret_prob etc. keep working
        return getattr(self.lens, name)


def _free_bytes(dev) -> int:
    """Bytes a new tensor can still get: what the driver reports free plus what torch's caching allocator holds unused."""
    return int(torch.cuda.mem_get_info(dev)[0] + torch.cuda.memory_reserved(dev) - torch.cuda.memory_allocated(dev))


class InterventionEngine:
    """Holds one SAE + one lens (norm weight and unembedding shard) and runs sweeps over them."""

    def __init__(self, sae: ops.SaeWeights, lens: ops.LensWeights, *, softcap: float = ops.GEMMA2_SOFTCAP,
                 cta_pair=5, n_chunks: Optional[int] = None, encode_cap: int = 128):
        self.sae, self.lens = sae, lens
        self.softcap, self.cta_pair, self.n_chunks, self.encode_cap = softcap, cta_pair, n_chunks, encode_cap
        self.stats = SweepStats()
        self.time_lens = False
        self.ll_fused = True         # nested sweeps: one GEMM sweep per condition after the first
        # K1b / K6 on a side stream under the lens sweeps (on the 16 SMs a grid of 4-CTA clusters leaves
        # free). Measured: no gain (574 k vs 579 k token-conditions/s) -- the side kernels share the power
        # budget and can sit on SMs the next persistent lens grid wants -- so it is off by default.
        self.overlap = False
        self._side = None
        # A condition without a predecessor (the first budget of a chain) gets its exact LL-Top-k from the logits the
        # lens kernel writes in its one sweep (row sub-batches of ll_dense_rows rows, ops.lens_ll_from_logits) instead
        # of a second sweep of the GEMM (K5): rows * V * 4 bytes of HBM traffic each way instead of 2 d V FLOP per row.
        self.ll_dense_first = True
        self.ll_dense_rows = 65536
        self._dense_logits = None
        self._dense_logits2 = None
        self._update_matrix = None
        self.use_ranked = True       # nested budgets read prefixes of rank-ordered lists (ops.rank_lists) in K1b
        self.multi_budget = True     # ... and all budgets of a chain are produced in one pass over the rows when they fit
        self.ll_alpha = 0.5          # candidate threshold = ll_alpha * previous condition's k-th best sum
        self.ll_cand_cap = 256       # candidate slots per prompt

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

    def ll_topk(self, h: torch.Tensor, lse: torch.Tensor, ll: LLSpec):
        """LL-Top-k of an already normalised operand, given the rows' logsumexp (second sweep)."""
        if ll.seg is None and ll.rows_per_prompt == ops.LL_SLOTS:
            hp, lp, cp, pp = h, lse, ll.cur_id, ll.prev_id
        else:
            seg = ll.seg if ll.seg is not None else torch.arange(0, h.shape[0] + 1, ll.rows_per_prompt)
            hp = ops.pad_prompt_rows(h, seg, 0.0)
            lp = ops.pad_prompt_rows(lse, seg, float("inf"))
            cp = ops.pad_prompt_rows(ll.cur_id, seg, -1) if ll.cur_id is not None else None
            pp = ops.pad_prompt_rows(ll.prev_id, seg, -1) if ll.prev_id is not None else None
        if self.time_lens:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        out = ops.ll_topk(hp, self.lens.w_u, lp, cp, pp, k=ll.k, softcap=self.softcap,
                          vocab_offset=self.lens.vocab_offset, cta_pair=self.cta_pair)
        if self.time_lens:
            e1.record()
            self.stats.ll_events.append((e0, e1, 2.0 * hp.shape[0] * hp.shape[1] * self.lens.w_u.shape[0]))
        self.stats.launches += 2
        return out

    def read_fused(self, h: torch.Tensor, secret_id: Optional[torch.Tensor], ll: LLSpec, prev: LLState):
        """Lens readout + LL-Top-k in ONE sweep (K3F + K4 + K6), using the previous nested condition's
        logsumexp as the softmax reference. Prompts whose top-k the certificate does not cover are
        redone exactly with the two-sweep kernel; the result is exact either way.
        Needs the 128-rows-per-prompt layout. Returns (Readout, LLState for the next condition)."""
        tau = (prev.theta * self.ll_alpha).clamp_min(1e-30)
        if self.time_lens:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        p, cnt, cid = ops.lens_ll_partials(h, self.lens.w_u, prev.lse, tau, ll.cur_id, ll.prev_id,
                                           softcap=self.softcap, secret_id=secret_id,
                                           vocab_offset=self.lens.vocab_offset, n_chunks=self.n_chunks,
                                           cta_pair=self.cta_pair, cand_cap=self.ll_cand_cap)
        if self.time_lens:
            e1.record()
            self.stats.lens_events.append((e0, e1, 2.0 * h.shape[0] * h.shape[1] * self.lens.w_u.shape[0]))
        out = ops.lens_merge(p, finalize=True)
        val, ids, theta, flag = ops.ll_candidates_exact(h, self.lens.w_u, out.lse, prev.lse, tau, ll.cur_id,
                                                        ll.prev_id, cnt, cid, k=ll.k, softcap=self.softcap,
                                                        vocab_offset=self.lens.vocab_offset)
        self.stats.launches += 3
        n_cand = cnt.sum()
        bad = torch.nonzero(flag == 0).flatten()          # host sync: a few bytes per condition
        self.stats.ll_candidates += int(n_cand.item())
        self.stats.ll_fallback_prompts += int(bad.numel())
        self.stats.ll_fused_prompts += int(flag.numel())
        if bad.numel():
            rows = (bad[:, None] * ops.LL_SLOTS + torch.arange(ops.LL_SLOTS, device=h.device)[None]).flatten()
            sub = LLSpec(ll.cur_id[rows].contiguous() if ll.cur_id is not None else None,
                         ll.prev_id[rows].contiguous() if ll.prev_id is not None else None,
                         rows_per_prompt=ops.LL_SLOTS, k=ll.k)
            bv, bi = self.ll_topk(h[rows].contiguous(), out.lse[rows].contiguous(), sub)
            val[bad], ids[bad], theta[bad] = bv, bi, bv[:, ll.k - 1]
        return Readout(out, val, ids), LLState(out.lse, theta)

    def release_scratch(self) -> None:
        """Drop the logits scratch of ``read_dense_ll`` (tens of GB) and the rank-update matrix."""
        self._dense_logits = None
        self._dense_logits2 = None
        self._update_matrix = None

    def read_dense_ll(self, h: torch.Tensor, secret_id: Optional[torch.Tensor], ll: LLSpec) -> Readout:
        """Lens readout + exact LL-Top-k of a condition with no predecessor, one GEMM sweep (see ``ll_dense_first``)."""
        n, V = h.shape[0], self.lens.w_u.shape[0]
        step = max(ops.LL_SLOTS, min(self.ll_dense_rows, n) // ops.LL_SLOTS * ops.LL_SLOTS)
        free = _free_bytes(h.device)
        have = 0 if self._dense_logits is None else self._dense_logits.numel() * 4
        while step > 8 * ops.LL_SLOTS and step * V * 4 > have + 0.8 * free:    # do not take more than most of what is free
            step = step // 2 // ops.LL_SLOTS * ops.LL_SLOTS
        if self._dense_logits is None or self._dense_logits.shape[0] < step or self._dense_logits.shape[1] != V:
            self._dense_logits = None
            self._dense_logits = torch.empty(step, V, dtype=torch.float32, device=h.device)
        outs, sums, ids = [], [], []
        for r0 in range(0, n, step):
            r1 = min(n, r0 + step)
            sl = lambda t: None if t is None else t[r0:r1]      # noqa: E731
            if self.time_lens:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
            o, sv, si = ops.lens_ll_from_logits(h[r0:r1], self.lens.w_u, sl(ll.cur_id), sl(ll.prev_id), self._dense_logits,
                                                k=ll.k, softcap=self.softcap, secret_id=sl(secret_id), cta_pair=self.cta_pair)
            if self.time_lens:
                e1.record()      # the sweep, its merge and the reduction of the stored logits
                self.stats.lens_events.append((e0, e1, 2.0 * (r1 - r0) * h.shape[1] * V))
            outs.append(o)
            sums.append(sv)
            ids.append(si)
            self.stats.launches += 4
        lens = ops.LensReadout(lse=torch.cat([o.lse for o in outs]), secret_prob=torch.cat([o.secret_prob for o in outs]),
                               top_prob=torch.cat([o.top_prob for o in outs]), top_id=torch.cat([o.top_id for o in outs]))
        return Readout(lens, torch.cat(sums), torch.cat(ids))

    def read(self, h: torch.Tensor, secret_id: Optional[torch.Tensor], finalize: bool = True,
             ll: Optional[LLSpec] = None):
        """Lens readout of an already normalised operand (+ LL-Top-k when ``ll`` is given)."""
        if (ll is not None and finalize and self.ll_dense_first and self.lens.vocab_offset == 0 and ll.seg is None
                and ll.rows_per_prompt == ops.LL_SLOTS and h.shape[0] % ops.LL_SLOTS == 0 and h.shape[0] > 0):
            return self.read_dense_ll(h, secret_id, ll)
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
        if ll is None or not finalize:
            return out
        ll_sum, ll_id = self.ll_topk(h, out.lse, ll)
        return Readout(out, ll_sum, ll_id)

    def ablate_and_read(self, x, acts, rank, budget: int, *, rank_rows: int = 0, row_table=None,
                        secret_id=None, mode: str = "error", finalize: bool = True, ll: Optional[LLSpec] = None,
                        ll_state: Optional[LLState] = None, ranked: Optional[ops.RankedActs] = None):
        """With ``ll_state`` (from the previous nested condition) returns (Readout, next LLState).
        ``ranked`` (``ops.rank_lists`` of the same acts and ranking; error-term mode only) lets the budget read
        just the leading entries of rank-ordered lists instead of looking every active latent up in the table."""
        if self.time_lens:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        if ranked is not None and mode == "error" and row_table is None:
            _, h = ops.sae_ablate_ranked(x, ranked, self.sae, budget, w_norm=self.lens.w_norm, eps=self.lens.eps,
                                         want_x=False, want_h=True)
        else:
            _, h = ops.sae_ablate_decode(x, acts, self.sae, rank, budget, mode=mode, rank_rows=rank_rows,
                                         row_table=row_table, w_norm=self.lens.w_norm, eps=self.lens.eps,
                                         want_x=False, want_h=True)
        if self.time_lens:
            e1.record()      # algorithmic traffic: x in, h out
            self.stats.row_events.append((e0, e1, 4.0 * x.shape[0] * x.shape[1]))
        self.stats.launches += 1
        if ll_state is not None:
            return self.read_fused(h, secret_id, ll, ll_state)
        return self.read(h, secret_id, finalize, ll)

    def project_and_read(self, x, U, *, secret_id=None, finalize: bool = True, ll: Optional[LLSpec] = None,
                         ll_state: Optional[LLState] = None):
        """With ``ll_state`` (from the previous, smaller rank) returns (Readout, next LLState)."""
        if self.time_lens:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        _, h = ops.lowrank_remove(x, U, w_norm=self.lens.w_norm, eps=self.lens.eps, want_x=False, want_h=True)
        if self.time_lens:
            e1.record()
            self.stats.row_events.append((e0, e1, 4.0 * x.shape[0] * x.shape[1]))
        self.stats.launches += 1
        if ll_state is not None:
            return self.read_fused(h, secret_id, ll, ll_state)
        return self.read(h, secret_id, finalize, ll)

    # ---- sweeps -----------------------------------------------------------------------------------
    def _nested_budgets(self, x, acts, rank, rank_rows: int, budgets: Sequence[int], secret_id, mode: str,
                        ll: Optional[LLSpec], state: Optional[LLState] = None):
        """Budgets that are prefixes of ONE rank table, ascending: every condition's exact lse is the next
        one's softmax reference, so with LL-Top-k only the first condition needs two sweeps (and not even
        that one when ``state`` from a closely related earlier chain is passed in). Returns
        ({budget: readout}, last LLState)."""
        fusable = (ll is not None and self.ll_fused and ll.seg is None and ll.rows_per_prompt == ops.LL_SLOTS
                   and self.lens.vocab_offset == 0)
        if fusable and self.overlap and len(budgets) >= 3:
            return self._nested_budgets_overlapped(x, acts, rank, rank_rows, budgets, secret_id, mode, ll, state)
        out = {}
        ranked = None
        if self.use_ranked and mode == "error" and len(budgets) > 1:
            # one pass over the active lists for the whole chain: each row's ranked latents in rank order
            if self.time_lens:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
            ranked = ops.rank_lists(acts, rank, rank_rows)
            if self.time_lens:
                e1.record()      # part of K1b's cost, no algorithmic bytes of its own
                self.stats.row_events.append((e0, e1, 0.0))
            self.stats.launches += 1
        bs = sorted(set(int(v) for v in budgets))
        h_all = None
        if ranked is not None and self.multi_budget and 1 < len(bs) <= 16:
            # all budgets' lens operands in ONE pass over the rows (x read once, every W_dec row gathered once): needs
            # len(bs) operands of rows * d bf16 at once
            need = len(bs) * x.shape[0] * x.shape[1] * 2
            if need <= 0.45 * _free_bytes(x.device):
                if self.time_lens:
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                _, h_all, _ = ops.sae_ablate_ranked_multi(x, ranked, self.sae, bs, w_norm=self.lens.w_norm, eps=self.lens.eps)
                if self.time_lens:
                    e1.record()      # algorithmic traffic: x in, h out, per budget
                    self.stats.row_events.append((e0, e1, 4.0 * x.shape[0] * x.shape[1] * len(bs)))
                self.stats.launches += 1
        for i, b in enumerate(bs):
            if h_all is not None:
                if fusable and state is not None:
                    out[b], state = self.read_fused(h_all[i], secret_id, ll, state)
                    continue
                r = self.read(h_all[i], secret_id, True, ll)
            elif fusable and state is not None:
                out[b], state = self.ablate_and_read(x, acts, rank, b, rank_rows=rank_rows, secret_id=secret_id,
                                                     mode=mode, ll=ll, ll_state=state, ranked=ranked)
                continue
            else:
                r = self.ablate_and_read(x, acts, rank, b, rank_rows=rank_rows, secret_id=secret_id, mode=mode, ll=ll,
                                         ranked=ranked)
            out[b] = r
            if fusable:
                state = LLState(r.lens.lse, r.ll_sum[:, ll.k - 1].contiguous())
        return out, state

    def _nested_budgets_overlapped(self, x, acts, rank, rank_rows: int, budgets: Sequence[int], secret_id, mode: str,
                                   ll: LLSpec, state: Optional[LLState]):
        """The same chain with the small kernels taken off the critical path. The lens sweeps cannot use
        all SMs (clusters of four CTAs fill 132 of 148), so while sweep i runs on the main stream a side
        stream produces the operand of the next condition (K1b of budget i+1, i+2) and evaluates the
        LL candidates of the previous one (K6 of budget i-1) on the SMs the cluster grid leaves free.
        What changes numerically: nothing that is returned -- the candidate threshold of condition i
        comes from the k-th best sum of condition i-2 instead of i-1 (K6 of i-1 is still running), the
        certificate is unchanged, and prompts it does not cover are redone with the two-sweep kernel
        once the chain is through."""
        main = torch.cuda.current_stream()
        if self._side is None:
            self._side = torch.cuda.Stream()
        side = self._side
        bs = sorted(int(v) for v in budgets)
        n, k = len(bs), ll.k
        h = [None] * n
        h_ready = [None] * n

        def produce(i):                      # K1b for budget i on the side stream
            with torch.cuda.stream(side):
                _, hi = ops.sae_ablate_decode(x, acts, self.sae, rank, bs[i], mode=mode, rank_rows=rank_rows,
                                              w_norm=self.lens.w_norm, eps=self.lens.eps, want_x=False, want_h=True)
                ev = torch.cuda.Event()
                ev.record(side)
            h[i], h_ready[i] = hi, ev
            self.stats.launches += 1

        side.wait_stream(main)               # x, acts, rank were produced on the main stream
        produce(0)
        if n > 1:
            produce(1)
        out, fused = {}, []                  # fused: per one-sweep condition, what K6 left on the side stream
        theta_done = None                    # (theta tensor, event) of the latest K6 whose result may be used
        theta_now = None if state is None else state.theta
        for i, b in enumerate(bs):
            main.wait_event(h_ready[i])
            if state is None:                # no reference yet: K3 + K5 on the main stream
                r = self.read(h[i], secret_id, True, ll)
                out[b] = r
                state = LLState(r.lens.lse, r.ll_sum[:, k - 1].contiguous())
                theta_now = state.theta
            else:
                if theta_done is not None:   # K6 of condition i-2 finished during sweep i-1
                    main.wait_event(theta_done[1])
                    theta_now = theta_done[0]
                tau = (theta_now * self.ll_alpha).clamp_min(1e-30)
                ref = state.lse
                if self.time_lens:
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                p, cnt, cid = ops.lens_ll_partials(h[i], self.lens.w_u, ref, tau, ll.cur_id, ll.prev_id,
                                                   softcap=self.softcap, secret_id=secret_id,
                                                   vocab_offset=self.lens.vocab_offset, n_chunks=self.n_chunks,
                                                   cta_pair=self.cta_pair, cand_cap=self.ll_cand_cap)
                if self.time_lens:
                    e1.record()
                    self.stats.lens_events.append((e0, e1, 2.0 * h[i].shape[0] * h[i].shape[1] * self.lens.w_u.shape[0]))
                lens_out = ops.lens_merge(p, finalize=True)
                swept = torch.cuda.Event()
                swept.record(main)
                if fused:                    # the previous condition's K6 may be used from the next sweep on
                    theta_done = (fused[-1]["theta"], fused[-1]["done"])
                with torch.cuda.stream(side):
                    side.wait_event(swept)
                    val, ids, theta, flag = ops.ll_candidates_exact(h[i], self.lens.w_u, lens_out.lse, ref, tau,
                                                                    ll.cur_id, ll.prev_id, cnt, cid, k=k,
                                                                    softcap=self.softcap,
                                                                    vocab_offset=self.lens.vocab_offset)
                    done = torch.cuda.Event()
                    done.record(side)
                fused.append(dict(i=i, b=b, lens=lens_out, val=val, ids=ids, theta=theta, flag=flag, cnt=cnt, done=done))
                self.stats.launches += 3
                state = LLState(lens_out.lse, theta)
            if i + 2 < n:
                produce(i + 2)
        main.wait_stream(side)
        # resolve the certificates (one host sync for the whole chain) and redo what they do not cover
        if fused:
            flags = torch.stack([f["flag"] for f in fused])
            counts = torch.stack([f["cnt"].sum() for f in fused]).sum()
            bad_all = torch.nonzero(flags == 0)
            self.stats.ll_candidates += int(counts.item())
            self.stats.ll_fused_prompts += int(flags.numel())
            self.stats.ll_fallback_prompts += int(bad_all.shape[0])
            for j, f in enumerate(fused):
                bad = bad_all[bad_all[:, 0] == j, 1]
                if bad.numel():
                    rows = (bad[:, None] * ops.LL_SLOTS + torch.arange(ops.LL_SLOTS, device=x.device)[None]).flatten()
                    sub = LLSpec(ll.cur_id[rows].contiguous() if ll.cur_id is not None else None,
                                 ll.prev_id[rows].contiguous() if ll.prev_id is not None else None,
                                 rows_per_prompt=ops.LL_SLOTS, k=k)
                    bv, bi = self.ll_topk(h[f["i"]][rows].contiguous(), f["lens"].lse[rows].contiguous(), sub)
                    f["val"][bad], f["ids"][bad], f["theta"][bad] = bv, bi, bv[:, k - 1]
                for t in (f["val"], f["ids"], f["theta"]):
                    t.record_stream(main)    # allocated on the side stream, handed to the caller's stream
                out[f["b"]] = Readout(f["lens"], f["val"], f["ids"])
        return out, state

    def targeted_budget_sweep(self, x: torch.Tensor, rows_per_prompt: int, secret_id: Optional[torch.Tensor],
                              budgets: Sequence[int] = DEFAULT_BUDGETS, mode: str = "error",
                              ll: Optional[LLSpec] = None) -> Dict[int, ops.LensReadout]:
        """BASELINE.json configs[1]: encode once, target once (top-1024 of the per-prompt mean),
        then one ablate+decode+lens pass per budget (budgets are prefixes of the sorted list)."""
        acts = self.encode(x)
        _, _, rank = self.targeted_rank(acts, rows_per_prompt, m=min(max(budgets), TB_TOPM_MAX))
        return self._nested_budgets(x, acts, rank, rows_per_prompt, budgets, secret_id, mode, ll)[0]

    def update_matrix(self) -> torch.Tensor:
        """D = (W_dec (.) (1 + w)) W_U^T, bf16, tile-major, computed on first use and kept (8.4 GB for Gemma-2 / 16k)."""
        if self._update_matrix is None:
            self._update_matrix = ops.lens_update_matrix(self.sae, self.lens)
        return self._update_matrix

    def targeted_budget_sweep_rank_update(self, x: torch.Tensor, secret_id: Optional[torch.Tensor],
                                          budgets: Sequence[int] = DEFAULT_BUDGETS, *, split: bool = False,
                                          sub_rows: int = 65536, ll: Optional[LLSpec] = None):
        """configs[1] with the nested budgets as RANK UPDATES of stored base logits: per row sub-batch one
        sweep of L0 = (x (.) (1 + w)) W_U^T, then per budget rstd_m * (L0 - A_m D[ids]) with a product of depth m instead
        of d. Error-term ablation, 128 rows per prompt, single shard. With ``ll`` every budget's soft-capped logits are
        also stored (a second scratch of the same size) and reduced to the LL-Top-k by the HBM-bound column-sum kernel of
        the first-condition path: exact for this path's logits, no candidate lists. Returns LensReadout per budget, or
        Readout (lens, ll_sum, ll_id) with ``ll``. NOT the arithmetic of the reference path -- the normalised residual is
        not rounded to bf16 per budget and D is a rounded product -- so this is a workload of its own, reported beside the
        dense sweep."""
        T = ops.LL_SLOTS
        n, V = x.shape[0], self.lens.w_u.shape[0]
        if n % T or self.lens.vocab_offset != 0 or V % 32:
            raise ops.TabooKernelError("rank-update sweep: whole prompts of 128 rows, unsharded vocabulary, V % 32 == 0",
                                       ops._lib.TB_ERR_INVALID)
        bs = sorted(int(v) for v in budgets)
        D = self.update_matrix()
        acts = self.encode(x)
        ids, _, rank = self.targeted_rank(acts, T, m=min(max(bs), TB_TOPM_MAX))
        ranked = ops.rank_lists(acts, rank, T)
        A = ops.rank_acts_dense(ranked, n, split)
        if len(bs) <= 16:               # K1b without an output row: only the RMSNorm factors of the rounded x', one pass
            _, _, rstd = ops.sae_ablate_ranked_multi(x, ranked, self.sae, bs, w_norm=self.lens.w_norm, eps=self.lens.eps,
                                                     want_h=False, want_rstd=True)
        else:
            rstd = torch.empty(len(bs), n, dtype=torch.float32, device=x.device)
            for i, b in enumerate(bs):
                ops.sae_ablate_ranked(x, ranked, self.sae, b, w_norm=self.lens.w_norm, eps=self.lens.eps, want_x=False,
                                      want_h=False, out_rstd=rstd[i])
        self.stats.launches += 4
        if ll is not None and (ll.seg is not None or ll.rows_per_prompt != T):
            raise ops.TabooKernelError("rank-update sweep: LL-Top-k needs 128 row slots per prompt", ops._lib.TB_ERR_INVALID)
        n_scratch = 2 if ll is not None else 1
        step = max(T, min(sub_rows, n) // T * T)
        free = _free_bytes(x.device)
        have = sum(0 if t is None else t.numel() * 4 for t in (self._dense_logits, self._dense_logits2)[:n_scratch])
        while step > 8 * T and n_scratch * step * V * 4 > have + 0.8 * free:
            step = step // 2 // T * T
        if self._dense_logits is None or self._dense_logits.numel() < step * V:
            self._dense_logits = None
            self._dense_logits = torch.empty(step, V, dtype=torch.float32, device=x.device)
        if ll is not None and (self._dense_logits2 is None or self._dense_logits2.numel() < step * V):
            self._dense_logits2 = None
            self._dense_logits2 = torch.empty(step * V, dtype=torch.float32, device=x.device)
        l0 = self._dense_logits
        stored = self._dense_logits2 if ll is not None else None
        outs = {b: [] for b in bs}
        lls = {b: [] for b in bs}
        for r0 in range(0, n, step):
            r1 = min(n, r0 + step)
            xw = ops.scale_cols(x[r0:r1], self.lens.w_norm)
            if self.time_lens:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
            ops.lens_base_logits(xw, self.lens.w_u, l0, cta_pair=self.cta_pair)
            if self.time_lens:
                e1.record()
                self.stats.lens_events.append((e0, e1, 2.0 * (r1 - r0) * x.shape[1] * V))
            for i, b in enumerate(bs):
                if self.time_lens:
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                o = ops.lens_rank_update(A[r0:r1], ids[r0 // T:r1 // T], D, l0, rstd[i, r0:r1], b, V=V, softcap=self.softcap,
                                         secret_id=None if secret_id is None else secret_id[r0:r1], logits_out=stored)
                if ll is not None:
                    lls[b].append(ops.ll_topk_from_tiles(stored, o.lse, None if ll.cur_id is None else ll.cur_id[r0:r1],
                                                         None if ll.prev_id is None else ll.prev_id[r0:r1], V, ll.k))
                if self.time_lens:
                    e1.record()      # FLOPs actually done: depth = the budget rounded up to 64 (x 2 with remainders)
                    depth = -(-b // 64) * 64 * (2 if split else 1)
                    self.stats.ll_events.append((e0, e1, 2.0 * (r1 - r0) * depth * V))
                outs[b].append(o)
            self.stats.launches += 2 + 2 * len(bs)
        cat = lambda xs, f: torch.cat([getattr(o, f) for o in xs])      # noqa: E731
        lens_out = {b: ops.LensReadout(lse=cat(v, "lse"), secret_prob=cat(v, "secret_prob"), top_prob=cat(v, "top_prob"),
                                       top_id=cat(v, "top_id")) for b, v in outs.items()}
        if ll is None:
            return lens_out
        return {b: Readout(lens_out[b], torch.cat([s_ for s_, _ in lls[b]]), torch.cat([i_ for _, i_ in lls[b]])) for b in bs}

    def random_budget_sweep(self, x: torch.Tensor, secret_id, budgets: Sequence[int] = DEFAULT_BUDGETS,
                            seeds: Iterable[int] = range(32), mode: str = "error",
                            acts: Optional[ops.SparseActs] = None, ll: Optional[LLSpec] = None
                            ) -> Dict[tuple, ops.LensReadout]:
        """BASELINE.json configs[2]: random-latent control, (budget, seed) conditions. One rank table per
        seed (shared by all prompts); budgets are nested within a seed. The smallest budget of every seed
        after the first starts from the previous seed's smallest-budget lse (one random latent ablated
        either way), so only the very first condition of the sweep pays the second LL-Top-k sweep."""
        acts = acts if acts is not None else self.encode(x)
        out, first_state = {}, None
        for seed in seeds:
            rank = random_rank_table(self.sae.S, seed, min(max(budgets), TB_TOPM_MAX), x.device)
            res, _ = self._nested_budgets(x, acts, rank, 0, budgets, secret_id, mode, ll, state=first_state)
            if first_state is None and ll is not None and self.ll_fused:
                r0 = res[min(int(v) for v in budgets)]
                if isinstance(r0, Readout) and r0.ll_sum is not None and ll.rows_per_prompt == ops.LL_SLOTS:
                    first_state = LLState(r0.lens.lse, r0.ll_sum[:, ll.k - 1].contiguous())
            for b, r in res.items():
                out[(b, int(seed))] = r
        return out

    def projection_sweep(self, x: torch.Tensor, U: torch.Tensor, secret_id,
                         ranks: Sequence[int] = DEFAULT_RANKS, ll: Optional[LLSpec] = None
                         ) -> Dict[int, ops.LensReadout]:
        """BASELINE.json configs[3]: remove the first r columns of U for each r. The ranks are nested
        subspaces, so with LL-Top-k every rank after the first reads in one sweep with the previous rank's
        exact lse as reference (like nested ablation budgets; results are exact either way)."""
        fusable = (ll is not None and self.ll_fused and ll.seg is None and ll.rows_per_prompt == ops.LL_SLOTS)
        out, state = {}, None
        rs = sorted(set(int(v) for v in ranks))
        # the tensor-core ranks share ONE coefficient pass (the coefficients of the largest rank: rank r uses its first r)
        tc = [r for r in rs if r >= ops.LOWRANK_MULTI_MIN]
        h_multi = {}
        d = x.shape[1]
        if (self.multi_budget and len(tc) >= 2 and d % 256 == 0 and d <= 4096
                and len(tc) * x.shape[0] * d * 2 <= 0.45 * _free_bytes(x.device)):
            if self.time_lens:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
            _, hm = ops.lowrank_remove_multi(x, U, tc, w_norm=self.lens.w_norm, eps=self.lens.eps)
            if self.time_lens:
                e1.record()
                self.stats.row_events.append((e0, e1, 4.0 * x.shape[0] * d * len(tc)))
            self.stats.launches += 1
            h_multi = {r: hm[i] for i, r in enumerate(tc)}
        for r in rs:
            if r in h_multi:
                if fusable and state is not None:
                    out[r], state = self.read_fused(h_multi[r], secret_id, ll, state)
                    continue
                out[r] = self.read(h_multi[r], secret_id, True, ll)
            elif fusable and state is not None:
                out[r], state = self.project_and_read(x, U[:, :r].contiguous(), secret_id=secret_id, ll=ll, ll_state=state)
                continue
            else:
                out[r] = self.project_and_read(x, U[:, :r].contiguous(), secret_id=secret_id, ll=ll)
            if fusable:
                state = LLState(out[r].lens.lse, out[r].ll_sum[:, ll.k - 1].contiguous())
        return out
