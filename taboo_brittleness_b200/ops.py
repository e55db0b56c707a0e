"""Tensor-level wrappers over the C ABI (include/tb_b200.h).

Every function takes CUDA torch tensors, allocates its outputs (and scratch) with torch, and
launches on torch's current stream. Nothing here computes on the CPU, and nothing falls back:
a missing library or a failed call raises ``TabooKernelError``.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import TB_TOPK, TB_TOPM_MAX, TabooKernelError, check

GEMMA2_EPS = 1e-6          # Gemma2Config.rms_norm_eps
GEMMA2_SOFTCAP = 30.0      # Gemma2Config.final_logit_softcapping
BLOCK_M = 128
BLOCK_N = 256


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _need(t: torch.Tensor, dtype: torch.dtype, name: str, ndim: Optional[int] = None) -> torch.Tensor:
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise TabooKernelError(f"{name} must be a CUDA tensor (no CPU path exists)", _lib.TB_ERR_INVALID)
    if t.dtype != dtype:
        raise TabooKernelError(f"{name} must be {dtype}, got {t.dtype}", _lib.TB_ERR_INVALID)
    if ndim is not None and t.dim() != ndim:
        raise TabooKernelError(f"{name} must be {ndim}-D, got shape {tuple(t.shape)}", _lib.TB_ERR_INVALID)
    if not t.is_contiguous():
        raise TabooKernelError(f"{name} must be contiguous", _lib.TB_ERR_INVALID)
    # kernels launch on torch's CURRENT device and stream: a tensor of another GPU would be read through a
    # peer mapping (or fault), and the per-device scratch caches would be mixed up
    if t.device.index != torch.cuda.current_device():
        raise TabooKernelError(f"{name} lives on cuda:{t.device.index} but the current device is "
                               f"cuda:{torch.cuda.current_device()} (use torch.cuda.device(...) or set_device)",
                               _lib.TB_ERR_INVALID)
    return t


def device_check() -> None:
    """Raise unless the current device can run the kernels (sm_100)."""
    check(_lib.load().tb_device_check(), "tb_device_check")


# ------------------------------------------------------------------------------------------
# weights
# ------------------------------------------------------------------------------------------
@dataclass
class SaeWeights:
    """Gemma Scope JumpReLU SAE parameters in the layout the kernels read.

    sae_lens stores W_enc [d, S], W_dec [S, d] (fp32); the encode GEMM wants the K-major
    transpose W_enc^T [S, d]. bf16 for the two matrices, fp32 for the vectors.
    """
    w_enc_t: torch.Tensor   # bf16 [S, d]
    b_enc: torch.Tensor     # f32 [S]
    theta: torch.Tensor     # f32 [S]
    w_dec: torch.Tensor     # bf16 [S, d]
    b_dec: torch.Tensor     # f32 [d]

    @property
    def d(self) -> int:
        return self.w_dec.shape[1]

    @property
    def S(self) -> int:
        return self.w_dec.shape[0]

    @staticmethod
    def from_sae_lens_layout(W_enc, b_enc, threshold, W_dec, b_dec, device="cuda") -> "SaeWeights":
        """W_enc [d, S], W_dec [S, d] as sae_lens / Gemma Scope params.npz hold them."""
        dev = torch.device(device)
        return SaeWeights(
            w_enc_t=W_enc.to(dev).t().contiguous().to(torch.bfloat16),
            b_enc=b_enc.to(dev, torch.float32).contiguous(),
            theta=threshold.to(dev, torch.float32).contiguous(),
            w_dec=W_dec.to(dev).contiguous().to(torch.bfloat16),
            b_dec=b_dec.to(dev, torch.float32).contiguous(),
        )


@dataclass
class LensWeights:
    """Final RMSNorm weight (f32 [d]) and unembedding shard W_U (bf16 [V_shard, d])."""
    w_norm: torch.Tensor
    w_u: torch.Tensor
    vocab_offset: int = 0
    eps: float = GEMMA2_EPS


@dataclass
class SparseActs:
    """Row-wise lists of the non-zero SAE latents (what replaces the dense [rows, S] acts)."""
    count: torch.Tensor                 # i32 [rows]
    idx: torch.Tensor                   # i32 flat
    val: torch.Tensor                   # f32 flat
    S: int
    cap: int = 0                        # fixed stride when row_start is None
    row_start: Optional[torch.Tensor] = None  # i64 [rows] (CSR)

    @property
    def n_rows(self) -> int:
        return self.count.shape[0]

    def to_dense(self) -> torch.Tensor:
        """Dense [rows, S] f32 view for tests (small inputs only)."""
        n = self.n_rows
        dense = torch.zeros(n, self.S, dtype=torch.float32, device=self.count.device)
        cnt = self.count.to(torch.int64)
        if self.row_start is None:
            cnt = cnt.clamp(max=self.cap)
            start = torch.arange(n, device=cnt.device, dtype=torch.int64) * self.cap
        else:
            start = self.row_start
        rows = torch.repeat_interleave(torch.arange(n, device=cnt.device), cnt)
        offs = torch.arange(int(cnt.sum().item()), device=cnt.device) - torch.repeat_interleave(
            torch.cumsum(cnt, 0) - cnt, cnt)
        pos = torch.repeat_interleave(start, cnt) + offs
        dense[rows, self.idx[pos].to(torch.int64)] = self.val[pos]
        return dense


# ------------------------------------------------------------------------------------------
# validation GEMM
# ------------------------------------------------------------------------------------------
def gemm_nt(a: torch.Tensor, b: torch.Tensor, n_chunks: int = 1, cta_pair: bool = False) -> torch.Tensor:
    """a [M, K] @ b [N, K]^T -> f32 [M, N] through the TMA/tcgen05 main loop (tests only)."""
    _need(a, torch.bfloat16, "a", 2)
    _need(b, torch.bfloat16, "b", 2)
    M, K = a.shape
    N = b.shape[0]
    out = torch.empty(M, N, dtype=torch.float32, device=a.device)
    check(_lib.load().tb_gemm_bf16_nt(_ptr(a), _ptr(b), M, N, K, n_chunks, int(cta_pair), _ptr(out),
                                      _stream()), "tb_gemm_bf16_nt")
    return out


# ------------------------------------------------------------------------------------------
# K1 SAE encode
# ------------------------------------------------------------------------------------------
def sae_encode_raw(x: torch.Tensor, sae: SaeWeights, *, cap: int = 0,
                   row_start: Optional[torch.Tensor] = None, row_cap: Optional[torch.Tensor] = None,
                   idx: Optional[torch.Tensor] = None, val: Optional[torch.Tensor] = None,
                   rank: Optional[torch.Tensor] = None, rank_rows: int = 0, budget: int = 0
                   ) -> torch.Tensor:
    """One launch of tb_sae_encode; returns count i32 [rows]. idx/val None = count only."""
    _need(x, torch.bfloat16, "x", 2)
    n, d = x.shape
    if d != sae.d:
        raise TabooKernelError(f"x has d={d}, SAE has d={sae.d}", _lib.TB_ERR_INVALID)
    count = torch.empty(n, dtype=torch.int32, device=x.device)
    if n == 0:
        return count
    if rank is not None:
        _need(rank, torch.uint16, "rank", 2)
    check(_lib.load().tb_sae_encode(
        _ptr(x), _ptr(sae.w_enc_t), _ptr(sae.b_enc), _ptr(sae.theta), n, d, sae.S,
        _ptr(row_start), _ptr(row_cap), cap, _ptr(idx), _ptr(val), _ptr(count),
        _ptr(rank), rank_rows, budget, _stream()), "tb_sae_encode")
    return count


def sae_encode(x: torch.Tensor, sae: SaeWeights, cap: int = 128, *, rank: Optional[torch.Tensor] = None,
               rank_rows: int = 0, budget: int = 0) -> SparseActs:
    """JumpReLU encode to sparse row lists (replaces ``sae.encode``, 02_run_sae_baseline.py:67).

    First pass writes fixed-stride lists of ``cap`` slots per row. Rows with more active latents than
    that (the <bos>-scale row of every prompt has thousands) are encoded once more, alone, into a region
    appended to the lists, and ``row_start`` then points each row at its list -- no active latent is ever
    dropped, and the second pass costs the overflowing rows only (about 1 in 128), not a second sweep.
    """
    n = x.shape[0]
    dev = x.device
    idx = torch.empty(max(n * cap, 1), dtype=torch.int32, device=dev)
    val = torch.empty(max(n * cap, 1), dtype=torch.float32, device=dev)
    n_sub = ENCODE_SUB_LISTS
    if (rank is None and n_sub > 1 and n >= ENCODE_CHUNKED_MIN_ROWS and cap % n_sub == 0
            and (-(-sae.S // BLOCK_N)) % n_sub == 0):
        # large batches: the latent sweep in n_sub chunks, so that the row blocks in flight fit L2
        _need(x, torch.bfloat16, "x", 2)
        sub = torch.empty(n_sub * n, dtype=torch.int32, device=dev)
        count = torch.empty(n, dtype=torch.int32, device=dev)
        trunc = torch.empty(n, dtype=torch.int32, device=dev)
        check(_lib.load().tb_sae_encode_chunked(_ptr(x), _ptr(sae.w_enc_t), _ptr(sae.b_enc), _ptr(sae.theta), n, x.shape[1],
                                                sae.S, cap, n_sub, _ptr(idx), _ptr(val), _ptr(sub), _ptr(count), _ptr(trunc),
                                                _stream()), "tb_sae_encode_chunked")
        over = torch.nonzero((trunc != 0) | (count > cap)).flatten()
        if over.numel() == 0:
            return SparseActs(count=count, idx=idx, val=val, S=sae.S, cap=cap)
        return _reencode_overflowing(x, sae, cap, idx, val, count, over)
    count = sae_encode_raw(x, sae, cap=cap, idx=idx, val=val, rank=rank, rank_rows=rank_rows, budget=budget)
    if n == 0 or int(count.max().item()) <= cap:
        return SparseActs(count=count, idx=idx, val=val, S=sae.S, cap=cap)
    if rank is not None:
        # the epilogue's rank mask is looked up by row number: keep all rows in place and redo them all
        c64 = count.to(torch.int64)
        row_start = torch.cumsum(c64, 0) - c64
        nnz = int(c64.sum().item())
        idx = torch.empty(nnz, dtype=torch.int32, device=dev)
        val = torch.empty(nnz, dtype=torch.float32, device=dev)
        count2 = sae_encode_raw(x, sae, row_start=row_start, row_cap=count, idx=idx, val=val,
                                rank=rank, rank_rows=rank_rows, budget=budget)
        return SparseActs(count=count2, idx=idx, val=val, S=sae.S, cap=0, row_start=row_start)
    return _reencode_overflowing(x, sae, cap, idx, val, count, torch.nonzero(count > cap).flatten())


ENCODE_SUB_LISTS = 2               # latent chunks of the large-batch encode (1 = one sweep per row block)
ENCODE_CHUNKED_MIN_ROWS = 128 * 148


def _reencode_overflowing(x, sae, cap, idx, val, count, over) -> SparseActs:
    """Rows whose list did not fit are encoded once more, alone, into a region appended to the lists."""
    n = x.shape[0]
    dev = x.device
    c_over = count[over].to(torch.int64)
    extra = int(c_over.sum().item())
    idx = torch.cat([idx[:n * cap], torch.empty(extra, dtype=torch.int32, device=dev)])
    val = torch.cat([val[:n * cap], torch.empty(extra, dtype=torch.float32, device=dev)])
    row_start = torch.arange(n, device=dev, dtype=torch.int64) * cap
    start_over = n * cap + torch.cumsum(c_over, 0) - c_over
    row_start[over] = start_over
    count_over = sae_encode_raw(x[over].contiguous(), sae, row_start=start_over, row_cap=count[over].contiguous(),
                                idx=idx, val=val)
    count = count.clone()
    count[over] = count_over
    return SparseActs(count=count, idx=idx, val=val, S=sae.S, cap=0, row_start=row_start)


# ------------------------------------------------------------------------------------------
# a2 targeting
# ------------------------------------------------------------------------------------------
def latent_mean(acts: SparseActs, rows_per_prompt: int = 0, seg: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Mean activation per prompt and latent, f32 [P, S] (02_run_sae_baseline.py:70).

    ``seg`` (i64 [P+1]) gives each prompt's row window (the response slice, :64); otherwise
    prompts are consecutive blocks of ``rows_per_prompt`` rows.
    """
    if seg is not None:
        _need(seg, torch.int64, "seg", 1)
        n_prompts = seg.shape[0] - 1
    else:
        if rows_per_prompt <= 0 or acts.n_rows % rows_per_prompt:
            raise TabooKernelError("rows_per_prompt must divide the row count", _lib.TB_ERR_INVALID)
        n_prompts = acts.n_rows // rows_per_prompt
    out = torch.empty(n_prompts, acts.S, dtype=torch.float32, device=acts.count.device)
    check(_lib.load().tb_sae_latent_mean(
        _ptr(acts.row_start), acts.cap, _ptr(acts.count), _ptr(acts.idx), _ptr(acts.val), _ptr(seg),
        rows_per_prompt, n_prompts, acts.S, _ptr(out), _stream()), "tb_sae_latent_mean")
    return out


def latent_wsum(acts: SparseActs, rows_per_prompt: int = 0, seg: Optional[torch.Tensor] = None, *,
                row_weight: Optional[torch.Tensor] = None, power: int = 1) -> torch.Tensor:
    """f32 [P, S]: sum over a prompt's rows of w[t] * a[t, s]**power (w None = 1)."""
    if seg is not None:
        _need(seg, torch.int64, "seg", 1)
        n_prompts = seg.shape[0] - 1
    else:
        if rows_per_prompt <= 0 or acts.n_rows % rows_per_prompt:
            raise TabooKernelError("rows_per_prompt must divide the row count", _lib.TB_ERR_INVALID)
        n_prompts = acts.n_rows // rows_per_prompt
    if row_weight is not None:
        _need(row_weight, torch.float32, "row_weight", 1)
    out = torch.empty(n_prompts, acts.S, dtype=torch.float32, device=acts.count.device)
    check(_lib.load().tb_sae_latent_wsum(
        _ptr(acts.row_start), acts.cap, _ptr(acts.count), _ptr(acts.idx), _ptr(acts.val), _ptr(seg), rows_per_prompt,
        n_prompts, acts.S, _ptr(row_weight), power, _ptr(out), _stream()), "tb_sae_latent_wsum")
    return out


def topm_select(mean: torch.Tensor, m: int, want_rank: bool = True) -> Tuple[torch.Tensor, torch.Tensor, Optional[torch.Tensor]]:
    """Top-m latents per prompt, sorted descending: (idx i32 [P,m], val f32 [P,m], rank u16 [P,S] or None)."""
    _need(mean, torch.float32, "mean", 2)
    P, S = mean.shape
    idx = torch.empty(P, m, dtype=torch.int32, device=mean.device)
    val = torch.empty(P, m, dtype=torch.float32, device=mean.device)
    rank = torch.empty(P, S, dtype=torch.uint16, device=mean.device) if want_rank else None
    check(_lib.load().tb_topm_select(_ptr(mean), P, S, m, _ptr(idx), _ptr(val), _ptr(rank), _stream()),
          "tb_topm_select")
    return idx, val, rank


def rank_table_from_indices(indices: torch.Tensor, S: int) -> torch.Tensor:
    """u16 [n_tables, S] rank table from index lists [n_tables, m] (e.g. seeded random control:
    ``torch.randperm(S, generator=cpu_gen(seed))[:m]``)."""
    if indices.dim() == 1:
        indices = indices[None]
    n, m = indices.shape
    if m > TB_TOPM_MAX:
        raise TabooKernelError(f"at most {TB_TOPM_MAX} latents per table", _lib.TB_ERR_INVALID)
    table = torch.full((n, S), _lib.TB_RANK_NONE, dtype=torch.int32)
    src = torch.arange(m, dtype=torch.int32).expand(n, m)
    table.scatter_(1, indices.to(torch.int64).cpu(), src)
    return table.to(torch.uint16)


# ------------------------------------------------------------------------------------------
# K1b / K2 / a6
# ------------------------------------------------------------------------------------------
def sae_ablate_decode(x: Optional[torch.Tensor], acts: SparseActs, sae: SaeWeights, rank: Optional[torch.Tensor],
                      budget: int, *, mode: str = "error", rank_rows: int = 0,
                      row_table: Optional[torch.Tensor] = None, w_norm: Optional[torch.Tensor] = None,
                      eps: float = GEMMA2_EPS, want_x: bool = True, want_h: bool = False
                      ) -> Tuple[Optional[torch.Tensor], Optional[torch.Tensor]]:
    """Zero the latents with rank < budget and decode back to the residual (Plan:126).

    mode "error": x' = x - sum_{ablated} a_i W_dec[i]  (keeps the SAE reconstruction error);
    mode "plain": x' = sum_{kept} a_i W_dec[i] + b_dec.  Returns (x' bf16, h bf16) as requested.
    """
    n = acts.n_rows
    d = sae.d
    dev = acts.count.device
    imode = {"error": 0, "plain": 1}[mode]
    if imode == 0:
        _need(x, torch.bfloat16, "x", 2)
    if rank is not None:
        _need(rank, torch.uint16, "rank", 2)
    if row_table is not None:
        _need(row_table, torch.int32, "row_table", 1)
    out_x = torch.empty(n, d, dtype=torch.bfloat16, device=dev) if want_x else None
    out_h = torch.empty(n, d, dtype=torch.bfloat16, device=dev) if want_h else None
    if n == 0:
        return out_x, out_h
    check(_lib.load().tb_sae_ablate_decode(
        _ptr(x), _ptr(acts.row_start), acts.cap, _ptr(acts.count), _ptr(acts.idx), _ptr(acts.val),
        _ptr(sae.w_dec), _ptr(sae.b_dec), _ptr(rank), _ptr(row_table), rank_rows, budget, imode,
        n, d, sae.S, _ptr(w_norm), eps, _ptr(out_x), _ptr(out_h), _stream()), "tb_sae_ablate_decode")
    return out_x, out_h


@dataclass
class RankedActs:
    """Per row, the active latents that appear in one ranking, ordered by rank (``tb_sae_rank_lists``): what the
    nested budgets of a sweep share. Same list layout as the ``SparseActs`` it was made from."""
    count: torch.Tensor                 # i32 [rows]
    key: torch.Tensor                   # i32 flat: rank << 20 | latent
    val: torch.Tensor                   # f32 flat
    cap: int = 0
    row_start: Optional[torch.Tensor] = None

    @property
    def n_rows(self) -> int:
        return self.count.shape[0]


def rank_lists(acts: SparseActs, rank: torch.Tensor, rank_rows: int = 0) -> RankedActs:
    """Order every row's ranked active latents by rank, once per ranking (per prompt: ``rank_rows`` rows share a
    table; 0: one table for all rows)."""
    _need(rank, torch.uint16, "rank", 2)
    dev = acts.count.device
    key = torch.empty_like(acts.idx)
    val = torch.empty_like(acts.val)
    cnt = torch.empty(acts.n_rows, dtype=torch.int32, device=dev)
    if acts.n_rows:
        check(_lib.load().tb_sae_rank_lists(_ptr(acts.row_start), acts.cap, _ptr(acts.count), _ptr(acts.idx), _ptr(acts.val),
                                            _ptr(rank), rank_rows, acts.n_rows, acts.S, _ptr(key), _ptr(val), _ptr(cnt),
                                            _stream()), "tb_sae_rank_lists")
    return RankedActs(count=cnt, key=key, val=val, cap=acts.cap, row_start=acts.row_start)


def sae_ablate_ranked(x: torch.Tensor, ranked: RankedActs, sae: SaeWeights, budget: int, *,
                      w_norm: Optional[torch.Tensor] = None, eps: float = GEMMA2_EPS, want_x: bool = True,
                      want_h: bool = False, out_rstd: Optional[torch.Tensor] = None
                      ) -> Tuple[Optional[torch.Tensor], Optional[torch.Tensor]]:
    """x' = x - sum over the row's active latents with rank < budget of a_i W_dec[i] (the error-term form of
    ``sae_ablate_decode``), from rank-ordered lists: a budget reads only its prefix. ``out_rstd`` (f32 [rows]) receives
    the final RMSNorm's 1 / sqrt(mean(x'^2) + eps) of the rounded x'."""
    _need(x, torch.bfloat16, "x", 2)
    n, d = x.shape
    out_x = torch.empty(n, d, dtype=torch.bfloat16, device=x.device) if want_x else None
    out_h = torch.empty(n, d, dtype=torch.bfloat16, device=x.device) if want_h else None
    if out_rstd is not None:
        _need(out_rstd, torch.float32, "out_rstd", 1)
        if out_rstd.shape[0] != n:
            raise TabooKernelError("out_rstd must hold one value per row", _lib.TB_ERR_INVALID)
    if n == 0:
        return out_x, out_h
    check(_lib.load().tb_sae_ablate_ranked(_ptr(x), _ptr(ranked.row_start), ranked.cap, _ptr(ranked.count), _ptr(ranked.key),
                                           _ptr(ranked.val), _ptr(sae.w_dec), int(budget), n, d, _ptr(w_norm), eps,
                                           _ptr(out_x), _ptr(out_h), _ptr(out_rstd), _stream()), "tb_sae_ablate_ranked")
    return out_x, out_h


def sae_ablate_ranked_multi(x: torch.Tensor, ranked: RankedActs, sae: SaeWeights, budgets: Sequence[int], *,
                            w_norm: Optional[torch.Tensor] = None, eps: float = GEMMA2_EPS, want_x: bool = False,
                            want_h: bool = True, want_rstd: bool = False):
    """All nested budgets in one pass over the rows (``tb_sae_ablate_ranked_multi``): returns (x' bf16 [B, rows, d] or None,
    h bf16 [B, rows, d] or None, rstd f32 [B, rows] or None) for the ascending ``budgets``; the same bits as one
    ``sae_ablate_ranked`` call per budget."""
    _need(x, torch.bfloat16, "x", 2)
    n, d = x.shape
    bs = [int(b) for b in budgets]
    if bs != sorted(set(bs)) or not bs:
        raise TabooKernelError("budgets must be strictly ascending", _lib.TB_ERR_INVALID)
    B = len(bs)
    out_x = torch.empty(B, n, d, dtype=torch.bfloat16, device=x.device) if want_x else None
    out_h = torch.empty(B, n, d, dtype=torch.bfloat16, device=x.device) if want_h else None
    rstd = torch.empty(B, n, dtype=torch.float32, device=x.device) if want_rstd else None
    if n == 0:
        return out_x, out_h, rstd
    import ctypes
    arr = (ctypes.c_int32 * B)(*bs)
    check(_lib.load().tb_sae_ablate_ranked_multi(_ptr(x), _ptr(ranked.row_start), ranked.cap, _ptr(ranked.count), _ptr(ranked.key),
                                                 _ptr(ranked.val), _ptr(sae.w_dec), ctypes.cast(arr, ctypes.c_void_p), B, n, d,
                                                 _ptr(w_norm), eps, _ptr(out_x), _ptr(out_h), _ptr(rstd), n * d, _stream()),
          "tb_sae_ablate_ranked_multi")
    return out_x, out_h, rstd


LOWRANK_SMALL = 8     # ranks up to this can use the kernel that keeps U^T in shared memory
LOWRANK_TC_MIN = 8    # ... and from this rank on the tensor-core path is faster (measured: 1.88 vs 2.22 ms at r = 8)
LOWRANK_MULTI_MIN = 4  # in a sweep of nested ranks the coefficient pass is shared: the removal + norm pass alone (1.1 ms per
                       # 262,144 rows) beats the CUDA-core kernel from rank 4 on (1.65 ms)


_PROJ_SCRATCH: dict = {}


def lowrank_remove(x: torch.Tensor, U: torch.Tensor, *, w_norm: Optional[torch.Tensor] = None,
                   eps: float = GEMMA2_EPS, want_x: bool = True, want_h: bool = False, one_pass: bool = False,
                   dbg_c: Optional[torch.Tensor] = None
                   ) -> Tuple[Optional[torch.Tensor], Optional[torch.Tensor]]:
    """x' = x - (x U) U^T for orthonormal U [d, r], r <= 64 (Plan:148). Small ranks run in one
    HBM-bound CUDA-core kernel (up to 7 by default); from 8 on both thin products go to the tensor cores, in three
    passes over the rows (``tb_lowrank_remove_tc``). ``one_pass=True`` (d a multiple of 256, at most 4096) takes
    ``tb_lowrank_remove_fused`` instead: x read once and h written once by a cluster of d / 256 CTAs that keeps
    each 128-row block in shared memory -- same results, 14 KB instead of 35 KB of HBM traffic per row, but
    measured SLOWER than the three passes (the per-block chain of cluster barriers and
    distributed-shared-memory exchanges is six times the block's HBM time), so it is not the default."""
    _need(x, torch.bfloat16, "x", 2)
    _need(U, torch.bfloat16, "U", 2)
    n, d = x.shape
    if U.shape[0] != d:
        raise TabooKernelError(f"U must be [d={d}, r], got {tuple(U.shape)}", _lib.TB_ERR_INVALID)
    r = U.shape[1]
    ut = U.t().contiguous()
    out_x = torch.empty(n, d, dtype=torch.bfloat16, device=x.device) if want_x else None
    out_h = torch.empty(n, d, dtype=torch.bfloat16, device=x.device) if want_h else None
    if n == 0:
        return out_x, out_h
    lib = _lib.load()
    if r >= LOWRANK_TC_MIN and one_pass and lib.tb_lowrank_fused_supported(d, r):
        # one pass over the rows: a cluster of d / 256 CTAs keeps each 128-row block in shared memory
        check(lib.tb_lowrank_remove_fused(_ptr(x), _ptr(ut), r, n, d, _ptr(w_norm), eps, _ptr(out_x), _ptr(out_h),
                                          _ptr(dbg_c), _stream()), "tb_lowrank_remove_fused")
        return out_x, out_h
    if r >= LOWRANK_TC_MIN:
        need = int(lib.tb_lowrank_tc_scratch_bytes(n, d, r))
        if need < 0:
            raise TabooKernelError(f"tb_lowrank_tc_scratch_bytes({n}, {d}, {r}) failed", _lib.TB_ERR_INVALID)
        key = (x.device.index, torch.cuda.current_stream().cuda_stream)
        buf = _PROJ_SCRATCH.get(key)
        if buf is None or buf.numel() < need:      # grown on demand, reused by later calls on this stream
            buf = _PROJ_SCRATCH[key] = torch.empty(need, dtype=torch.uint8, device=x.device)
        check(lib.tb_lowrank_remove_tc(_ptr(x), _ptr(ut), r, n, d, _ptr(w_norm), eps, _ptr(buf), buf.numel(),
                                       _ptr(out_x), _ptr(out_h), _stream()), "tb_lowrank_remove_tc")
        return out_x, out_h
    check(lib.tb_lowrank_remove(_ptr(x), _ptr(ut), r, n, d, _ptr(w_norm), eps, _ptr(out_x),
                                _ptr(out_h), _stream()), "tb_lowrank_remove")
    return out_x, out_h


def lowrank_remove_multi(x: torch.Tensor, U: torch.Tensor, ranks: Sequence[int], *, w_norm: Optional[torch.Tensor] = None,
                         eps: float = GEMMA2_EPS, want_x: bool = False, want_h: bool = True
                         ) -> Tuple[Optional[torch.Tensor], Optional[torch.Tensor]]:
    """x - (x U_r) U_r^T for nested ranks r (the first r columns of U [d, >= max r]) in one call: one coefficient pass
    for the largest rank, then one removal + norm pass per rank (``tb_lowrank_remove_tc_multi``; tensor-core path, the
    same bits as ``lowrank_remove`` per rank for ranks >= LOWRANK_TC_MIN). Returns ([R, rows, d] or None) x 2."""
    _need(x, torch.bfloat16, "x", 2)
    _need(U, torch.bfloat16, "U", 2)
    n, d = x.shape
    rs = [int(r) for r in ranks]
    if not rs or rs != sorted(set(rs)) or rs[-1] > U.shape[1] or U.shape[0] != d:
        raise TabooKernelError("ranks must be strictly ascending and covered by U [d, r_max]", _lib.TB_ERR_INVALID)
    R = len(rs)
    ut = U[:, :rs[-1]].t().contiguous()
    out_x = torch.empty(R, n, d, dtype=torch.bfloat16, device=x.device) if want_x else None
    out_h = torch.empty(R, n, d, dtype=torch.bfloat16, device=x.device) if want_h else None
    if n == 0:
        return out_x, out_h
    lib = _lib.load()
    need = int(lib.tb_lowrank_tc_scratch_bytes(n, d, rs[-1]))
    if need < 0:
        raise TabooKernelError(f"tb_lowrank_tc_scratch_bytes({n}, {d}, {rs[-1]}) failed", _lib.TB_ERR_INVALID)
    key = (x.device.index, torch.cuda.current_stream().cuda_stream)
    buf = _PROJ_SCRATCH.get(key)
    if buf is None or buf.numel() < need:
        buf = _PROJ_SCRATCH[key] = torch.empty(need, dtype=torch.uint8, device=x.device)
    import ctypes
    arr = (ctypes.c_int32 * R)(*rs)
    check(lib.tb_lowrank_remove_tc_multi(_ptr(x), _ptr(ut), ctypes.cast(arr, ctypes.c_void_p), R, n, d, _ptr(w_norm), eps, _ptr(buf),
                                         buf.numel(), _ptr(out_x), _ptr(out_h), n * d, _stream()), "tb_lowrank_remove_tc_multi")
    return out_x, out_h


def rmsnorm(x: torch.Tensor, w_norm: torch.Tensor, eps: float = GEMMA2_EPS) -> torch.Tensor:
    """Gemma2RMSNorm on a bf16 residual -> bf16 (models.py:135, ``model.model.norm``)."""
    _need(x, torch.bfloat16, "x", 2)
    _need(w_norm, torch.float32, "w_norm", 1)
    out = torch.empty_like(x)
    if x.shape[0] == 0:
        return out
    check(_lib.load().tb_rmsnorm(_ptr(x), _ptr(w_norm), eps, x.shape[0], x.shape[1], _ptr(out), _stream()),
          "tb_rmsnorm")
    return out


# ------------------------------------------------------------------------------------------
# K3 / K4 logit lens
# ------------------------------------------------------------------------------------------
@dataclass
class LensPartials:
    """(max, sumexp, secret logit, top-10) per (part, row); parts = vocab chunks or shards."""
    pm: torch.Tensor    # f32 [parts, rows]
    ps: torch.Tensor
    psec: torch.Tensor
    ptv: torch.Tensor   # f32 [parts, rows, 10]
    pti: torch.Tensor   # i32 [parts, rows, 10]

    @property
    def n_parts(self) -> int:
        return self.pm.shape[0]

    @property
    def n_rows(self) -> int:
        return self.pm.shape[1]

    def pack(self) -> torch.Tensor:
        """[parts, rows, 23] f32 (ids bit-cast) -- one tensor for a single all_gather."""
        return torch.cat([self.pm[..., None], self.ps[..., None], self.psec[..., None], self.ptv,
                          self.pti.view(torch.float32)], dim=-1).contiguous()

    @staticmethod
    def unpack(t: torch.Tensor) -> "LensPartials":
        t = t.reshape(-1, t.shape[-2], 23)
        return LensPartials(pm=t[..., 0].contiguous(), ps=t[..., 1].contiguous(), psec=t[..., 2].contiguous(),
                            ptv=t[..., 3:13].contiguous(), pti=t[..., 13:23].contiguous().view(torch.int32))


@dataclass
class LensReadout:
    lse: torch.Tensor          # f32 [rows]  log-sum-exp of the (soft-capped) logits
    secret_prob: torch.Tensor  # f32 [rows]
    top_prob: torch.Tensor     # f32 [rows, 10], descending
    top_id: torch.Tensor       # i32 [rows, 10]


# cluster modes of the lens kernels (the ``cta_pair`` argument of the C ABI):
#   0 single CTAs, 1 CTA pairs, 2 = 2x1 pairs (two row blocks share the W_U tiles through TMA
#   multicast), 3 = 1x2 pairs (two chunks share the row block), 4 = 2x2 pairs (both)
#   5 = 2x1 pairs as the PREFERRED cluster shape with plain pairs as the fallback: groups of four CTAs that
#   do not fit a GPC as one cluster run as two pairs, so all 148 SMs work (37 groups instead of 33 clusters)
CLUSTER_MODES = {0: (1, 1, 1), 1: (2, 1, 1), 2: (2, 2, 1), 3: (2, 1, 2), 4: (2, 2, 2), 5: (2, 2, 1)}   # (group, R, C)
# clusters that can be co-resident on a 148-SM B200 (cudaOccupancyMaxActiveClusters: a cluster must
# fit one GPC, so sizes 4 and 8 leave SMs idle); only used to rank schedules here
_B200_CLUSTERS = {1: 148, 2: 74, 4: 33, 6: 22, 8: 15}


def _resident_groups(cta_pair, n_sms: int = 148) -> int:
    """Co-resident clusters (mode 5: groups of four CTAs, whether they run as one cluster or two pairs)."""
    m = _mode(cta_pair)
    group, R, C = CLUSTER_MODES[m]
    size = group * R * C
    if m == 5:
        return max(n_sms // size, 1)
    return max(_B200_CLUSTERS.get(size, n_sms // size) * n_sms // 148, 1)


def _mode(cta_pair) -> int:
    m = int(cta_pair)
    if m not in CLUSTER_MODES:
        raise TabooKernelError(f"cluster mode {cta_pair!r} not in 0..5", _lib.TB_ERR_INVALID)
    return m


def pick_chunks(n_rows: int, V: int, n_sms: int = 148, cta_pair=False, *, sweep_a: bool = False) -> int:
    """How many vocab chunks to sweep: fill all clusters with little tail waste while keeping the
    stationary blocks of the resident clusters (1.835 MB each at d=3584) a small part of L2."""
    group, R, C = CLUSTER_MODES[_mode(cta_pair)]
    clusters = _resident_groups(cta_pair, n_sms)
    if sweep_a:
        blocks = -(-n_rows // BLOCK_N)
        n_tiles = -(-V // (BLOCK_M * group))
        max_tpc = LL_MAX_CHUNK_IDS // (BLOCK_M * group)
    else:
        blocks = -(-n_rows // (BLOCK_M * group))
        n_tiles = -(-V // BLOCK_N)
        max_tpc = n_tiles
    best, best_cost = None, None
    for c in range(C, min(n_tiles, 128) + 1, C):
        tpc = -(-n_tiles // c)
        if -(-n_tiles // tpc) != c or tpc > max_tpc:      # would leave an empty chunk / overflow the bitmap
            continue
        units = -(-blocks // R) * (c // C)
        waves = -(-units // clusters)
        cost = waves * (tpc + 1)          # time ~ waves * tiles per chunk (+1 tile of pipeline fill per unit)
        cost *= 1.0 + 0.004 * c           # fewer chunks = more clusters per W_U stream = less DRAM traffic
        resident = min(blocks, -(-clusters // (c // C)) * R)
        mb = resident * (1.835 if (group == 2 or sweep_a) else 0.92)
        if mb > 64.0:                     # stationary blocks of the resident clusters should stay in L2 (measured:
                                          # 33 blocks = 61 MB are fine, 38 blocks = 70 MB cost 2 %)
            cost *= 1.5
        if best_cost is None or cost < best_cost:
            best, best_cost = c, cost
    return best or C


def plan_row_splits(n_rows: int, V: int, cta_pair=5, *, sweep_a: bool = False, n_sms: int = 148):
    """[(row_start, row_stop, n_chunks), ...] for one lens sweep. The kernels hand out (row group, chunk)
    units round-robin to the resident clusters; when the unit count is just above a whole number of
    waves (e.g. 2048 units on 33 clusters = 62.06 waves) the last wave keeps two clusters busy for a
    full chunk while 31 idle. The row groups of that ragged wave are then swept in a second launch
    with many short chunks, one unit per cluster; partials of different chunkings merge as usual."""
    group, R, C = CLUSTER_MODES[_mode(cta_pair)]
    clusters = _resident_groups(cta_pair, n_sms)
    c = pick_chunks(n_rows, V, n_sms, cta_pair, sweep_a=sweep_a)
    rows_per_group = BLOCK_N * R if sweep_a else BLOCK_M * group * R
    groups = -(-n_rows // rows_per_group)
    per_group = c // C
    units = groups * per_group
    full, rem = divmod(units, clusters)
    if full < 4 or rem == 0 or rem > clusters // 2:
        return [(0, n_rows, c)]
    tail_groups = -(-rem // per_group)
    main_rows = (groups - tail_groups) * rows_per_group
    if main_rows <= 0 or main_rows >= n_rows:
        return [(0, n_rows, c)]
    # tail: as many chunks as give about one unit per cluster
    n_tiles = -(-V // (BLOCK_M * group)) if sweep_a else -(-V // BLOCK_N)
    want = max(C, (clusters // tail_groups) * C)
    c_tail = c
    for cand in range(min(want, n_tiles), c, -1):
        tpc = -(-n_tiles // cand)
        if cand % C == 0 and -(-n_tiles // tpc) == cand and tpc >= 8:
            c_tail = cand
            break
    if c_tail == c:
        return [(0, n_rows, c)]
    return [(0, main_rows, c), (main_rows, n_rows, c_tail)]


def _cat_partials(parts) -> "LensPartials":
    return LensPartials(pm=torch.cat([q.pm for q in parts], 1), ps=torch.cat([q.ps for q in parts], 1),
                        psec=torch.cat([q.psec for q in parts], 1), ptv=torch.cat([q.ptv for q in parts], 1),
                        pti=torch.cat([q.pti for q in parts], 1))


def _alloc_partials(n_chunks: int, n: int, dev) -> "LensPartials":
    parts = n_chunks * _lib.load().tb_lens_parts_per_chunk()     # one partial per (chunk, column half, row)
    return LensPartials(
        pm=torch.empty(parts, n, dtype=torch.float32, device=dev),
        ps=torch.empty(parts, n, dtype=torch.float32, device=dev),
        psec=torch.empty(parts, n, dtype=torch.float32, device=dev),
        ptv=torch.empty(parts, n, TB_TOPK, dtype=torch.float32, device=dev),
        pti=torch.empty(parts, n, TB_TOPK, dtype=torch.int32, device=dev))


def lens_partials(h: torch.Tensor, w_u: torch.Tensor, *, softcap: float = 0.0,
                  secret_id: Optional[torch.Tensor] = None, vocab_offset: int = 0,
                  n_chunks: Optional[int] = None, cta_pair=5) -> LensPartials:
    _need(h, torch.bfloat16, "h", 2)
    _need(w_u, torch.bfloat16, "w_u", 2)
    n, d = h.shape
    V = w_u.shape[0]
    if secret_id is not None:
        _need(secret_id, torch.int32, "secret_id", 1)
    if n_chunks is None:
        splits = plan_row_splits(n, V, cta_pair, n_sms=torch.cuda.get_device_properties(h.device).multi_processor_count)
        if len(splits) > 1:     # ragged last wave: sweep its rows separately (see plan_row_splits)
            return _cat_partials([lens_merge(lens_partials(
                h[r0:r1], w_u, softcap=softcap, secret_id=None if secret_id is None else secret_id[r0:r1],
                vocab_offset=vocab_offset, n_chunks=c, cta_pair=cta_pair), finalize=False) for r0, r1, c in splits])
        n_chunks = splits[0][2]
    p = _alloc_partials(n_chunks, n, h.device)
    if n == 0:
        return p
    check(_lib.load().tb_lens_partials(
        _ptr(h), _ptr(w_u), n, d, V, vocab_offset, float(softcap or 0.0), _ptr(secret_id), n_chunks,
        _mode(cta_pair), _ptr(p.pm), _ptr(p.ps), _ptr(p.psec), _ptr(p.ptv), _ptr(p.pti), _stream()),
        "tb_lens_partials")
    return p


def lens_merge(p: LensPartials, finalize: bool = True):
    """Merge parts per row. finalize -> LensReadout, else a one-part LensPartials."""
    n = p.n_rows
    dev = p.pm.device
    lib = _lib.load()
    if finalize:
        out = LensReadout(lse=torch.empty(n, dtype=torch.float32, device=dev),
                          secret_prob=torch.empty(n, dtype=torch.float32, device=dev),
                          top_prob=torch.empty(n, TB_TOPK, dtype=torch.float32, device=dev),
                          top_id=torch.empty(n, TB_TOPK, dtype=torch.int32, device=dev))
        if n == 0:
            return out
        check(lib.tb_lens_merge(_ptr(p.pm), _ptr(p.ps), _ptr(p.psec), _ptr(p.ptv), _ptr(p.pti), p.n_parts, n, 1,
                                None, None, None, None, None, _ptr(out.lse), _ptr(out.secret_prob),
                                _ptr(out.top_prob), _ptr(out.top_id), _stream()), "tb_lens_merge")
        return out
    o = LensPartials(pm=torch.empty(1, n, dtype=torch.float32, device=dev),
                     ps=torch.empty(1, n, dtype=torch.float32, device=dev),
                     psec=torch.empty(1, n, dtype=torch.float32, device=dev),
                     ptv=torch.empty(1, n, TB_TOPK, dtype=torch.float32, device=dev),
                     pti=torch.empty(1, n, TB_TOPK, dtype=torch.int32, device=dev))
    if n == 0:
        return o
    check(lib.tb_lens_merge(_ptr(p.pm), _ptr(p.ps), _ptr(p.psec), _ptr(p.ptv), _ptr(p.pti), p.n_parts, n, 0,
                            _ptr(o.pm), _ptr(o.ps), _ptr(o.psec), _ptr(o.ptv), _ptr(o.pti), None, None, None,
                            None, _stream()), "tb_lens_merge")
    return o


def lens_dense_probs(h: torch.Tensor, w_u: torch.Tensor, *, softcap: float = 0.0, n_chunks: int = 4,
                     cta_pair: bool = True) -> Tuple[torch.Tensor, "LensReadout"]:
    """(probs f32 [rows, V], readout) of an already normalised operand: softmax(lm_head(h)) as a dense array
    (src/models.py:135-138, what ``get_layer_logits`` returns per layer) from ONE sweep -- the lens kernel writes
    the logits while it accumulates the logsumexp, an elementwise pass normalises them in place. For the short
    sequences of the compatibility surface: rows * V * 4 bytes are materialised."""
    _need(h, torch.bfloat16, "h", 2)
    _need(w_u, torch.bfloat16, "w_u", 2)
    n, d = h.shape
    V = w_u.shape[0]
    n_tiles = -(-V // BLOCK_N)
    n_chunks = max(1, min(n_chunks, n_tiles))
    while -(-n_tiles // -(-n_tiles // n_chunks)) != n_chunks:      # no empty chunk
        n_chunks -= 1
    p = _alloc_partials(n_chunks, n, h.device)
    probs = torch.empty(n, V, dtype=torch.float32, device=h.device)
    lib = _lib.load()
    if n:
        check(lib.tb_lens_dense_partials(_ptr(h), _ptr(w_u), n, d, V, float(softcap or 0.0), None, n_chunks, int(bool(cta_pair)),
                                         _ptr(p.pm), _ptr(p.ps), _ptr(p.psec), _ptr(p.ptv), _ptr(p.pti), _ptr(probs),
                                         LOGITS_ROW_MAJOR, _stream()), "tb_lens_dense_partials")
    out = lens_merge(p, finalize=True)
    if n:
        check(lib.tb_probs_from_logits(_ptr(probs), _ptr(out.lse), n, V, _stream()), "tb_probs_from_logits")
    return probs, out


def lens_ll_from_logits(h: torch.Tensor, w_u: torch.Tensor, cur_id: Optional[torch.Tensor], prev_id: Optional[torch.Tensor],
                        logits: torch.Tensor, *, k: int = TB_TOPK, softcap: float = 0.0,
                        secret_id: Optional[torch.Tensor] = None, cta_pair=5, row_major: bool = False):
    """Lens readout AND exact LL-Top-k of one row sub-batch from ONE sweep of the GEMM, for a condition that has no
    predecessor to take a softmax reference from: the lens kernel writes its soft-capped logits into ``logits``
    (f32 [>= rows, V] scratch) while it accumulates the readout; once the logsumexp is merged, an HBM-bound kernel
    sums exp(logit - lse) over each prompt's 128 row slots with the masked ids zeroed, and the radix select ranks each
    prompt's V sums. Rows in the 128-slots-per-prompt layout. Returns (LensReadout, sums f32 [P, k], ids i32 [P, k])."""
    _need(h, torch.bfloat16, "h", 2)
    _need(w_u, torch.bfloat16, "w_u", 2)
    _need(logits, torch.float32, "logits", 2)
    n, d = h.shape
    V = w_u.shape[0]
    if n % LL_SLOTS or logits.shape[1] != V or logits.shape[0] < n:
        raise TabooKernelError(f"rows={n} must be a multiple of {LL_SLOTS}; logits scratch {tuple(logits.shape)} too small",
                               _lib.TB_ERR_INVALID)
    for name, t in (("cur_id", cur_id), ("prev_id", prev_id), ("secret_id", secret_id)):
        if t is not None:
            _need(t, torch.int32, name, 1)
    dev = h.device
    n_chunks = pick_chunks(n, V, torch.cuda.get_device_properties(dev).multi_processor_count, cta_pair)
    p = _alloc_partials(n_chunks, n, dev)
    lib = _lib.load()
    P = n // LL_SLOTS
    if n == 0:
        return lens_merge(p), torch.empty(0, k, device=dev), torch.empty(0, k, dtype=torch.int32, device=dev)
    # the scratch is only ever read by the column-sum kernel: prompt tiles (16 KB of consecutive memory per prompt and
    # 32 columns) let that kernel stream; the row-major form is kept for vocabularies that are no multiple of 32
    layout = LOGITS_PROMPT_TILES if (V % 32 == 0 and not row_major) else LOGITS_ROW_MAJOR
    check(lib.tb_lens_dense_partials(_ptr(h), _ptr(w_u), n, d, V, float(softcap or 0.0), _ptr(secret_id), n_chunks,
                                     _mode(cta_pair), _ptr(p.pm), _ptr(p.ps), _ptr(p.psec), _ptr(p.ptv), _ptr(p.pti),
                                     _ptr(logits), layout, _stream()), "tb_lens_dense_partials")
    out = lens_merge(p, finalize=True)
    acc = torch.empty(P, V, dtype=torch.float32, device=dev)
    check(lib.tb_logits_ll_colsum(_ptr(logits), _ptr(out.lse), _ptr(cur_id), _ptr(prev_id), n, V, _ptr(acc), layout,
                                  _stream()), "tb_logits_ll_colsum")
    ids, sums, _ = topm_select(acc, k, want_rank=False)
    return out, sums, ids


# ------------------------------------------------------------------------------------------
# nested budgets as rank updates of stored base logits (tb_b200.h, "rank updates")
# ------------------------------------------------------------------------------------------
LOGITS_PROMPT_TILES_RAW = 2
RANKED_MAX = TB_TOPM_MAX


def scale_cols(x: torch.Tensor, w_norm: torch.Tensor) -> torch.Tensor:
    """bf16(x * (1 + w)): the operand of the base sweep L0 = (x (.) (1 + w)) W_U^T."""
    _need(x, torch.bfloat16, "x", 2)
    _need(w_norm, torch.float32, "w_norm", 1)
    out = torch.empty_like(x)
    if x.shape[0]:
        check(_lib.load().tb_scale_cols(_ptr(x), _ptr(w_norm), x.shape[0], x.shape[1], _ptr(out), _stream()), "tb_scale_cols")
    return out


def lens_update_matrix(sae: SaeWeights, lens: LensWeights) -> torch.Tensor:
    """D = (W_dec (.) (1 + w)) W_U^T in bf16, TILE-MAJOR [ceil(V / 256), S, 256] (ids past V zero): once per weight set (a
    plain library GEMM, not part of any timed step). ``update_matrix_rows`` gives rows of the [S, V] form back."""
    wd = (sae.w_dec.float() * (1.0 + lens.w_norm.float())[None]).to(torch.bfloat16)
    S, V = wd.shape[0], lens.w_u.shape[0]
    n_tiles = -(-V // BLOCK_N)
    out = torch.zeros(n_tiles, S, BLOCK_N, dtype=torch.bfloat16, device=wd.device)
    for t0 in range(0, n_tiles, 64):           # 64 tiles at a time: no second copy of the whole matrix
        t1 = min(n_tiles, t0 + 64)
        v0, v1 = t0 * BLOCK_N, min(V, t1 * BLOCK_N)
        blk = torch.matmul(wd, lens.w_u[v0:v1].t())                      # [S, v1 - v0]
        if (v1 - v0) % BLOCK_N:
            blk = torch.nn.functional.pad(blk, (0, BLOCK_N - (v1 - v0) % BLOCK_N))
        out[t0:t1] = blk.view(S, t1 - t0, BLOCK_N).permute(1, 0, 2)
    return out


def update_matrix_rows(D_tiles: torch.Tensor, ids: torch.Tensor, V: int) -> torch.Tensor:
    """Rows ``ids`` of D in the plain [len(ids), V] form (tests, probes)."""
    return D_tiles[:, ids.long()].permute(1, 0, 2).reshape(ids.numel(), -1)[:, :V]


def rank_acts_dense(ranked: RankedActs, n_rows: int, split: bool = False) -> torch.Tensor:
    """The rank-ordered lists as a dense bf16 matrix [rows, 1024 (+ 1024 remainders)]: column r = the activation of the
    prompt's rank-r latent."""
    cols = RANKED_MAX * (2 if split else 1)
    out = torch.empty(n_rows, cols, dtype=torch.bfloat16, device=ranked.count.device)
    if n_rows:
        check(_lib.load().tb_rank_acts_dense(_ptr(ranked.row_start), ranked.cap, _ptr(ranked.count), _ptr(ranked.key),
                                             _ptr(ranked.val), n_rows, cols, RANKED_MAX if split else 0, _ptr(out), _stream()),
              "tb_rank_acts_dense")
    return out


def lens_base_logits(xw: torch.Tensor, w_u: torch.Tensor, l0: torch.Tensor, cta_pair=5) -> None:
    """L0 = xw W_U^T (raw products, f32) into ``l0`` (>= rows * V floats) in prompt tiles, one sweep, no readout."""
    _need(xw, torch.bfloat16, "xw", 2)
    _need(w_u, torch.bfloat16, "w_u", 2)
    n, d = xw.shape
    V = w_u.shape[0]
    if l0.dtype != torch.float32 or l0.numel() < n * V or not l0.is_cuda:
        raise TabooKernelError("l0 scratch too small", _lib.TB_ERR_INVALID)
    n_chunks = pick_chunks(n, V, torch.cuda.get_device_properties(xw.device).multi_processor_count, cta_pair)
    check(_lib.load().tb_lens_dense_partials(_ptr(xw), _ptr(w_u), n, d, V, 0.0, None, n_chunks, _mode(cta_pair), None, None, None,
                                             None, None, _ptr(l0), LOGITS_PROMPT_TILES_RAW, _stream()), "tb_lens_dense_partials")


def lens_rank_update(acts_dense: torch.Tensor, rank_ids: torch.Tensor, D: torch.Tensor, l0: torch.Tensor, rstd: torch.Tensor,
                     budget: int, *, V: Optional[int] = None, softcap: float = 0.0, secret_id: Optional[torch.Tensor] = None,
                     n_chunks: Optional[int] = None, logits_out: Optional[torch.Tensor] = None) -> "LensReadout":
    """Lens readout of one budget from the base logits: rstd * (L0 - acts[:, :budget] D[rank_ids[:budget]]) per prompt."""
    _need(acts_dense, torch.bfloat16, "acts_dense", 2)
    _need(rank_ids, torch.int32, "rank_ids", 2)
    _need(D, torch.bfloat16, "D", 3)
    _need(rstd, torch.float32, "rstd", 1)
    n, cols = acts_dense.shape
    V = int(V if V is not None else D.shape[0] * D.shape[2])
    if D.shape[2] != BLOCK_N or D.shape[0] != -(-V // BLOCK_N):
        raise TabooKernelError("D must be tile-major [ceil(V / 256), S, 256] (ops.lens_update_matrix)", _lib.TB_ERR_INVALID)
    if secret_id is not None:
        _need(secret_id, torch.int32, "secret_id", 1)
    if rank_ids.shape[0] * LL_SLOTS != n or rstd.shape[0] != n or l0.numel() < n * V:
        raise TabooKernelError("rank_ids / rstd / l0 do not match the rows", _lib.TB_ERR_INVALID)
    if logits_out is not None and (logits_out.dtype != torch.float32 or logits_out.numel() < n * V or not logits_out.is_cuda
                                   or logits_out.data_ptr() == l0.data_ptr()):
        raise TabooKernelError("logits_out: its own f32 scratch of rows * V", _lib.TB_ERR_INVALID)
    n_tiles = -(-V // BLOCK_N)
    sms = torch.cuda.get_device_properties(acts_dense.device).multi_processor_count
    if n_chunks is None:       # a dozen (prompt, chunk) units per SM: little tail waste, chunks still hundreds of tiles long
        n_chunks = max(1, min(max(1, n_tiles // 8), -(-12 * sms // max(1, n // LL_SLOTS))))
    n_chunks = max(1, min(n_chunks, n_tiles))
    while -(-n_tiles // -(-n_tiles // n_chunks)) != n_chunks:
        n_chunks -= 1
    lo_off = RANKED_MAX if cols > RANKED_MAX else 0
    parts = n_chunks * _lib.load().tb_lens_rank_update_parts(int(budget), lo_off)
    dev = acts_dense.device
    p = LensPartials(pm=torch.empty(parts, n, dtype=torch.float32, device=dev), ps=torch.empty(parts, n, dtype=torch.float32, device=dev),
                     psec=torch.empty(parts, n, dtype=torch.float32, device=dev),
                     ptv=torch.empty(parts, n, TB_TOPK, dtype=torch.float32, device=dev),
                     pti=torch.empty(parts, n, TB_TOPK, dtype=torch.int32, device=dev))
    check(_lib.load().tb_lens_rank_update_partials(_ptr(acts_dense), cols, lo_off, _ptr(rank_ids),
                                                   rank_ids.shape[1], _ptr(D), D.shape[1], _ptr(l0), _ptr(rstd), n, V, int(budget),
                                                   float(softcap or 0.0), _ptr(secret_id), n_chunks, _ptr(p.pm), _ptr(p.ps),
                                                   _ptr(p.psec), _ptr(p.ptv), _ptr(p.pti), _ptr(logits_out), _stream()),
          "tb_lens_rank_update_partials")
    return lens_merge(p, finalize=True)


def ll_topk_from_tiles(logits_tiles: torch.Tensor, lse: torch.Tensor, cur_id: Optional[torch.Tensor],
                       prev_id: Optional[torch.Tensor], V: int, k: int = TB_TOPK) -> Tuple[torch.Tensor, torch.Tensor]:
    """LL-Top-k (sums f32 [P, k], ids i32 [P, k]) from soft-capped logits stored in prompt tiles (log2 units)."""
    n = lse.shape[0]
    acc = torch.empty(n // LL_SLOTS, V, dtype=torch.float32, device=lse.device)
    check(_lib.load().tb_logits_ll_colsum(_ptr(logits_tiles), _ptr(lse), _ptr(cur_id), _ptr(prev_id), n, V, _ptr(acc),
                                          LOGITS_PROMPT_TILES, _stream()), "tb_logits_ll_colsum")
    ids, sums, _ = topm_select(acc, k, want_rank=False)
    return sums, ids


def logit_lens(x: torch.Tensor, lens: LensWeights, *, softcap: float = 0.0,
               secret_id: Optional[torch.Tensor] = None, n_chunks: Optional[int] = None,
               cta_pair=5, prenormed: bool = False) -> LensReadout:
    """norm -> lm_head -> softmax readouts of models.py:135-138 without the [rows, V] tensor.

    x: bf16 residual rows (or the already normalised operand h when ``prenormed``).
    """
    h = x if prenormed else rmsnorm(x, lens.w_norm, lens.eps)
    p = lens_partials(h, lens.w_u, softcap=softcap, secret_id=secret_id, vocab_offset=lens.vocab_offset,
                      n_chunks=n_chunks, cta_pair=cta_pair)
    return lens_merge(p, finalize=True)


# ------------------------------------------------------------------------------------------
# K5 LL-Top-k
# ------------------------------------------------------------------------------------------
LL_SLOTS = 128             # row slots per prompt in the K5 kernel
LOGITS_ROW_MAJOR, LOGITS_PROMPT_TILES = 0, 1     # tb_b200.h: layout of a stored-logits array
LL_MAX_CHUNK_IDS = 65536   # vocabulary ids one chunk may span (mask bitmap)


def pick_chunks_ll(n_rows: int, V: int, n_sms: int = 148, cta_pair=5) -> int:
    """Vocab chunks for the K5 sweep (bitmap limit as a floor on the count)."""
    return pick_chunks(n_rows, V, n_sms, cta_pair, sweep_a=True)


def pad_prompt_rows(t: torch.Tensor, seg: torch.Tensor, fill) -> torch.Tensor:
    """Scatter rows [seg[p], seg[p+1]) of ``t`` to row slots [128 p, 128 p + len) of a new tensor
    filled with ``fill`` elsewhere (the K5 row layout). seg: i64 [P+1] on the host or device."""
    seg = seg.to(t.device, torch.int64)
    P = seg.shape[0] - 1
    lens = seg[1:] - seg[:-1]
    if int(lens.max().item()) > LL_SLOTS:
        raise TabooKernelError(f"a prompt has more than {LL_SLOTS} positions", _lib.TB_ERR_INVALID)
    out = torch.full((P * LL_SLOTS,) + tuple(t.shape[1:]), fill, dtype=t.dtype, device=t.device)
    rows = torch.arange(int(seg[0].item()), int(seg[-1].item()), device=t.device)
    prompt = torch.repeat_interleave(torch.arange(P, device=t.device), lens)
    dst = prompt * LL_SLOTS + (rows - seg[:-1][prompt])
    out[dst] = t[rows]
    return out


def ll_topk(h: torch.Tensor, w_u: torch.Tensor, lse: torch.Tensor, cur_id: Optional[torch.Tensor] = None,
            prev_id: Optional[torch.Tensor] = None, *, k: int = TB_TOPK, softcap: float = 0.0,
            vocab_offset: int = 0, n_chunks: Optional[int] = None, cta_pair=5,
            finalize: bool = True):
    """LL-Top-k per prompt (01_reproduce_logit_lens.py:35-71, 147-149) on rows laid out 128 slots
    per prompt (``pad_prompt_rows``): h bf16 [P*128, d] normalised operand, lse f32 [P*128] from the
    lens readout (+inf in empty slots), cur_id / prev_id i32 [P*128] ids zeroed per position.

    Returns (sums f32 [P, k], ids i32 [P, k]) sorted descending; with ``finalize=False`` the merged
    one-part lists (f32 [1, P, 10], i32 [1, P, 10]) for an all_gather across vocabulary shards.
    """
    _need(h, torch.bfloat16, "h", 2)
    _need(w_u, torch.bfloat16, "w_u", 2)
    _need(lse, torch.float32, "lse", 1)
    n, d = h.shape
    V = w_u.shape[0]
    if n % LL_SLOTS or lse.shape[0] != n:
        raise TabooKernelError(f"K5 wants {LL_SLOTS} row slots per prompt: rows={n}, lse={lse.shape[0]}",
                               _lib.TB_ERR_INVALID)
    for name, t in (("cur_id", cur_id), ("prev_id", prev_id)):
        if t is not None:
            _need(t, torch.int32, name, 1)
    P = n // LL_SLOTS
    dev = h.device
    if n_chunks is None:
        splits = plan_row_splits(n, V, cta_pair, sweep_a=True,
                                 n_sms=torch.cuda.get_device_properties(dev).multi_processor_count)
        if len(splits) > 1:     # ragged last wave: sweep its row slots separately (see plan_row_splits)
            lists = [ll_topk(h[r0:r1], w_u, lse[r0:r1], None if cur_id is None else cur_id[r0:r1],
                             None if prev_id is None else prev_id[r0:r1], k=k, softcap=softcap,
                             vocab_offset=vocab_offset, n_chunks=c, cta_pair=cta_pair, finalize=False)
                     for r0, r1, c in splits]
            pv, pi = torch.cat([q[0] for q in lists], 1), torch.cat([q[1] for q in lists], 1)
            return (pv, pi) if not finalize else ll_topk_merge(pv, pi, k=k)
        n_chunks = splits[0][2]
    parts = n_chunks * CLUSTER_MODES[_mode(cta_pair)][0]
    pv = torch.empty(parts, P, TB_TOPK, dtype=torch.float32, device=dev)
    pi = torch.empty(parts, P, TB_TOPK, dtype=torch.int32, device=dev)
    lib = _lib.load()
    if P:
        check(lib.tb_ll_topk_partials(_ptr(h), _ptr(w_u), n, d, V, vocab_offset, float(softcap or 0.0), _ptr(lse),
                                      _ptr(cur_id), _ptr(prev_id), n_chunks, _mode(cta_pair), _ptr(pv), _ptr(pi),
                                      _stream()), "tb_ll_topk_partials")
    return ll_topk_merge(pv, pi, k=k, finalize=finalize)


def ll_topk_merge(pv: torch.Tensor, pi: torch.Tensor, *, k: int = TB_TOPK, finalize: bool = True):
    """Merge sorted (sum, id) lists [parts, P, 10] -> ([P, k], [P, k]) (or [1, P, 10] pairs)."""
    _need(pv, torch.float32, "part_val", 3)
    _need(pi, torch.int32, "part_id", 3)
    parts, P, _ = pv.shape
    kk = k if finalize else TB_TOPK
    ov = torch.empty(P, kk, dtype=torch.float32, device=pv.device)
    oi = torch.empty(P, kk, dtype=torch.int32, device=pv.device)
    if P:
        check(_lib.load().tb_ll_topk_merge(_ptr(pv), _ptr(pi), parts, P, kk, _ptr(ov), _ptr(oi), _stream()),
              "tb_ll_topk_merge")
    return (ov, oi) if finalize else (ov[None], oi[None])


# ------------------------------------------------------------------------------------------
# K3F + K6: lens readout and LL-Top-k candidates in one sweep
# ------------------------------------------------------------------------------------------
def lens_ll_partials(h: torch.Tensor, w_u: torch.Tensor, ref_lse: torch.Tensor, tau: torch.Tensor,
                     cur_id: Optional[torch.Tensor] = None, prev_id: Optional[torch.Tensor] = None, *,
                     softcap: float = 0.0, secret_id: Optional[torch.Tensor] = None, vocab_offset: int = 0,
                     n_chunks: Optional[int] = None, cta_pair=5, cand_cap: int = 256):
    """K3 readout partials with ``ref_lse`` as the softmax reference, plus per prompt (128 row slots)
    the ids whose masked column sum of exp(logit - ref_lse) reaches ``tau`` [P].
    Returns (LensPartials, cand_count i32 [P], cand_id i32 [P, cand_cap])."""
    _need(h, torch.bfloat16, "h", 2)
    _need(w_u, torch.bfloat16, "w_u", 2)
    _need(ref_lse, torch.float32, "ref_lse", 1)
    _need(tau, torch.float32, "tau", 1)
    n, d = h.shape
    V = w_u.shape[0]
    if n % LL_SLOTS or ref_lse.shape[0] != n or tau.shape[0] != n // LL_SLOTS:
        raise TabooKernelError(f"K3F wants {LL_SLOTS} row slots per prompt: rows={n}, ref={ref_lse.shape[0]}, "
                               f"tau={tau.shape[0]}", _lib.TB_ERR_INVALID)
    for name, t in (("cur_id", cur_id), ("prev_id", prev_id), ("secret_id", secret_id)):
        if t is not None:
            _need(t, torch.int32, name, 1)
    dev = h.device
    P = n // LL_SLOTS
    if n_chunks is None:
        splits = plan_row_splits(n, V, cta_pair, n_sms=torch.cuda.get_device_properties(dev).multi_processor_count)
        if len(splits) > 1:     # ragged last wave: sweep its rows separately (see plan_row_splits)
            outs = []
            for r0, r1, c in splits:
                sl = lambda t: None if t is None else t[r0:r1]      # noqa: E731
                q, cc, ci = lens_ll_partials(h[r0:r1], w_u, ref_lse[r0:r1], tau[r0 // LL_SLOTS:r1 // LL_SLOTS], sl(cur_id),
                                             sl(prev_id), softcap=softcap, secret_id=sl(secret_id),
                                             vocab_offset=vocab_offset, n_chunks=c, cta_pair=cta_pair, cand_cap=cand_cap)
                outs.append((lens_merge(q, finalize=False), cc, ci))
            return (_cat_partials([o[0] for o in outs]), torch.cat([o[1] for o in outs]),
                    torch.cat([o[2] for o in outs]))
        n_chunks = splits[0][2]
    p = _alloc_partials(n_chunks, n, dev)
    cand_count = torch.zeros(P, dtype=torch.int32, device=dev)      # the kernel appends with atomics
    cand_id = torch.empty(P, cand_cap, dtype=torch.int32, device=dev)
    if n:
        check(_lib.load().tb_lens_ll_partials(
            _ptr(h), _ptr(w_u), n, d, V, vocab_offset, float(softcap or 0.0), _ptr(secret_id), _ptr(ref_lse),
            _ptr(tau), _ptr(cur_id), _ptr(prev_id), n_chunks, _mode(cta_pair), _ptr(p.pm), _ptr(p.ps), _ptr(p.psec),
            _ptr(p.ptv), _ptr(p.pti), _ptr(cand_count), _ptr(cand_id), cand_cap, _stream()), "tb_lens_ll_partials")
    return p, cand_count, cand_id


_CAND_SCRATCH = {}     # (device, rows, d) -> (wc bf16 [rows, d], sums f32 [P, 128]); reused across conditions


def ll_candidates_exact(h: torch.Tensor, w_u: torch.Tensor, lse: torch.Tensor, ref_lse: torch.Tensor,
                        tau: torch.Tensor, cur_id: Optional[torch.Tensor], prev_id: Optional[torch.Tensor],
                        cand_count: torch.Tensor, cand_id: torch.Tensor, *, k: int = TB_TOPK, softcap: float = 0.0,
                        vocab_offset: int = 0):
    """Exact masked column sums of the listed candidates -> (sums f32 [P, k], ids i32 [P, k],
    theta f32 [P] = k-th best sum, flag i32 [P] = 1 where the top-k is certified exact)."""
    n, d = h.shape
    P = n // LL_SLOTS
    dev = h.device
    val = torch.empty(P, k, dtype=torch.float32, device=dev)
    ids = torch.empty(P, k, dtype=torch.int32, device=dev)
    theta = torch.empty(P, dtype=torch.float32, device=dev)
    flag = torch.empty(P, dtype=torch.int32, device=dev)
    if P:
        # gather the candidates' W_U rows, block-diagonal tcgen05 GEMM, rank + certify
        key = (dev, n, d)
        if key not in _CAND_SCRATCH:
            _CAND_SCRATCH.clear()
            _CAND_SCRATCH[key] = (torch.empty(n, d, dtype=torch.bfloat16, device=dev),
                                  torch.empty(P, LL_SLOTS, dtype=torch.float32, device=dev))
        wc, sums = _CAND_SCRATCH[key]
        check(_lib.load().tb_ll_candidates_exact_tc(
            _ptr(h), _ptr(w_u), n, d, w_u.shape[0], vocab_offset, float(softcap or 0.0), _ptr(lse), _ptr(ref_lse),
            _ptr(tau), _ptr(cur_id), _ptr(prev_id), _ptr(cand_count), _ptr(cand_id), cand_id.shape[1], k, _ptr(wc),
            _ptr(sums), _ptr(val), _ptr(ids), _ptr(theta), _ptr(flag), _stream()), "tb_ll_candidates_exact_tc")
    return val, ids, theta, flag


# ------------------------------------------------------------------------------------------
# a11 on cached dense probabilities
# ------------------------------------------------------------------------------------------
def probs_ll_colsum(probs: torch.Tensor, cur_id: Optional[torch.Tensor] = None,
                    prev_id: Optional[torch.Tensor] = None) -> torch.Tensor:
    """acc f32 [V] = sum over positions of probs [T, V] with probs[i, cur_id[i]] and
    probs[i, prev_id[i]] zeroed (aggregate_response_logits, 01_reproduce_logit_lens.py:35-71)."""
    _need(probs, torch.float32, "probs", 2)
    for name, t in (("cur_id", cur_id), ("prev_id", prev_id)):
        if t is not None:
            _need(t, torch.int32, name, 1)
    T, V = probs.shape
    acc = torch.empty(V, dtype=torch.float32, device=probs.device)
    check(_lib.load().tb_probs_ll_colsum(_ptr(probs), T, V, _ptr(cur_id), _ptr(prev_id), _ptr(acc), _stream()),
          "tb_probs_ll_colsum")
    return acc


def probs_ll_topk(probs: torch.Tensor, cur_id: Optional[torch.Tensor], prev_id: Optional[torch.Tensor], k: int):
    """(sums f32 [k], ids i32 [k]) of the k largest masked column sums (01:147-149 before decoding)."""
    acc = probs_ll_colsum(probs, cur_id, prev_id)
    idx, val, _ = topm_select(acc[None], k)
    return val[0], idx[0]
