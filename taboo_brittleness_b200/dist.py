"""Multi-GPU plumbing of the intervention-and-readout path: one process per GPU, ``torch.distributed``
(NCCL over NVLink on the GPUs; the CPU tests drive the same code over gloo).

The reference is single-process, single-device (no collective anywhere), so
there is no reference interface to mirror here; the two patterns are the ones BASELINE.json's
north_star names:

* **sweep sharding** (configs[1], configs[2]): prompts -- and, for the random control, (seed, prompt)
  conditions -- are independent units. Every rank holds full weight replicas and works on a
  contiguous range of units; nothing is exchanged inside the loop, per-row readouts are gathered
  once at the end (``gather_rows``).
* **vocabulary-sharded lens** (configs[3]): rank g owns rows ``vocab_shard(V, g, G)`` of W_U, every
  rank sees all residual rows. One exchange: an all_gather of the merged per-row partials
  (max, sumexp, secret logit, top-10 = 23 words = 92 bytes per row and rank), then the same merge
  kernel (K4) that merges vocabulary chunks on one GPU. LL-Top-k lists are exchanged the same way.

Only the exchange and the index arithmetic live here; the kernels are called through ``ops``.
The compute callables are parameters (``partials_fn`` / ``merge_fn``) so that the CPU tests can run
the plumbing with the oracle's restatement of a partial; the defaults are the CUDA kernels and there
is no other built-in path.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence, Tuple

import torch
import torch.distributed as td

from . import ops

TILE_V = 256   # vocabulary shards are aligned to the lens kernel's N tile


def world(group=None) -> Tuple[int, int]:
    """(rank, world_size); (0, 1) when torch.distributed is not initialised."""
    if td.is_available() and td.is_initialized():
        return td.get_rank(group), td.get_world_size(group)
    return 0, 1


def shard_range(n: int, rank: int, world_size: int) -> Tuple[int, int]:
    """Contiguous, balanced split of n units: the first n % world_size ranks get one more."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError(f"rank {rank} / world {world_size}")
    base, extra = divmod(n, world_size)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def shard_conditions(n_prompts: int, n_seeds: int, rank: int, world_size: int) -> List[Tuple[int, int, int]]:
    """Random-control sweep (configs[2]): (seed, prompt) pairs are the units, laid out seed-major.
    Returns this rank's work as (seed, prompt_start, prompt_stop) runs, so that a rank touches as
    few seeds (rank tables) as possible and the encode of a prompt range is reused across them."""
    lo, hi = shard_range(n_prompts * n_seeds, rank, world_size)
    runs = []
    u = lo
    while u < hi:
        seed, p0 = divmod(u, n_prompts)
        p1 = min(n_prompts, p0 + (hi - u))
        runs.append((seed, p0, p1))
        u += p1 - p0
    return runs


def condition_unit_counts(n_prompts: int, n_seeds: int, world_size: int) -> List[int]:
    """(seed, prompt) units per rank under ``shard_conditions``, in rank order."""
    return [hi - lo for lo, hi in (shard_range(n_prompts * n_seeds, r, world_size) for r in range(world_size))]


def run_sharded_conditions(n_prompts: int, n_seeds: int, sweep_fn: Callable, group=None) -> torch.Tensor:
    """The random-control sweep (configs[2]) over the ranks of ``group``: this rank's (seed, prompt) units are
    grouped by prompt range, ``sweep_fn(p0, p1, seeds)`` is called once per range (so the encode of the range
    is shared by its seeds) and returns ``{seed: tensor [p1 - p0, ...]}`` -- one row of readouts per prompt;
    the rows of all units are then gathered to every rank in global unit order (seed-major):
    returns [n_seeds * n_prompts, ...]. No collective inside the loop, one ragged gather at the end."""
    r, w = world(group)
    runs = shard_conditions(n_prompts, n_seeds, r, w)
    ranges: dict = {}
    for seed, p0, p1 in runs:
        ranges.setdefault((p0, p1), []).append(seed)
    per_unit = {}
    for (p0, p1), seeds in ranges.items():
        res = sweep_fn(p0, p1, seeds)
        for seed in seeds:
            per_unit[(seed, p0)] = res[seed]
    if not runs:
        raise ValueError("a rank without units cannot know the readout shape: use at most n_prompts * n_seeds ranks")
    local = torch.cat([per_unit[(seed, p0)] for seed, p0, _ in runs]).contiguous()
    return gather_rows(local, condition_unit_counts(n_prompts, n_seeds, w), group)


def vocab_shard(V: int, rank: int, world_size: int, align: int = TILE_V) -> Tuple[int, int]:
    """Rows of W_U owned by ``rank``: equal tile-aligned shards (32,000 per GPU at V = 256000, G = 8);
    the last rank takes the remainder."""
    tiles = -(-V // align)
    t0, t1 = shard_range(tiles, rank, world_size)
    return min(V, t0 * align), min(V, t1 * align)


def vocab_shards_weighted(V: int, weights: Sequence[float], align: int = TILE_V) -> List[Tuple[int, int]]:
    """Tile-aligned vocabulary shards whose sizes follow ``weights`` (one per rank, e.g. each GPU's measured
    lens-GEMM rate: the GPUs of one box settle at different clocks under their power caps, and with equal shards
    every exchange waits for the slowest). Every rank gets at least one tile; the shards cover [0, V) in rank order."""
    w = [max(float(x), 0.0) for x in weights]
    if not w or sum(w) <= 0.0:
        raise ValueError("weights must be positive")
    tiles = -(-V // align)
    n = len(w)
    if tiles < n:
        raise ValueError(f"{tiles} tiles cannot be split over {n} ranks")
    total = sum(w)
    bounds, acc = [0], 0.0
    for r in range(n):
        acc += w[r]
        edge = int(round(tiles * acc / total))
        edge = max(edge, bounds[-1] + 1)                  # at least one tile each ...
        edge = min(edge, tiles - (n - 1 - r))             # ... and room for the ranks that follow
        bounds.append(edge)
    bounds[-1] = tiles
    return [(min(V, bounds[r] * align), min(V, bounds[r + 1] * align)) for r in range(n)]


# ------------------------------------------------------------------------------------------------
# exchanges
# ------------------------------------------------------------------------------------------------
def all_gather_cat(t: torch.Tensor, group=None) -> torch.Tensor:
    """all_gather of equally shaped tensors, concatenated along dim 0 in rank order."""
    _, w = world(group)
    if w == 1:
        return t
    t = t.contiguous()
    out = torch.empty((w * t.shape[0],) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    td.all_gather_into_tensor(out, t, group=group)
    return out


def gather_rows(t: torch.Tensor, counts: Sequence[int], group=None) -> torch.Tensor:
    """Gather per-rank row blocks of different lengths (``counts[r]`` rows on rank r) to every rank,
    in rank order: pad to the longest, one all_gather, strip the padding."""
    r, w = world(group)
    if w == 1:
        return t
    if len(counts) != w or t.shape[0] != counts[r]:
        raise ValueError(f"rank {r} holds {t.shape[0]} rows, counts say {list(counts)}")
    longest = max(counts)
    padded = torch.zeros((longest,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    padded[:t.shape[0]] = t
    full = all_gather_cat(padded, group).view((w, longest) + tuple(t.shape[1:]))
    return torch.cat([full[i, :counts[i]] for i in range(w)], dim=0)


# ------------------------------------------------------------------------------------------------
# vocabulary-sharded lens
# ------------------------------------------------------------------------------------------------
def _kernel_partials(h, lens: ops.LensWeights, softcap, secret_id, n_chunks, cta_pair) -> torch.Tensor:
    p = ops.lens_partials(h, lens.w_u, softcap=softcap, secret_id=secret_id, vocab_offset=lens.vocab_offset,
                          n_chunks=n_chunks, cta_pair=cta_pair)
    return ops.lens_merge(p, finalize=False).pack()            # [1, rows, 23]


def _kernel_merge(packed: torch.Tensor) -> ops.LensReadout:
    return ops.lens_merge(ops.LensPartials.unpack(packed), finalize=True)


def lens_readout_vocab_sharded(h: torch.Tensor, lens_shard: ops.LensWeights, *, softcap: float = 0.0,
                               secret_id: Optional[torch.Tensor] = None, group=None,
                               n_chunks: Optional[int] = None, cta_pair=5,
                               partials_fn: Optional[Callable] = None, merge_fn: Optional[Callable] = None,
                               timers: Optional[dict] = None):
    """Lens readout with W_U split by rows across the ranks of ``group``.

    h: normalised operand, all rows on every rank; ``lens_shard.w_u`` = this rank's rows of W_U,
    ``lens_shard.vocab_offset`` = global id of its first row. Every rank returns the full readout.
    ``timers`` (a dict of lists, CUDA only) receives (start, end) event pairs of the three phases on the
    current stream: "partials" (shard GEMM + local merge), "gather" (the exchange), "merge" (K4 over ranks).
    """
    partials_fn = partials_fn or _kernel_partials
    merge_fn = merge_fn or _kernel_merge

    def mark():
        if timers is None:
            return None
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        return e

    t0 = mark()
    local = partials_fn(h, lens_shard, softcap, secret_id, n_chunks, cta_pair)   # [1, rows, 23]
    t1 = mark()
    gathered = all_gather_cat(local, group)                                       # [world, rows, 23]
    t2 = mark()
    out = merge_fn(gathered)
    t3 = mark()
    if timers is not None:
        timers.setdefault("partials", []).append((t0, t1))
        timers.setdefault("gather", []).append((t1, t2))
        timers.setdefault("merge", []).append((t2, t3))
    return out


def ll_topk_vocab_sharded(h: torch.Tensor, lens_shard: ops.LensWeights, lse: torch.Tensor,
                          cur_id: Optional[torch.Tensor], prev_id: Optional[torch.Tensor], *, k: int = 10,
                          softcap: float = 0.0, group=None, n_chunks: Optional[int] = None, cta_pair=5,
                          partials_fn: Optional[Callable] = None, merge_fn: Optional[Callable] = None):
    """LL-Top-k with W_U split across ranks: ``lse`` must be the full-vocabulary logsumexp (from
    ``lens_readout_vocab_sharded``); each rank reduces its shard to one sorted list of 10 per
    prompt, the lists are all-gathered (80 bytes per prompt and rank) and merged."""
    if partials_fn is None:
        def partials_fn(h_, lens_, lse_, cur_, prev_):
            return ops.ll_topk(h_, lens_.w_u, lse_, cur_, prev_, softcap=softcap, vocab_offset=lens_.vocab_offset,
                               n_chunks=n_chunks, cta_pair=cta_pair, finalize=False)
    if merge_fn is None:
        def merge_fn(pv, pi):
            return ops.ll_topk_merge(pv, pi, k=k)
    pv, pi = partials_fn(h, lens_shard, lse, cur_id, prev_id)                      # [1, P, 10] each
    return merge_fn(all_gather_cat(pv, group), all_gather_cat(pi, group))
