"""world_size-2 (and 3) CPU tests of dist.py over gloo: sharding arithmetic, the all_gather
exchange of lens partials / LL-Top-k lists with global vocabulary ids, and ragged row gathers. The
per-shard compute is the oracle's restatement of a partial (the kernels need a GPU; test_gpu_lens
checks that K3/K4 produce exactly these partials), so what is tested here is the multi-process path."""
import os
import socket

import pytest
import torch
import torch.distributed as td
import torch.multiprocessing as mp

from _common import oracle, synthetic
from taboo_brittleness_b200 import dist as tbd
from taboo_brittleness_b200 import ops


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _spawn(fn, world, *args):
    port = _free_port()
    mp.spawn(_entry, args=(world, port, fn, args), nprocs=world, join=True)


def _entry(rank, world, port, fn, args):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    td.init_process_group("gloo", rank=rank, world_size=world)
    try:
        fn(rank, world, *args)
    finally:
        td.destroy_process_group()


def test_shard_arithmetic():
    for n, w in [(4096, 8), (10, 3), (2, 4), (0, 2)]:
        spans = [tbd.shard_range(n, r, w) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        sizes = [b - a for a, b in spans]
        assert max(sizes) - min(sizes) <= 1
    assert tbd.vocab_shard(256000, 3, 8) == (96000, 128000)          # 32k per GPU (BASELINE configs[3])
    assert [tbd.vocab_shard(1000, r, 3) for r in range(3)] == [(0, 512), (512, 768), (768, 1000)]
    # random control: 4 prompts x 3 seeds over 2 ranks -> 6 units each, seed-major runs
    runs = [tbd.shard_conditions(4, 3, r, 2) for r in range(2)]
    assert runs[0] == [(0, 0, 4), (1, 0, 2)] and runs[1] == [(1, 2, 4), (2, 0, 4)]
    covered = sorted((s, p) for rr in runs for s, a, b in rr for p in range(a, b))
    assert covered == [(s, p) for s in range(3) for p in range(4)]
    # speed-weighted shards: tile-aligned, contiguous, cover the vocabulary, follow the weights
    eq = tbd.vocab_shards_weighted(256000, [1.0] * 8)
    assert eq == [tbd.vocab_shard(256000, r, 8) for r in range(8)]
    sh = tbd.vocab_shards_weighted(256000, [1.0, 1.0, 0.9, 1.0, 1.1, 1.0, 1.0, 1.0])
    assert sh[0][0] == 0 and sh[-1][1] == 256000 and all(a[1] == b[0] for a, b in zip(sh, sh[1:]))
    assert all(v0 % 256 == 0 for v0, _ in sh) and sh[2][1] - sh[2][0] < sh[4][1] - sh[4][0]
    tiny = tbd.vocab_shards_weighted(1000, [1e-9, 1.0, 1e-9])
    assert [b - a for a, b in tiny] == [256, 512, 232]


def _lens_worker(rank, world, softcap):
    torch.manual_seed(0)
    P, T, d, V = 2, 16, 64, 1000
    x = synthetic.make_residuals(P, T, d, seed=7)
    lens = synthetic.make_lens(d, V, seed=7)
    sid = synthetic.make_secret_ids(P, T, V, seed=7)
    logits = oracle.lens_logits(x, lens.w_norm, lens.W_U, softcap)
    v0, v1 = tbd.vocab_shard(V, rank, world)
    shard = ops.LensWeights(w_norm=lens.w_norm, w_u=lens.W_U[v0:v1], vocab_offset=v0)

    def partials_fn(h, lw, cap, secret, n_chunks, pair):
        return oracle.lens_partial_packed(logits[:, lw.vocab_offset:lw.vocab_offset + lw.w_u.shape[0]],
                                          lw.vocab_offset, secret)

    out = tbd.lens_readout_vocab_sharded(x, shard, softcap=softcap or 0.0, secret_id=sid,
                                         partials_fn=partials_fn, merge_fn=oracle.merge_partials_packed)
    probs = torch.softmax(logits.float(), -1)
    assert torch.allclose(out["lse"], torch.logsumexp(logits.float(), -1), atol=1e-5)
    assert torch.allclose(out["secret_prob"], oracle.secret_prob(probs, sid), atol=1e-6)
    tp, ti = torch.topk(probs, 10, dim=-1)
    assert torch.equal(out["top_id"].to(torch.int64), ti)
    assert torch.allclose(out["top_prob"], tp, atol=1e-6)

    # LL-Top-k lists: per-shard lists of global ids, merged after the gather
    cur = torch.randint(0, V, (P * T,), generator=torch.Generator().manual_seed(1)).tolist()
    acc = torch.stack([oracle.aggregate_response_probs(probs[p * T:(p + 1) * T], cur[p * T:(p + 1) * T])
                       for p in range(P)])
    vals, ids = tbd.ll_topk_vocab_sharded(
        x, shard, out["lse"], None, None, k=10,
        partials_fn=lambda h, lw, lse, c, pv: oracle.ll_topk_lists(acc[:, v0:v1], v0),
        merge_fn=oracle.ll_topk_merge_lists)
    rv, ri = torch.topk(acc, 10, dim=-1)
    assert torch.equal(ids.to(torch.int64), ri) and torch.allclose(vals, rv)


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("softcap", [None, 30.0])
def test_vocab_sharded_lens_over_gloo(world, softcap):
    _spawn(_lens_worker, world, softcap)


def _gather_worker(rank, world):
    counts = [3, 0, 5][:world]
    t = torch.arange(counts[rank] * 2, dtype=torch.float32).view(counts[rank], 2) + 100 * rank
    full = tbd.gather_rows(t, counts)
    want = torch.cat([torch.arange(c * 2, dtype=torch.float32).view(c, 2) + 100 * r for r, c in enumerate(counts)])
    assert torch.equal(full, want)
    same = tbd.all_gather_cat(torch.full((2, 3), float(rank)))
    assert same.shape == (2 * world, 3) and same[2 * rank, 0] == rank


@pytest.mark.parametrize("world", [2, 3])
def test_ragged_row_gather_over_gloo(world):
    _spawn(_gather_worker, world)


def _conditions_worker(rank, world, n_prompts, n_seeds):
    calls = []

    def sweep_fn(p0, p1, seeds):
        calls.append((p0, p1, tuple(seeds)))
        # readout of unit (seed, prompt) = [seed, prompt, seed * 1000 + prompt]
        return {s: torch.stack([torch.tensor([s, p, s * 1000 + p], dtype=torch.float32) for p in range(p0, p1)])
                for s in seeds}
    full = tbd.run_sharded_conditions(n_prompts, n_seeds, sweep_fn)
    want = torch.tensor([[s, p, s * 1000 + p] for s in range(n_seeds) for p in range(n_prompts)], dtype=torch.float32)
    assert torch.equal(full, want)
    # every prompt range is swept once per rank, whatever the number of seeds that share it
    assert len(calls) == len({(a, b) for a, b, _ in calls})
    mine = sum((b - a) * len(sd) for a, b, sd in calls)
    assert mine == tbd.condition_unit_counts(n_prompts, n_seeds, world)[rank]


@pytest.mark.parametrize("world,n_prompts,n_seeds", [(2, 6, 32), (3, 5, 4), (2, 4, 3)])
def test_random_control_units_over_gloo(world, n_prompts, n_seeds):
    """configs[2]: (seed, prompt) units split over the ranks, encode shared per prompt range, ragged gather
    back into global seed-major order (32 seeds over 2 ranks: whole seeds; 4 seeds over 3: cut ranges)."""
    _spawn(_conditions_worker, world, n_prompts, n_seeds)


def test_single_process_is_the_identity():
    assert tbd.world() == (0, 1)
    t = torch.ones(3, 2)
    assert tbd.all_gather_cat(t) is t and tbd.gather_rows(t, [3]) is t
