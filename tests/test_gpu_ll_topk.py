"""K5 parity: LL-Top-k per prompt (masked column sums of the lens probabilities -> top-k ids)
against the oracle's restatement of aggregate_response_logits + torch.topk
(src/01_reproduce_logit_lens.py:35-71, 147-149), which tests/test_oracle_golden.py pins on the
reference's own function body and real cached fixtures. Id sets must be equal except at
near-ties of the k-th sum; the sums agree to 2e-3 relative (bf16 operand, fp32 accumulate)."""
import pytest
import torch

from _common import BF16, cuda_bf16, oracle, synthetic

pytestmark = pytest.mark.gpu

SUM_REL_TOL = 2e-3
SUM_ABS_TOL = 2e-4
TIE_REL_GAP = 2e-3


def _problem(P, T, d, V, seed=0):
    x = synthetic.make_residuals(P, T, d, seed=1234 + seed)
    lens = synthetic.make_lens(d, V, seed=42 + seed)
    sid = synthetic.make_secret_ids(P, T, V, seed=1234 + seed)
    synthetic.plant_secret_direction(lens, x, sid)
    return x, lens, sid


def _mask_ids(probs, P, T, V, seed):
    """cur/prev ids per position: random ids, except that on every other position the current
    id is the prompt's strongest token and on every third the previous id is its runner-up, so
    the masking really changes the ranking."""
    g = torch.Generator().manual_seed(seed)
    cur = torch.randint(0, V, (P * T,), generator=g, dtype=torch.int32)
    prev = torch.randint(0, V, (P * T,), generator=g, dtype=torch.int32)
    for p in range(P):
        acc = probs[p * T:(p + 1) * T].sum(0)
        top = torch.topk(acc, 2).indices.to(torch.int32)
        cur[p * T:(p + 1) * T:2] = top[0]
        prev[p * T + 1:(p + 1) * T:3] = top[1]
        prev[p * T] = -1          # position 0 has no previous token (01:64: "if i > 0")
    return cur, prev


def _oracle_lists(probs, cur, prev, seg, k):
    out = []
    for p in range(len(seg) - 1):
        sl = slice(seg[p], seg[p + 1])
        # (the oracle, like the reference, ignores prev at position 0; _mask_ids puts -1 there)
        acc = oracle.aggregate_response_probs(probs[sl], cur[sl].tolist(), prev[sl].tolist())
        vals, ids = torch.topk(acc, k + 1)
        out.append((acc, vals, ids))
    return out


def _compare(vals, ids, ref, k):
    vals, ids = vals.cpu(), ids.cpu()
    n_tie = 0
    for p, (acc, rv, ri) in enumerate(ref):
        got = [int(i) for i in ids[p].tolist()]
        assert len(set(got)) == k and min(got) >= 0
        # sums of the ids the kernel picked agree with the oracle's sums for those ids
        want = acc[torch.tensor(got)]
        assert torch.allclose(vals[p], want, rtol=SUM_REL_TOL, atol=SUM_ABS_TOL), (p, vals[p], want)
        assert (vals[p][:-1] >= vals[p][1:]).all()
        if set(got) != set(ri[:k].tolist()):
            kth = float(rv[k - 1])
            for i in set(got) ^ set(ri[:k].tolist()):
                assert abs(float(acc[i]) - kth) <= TIE_REL_GAP * kth, f"prompt {p}: id {i} is not a near-tie"
            n_tie += 1
    return n_tie


def _run(x, lens, P, T, V, softcap, chunks, pair, k, seed, masks=True):
    from taboo_brittleness_b200 import ops
    probs = oracle.lens_probs(x, lens.w_norm, lens.W_U, softcap, residual_dtype=BF16)
    cur, prev = _mask_ids(probs, P, T, V, seed) if masks else (None, None)
    lw = ops.LensWeights(lens.w_norm.cuda(), cuda_bf16(lens.W_U))
    h = ops.rmsnorm(cuda_bf16(x), lw.w_norm)
    lse = ops.lens_merge(ops.lens_partials(h, lw.w_u, softcap=softcap or 0.0)).lse
    seg = torch.arange(0, P * T + 1, T)
    if T != ops.LL_SLOTS:
        h_p = ops.pad_prompt_rows(h, seg, 0.0)
        lse_p = ops.pad_prompt_rows(lse, seg, float("inf"))
        cur_p = ops.pad_prompt_rows(cur.cuda(), seg, -1) if masks else None
        prev_p = ops.pad_prompt_rows(prev.cuda(), seg, -1) if masks else None
    else:
        h_p, lse_p = h, lse
        cur_p, prev_p = (cur.cuda(), prev.cuda()) if masks else (None, None)
    vals, ids = ops.ll_topk(h_p, lw.w_u, lse_p, cur_p, prev_p, k=k, softcap=softcap or 0.0, n_chunks=chunks,
                            cta_pair=pair)
    torch.cuda.synchronize()
    if not masks:
        cur = torch.full((P * T,), -1, dtype=torch.int32)
        prev = cur.clone()
    ref = _oracle_lists(probs, cur, prev, seg.tolist(), k)
    return _compare(vals, ids, ref, k)


@pytest.mark.parametrize("softcap", [None, 30.0])
@pytest.mark.parametrize("pair", [False, True])
@pytest.mark.parametrize("chunks", [1, 3, None])
def test_ll_topk_small(cuda_lib, softcap, pair, chunks):
    x, lens, _ = _problem(4, 128, 512, 8192)
    n_tie = _run(x, lens, 4, 128, 8192, softcap, chunks, pair, 10, seed=1)
    assert n_tie <= 1


@pytest.mark.parametrize("mode", [2, 3, 4, 5])
def test_ll_topk_multicast_clusters(cuda_lib, mode):
    """Cluster modes of the K5 sweep: 5 prompts = 3 blocks of 256 row slots (odd), V = 4800."""
    x, lens, _ = _problem(5, 128, 264, 4800, seed=6)
    _run(x, lens, 5, 128, 4800, 30.0, 4, mode, 10, seed=6)


@pytest.mark.parametrize("T", [1, 37, 100])
def test_ll_topk_short_prompts_are_padded(cuda_lib, T):
    """T < 128: rows are scattered to 128 slots per prompt; empty slots carry lse = +inf.
    3 prompts -> the last 256-row block is half empty; V = 5000 is a ragged last tile."""
    x, lens, _ = _problem(3, T, 264, 5000, seed=2)
    _run(x, lens, 3, T, 5000, 30.0, 2, True, 5, seed=2)


def test_ll_topk_without_masks(cuda_lib):
    x, lens, _ = _problem(2, 128, 512, 4096, seed=3)
    _run(x, lens, 2, 128, 4096, None, 2, True, 10, seed=3, masks=False)


def test_ll_topk_gemma_shapes_full_vocab(cuda_lib):
    x, lens, _ = _problem(2, 128, 3584, 256000, seed=4)
    _run(x, lens, 2, 128, 256000, 30.0, None, True, 10, seed=4)


def test_ll_topk_vocab_shards_merge_to_the_single_pass_answer(cuda_lib):
    from taboo_brittleness_b200 import ops
    x, lens, _ = _problem(2, 128, 512, 4096, seed=5)
    wu = cuda_bf16(lens.W_U)
    h = ops.rmsnorm(cuda_bf16(x), lens.w_norm.cuda())
    lse = ops.lens_merge(ops.lens_partials(h, wu, softcap=30.0)).lse
    g = torch.Generator().manual_seed(5)
    cur = torch.randint(0, 4096, (256,), generator=g, dtype=torch.int32).cuda()
    full_v, full_i = ops.ll_topk(h, wu, lse, cur, None, softcap=30.0, n_chunks=2)
    parts = [ops.ll_topk(h, wu[s * 2048:(s + 1) * 2048].contiguous(), lse, cur, None, softcap=30.0,
                         vocab_offset=s * 2048, n_chunks=1, finalize=False) for s in range(2)]
    mv, mi = ops.ll_topk_merge(torch.cat([p[0] for p in parts]), torch.cat([p[1] for p in parts]))
    torch.cuda.synchronize()
    assert torch.equal(mi, full_i)
    assert torch.allclose(mv, full_v, rtol=1e-6, atol=1e-7)


def test_ll_topk_rejects_long_prompts_and_wide_chunks(cuda_lib):
    from taboo_brittleness_b200 import ops
    from taboo_brittleness_b200._lib import TabooKernelError
    h = torch.zeros(300, 64, dtype=BF16, device="cuda")
    with pytest.raises(TabooKernelError):
        ops.pad_prompt_rows(h, torch.tensor([0, 300]), 0.0)
    wu = torch.zeros(70000, 64, dtype=BF16, device="cuda")
    hh = torch.zeros(128, 64, dtype=BF16, device="cuda")
    with pytest.raises(TabooKernelError, match="bitmap"):
        ops.ll_topk(hh, wu, torch.zeros(128, device="cuda"), n_chunks=1)


def test_sweep_with_ll_topk_matches_oracle_path(cuda_lib):
    """encode -> targeted ablation -> decode -> lens -> LL-Top-k through InterventionEngine, T = 64
    (padded to the 128-slot layout inside), against the oracle's whole path."""
    from _common import sae_to_cuda
    from taboo_brittleness_b200 import interventions, ops
    P, T, d, S, V, budget = 3, 64, 512, 2048, 4096, 8
    x = synthetic.make_residuals(P, T, d, seed=11)
    sae = synthetic.make_sae(d, S, seed=11, target_l0=40.0)
    lens = synthetic.make_lens(d, V, seed=11)
    sid = synthetic.make_secret_ids(P, T, V, seed=11)
    synthetic.plant_secret_direction(lens, x, sid)
    ref = oracle.run_path(x, T, sae.as_dict(), lens.w_norm, lens.W_U, sid, budget, 30.0, mode="error",
                          residual_dtype=BF16)
    cur, prev = _mask_ids(ref["lse_free_probs"], P, T, V, seed=11)
    eng = interventions.InterventionEngine(sae_to_cuda(sae), ops.LensWeights(lens.w_norm.cuda(), cuda_bf16(lens.W_U)),
                                           softcap=30.0, cta_pair=4)
    ll = interventions.LLSpec(cur.cuda(), prev.cuda(), rows_per_prompt=T, k=5)
    out = eng.targeted_budget_sweep(cuda_bf16(x), T, sid.cuda(), [budget], ll=ll)[budget]
    torch.cuda.synchronize()
    assert (out.secret_prob.cpu() - ref["secret_prob"]).abs().max().item() <= 1e-3
    lists = _oracle_lists(ref["lse_free_probs"], cur, prev, list(range(0, P * T + 1, T)), 5)
    _compare(out.ll_sum, out.ll_id, lists, 5)
