"""K3F + K6: lens readout and LL-Top-k in one sweep, with the previous condition's logsumexp as the
softmax reference. Checked against the oracle (same comparators as test_gpu_ll_topk.py) and against
the two-sweep kernels: the per-row readouts must agree with K3 + K4, certified prompts must carry the
oracle's top-k, uncertified ones are redone by the engine, so a sweep is exact either way."""
import pytest
import torch

from _common import BF16, cuda_bf16, oracle, sae_to_cuda, synthetic
from test_gpu_ll_topk import _compare, _mask_ids, _oracle_lists, _problem

pytestmark = pytest.mark.gpu


def _setup(P, d, V, seed, softcap):
    from taboo_brittleness_b200 import ops
    T = 128
    x, lens, sid = _problem(P, T, d, V, seed=seed)
    probs = oracle.lens_probs(x, lens.w_norm, lens.W_U, softcap, residual_dtype=BF16)
    cur, prev = _mask_ids(probs, P, T, V, seed)
    w_u = cuda_bf16(lens.W_U)
    h = ops.rmsnorm(cuda_bf16(x), lens.w_norm.cuda())
    plain = ops.lens_merge(ops.lens_partials(h, w_u, softcap=softcap or 0.0, secret_id=sid.cuda(), cta_pair=1))
    ref_lists = _oracle_lists(probs, cur, prev, list(range(0, P * T + 1, T)), 10)
    return h, w_u, sid.cuda(), cur.cuda(), prev.cuda(), plain, ref_lists


@pytest.mark.parametrize("mode", [1, 4, 5])
@pytest.mark.parametrize("softcap", [None, 30.0])
@pytest.mark.parametrize("noise", [0.0, 0.3])
def test_fused_sweep_readouts_and_certified_topk(cuda_lib, mode, softcap, noise):
    from taboo_brittleness_b200 import ops
    P, d, V = 5, 512, 8192
    h, w_u, sid, cur, prev, plain, ref_lists = _setup(P, d, V, 3, softcap)
    g = torch.Generator().manual_seed(1)
    ref_lse = plain.lse + noise * torch.randn(P * 128, generator=g).cuda()      # an imperfect reference
    theta_true = torch.tensor([float(r[1][9]) for r in ref_lists]).cuda()
    tau = (0.5 * theta_true).contiguous()
    p, cnt, cid = ops.lens_ll_partials(h, w_u, ref_lse, tau, cur, prev, softcap=softcap or 0.0, secret_id=sid,
                                       n_chunks=4, cta_pair=mode)
    out = ops.lens_merge(p)
    val, ids, theta, flag = ops.ll_candidates_exact(h, w_u, out.lse, ref_lse, tau, cur, prev, cnt, cid, k=10,
                                                    softcap=softcap or 0.0)
    torch.cuda.synchronize()
    # per-row readouts: same as the running-maximum kernel up to fp32 summation order
    assert torch.allclose(out.lse, plain.lse, atol=2e-4)
    assert torch.allclose(out.secret_prob, plain.secret_prob, atol=1e-4)
    assert torch.equal(out.top_id, plain.top_id)
    assert torch.allclose(out.top_prob, plain.top_prob, atol=1e-4)
    # certified prompts carry the oracle's top-10; with a good reference all of them certify
    ok = flag.cpu().bool()
    assert ok.any()
    if noise == 0.0:
        assert ok.all()
    keep = [i for i in range(P) if ok[i]]
    _compare(val[ok], ids[ok], [ref_lists[i] for i in keep], 10)
    assert torch.allclose(theta[ok], val[ok][:, 9])
    assert (cnt.cpu() >= 10).all() and (cnt.cpu() <= 256).all()


def test_fused_overflow_is_flagged(cuda_lib):
    """A threshold far too low lists more ids than the buffer holds: flag 0, never a wrong answer."""
    from taboo_brittleness_b200 import ops
    h, w_u, sid, cur, prev, plain, _ = _setup(2, 264, 4800, 4, 30.0)
    tau = torch.full((2,), 1e-12, device="cuda")
    p, cnt, cid = ops.lens_ll_partials(h, w_u, plain.lse, tau, cur, prev, softcap=30.0, n_chunks=4, cta_pair=4,
                                       cand_cap=64)
    out = ops.lens_merge(p)
    _, _, _, flag = ops.ll_candidates_exact(h, w_u, out.lse, plain.lse, tau, cur, prev, cnt, cid, softcap=30.0)
    torch.cuda.synchronize()
    assert (cnt.cpu() > 64).all() and (flag.cpu() == 0).all()


@pytest.mark.parametrize("overlap", [False, True])
@pytest.mark.parametrize("alpha", [0.5, 1e-9])
def test_nested_budget_sweep_one_sweep_per_condition_is_exact(cuda_lib, alpha, overlap):
    """InterventionEngine.targeted_budget_sweep with LL-Top-k: first budget two sweeps, the rest one
    sweep each (alpha = 1e-9 forces every prompt through the fallback) -- both must equal the oracle."""
    from taboo_brittleness_b200 import interventions, ops
    P, T, d, S, V = 3, 128, 512, 2048, 4096
    budgets = [1, 4, 16, 64]
    x = synthetic.make_residuals(P, T, d, seed=12)
    sae = synthetic.make_sae(d, S, seed=12, target_l0=40.0)
    lens = synthetic.make_lens(d, V, seed=12)
    sid = synthetic.make_secret_ids(P, T, V, seed=12)
    synthetic.plant_secret_direction(lens, x, sid)
    g = torch.Generator().manual_seed(12)
    cur = torch.randint(0, V, (P * T,), generator=g, dtype=torch.int32)
    prev = torch.randint(0, V, (P * T,), generator=g, dtype=torch.int32)
    prev[::T] = -1
    eng = interventions.InterventionEngine(sae_to_cuda(sae), ops.LensWeights(lens.w_norm.cuda(), cuda_bf16(lens.W_U)),
                                           softcap=30.0, cta_pair=4)
    eng.ll_alpha = alpha
    eng.overlap = overlap        # side-stream variant of the chain: same results
    ll = interventions.LLSpec(cur.cuda(), prev.cuda(), rows_per_prompt=T, k=10)
    res = eng.targeted_budget_sweep(cuda_bf16(x), T, sid.cuda(), budgets, ll=ll)
    torch.cuda.synchronize()
    assert eng.stats.ll_fused_prompts == P * (len(budgets) - 1)
    if alpha < 1e-6:
        assert eng.stats.ll_fallback_prompts == eng.stats.ll_fused_prompts
    else:
        assert eng.stats.ll_fallback_prompts < eng.stats.ll_fused_prompts
    for b in budgets:
        ref = oracle.run_path(x, T, sae.as_dict(), lens.w_norm, lens.W_U, sid, b, 30.0, mode="error",
                              residual_dtype=BF16)
        assert (res[b].secret_prob.cpu() - ref["secret_prob"]).abs().max().item() <= 1e-3
        lists = _oracle_lists(ref["lse_free_probs"], cur, prev, list(range(0, P * T + 1, T)), 10)
        _compare(res[b].ll_sum, res[b].ll_id, lists, 10)


def test_random_control_sweep_with_ll_topk_is_exact(cuda_lib):
    """configs[2] with LL-Top-k: budgets nested within a seed, the next seed's first budget seeded by the
    first seed's; every (budget, seed) condition must equal the oracle."""
    from taboo_brittleness_b200 import interventions, ops
    P, T, d, S, V = 2, 128, 512, 2048, 4096
    budgets, seeds = [2, 32], [0, 5]
    x = synthetic.make_residuals(P, T, d, seed=13)
    sae = synthetic.make_sae(d, S, seed=13, target_l0=40.0)
    lens = synthetic.make_lens(d, V, seed=13)
    sid = synthetic.make_secret_ids(P, T, V, seed=13)
    synthetic.plant_secret_direction(lens, x, sid)
    g = torch.Generator().manual_seed(13)
    cur = torch.randint(0, V, (P * T,), generator=g, dtype=torch.int32)
    prev = torch.cat([torch.tensor([-1], dtype=torch.int32), cur[:-1]])
    prev[::T] = -1
    eng = interventions.InterventionEngine(sae_to_cuda(sae), ops.LensWeights(lens.w_norm.cuda(), cuda_bf16(lens.W_U)),
                                           softcap=30.0)
    ll = interventions.LLSpec(cur.cuda(), prev.cuda(), rows_per_prompt=T, k=10)
    res = eng.random_budget_sweep(cuda_bf16(x), sid.cuda(), budgets, seeds, ll=ll)
    torch.cuda.synchronize()
    assert eng.stats.ll_fused_prompts == P * (len(budgets) * len(seeds) - 1)     # only the first condition took two sweeps
    acts = oracle.jumprelu_encode(x, sae.W_enc, sae.b_enc, sae.threshold)
    for seed in seeds:
        idx = oracle.random_latents(S, seed, max(budgets))
        for b in budgets:
            x1 = oracle.ablate_decode(x, acts, idx[:b], sae.W_dec, sae.b_dec, "error", residual_dtype=BF16)
            probs = oracle.lens_probs(x1, lens.w_norm, lens.W_U, 30.0, residual_dtype=BF16)
            lists = _oracle_lists(probs, cur, prev, list(range(0, P * T + 1, T)), 10)
            _compare(res[(b, seed)].ll_sum, res[(b, seed)].ll_id, lists, 10)
            assert (res[(b, seed)].secret_prob.cpu() - oracle.secret_prob(probs, sid)).abs().max().item() <= 1e-3


def test_row_splits_with_different_chunkings_give_the_same_readouts(cuda_lib, monkeypatch):
    """ops.plan_row_splits may sweep the rows of a ragged last wave in a second launch with another
    chunk count: K3, K3F + K6 and K5 must return what one launch returns (forced split here)."""
    from taboo_brittleness_b200 import ops
    P, d, V = 5, 512, 8192
    h, w_u, sid, cur, prev, plain, ref_lists = _setup(P, d, V, 7, 30.0)
    tau = (0.5 * torch.tensor([float(r[1][9]) for r in ref_lists])).cuda().contiguous()
    one = ops.lens_merge(ops.lens_partials(h, w_u, softcap=30.0, secret_id=sid, n_chunks=2))
    p1, c1, i1 = ops.lens_ll_partials(h, w_u, plain.lse, tau, cur, prev, softcap=30.0, secret_id=sid, n_chunks=2)
    v1, id1 = ops.ll_topk(h, w_u, plain.lse, cur, prev, softcap=30.0, n_chunks=2)
    monkeypatch.setattr(ops, "plan_row_splits", lambda n, V_, mode=5, **kw: [(0, 256, 2), (256, n, 4)])
    two = ops.lens_merge(ops.lens_partials(h, w_u, softcap=30.0, secret_id=sid))
    p2, c2, i2 = ops.lens_ll_partials(h, w_u, plain.lse, tau, cur, prev, softcap=30.0, secret_id=sid)
    v2, id2 = ops.ll_topk(h, w_u, plain.lse, cur, prev, softcap=30.0)
    torch.cuda.synchronize()
    assert p2.n_parts == 1 and p2.n_rows == P * 128
    assert torch.equal(two.top_id, one.top_id) and torch.allclose(two.lse, one.lse, atol=1e-5)
    assert torch.allclose(two.secret_prob, one.secret_prob, atol=1e-6)
    f1, f2 = ops.lens_merge(p1), ops.lens_merge(p2)
    assert torch.equal(f1.top_id, f2.top_id) and torch.allclose(f1.lse, f2.lse, atol=1e-5)
    assert torch.equal(c1, c2)
    for q in range(P):      # same candidate sets (order within a list depends on the launch)
        n = int(c1[q])
        assert set(i1[q, :n].tolist()) == set(i2[q, :n].tolist())
    assert torch.equal(id1, id2) and torch.allclose(v1, v2, rtol=1e-5, atol=1e-7)


@pytest.mark.parametrize("P", [5, 2])
def test_candidate_sums_on_tensor_cores_match_the_oracle_lists(cuda_lib, P):
    """K6: gather + block-diagonal tcgen05 GEMM + finalize against the oracle's LL lists (odd and even
    prompt counts: the last tile of the odd case is half empty)."""
    from taboo_brittleness_b200 import ops
    h, w_u, sid, cur, prev, plain, ref_lists = _setup(P, 512, 8192, 9, 30.0)
    g = torch.Generator().manual_seed(2)
    ref_lse = plain.lse + 0.1 * torch.randn(P * 128, generator=g).cuda()
    tau = (0.4 * torch.tensor([float(r[1][9]) for r in ref_lists])).cuda().contiguous()
    p, cnt, cid = ops.lens_ll_partials(h, w_u, ref_lse, tau, cur, prev, softcap=30.0, n_chunks=4)
    out = ops.lens_merge(p)
    b = ops.ll_candidates_exact(h, w_u, out.lse, ref_lse, tau, cur, prev, cnt, cid, softcap=30.0)
    torch.cuda.synchronize()
    assert (cnt.cpu() <= 128).all()
    assert b[3].bool().any()
    _compare(b[0][b[3].bool()], b[1][b[3].bool()], [r for r, f in zip(ref_lists, b[3].tolist()) if f], 10)


@pytest.mark.parametrize("softcap", [None, 30.0])
def test_first_condition_ll_from_stored_logits_equals_second_sweep(cuda_lib, softcap):
    """A condition without a predecessor: LL-Top-10 from the logits the lens kernel writes in its one sweep
    (ops.lens_ll_from_logits: dense store + per-prompt column sums + radix select, in row sub-batches) against the
    second GEMM sweep (K5) and the oracle's lists; the lens readout must be the plain kernel's."""
    from taboo_brittleness_b200 import interventions, ops
    P, d, V = 5, 512, 8192
    h, w_u, sid, cur, prev, plain, ref_lists = _setup(P, d, V, 7, softcap)
    scratch = torch.empty(2 * 128, V, dtype=torch.float32, device="cuda")
    outs = [ops.lens_ll_from_logits(h[r0:r0 + 256], w_u, cur[r0:r0 + 256], prev[r0:r0 + 256], scratch, softcap=softcap or 0.0,
                                    secret_id=sid[r0:r0 + 256], cta_pair=m) for r0, m in ((0, 5), (256, 1), (512, 4))]
    v5, i5 = ops.ll_topk(h, w_u, plain.lse, cur, prev, softcap=softcap or 0.0)
    torch.cuda.synchronize()
    lse = torch.cat([o[0].lse for o in outs])
    top = torch.cat([o[0].top_id for o in outs])
    sums = torch.cat([o[1] for o in outs])
    ids = torch.cat([o[2] for o in outs])
    assert torch.allclose(lse, plain.lse, atol=2e-4) and torch.equal(top, plain.top_id)
    assert torch.allclose(torch.cat([o[0].secret_prob for o in outs]), plain.secret_prob, atol=1e-5)
    assert torch.allclose(sums, v5, rtol=1e-4, atol=1e-7)
    same = (ids.sort(1).values == i5.sort(1).values).all(1)
    for q in torch.nonzero(~same).flatten().tolist():
        assert abs(float(sums[q, 9] - v5[q, 9])) <= 2e-3 * float(v5[q, 9])
    _compare(sums, ids, ref_lists, 10)
    # through the engine: the first budget of a chain takes this path, the second the one-sweep K3F + K6 path
    x = synthetic.make_residuals(P, 128, d, seed=71)
    lens = synthetic.make_lens(d, V, seed=71)
    sae = synthetic.make_sae(d, 1024, seed=71, target_l0=20.0)
    sidc = synthetic.make_secret_ids(P, 128, V, seed=71)

    def mk(dense):
        e = interventions.InterventionEngine(sae_to_cuda(sae), ops.LensWeights(lens.w_norm.cuda(), cuda_bf16(lens.W_U)),
                                             softcap=softcap or 0.0)
        e.ll_dense_first = dense
        e.ll_dense_rows = 256
        return e
    ll = interventions.LLSpec(cur, prev, rows_per_prompt=128, k=10)
    a = mk(True).targeted_budget_sweep(cuda_bf16(x), 128, sidc.cuda(), [2, 16], ll=ll)
    b = mk(False).targeted_budget_sweep(cuda_bf16(x), 128, sidc.cuda(), [2, 16], ll=ll)
    torch.cuda.synchronize()
    for bud in (2, 16):
        assert torch.equal(a[bud].top_id, b[bud].top_id) and torch.allclose(a[bud].lens.lse, b[bud].lens.lse, atol=2e-4)
        eq = (a[bud].ll_id.sort(1).values == b[bud].ll_id.sort(1).values).all(1)
        for q in torch.nonzero(~eq).flatten().tolist():
            assert abs(float(a[bud].ll_sum[q, 9] - b[bud].ll_sum[q, 9])) <= 2e-3 * float(b[bud].ll_sum[q, 9])
        assert torch.allclose(a[bud].ll_sum[:, 0], b[bud].ll_sum[:, 0], rtol=2e-3)


@pytest.mark.parametrize("mode", [1, 5])
def test_stored_logits_prompt_tiles_are_the_row_major_logits(cuda_lib, mode):
    """The two layouts tb_lens_dense_partials writes (tb_b200.h: TB_LOGITS_ROW_MAJOR, TB_LOGITS_PROMPT_TILES) hold the same
    logits (prompt tiles in log2 units), the column sums of either agree to rounding, and ops.lens_ll_from_logits returns the same lists from both."""
    from taboo_brittleness_b200 import _lib, ops
    P, d, V = 3, 264, 4096 + 96              # 17 column tiles of 256: the last one ragged, V % 32 == 0
    h, w_u, sid, cur, prev, plain, _ = _setup(P, d, V, 11, 30.0)
    n = P * 128
    lib = _lib.load()
    st = torch.cuda.current_stream().cuda_stream
    bufs, accs = {}, {}
    for layout in (ops.LOGITS_ROW_MAJOR, ops.LOGITS_PROMPT_TILES):
        c = ops.pick_chunks(n, V, 148, mode)
        p = ops._alloc_partials(c, n, h.device)
        buf = torch.full((n, V), float("nan"), dtype=torch.float32, device="cuda")
        _lib.check(lib.tb_lens_dense_partials(h.data_ptr(), w_u.data_ptr(), n, d, V, 30.0, sid.data_ptr(), c, mode, p.pm.data_ptr(),
                                              p.ps.data_ptr(), p.psec.data_ptr(), p.ptv.data_ptr(), p.pti.data_ptr(), buf.data_ptr(),
                                              layout, st), "dense")
        out = ops.lens_merge(p)
        acc = torch.empty(P, V, dtype=torch.float32, device="cuda")
        _lib.check(lib.tb_logits_ll_colsum(buf.data_ptr(), out.lse.data_ptr(), cur.data_ptr(), prev.data_ptr(), n, V, acc.data_ptr(),
                                           layout, st), "colsum")
        bufs[layout], accs[layout] = buf, acc
        assert torch.equal(out.top_id, plain.top_id)
    torch.cuda.synchronize()
    tiles = bufs[ops.LOGITS_PROMPT_TILES].view(P, V // 32, 128, 32).permute(0, 2, 1, 3).reshape(n, V)
    assert not torch.isnan(tiles).any()
    ln2 = torch.tensor(0.6931471805599453, dtype=torch.float32, device="cuda")      # tiles are in log2 units
    assert torch.equal(tiles * ln2, bufs[ops.LOGITS_ROW_MAJOR])
    assert torch.allclose(accs[ops.LOGITS_PROMPT_TILES], accs[ops.LOGITS_ROW_MAJOR], rtol=1e-5, atol=1e-12)
    scratch = torch.empty(n, V, dtype=torch.float32, device="cuda")
    a = ops.lens_ll_from_logits(h, w_u, cur, prev, scratch, softcap=30.0, secret_id=sid, cta_pair=mode)
    b = ops.lens_ll_from_logits(h, w_u, cur, prev, scratch, softcap=30.0, secret_id=sid, cta_pair=mode, row_major=True)
    torch.cuda.synchronize()
    assert torch.allclose(a[1], b[1], rtol=1e-5)
    same = (a[2].sort(1).values == b[2].sort(1).values).all(1)
    for q in torch.nonzero(~same).flatten().tolist():       # a tie at the k-th place to within rounding
        assert abs(float(a[1][q, 9] - b[1][q, 9])) <= 1e-5 * float(b[1][q, 9])
    with pytest.raises(ops.TabooKernelError):               # prompt tiles want whole prompts and V % 32 == 0
        _lib.check(lib.tb_logits_ll_colsum(scratch.data_ptr(), a[0].lse.data_ptr(), None, None, n - 1, V, accs[0].data_ptr(),
                                           ops.LOGITS_PROMPT_TILES, st), "colsum")
