"""The sweeps of BASELINE.json's configs through InterventionEngine, against the oracle's whole path:
configs[2] random-latent control (budgets x seeds, one rank table per seed shared by all prompts),
configs[3] low-rank direction removal at several ranks, and configs[4]'s 131k-wide SAE encode."""
import pytest
import torch

from _common import BF16, cuda_bf16, oracle, sae_to_cuda, synthetic, topk_sets_match

pytestmark = pytest.mark.gpu

P, T, D, S, V = 3, 64, 512, 2048, 4096


@pytest.fixture(scope="module")
def setup(cuda_lib):
    from taboo_brittleness_b200 import interventions, ops
    x = synthetic.make_residuals(P, T, D, seed=31)
    sae = synthetic.make_sae(D, S, seed=31, target_l0=40.0)
    lens = synthetic.make_lens(D, V, seed=31)
    sid = synthetic.make_secret_ids(P, T, V, seed=31)
    synthetic.plant_secret_direction(lens, x, sid)
    eng = interventions.InterventionEngine(sae_to_cuda(sae), ops.LensWeights(lens.w_norm.cuda(), cuda_bf16(lens.W_U)),
                                           softcap=30.0)
    return dict(x=x, sae=sae, lens=lens, sid=sid, eng=eng)


def _check_readout(out, x1, lens, sid):
    probs = oracle.lens_probs(x1, lens.w_norm, lens.W_U, 30.0, residual_dtype=BF16)
    assert (out.secret_prob.cpu() - oracle.secret_prob(probs, sid)).abs().max().item() <= 1e-3
    n_exact, n_tie, bad = topk_sets_match(out.top_id.cpu(), probs, 10)
    assert not bad and n_exact >= 0.97 * x1.shape[0]


def test_random_control_sweep(setup):
    """Seeded random latents (torch.randperm on the host): the kernels and the oracle
    ablate the same indices; readouts agree for every (budget, seed)."""
    eng, x, sae, lens, sid = (setup[k] for k in ("eng", "x", "sae", "lens", "sid"))
    budgets, seeds = [4, 64, 1024], [0, 7]
    res = eng.random_budget_sweep(cuda_bf16(x), sid.cuda(), budgets, seeds)
    torch.cuda.synchronize()
    acts = oracle.jumprelu_encode(x, sae.W_enc, sae.b_enc, sae.threshold)
    for seed in seeds:
        idx = oracle.random_latents(S, seed, max(budgets))
        for b in budgets:
            x1 = oracle.ablate_decode(x, acts, idx[:b], sae.W_dec, sae.b_dec, "error", residual_dtype=BF16)
            _check_readout(res[(b, seed)], x1, lens, sid)
    # different seeds really are different conditions
    assert not torch.equal(res[(1024, 0)].secret_prob, res[(1024, 7)].secret_prob)


def test_projection_sweep(setup):
    """x - (x U_r) U_r^T for nested r (first r columns of one orthonormal U), then the lens."""
    eng, x, lens, sid = (setup[k] for k in ("eng", "x", "lens", "sid"))
    U = synthetic.bf16_exact(oracle.random_orthonormal(D, 64, seed=3))
    ranks = [1, 8, 64]
    res = eng.projection_sweep(cuda_bf16(x), cuda_bf16(U), sid.cuda(), ranks)
    torch.cuda.synchronize()
    for r in ranks:
        x1 = oracle.lowrank_remove(x, U[:, :r], residual_dtype=BF16)
        _check_readout(res[r], x1, lens, sid)


def test_projection_sweep_ll_topk_chained_over_ranks(cuda_lib):
    """With LL-Top-k and the 128-row layout, every rank after the first reads in ONE sweep (previous rank's
    lse as reference). The lists must equal those of the two-sweep path, rank by rank."""
    from taboo_brittleness_b200 import interventions, ops
    p_, t_, d_, v_ = 4, 128, 512, 8192
    x = synthetic.make_residuals(p_, t_, d_, seed=41)
    lens = synthetic.make_lens(d_, v_, seed=41)
    sid = synthetic.make_secret_ids(p_, t_, v_, seed=41)
    synthetic.plant_secret_direction(lens, x, sid)
    sae = synthetic.make_sae(d_, 1024, seed=41, target_l0=20.0)
    U = synthetic.bf16_exact(oracle.random_orthonormal(d_, 16, seed=4))
    g = torch.Generator().manual_seed(41)
    cur = torch.randint(0, v_, (p_ * t_,), generator=g, dtype=torch.int32).cuda()
    ll = interventions.LLSpec(cur, None, rows_per_prompt=t_, k=10)
    def mk(fused):
        e = interventions.InterventionEngine(sae_to_cuda(sae), ops.LensWeights(lens.w_norm.cuda(), cuda_bf16(lens.W_U)),
                                             softcap=30.0)
        e.ll_fused = fused
        return e
    ranks = [1, 4, 16]
    e1, e2 = mk(True), mk(False)
    one = e1.projection_sweep(cuda_bf16(x), cuda_bf16(U), sid.cuda(), ranks, ll=ll)
    two = e2.projection_sweep(cuda_bf16(x), cuda_bf16(U), sid.cuda(), ranks, ll=ll)
    torch.cuda.synchronize()
    assert e1.stats.ll_fused_prompts == 2 * p_ and e2.stats.ll_fused_prompts == 0
    for r in ranks:
        assert torch.allclose(one[r].lens.lse, two[r].lens.lse, atol=2e-4)
        assert torch.equal(one[r].top_id, two[r].top_id)
        same = (one[r].ll_id.sort(1).values == two[r].ll_id.sort(1).values).all(1)
        for q in torch.nonzero(~same).flatten().tolist():      # only near-ties of the 10th sum may differ
            a, b = one[r].ll_sum[q], two[r].ll_sum[q]
            assert abs(float(a[9] - b[9])) <= 2e-3 * float(b[9])
        assert torch.allclose(one[r].ll_sum[:, 0], two[r].ll_sum[:, 0], rtol=2e-3)


def test_projection_removes_the_planted_secret(setup):
    """Removing the direction that carries a prompt's secret drops its secret probability (the point
    of the intervention), and the kernel reports the same drop as the oracle."""
    eng, x, lens, sid = (setup[k] for k in ("eng", "x", "lens", "sid"))
    # direction = unembedding row of the first prompt's secret (normalised), as the only column of U
    u = lens.W_U[int(sid[0])].float()
    U = synthetic.bf16_exact((u / u.norm())[:, None])
    base = eng.read(eng_h(eng, x), sid.cuda())
    out = eng.project_and_read(cuda_bf16(x), cuda_bf16(U), secret_id=sid.cuda())
    torch.cuda.synchronize()
    x1 = oracle.lowrank_remove(x, U, residual_dtype=BF16)
    ref = oracle.secret_prob(oracle.lens_probs(x1, lens.w_norm, lens.W_U, 30.0, residual_dtype=BF16), sid)
    assert (out.secret_prob.cpu() - ref).abs().max().item() <= 1e-3
    assert out.secret_prob[:T].max().item() < base.secret_prob[:T].max().item()


def eng_h(eng, x):
    from taboo_brittleness_b200 import ops
    return ops.rmsnorm(cuda_bf16(x), eng.lens.w_norm, eng.lens.eps)


def test_encode_131k_latents(cuda_lib):
    """configs[4]'s Gemma Scope width_131k: S = 131072 (512 tiles), d = 3584, weights made on the device.
    Active sets against a dense fp32 product of the same bf16 operands."""
    from taboo_brittleness_b200 import ops
    d, S_big, n = 3584, 131072, 256
    g = torch.Generator(device="cuda").manual_seed(5)
    w_enc_t = (torch.randn(S_big, d, generator=g, device="cuda") * d ** -0.5).to(BF16)
    b_enc = torch.randn(S_big, generator=g, device="cuda") * 0.1
    x = synthetic.make_residuals(2, 128, d, seed=5, bos_rows=0, device="cuda").to(BF16)
    pre = x.float() @ w_enc_t.float().t() + b_enc
    theta = torch.quantile(pre[:, :4096].flatten()[:2_000_000], 0.9995).expand(S_big).contiguous()  # L0 ~ 65
    count = torch.empty(n, dtype=torch.int32, device="cuda")
    cap = 512
    idx = torch.empty(n * cap, dtype=torch.int32, device="cuda")
    val = torch.empty(n * cap, dtype=torch.float32, device="cuda")
    from taboo_brittleness_b200 import _lib
    _lib.check(cuda_lib.tb_sae_encode(x.data_ptr(), w_enc_t.data_ptr(), b_enc.data_ptr(), theta.data_ptr(), n, d, S_big,
                                      None, None, cap, idx.data_ptr(), val.data_ptr(), count.data_ptr(), None, 0, 0,
                                      torch.cuda.current_stream().cuda_stream), "tb_sae_encode")
    torch.cuda.synchronize()
    want = (pre > theta) & (pre > 0)
    assert int(count.max()) <= cap
    got = torch.zeros_like(want)
    rows = torch.repeat_interleave(torch.arange(n, device="cuda"), count.long())
    offs = torch.arange(int(count.sum()), device="cuda") - torch.repeat_interleave(torch.cumsum(count.long(), 0) - count.long(), count.long())
    cols = idx[rows * cap + offs].long()
    got[rows, cols] = True
    diff = got ^ want
    # only pre-activations within bf16-GEMM accumulation noise of the threshold may differ
    assert (pre[diff] - theta[0]).abs().max().item() <= 2e-3 * float(theta[0].abs()) + 1e-3 if diff.any() else True
    assert diff.sum().item() <= 8
    assert 20 < want.sum(1).float().mean().item() < 200
