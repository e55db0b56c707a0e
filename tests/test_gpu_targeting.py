"""targeting.py (the Execution Plan's spike tokens, secret score, PCA directions; prose only in the
reference) against the oracle's dense restatement, and the spike-restricted ablation end to end."""
import pytest
import torch

from _common import BF16, cuda_bf16, oracle, sae_to_cuda, synthetic

pytestmark = pytest.mark.gpu

P, T, D, S, V = 3, 64, 512, 2048, 4096


@pytest.fixture(scope="module")
def setup(cuda_lib):
    from taboo_brittleness_b200 import ops
    x = synthetic.make_residuals(P, T, D, seed=41)
    sae = synthetic.make_sae(D, S, seed=41, target_l0=40.0)
    lens = synthetic.make_lens(D, V, seed=41)
    sid = synthetic.make_secret_ids(P, T, V, seed=41)
    synthetic.plant_secret_direction(lens, x, sid)
    w = sae_to_cuda(sae)
    acts = ops.sae_encode(cuda_bf16(x), w, cap=256)
    out = ops.logit_lens(cuda_bf16(x), ops.LensWeights(lens.w_norm.cuda(), cuda_bf16(lens.W_U)), softcap=30.0,
                         secret_id=sid.cuda())
    return dict(x=x, sae=sae, w=w, lens=lens, sid=sid, acts=acts, secret_prob=out.secret_prob)


def test_spikes_and_secret_scores(setup):
    from taboo_brittleness_b200 import targeting
    sp = setup["secret_prob"]
    spikes = targeting.spike_positions(sp, T, 4)
    assert torch.equal(spikes.cpu(), oracle.spike_positions(sp.cpu(), T, 4))
    scores = targeting.secret_scores(setup["acts"], sp, T, spikes)
    ref = oracle.secret_scores(setup["acts"].to_dense().cpu(), sp.cpu(), T, spikes.cpu())
    assert torch.allclose(scores.cpu(), ref, rtol=2e-3, atol=1e-4 * float(ref.max()))
    assert (scores >= 0).all() and float(scores.max()) > 0
    idx, rank = targeting.score_rank_table(scores, 16)
    want = torch.topk(ref, 16, dim=1).indices
    for p in range(P):      # same top-16 sets outside near-ties of the fp32 vs fp64 scores
        diff = set(idx[p].tolist()) ^ set(want[p].tolist())
        kth = float(torch.topk(ref[p], 16).values[-1])
        assert all(abs(float(ref[p, i]) - kth) <= 2e-3 * max(kth, 1e-12) for i in diff)
    excl = torch.zeros(P * T, dtype=torch.bool, device="cuda")
    excl[spikes[0, 0]] = True
    assert int(targeting.spike_positions(sp, T, 4, exclude=excl)[0, 0]) != int(spikes[0, 0])


def test_ablation_restricted_to_spike_rows(setup):
    """Ablating the top-8 scored latents only at the spike rows: those rows change exactly as a full
    ablation would change them, all other rows are the unablated (bf16-rounded) residual."""
    from taboo_brittleness_b200 import ops, targeting
    sp, acts, w = setup["secret_prob"], setup["acts"], setup["w"]
    xs = cuda_bf16(setup["x"])
    spikes = targeting.spike_positions(sp, T, 4)
    _, rank = targeting.score_rank_table(targeting.secret_scores(acts, sp, T, spikes), 8)
    full, _ = ops.sae_ablate_decode(xs, acts, w, rank, 8, rank_rows=T)
    only, _ = ops.sae_ablate_decode(xs, acts, w, targeting.append_null_table(rank), 8,
                                    row_table=targeting.spike_row_table(spikes, T))
    mask = targeting.spike_mask(spikes, T)
    assert torch.equal(only[mask], full[mask])
    assert torch.equal(only[~mask], xs[~mask])
    assert not torch.equal(full[~mask], xs[~mask])       # the full ablation does touch other rows


def test_pca_directions(setup):
    from taboo_brittleness_b200 import ops, targeting
    g = torch.Generator().manual_seed(4)
    basis = torch.linalg.qr(torch.randn(D, 3, generator=g))[0]
    x = torch.randn(200, 3, generator=g) * torch.tensor([30.0, 10.0, 4.0]) @ basis.t() + 0.05 * torch.randn(200, D, generator=g) + 5.0
    U = targeting.pca_directions(x.cuda(), 3).cpu()
    ref = oracle.pca_directions(x, 3)
    assert torch.allclose(U.t() @ U, torch.eye(3), atol=1e-4)
    assert torch.allclose(U, ref, atol=2e-3)
    assert (U.t() @ basis).abs().max(1).values.min().item() > 0.99       # recovers the planted subspace axes
    # and the directions feed K2
    xb = cuda_bf16(synthetic.bf16_exact(x))
    out, _ = ops.lowrank_remove(xb, cuda_bf16(U))
    want = oracle.lowrank_remove(synthetic.bf16_exact(x), synthetic.bf16_exact(U), residual_dtype=BF16)
    assert torch.allclose(out.float().cpu(), want, rtol=0.008, atol=0.05)
