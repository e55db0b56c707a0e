"""Host-side pieces of the rank-update workload that need no GPU: the tile-major layout of the update matrix
D = (W_dec (.) (1 + w)) W_U^T (ops.lens_update_matrix / ops.update_matrix_rows) and the algebra the workload rests on,
restated with the oracle's fp32 path: rstd * (L0 - A D) are the lens logits of the ablated residual."""
import torch

from _common import oracle, synthetic


def test_update_matrix_is_tile_major_and_zero_padded():
    from taboo_brittleness_b200 import ops
    d, S, V = 64, 96, 256 * 2 + 40            # three tiles of 256 ids, the last one ragged
    sae = synthetic.make_sae(d, S, seed=3, target_l0=8.0, calib_rows=64)
    lens = synthetic.make_lens(d, V, seed=3)
    w = ops.SaeWeights.from_sae_lens_layout(sae.W_enc, sae.b_enc, sae.threshold, sae.W_dec, sae.b_dec, device="cpu")
    lw = ops.LensWeights(w_norm=lens.w_norm, w_u=lens.W_U.to(torch.bfloat16))
    D = ops.lens_update_matrix(w, lw)
    assert D.shape == (3, S, 256) and D.dtype == torch.bfloat16
    wd = (w.w_dec.float() * (1.0 + lens.w_norm.float())[None]).to(torch.bfloat16)
    want = torch.matmul(wd, lw.w_u.t())
    ids = torch.tensor([5, 0, 95, 17], dtype=torch.int32)
    assert torch.equal(ops.update_matrix_rows(D, ids, V), want[ids.long()])
    assert torch.equal(D[2, :, V - 512:], torch.zeros(S, 256 - (V - 512), dtype=torch.bfloat16))   # ids past V


def test_rank_update_algebra_against_the_oracle_path():
    """x' = x - sum a_j W_dec[j]  =>  RMSNorm(x') W_U^T = rstd(x') * ((x (.) (1+w)) W_U^T - sum a_j D[j]) exactly in real
    arithmetic; in fp32 without intermediate rounding the two agree to rounding."""
    d, S, V, n = 64, 128, 300, 20
    sae = synthetic.make_sae(d, S, seed=9, target_l0=10.0, calib_rows=64)
    lens = synthetic.make_lens(d, V, seed=9)
    x = synthetic.make_residuals(1, n, d, seed=9)
    acts = oracle.jumprelu_encode(x, sae.W_enc, sae.b_enc, sae.threshold)
    idx = acts.mean(0).topk(16).indices
    x1 = oracle.ablate_decode(x, acts, idx, sae.W_dec, sae.b_dec, "error", residual_dtype=torch.float32)
    w1 = 1.0 + lens.w_norm.float()
    rstd = torch.rsqrt(x1.pow(2).mean(1) + 1e-6)
    direct = (x1 * rstd[:, None] * w1[None]) @ lens.W_U.float().t()
    L0 = (x * w1[None]) @ lens.W_U.float().t()
    Dm = (sae.W_dec.float() * w1[None]) @ lens.W_U.float().t()
    update = rstd[:, None] * (L0 - acts[:, idx] @ Dm[idx])
    assert torch.allclose(update, direct, rtol=1e-4, atol=1e-3 * direct.abs().max().item())
