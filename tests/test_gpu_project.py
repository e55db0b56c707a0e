"""K2 parity: low-rank direction removal x' = x - (xU)U^T (Execution Plan:148) vs the oracle."""
import pytest
import torch

from _common import BF16, cuda_bf16, oracle, synthetic

pytestmark = pytest.mark.gpu

RESID_REL_TOL = 2e-2


@pytest.mark.parametrize("r", [1, 2, 5, 8, 9, 33, 64])
@pytest.mark.parametrize("n_rows", [1, 37, 512])
def test_lowrank_remove(cuda_lib, r, n_rows):
    from taboo_brittleness_b200 import ops
    d = 3584
    x = synthetic.make_residuals(1, n_rows, d, seed=77 + r)
    U = synthetic.bf16_exact(oracle.random_orthonormal(d, r, seed=r))
    w = synthetic.make_lens(d, 8, seed=1).w_norm
    out_x, out_h = ops.lowrank_remove(cuda_bf16(x), cuda_bf16(U), w_norm=w.cuda(), want_x=True, want_h=True)
    torch.cuda.synchronize()
    got = out_x.float().cpu()
    ref = oracle.lowrank_remove(x, U)
    rel = (got - ref).norm(dim=1) / ref.norm(dim=1)
    assert rel.max().item() <= RESID_REL_TOL
    assert ((got - ref).abs() <= 0.0079 * ref.abs() + 1e-3 * ref.abs().max()).all()
    # what is left along U is bf16 rounding noise only
    resid = (got @ U).abs().max().item()
    before = (x @ U).abs().max().item()
    assert resid <= 0.02 * before + 1.0
    # fused norm == standalone norm of the rounded residual up to the order of the fp32 sum of squares
    ref_h, got_h = ops.rmsnorm(out_x, w.cuda()).float(), out_h.float()
    assert (got_h != ref_h).float().mean().item() < 2e-3
    assert ((got_h - ref_h).abs() <= 0.0079 * ref_h.abs() + 1e-6).all()


def test_lowrank_remove_small_d_and_rank_zero(cuda_lib):
    from taboo_brittleness_b200 import ops
    x = synthetic.make_residuals(2, 9, 264, seed=5)
    U = synthetic.bf16_exact(oracle.random_orthonormal(264, 4, seed=2))
    out_x, _ = ops.lowrank_remove(cuda_bf16(x), cuda_bf16(U))
    ref = oracle.lowrank_remove(x, U, residual_dtype=BF16)
    assert torch.allclose(out_x.float().cpu(), ref, rtol=0.008, atol=0.05)
    same, _ = ops.lowrank_remove(cuda_bf16(x), torch.empty(264, 0, dtype=BF16, device="cuda"))
    assert torch.equal(same, cuda_bf16(x))


def test_lowrank_rejects_rank_above_64(cuda_lib):
    from taboo_brittleness_b200 import ops
    from taboo_brittleness_b200._lib import TabooKernelError
    x = torch.zeros(4, 64, dtype=BF16, device="cuda")
    with pytest.raises(TabooKernelError):
        ops.lowrank_remove(x, torch.zeros(64, 65, dtype=BF16, device="cuda"))
