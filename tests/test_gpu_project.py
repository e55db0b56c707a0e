"""K2 parity: low-rank direction removal x' = x - (xU)U^T (Execution Plan:148) vs the oracle."""
import pytest
import torch

from _common import BF16, cuda_bf16, oracle, synthetic

pytestmark = pytest.mark.gpu

RESID_REL_TOL = 2e-2


@pytest.mark.parametrize("r", [1, 2, 3, 4, 5, 8, 9, 33, 64])
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


@pytest.mark.parametrize("r,d,n_rows", [(9, 3584, 700), (16, 3584, 4096), (40, 520, 300), (64, 3584, 1300)])
def test_lowrank_tensor_core_path_h_alone_equals_h_next_to_x(cuda_lib, r, d, n_rows):
    """Ranks above 8: h alone (x' written into the h buffer and scaled in place) equals h next to x',
    and x' matches the oracle (fp32 coefficients; the kernel splits them into bf16 head + remainder)."""
    from taboo_brittleness_b200 import ops
    xc = synthetic.make_residuals(1, n_rows, d, seed=5 + r)
    Uc = synthetic.bf16_exact(oracle.random_orthonormal(d, r, seed=r))
    x, U = cuda_bf16(xc), cuda_bf16(Uc)
    w = synthetic.make_lens(d, 8, seed=1).w_norm.cuda()
    tx, th = ops.lowrank_remove(x, U, w_norm=w, want_x=True, want_h=True)
    _, th2 = ops.lowrank_remove(x, U, w_norm=w, want_x=False, want_h=True)
    torch.cuda.synchronize()
    assert torch.equal(th, th2)
    ref = oracle.lowrank_remove(xc, Uc, residual_dtype=BF16)
    got = tx.float().cpu()
    assert (got != ref).float().mean().item() < 1e-2
    assert ((got - ref).abs() <= 0.0079 * ref.abs() + 1e-3).all()


@pytest.mark.parametrize("r,d,n_rows", [(8, 256, 128), (9, 512, 300), (40, 512, 1000), (16, 3584, 700), (64, 3584, 4096),
                                        (33, 3584, 5000)])
def test_lowrank_one_pass_cluster_kernel(cuda_lib, r, d, n_rows):
    """tb_lowrank_remove_fused (a cluster of d / 256 CTAs holds each 128-row block; split-K coefficients reduced
    through distributed shared memory in rank order): the reduced coefficients equal x U to fp32 accuracy, x' and h
    equal the three-pass path except where a different fp32 summation order flips a bf16 rounding, h alone equals
    h next to x', and the result is deterministic."""
    from taboo_brittleness_b200 import ops
    xc = synthetic.make_residuals(-(-n_rows // 128), 128, d, seed=5 + r)[:n_rows]
    Uc = synthetic.bf16_exact(oracle.random_orthonormal(d, r, seed=r))
    x, U = cuda_bf16(xc), cuda_bf16(Uc)
    w = synthetic.make_lens(d, 8, seed=1).w_norm.cuda()
    rp = 32 * ((r + 31) // 32)
    dbg = torch.full((n_rows, rp), float("nan"), device="cuda")
    fx, fh = ops.lowrank_remove(x, U, w_norm=w, want_x=True, want_h=True, one_pass=True, dbg_c=dbg)
    _, fh2 = ops.lowrank_remove(x, U, w_norm=w, want_x=False, want_h=True, one_pass=True)
    fx3, fh3 = ops.lowrank_remove(x, U, w_norm=w, want_x=True, want_h=True, one_pass=True)
    tx, th = ops.lowrank_remove(x, U, w_norm=w, want_x=True, want_h=True, one_pass=False)
    torch.cuda.synchronize()
    c_ref = xc @ Uc
    assert (dbg[:, :r].cpu() - c_ref).abs().max().item() <= 5e-6 * c_ref.abs().max().item()
    assert rp == r or dbg[:, r:].abs().max().item() == 0.0
    assert torch.equal(fh, fh2) and torch.equal(fx, fx3) and torch.equal(fh, fh3)
    assert (fx != tx).float().mean().item() < 5e-3 and (fh != th).float().mean().item() < 5e-3
    ref = oracle.lowrank_remove(xc, Uc, residual_dtype=BF16)
    got = fx.float().cpu()
    assert (got != ref).float().mean().item() < 5e-3
    assert ((got - ref).abs() <= 0.0079 * ref.abs() + 1e-3).all()
    ref_h, got_h = ops.rmsnorm(fx, w).float(), fh.float()
    assert (got_h != ref_h).float().mean().item() < 2e-3
    assert ((got_h - ref_h).abs() <= 0.0079 * ref_h.abs() + 1e-6).all()


def test_lowrank_tc_scratch_is_checked(cuda_lib):
    from taboo_brittleness_b200 import _lib
    need = cuda_lib.tb_lowrank_tc_scratch_bytes(1000, 3584, 33)
    assert need >= 1000 * 128 * 2 + 3584 * 128 * 2 + 2 * 1000 * 4
    x = torch.zeros(1000, 3584, dtype=BF16, device="cuda")
    ut = torch.zeros(33, 3584, dtype=BF16, device="cuda")
    buf = torch.empty(need - 256, dtype=torch.uint8, device="cuda")
    rc = cuda_lib.tb_lowrank_remove_tc(x.data_ptr(), ut.data_ptr(), 33, 1000, 3584, None, 1e-6, buf.data_ptr(),
                                       buf.numel(), x.data_ptr(), None, torch.cuda.current_stream().cuda_stream)
    assert rc == _lib.TB_ERR_INVALID and "scratch" in _lib.last_error()


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


@pytest.mark.parametrize("r,n_rows", [(8, 129), (16, 700), (33, 1000), (64, 4096)])
def test_lowrank_cluster_apply_equals_separate_passes(cuda_lib, r, n_rows, monkeypatch):
    """Ranks 8..64 at d = 3584: the default (coefficient pass + one cluster kernel for apply and norm, x added into the
    accumulator by the tensor core) against the same path with the apply and norm passes separate
    (TB_K2_THREE_PASS=1): x' and h equal except for rare roundings that fall the other way."""
    from taboo_brittleness_b200 import ops
    d = 3584
    xc = synthetic.make_residuals(-(-n_rows // 128), 128, d, seed=9 + r)[:n_rows]
    x = cuda_bf16(xc)
    U = cuda_bf16(synthetic.bf16_exact(oracle.random_orthonormal(d, r, seed=r)))
    w = synthetic.make_lens(d, 8, seed=1).w_norm.cuda()
    ax, ah = ops.lowrank_remove(x, U, w_norm=w, want_x=True, want_h=True)
    _, ah2 = ops.lowrank_remove(x, U, w_norm=w, want_x=False, want_h=True)
    ox, none_h = ops.lowrank_remove(x, U)                      # x' alone, no norm weight
    assert none_h is None and torch.equal(ox, ax)
    monkeypatch.setenv("TB_K2_THREE_PASS", "1")
    tx, th = ops.lowrank_remove(x, U, w_norm=w, want_x=True, want_h=True)
    torch.cuda.synchronize()
    assert torch.equal(ah, ah2)
    assert (ax != tx).float().mean().item() < 1e-3 and (ah != th).float().mean().item() < 1e-3
    assert ((ax.float() - tx.float()).abs() <= 0.0079 * tx.float().abs() + 1e-3).all()


@pytest.mark.parametrize("d", [512, 3584])
def test_nested_ranks_share_one_coefficient_pass(cuda_lib, d):
    """tb_lowrank_remove_tc_multi: the coefficients of the largest rank computed once, every nested rank's removal + norm
    pass on its leading ones -- bit for bit what one tb_lowrank_remove_tc call per rank gives (ranks that are no multiple
    of 16 or 32 included: the rows of U^T past the rank are zero-filled by the tensor map)."""
    from taboo_brittleness_b200 import ops
    n = 700
    x = cuda_bf16(synthetic.make_residuals(1, n, d, seed=17))
    U = cuda_bf16(synthetic.bf16_exact(oracle.random_orthonormal(d, 64, seed=5)))
    w = synthetic.make_lens(d, 8, seed=2).w_norm.cuda()
    ranks = [8, 12, 16, 24, 32, 40, 64]
    mx, mh = ops.lowrank_remove_multi(x, U, ranks, w_norm=w, want_x=True, want_h=True)
    for i, r in enumerate(ranks):
        ox, oh = ops.lowrank_remove(x, U[:, :r].contiguous(), w_norm=w, want_x=True, want_h=True)
        torch.cuda.synchronize()
        assert torch.equal(mx[i], ox) and torch.equal(mh[i], oh), r
    # ranks below 8 are taken by a CUDA-core kernel when called alone: same residual to rounding, oracle tolerance either way
    lx, _ = ops.lowrank_remove_multi(x, U, [4, 8], w_norm=w, want_x=True, want_h=False)
    sx, _ = ops.lowrank_remove(x, U[:, :4].contiguous(), want_x=True)
    ref = oracle.lowrank_remove(x.float().cpu(), U[:, :4].float().cpu(), residual_dtype=BF16)
    torch.cuda.synchronize()
    assert ((lx[0].float().cpu() - ref).norm(dim=1) / ref.norm(dim=1)).max().item() <= 2e-2
    assert (lx[0] != sx).float().mean().item() < 2e-3 and torch.equal(lx[1], mx[0])
    one = ops.lowrank_remove_multi(x, U, [32], w_norm=w)
    assert one[0] is None and torch.equal(one[1][0], mh[ranks.index(32)])
    with pytest.raises(ops.TabooKernelError):
        ops.lowrank_remove_multi(x, U, [16, 8], w_norm=w)
