"""K1 / a2 / K1b parity: JumpReLU encode to sparse lists, mean-over-tokens top-m targeting, and
ablate + decode (both forms), against the oracle on the same bf16-rounded tensors.
BASELINE.json tolerances: active sets and ablated indices bit-exact except documented
near-threshold ties; reconstructed residuals within 2e-2 relative."""
import pytest
import torch

from _common import cuda_bf16, oracle, sae_to_cuda, synthetic

pytestmark = pytest.mark.gpu

RESID_REL_TOL = 2e-2   # BASELINE.json north_star
TIE_TAU = 2e-3         # |pre - theta| below this may flip (fp32 summation order on |pre| ~ 30)

D, S, T, P = 3584, 16384, 64, 3


@pytest.fixture(scope="module")
def problem():
    x = synthetic.make_residuals(P, T, D, seed=1234)
    sae = synthetic.make_sae(D, S, seed=42)
    pre = x @ sae.W_enc + sae.b_enc
    acts = oracle.jumprelu_encode(x, sae.W_enc, sae.b_enc, sae.threshold)
    return {"x": x, "sae": sae, "pre": pre, "acts": acts}


@pytest.fixture(scope="module")
def encoded(cuda_lib, problem):
    from taboo_brittleness_b200 import ops
    w = sae_to_cuda(problem["sae"])
    xs = cuda_bf16(problem["x"])
    acts = ops.sae_encode(xs, w, cap=128)
    torch.cuda.synchronize()
    return {"w": w, "x": xs, "acts": acts}


def test_encode_active_sets_and_values(problem, encoded):
    sp = encoded["acts"]
    ref = problem["acts"]
    dense = sp.to_dense().cpu()
    # <bos> rows have thousands of active latents: the CSR re-pass must have kicked in
    assert sp.row_start is not None and int(sp.count.max()) > 128
    mism = (dense > 0) != (ref > 0)
    near = (problem["pre"] - problem["sae"].threshold).abs() < TIE_TAU * problem["pre"].abs().clamp_min(1.0)
    assert not (mism & ~near).any(), "active set differs away from the threshold"
    assert mism.sum().item() <= 1e-5 * mism.numel()
    both = (dense > 0) & (ref > 0)
    # fp32 accumulation-order noise scales with the row: |pre| <~ |x|_2 (W_enc columns ~ unit norm)
    tol = (4e-6 * problem["x"].norm(dim=1, keepdim=True)).expand_as(ref)
    assert ((dense - ref).abs()[both] <= tol[both] + 2e-5 * ref.abs()[both]).all()
    # counts are the true number of active latents; lists are in increasing latent order
    assert torch.equal(sp.count.cpu().long(), (dense > 0).sum(1))
    c = sp.count.cpu().tolist()
    rs = sp.row_start.cpu().tolist()
    idx = sp.idx.cpu()
    for r in (0, 1, T, 2 * T + 5):
        row = idx[rs[r]:rs[r] + c[r]]
        assert (row[1:] > row[:-1]).all()
    # typical rows sit near the calibrated L0
    typical = torch.tensor([c[r] for r in range(P * T) if r % T != 0], dtype=torch.float32)
    assert 40 < typical.mean().item() < 120


def test_encode_fixed_stride_without_overflow(cuda_lib, problem, encoded):
    from taboo_brittleness_b200 import ops
    rows = [r for r in range(P * T) if r % T != 0][:100]
    xs = encoded["x"][rows].contiguous()
    sp = ops.sae_encode(xs, encoded["w"], cap=256)
    assert sp.row_start is None
    assert torch.equal(sp.to_dense().cpu(), encoded["acts"].to_dense().cpu()[rows])


def test_encode_count_only_and_fused_mask(cuda_lib, problem, encoded):
    from taboo_brittleness_b200 import ops
    cnt = ops.sae_encode_raw(encoded["x"], encoded["w"])
    assert torch.equal(cnt, encoded["acts"].count)
    # random-control mask fused in the encode epilogue == masking afterwards
    ridx = oracle.random_latents(S, seed=3, m=1024)
    rank = ops.rank_table_from_indices(ridx, S).cuda()
    for budget in (1, 64, 1024):
        sp = ops.sae_encode(encoded["x"], encoded["w"], cap=128, rank=rank, rank_rows=0, budget=budget)
        ref = encoded["acts"].to_dense().clone()
        ref[:, ridx[:budget].cuda()] = 0
        assert torch.equal(sp.to_dense(), ref)


def test_latent_mean_and_topm(problem, encoded):
    from taboo_brittleness_b200 import ops
    mean = ops.latent_mean(encoded["acts"], rows_per_prompt=T)
    # the reduction itself: against torch.mean of the kernel's own activations (tight) ...
    own = encoded["acts"].to_dense().cpu().view(P, T, S).mean(1)
    assert torch.allclose(mean.cpu(), own, rtol=2e-6, atol=1e-6)
    # ... and end to end against the oracle's activations (encode noise on the <bos> rows included)
    # (latents whose active set flipped at the threshold in some row of the prompt are excluded:
    #  one flip moves the mean by theta / T; those flips are pinned in the encode test)
    ref_mean = problem["acts"].view(P, T, S).mean(1)
    dense = encoded["acts"].to_dense().cpu()
    flipped = ((dense > 0) != (problem["acts"] > 0)).view(P, T, S).any(1)
    assert flipped.sum().item() <= 1e-4 * flipped.numel()
    err = (mean.cpu() - ref_mean).abs()
    assert (err[~flipped] <= 1e-4 * ref_mean.abs()[~flipped] + 2e-3).all()
    m = 1024
    idx, val, rank = ops.topm_select(mean, m)
    torch.cuda.synchronize()
    idx, val, rank, mean_c = idx.cpu().long(), val.cpu(), rank.cpu().to(torch.int32), mean.cpu()
    for p in range(P):
        tv, ti = torch.topk(mean_c[p], m)
        assert torch.equal(val[p], tv), "values must match torch.topk on the same input bit for bit"
        assert (val[p][:-1] >= val[p][1:]).all()
        # ids agree wherever the value is not tied with a neighbour
        untied = torch.ones(m, dtype=torch.bool)
        untied[1:] &= tv[1:] != tv[:-1]
        untied[:-1] &= tv[:-1] != tv[1:]
        assert torch.equal(idx[p][untied], ti[untied])
        assert torch.equal(mean_c[p][idx[p]], val[p])
        assert len(set(idx[p].tolist())) == m
        # rank table: position of each selected latent, 0xFFFF elsewhere
        assert torch.equal(rank[p][idx[p]], torch.arange(m, dtype=torch.int32))
        assert (rank[p] != 0xFFFF).sum().item() == m
    # against the oracle's own targeting (02:70-73) for a small budget: identical sets
    for p in range(P):
        _, oi = oracle.top_latents_of_mean(problem["acts"][p * T:(p + 1) * T], 10)
        assert set(oi.tolist()) == set(idx[p][:10].tolist())


def test_latent_mean_ragged_segments(cuda_lib, problem, encoded):
    from taboo_brittleness_b200 import ops
    seg = torch.tensor([3, 3, 20, 64, 190], dtype=torch.int64)   # empty, short, long windows
    mean = ops.latent_mean(encoded["acts"], seg=seg.cuda()).cpu()
    own = encoded["acts"].to_dense().cpu()
    for p in range(4):
        a = own[seg[p]:seg[p + 1]]
        ref = a.mean(0) if a.shape[0] else torch.zeros(S)
        assert torch.allclose(mean[p], ref, rtol=2e-6, atol=1e-6)


@pytest.mark.parametrize("mode", ["error", "plain"])
@pytest.mark.parametrize("budget", [1, 16, 1024])
def test_ablate_decode_targeted(problem, encoded, mode, budget):
    from taboo_brittleness_b200 import ops
    sae = problem["sae"]
    mean = ops.latent_mean(encoded["acts"], rows_per_prompt=T)
    idx, _, rank = ops.topm_select(mean, 1024)
    w_norm = synthetic.make_lens(D, 8, seed=42).w_norm
    out_x, out_h = ops.sae_ablate_decode(encoded["x"], encoded["acts"], encoded["w"], rank, budget, mode=mode,
                                         rank_rows=T, w_norm=w_norm.cuda(), want_x=True, want_h=True)
    torch.cuda.synchronize()
    got = out_x.float().cpu()
    # The decode is checked on the kernel's own active sets: JumpReLU jumps by ~theta at the
    # threshold, so the handful of documented near-threshold flips of the encode (pinned in
    # test_encode_active_sets_and_values) would otherwise show up here as O(theta) residual changes.
    acts = encoded["acts"].to_dense().cpu()
    for p in range(P):
        sl = slice(p * T, (p + 1) * T)
        ref32 = oracle.ablate_decode(problem["x"][sl], acts[sl], idx[p, :budget].cpu(), sae.W_dec,
                                     sae.b_dec, mode)
        rel = (got[sl] - ref32).norm(dim=1) / ref32.norm(dim=1).clamp_min(1e-6)
        assert rel.max().item() <= RESID_REL_TOL
        # elementwise: one bf16 rounding of the fp32 result (plus summation-order noise)
        assert ((got[sl] - ref32).abs() <= 0.0079 * ref32.abs() + 2e-3 * ref32.abs().max()).all()
    # fused RMSNorm output == standalone norm of the rounded residual: same formula, only the order of
    # the fp32 sum of squares differs (threads-per-row differ) -> rare 1-ulp bf16 flips, nothing more
    ref_h = ops.rmsnorm(out_x, w_norm.cuda()).float()
    got_h = out_h.float()
    assert (got_h != ref_h).float().mean().item() < 2e-3
    assert ((got_h - ref_h).abs() <= 0.0079 * ref_h.abs() + 1e-6).all()


def test_ablate_decode_random_control_and_zero_budget(problem, encoded):
    from taboo_brittleness_b200 import ops
    sae = problem["sae"]
    ridx = oracle.random_latents(S, seed=11, m=1024)
    rank = ops.rank_table_from_indices(ridx, S).cuda()
    out_x, _ = ops.sae_ablate_decode(encoded["x"], encoded["acts"], encoded["w"], rank, 256, mode="error")
    ref = oracle.ablate_decode(problem["x"], encoded["acts"].to_dense().cpu(), ridx[:256], sae.W_dec, sae.b_dec,
                               "error")
    rel = (out_x.float().cpu() - ref).norm(dim=1) / ref.norm(dim=1)
    assert rel.max().item() <= RESID_REL_TOL
    # budget 0 ablates nothing: the error-term form returns x unchanged
    same, _ = ops.sae_ablate_decode(encoded["x"], encoded["acts"], encoded["w"], rank, 0, mode="error")
    assert torch.equal(same, encoded["x"])


def test_encode_large_batches_on_cta_pairs_equal_small_batches(cuda_lib):
    """From 128 rows per SM upwards the encode runs on CTA pairs (UMMA 256 x 256), below on single CTAs.
    Same accumulation order either way: the lists of a big batch equal those of its pieces, bit for bit."""
    from taboo_brittleness_b200 import ops
    d, s_, t_, p_ = 264, 2048, 100, 200                       # 20,000 rows > 128 * 148; 200-row tail block
    x = cuda_bf16(synthetic.make_residuals(p_, t_, d, seed=77))
    w = sae_to_cuda(synthetic.make_sae(d, s_, seed=7, target_l0=20.0))
    big = ops.sae_encode(x, w, cap=128)                        # (<bos>-scale rows overflow 128 slots: CSR re-pass included)
    pieces = [ops.sae_encode(x[r0:r0 + 5000].contiguous(), w, cap=128) for r0 in range(0, x.shape[0], 5000)]
    torch.cuda.synchronize()
    assert torch.equal(big.count, torch.cat([q.count for q in pieces]))
    assert torch.equal(big.to_dense(), torch.cat([q.to_dense() for q in pieces]))
    assert int(big.count.sum()) > 10 * x.shape[0]


def test_ranked_lists_and_prefix_ablation(cuda_lib, problem, encoded):
    """Nested budgets from rank-ordered lists (tb_sae_rank_lists + tb_sae_ablate_ranked): the lists hold exactly a
    row's active latents with rank < 1024, in rank order; every budget gives the residual and the fused norm of
    tb_sae_ablate_decode (same terms, summed in rank order instead of latent order: rare bf16 rounding flips) and
    matches the oracle within the north_star tolerance."""
    from taboo_brittleness_b200 import ops
    acts, w, xs = encoded["acts"], encoded["w"], encoded["x"]
    mean = ops.latent_mean(acts, rows_per_prompt=T)
    idx, _, rank = ops.topm_select(mean, 1024)
    ranked = ops.rank_lists(acts, rank, T)
    torch.cuda.synchronize()
    cnt, rs = acts.count.cpu().tolist(), acts.row_start.cpu().tolist()
    rk = rank.cpu().to(torch.int32)
    a_idx, a_val = acts.idx.cpu(), acts.val.cpu()
    r_key, r_val, r_cnt = ranked.key.cpu(), ranked.val.cpu(), ranked.count.cpu().tolist()
    for r in (0, 1, 5, T, T + 3, 2 * T + 7):
        lat = a_idx[rs[r]:rs[r] + cnt[r]].long()
        ranks = rk[r // T][lat]
        keep = ranks < 1024
        want = sorted(zip(ranks[keep].tolist(), lat[keep].tolist(), a_val[rs[r]:rs[r] + cnt[r]][keep].tolist()))
        keys = r_key[rs[r]:rs[r] + r_cnt[r]]
        got = list(zip((keys >> 20).tolist(), (keys & 0xfffff).tolist(), r_val[rs[r]:rs[r] + r_cnt[r]].tolist()))
        assert got == want
    lens_w = synthetic.make_lens(D, 8, seed=1).w_norm.cuda()
    for budget in (1, 8, 100, 1024):
        ox, oh = ops.sae_ablate_decode(xs, acts, w, rank, budget, rank_rows=T, w_norm=lens_w, want_x=True, want_h=True)
        nx, nh = ops.sae_ablate_ranked(xs, ranked, w, budget, w_norm=lens_w, want_x=True, want_h=True)
        torch.cuda.synchronize()
        assert (nx != ox).float().mean().item() < 2e-3
        assert ((nx.float() - ox.float()).abs() <= 0.0079 * ox.float().abs() + 1e-3).all()
        assert (nh != oh).float().mean().item() < 4e-3
        ref = torch.cat([oracle.ablate_decode(problem["x"][p * T:(p + 1) * T], problem["acts"][p * T:(p + 1) * T],
                                              idx[p, :budget].cpu().long(), problem["sae"].W_dec, problem["sae"].b_dec,
                                              "error", residual_dtype=torch.bfloat16) for p in range(P)])
        rel = (nx.float().cpu() - ref).norm(dim=1) / ref.norm(dim=1)
        # (rows whose active set differs from the oracle's at a near-threshold latent are pinned in the encode test)
        clean = ~((acts.to_dense().cpu() > 0) != (problem["acts"] > 0)).any(1)
        assert rel[clean].max().item() <= RESID_REL_TOL


def test_ranked_lists_fixed_stride_random_table_and_empty(cuda_lib, problem, encoded):
    """The same kernels on fixed-stride lists (no overflowing row), one ranking shared by all rows (the random
    control: rank_rows = 0), a small d (one 16-byte slice per thread in the whole-CTA kernel), and an empty batch."""
    from taboo_brittleness_b200 import ops
    rows = [r for r in range(P * T) if r % T != 0][:150]
    xs = encoded["x"][rows].contiguous()
    sp = ops.sae_encode(xs, encoded["w"], cap=256)
    assert sp.row_start is None
    ridx = oracle.random_latents(S, seed=11, m=1024)
    rank = ops.rank_table_from_indices(ridx, S).cuda()
    ranked = ops.rank_lists(sp, rank, 0)
    w_norm = synthetic.make_lens(D, 8, seed=1).w_norm.cuda()
    for budget in (1, 300, 1024):
        ox, oh = ops.sae_ablate_decode(xs, sp, encoded["w"], rank, budget, rank_rows=0, w_norm=w_norm, want_x=True, want_h=True)
        nx, nh = ops.sae_ablate_ranked(xs, ranked, encoded["w"], budget, w_norm=w_norm, want_x=True, want_h=True)
        torch.cuda.synchronize()
        assert (nx != ox).float().mean().item() < 2e-3 and (nh != oh).float().mean().item() < 4e-3
        assert ((nx.float() - ox.float()).abs() <= 0.0079 * ox.float().abs() + 1e-3).all()
    dense = sp.to_dense()
    want_cnt = (dense[:, ridx.cuda()] > 0).sum(1)
    assert torch.equal(ranked.count.long(), want_cnt)
    # every row taken by the whole-CTA kernel (threshold 1) gives the row-group kernel's residual
    import os
    os.environ["TB_HEAVY_MIN"] = "1"
    try:
        hx, _ = ops.sae_ablate_ranked(xs, ranked, encoded["w"], 1024, w_norm=w_norm, want_x=True, want_h=True)
    finally:
        del os.environ["TB_HEAVY_MIN"]
    lx, _ = ops.sae_ablate_ranked(xs, ranked, encoded["w"], 1024, w_norm=w_norm, want_x=True, want_h=True)
    torch.cuda.synchronize()
    assert torch.equal(hx, lx)
    empty = ops.SparseActs(count=torch.empty(0, dtype=torch.int32, device="cuda"), idx=sp.idx[:1], val=sp.val[:1], S=S, cap=256)
    r0 = ops.rank_lists(empty, rank, 0)
    e0, _ = ops.sae_ablate_ranked(xs[:0], r0, encoded["w"], 4)
    assert r0.count.shape == (0,) and e0.shape == (0, D)


def test_all_budgets_in_one_pass_equal_one_launch_per_budget(cuda_lib, problem, encoded):
    """tb_sae_ablate_ranked_multi: the residual, the lens operand and the RMSNorm factor of every budget of a chain from one
    pass over the rows -- bit for bit what one tb_sae_ablate_ranked launch per budget gives (CSR lists with <bos>-scale rows
    for the whole-CTA kernel, fixed-stride lists with one shared ranking, every row forced through either kernel)."""
    import os
    from taboo_brittleness_b200 import ops
    acts, w, xs = encoded["acts"], encoded["w"], encoded["x"]
    mean = ops.latent_mean(acts, rows_per_prompt=T)
    _, _, rank = ops.topm_select(mean, 1024)
    ranked = ops.rank_lists(acts, rank, T)
    w_norm = synthetic.make_lens(D, 8, seed=1).w_norm.cuda()
    budgets = [1, 2, 4, 8, 16, 32, 64, 100, 256, 512, 1024]

    def check(x_rows, rk, same_kernel=True):
        mx, mh, mr = ops.sae_ablate_ranked_multi(x_rows, rk, w, budgets, w_norm=w_norm, want_x=True, want_h=True, want_rstd=True)
        for i, b in enumerate(budgets):
            rstd = torch.empty(x_rows.shape[0], dtype=torch.float32, device="cuda")
            ox, oh = ops.sae_ablate_ranked(x_rows, rk, w, b, w_norm=w_norm, want_x=True, want_h=True, out_rstd=rstd)
            torch.cuda.synchronize()
            assert torch.equal(mx[i], ox), b
            if same_kernel:
                assert torch.equal(mh[i], oh) and torch.equal(mr[i], rstd), b
            else:   # a long list sends a row to the whole-CTA kernel for ALL budgets here, per budget only from 48 entries
                    # on: the two kernels add the squares in a different order (512 vs 128 threads): last bit of rstd
                assert torch.allclose(mr[i], rstd, rtol=1e-6) and (mh[i] != oh).float().mean().item() < 1e-4, b

    check(xs, ranked, same_kernel=False)
    for hm in ("1", "100000"):          # every row through the whole-CTA kernel / through the row-group kernel
        os.environ["TB_HEAVY_MIN"] = hm
        try:
            check(xs, ranked)
        finally:
            del os.environ["TB_HEAVY_MIN"]
    rows = [r for r in range(P * T) if r % T != 0][:150]
    sp = ops.sae_encode(xs[rows].contiguous(), w, cap=256)
    rk2 = ops.rank_lists(sp, ops.rank_table_from_indices(oracle.random_latents(S, seed=11, m=1024), S).cuda(), 0)
    check(xs[rows].contiguous(), rk2, same_kernel=False)
    only_h = ops.sae_ablate_ranked_multi(xs, ranked, w, [4, 64], w_norm=w_norm)
    assert only_h[0] is None and only_h[2] is None and only_h[1].shape == (2, xs.shape[0], D)
    with pytest.raises(ops.TabooKernelError):
        ops.sae_ablate_ranked_multi(xs, ranked, w, [8, 4], w_norm=w_norm)


def test_encode_with_latent_chunks_equals_one_sweep(cuda_lib, problem):
    """Large batches encode with the latent sweep cut in two (tb_sae_encode_chunked: sub-lists per chunk, merged in
    place): same lists, counts and overflow handling as the one-sweep kernel (20,000 rows: ragged last block,
    <bos>-scale rows that overflow, rows whose first sub-list alone overflows its half of the slots)."""
    from taboo_brittleness_b200 import ops
    w = sae_to_cuda(problem["sae"])
    x = synthetic.make_residuals(157, 128, D, seed=99, device="cuda")[:20000].to(torch.bfloat16).contiguous()
    assert x.shape[0] >= ops.ENCODE_CHUNKED_MIN_ROWS
    chunked = ops.sae_encode(x, w, cap=128)
    old = ops.ENCODE_SUB_LISTS
    ops.ENCODE_SUB_LISTS = 1
    try:
        plain = ops.sae_encode(x, w, cap=128)
    finally:
        ops.ENCODE_SUB_LISTS = old
    torch.cuda.synchronize()
    assert torch.equal(chunked.count, plain.count)
    assert chunked.row_start is not None and plain.row_start is not None
    # same (latent, value) sequence per row; the two layouts may place overflowing rows differently
    for lo in range(0, 20000, 4000):
        rows = torch.arange(lo, lo + 4000, device="cuda")
        a = _dense_rows_for_test(chunked, rows)
        b = _dense_rows_for_test(plain, rows)
        assert torch.equal(a, b)
    c = chunked.count.cpu()
    rs = chunked.row_start.cpu()
    idx = chunked.idx.cpu()
    for r in (0, 1, 127, 128, 5000, 19999):
        row = idx[rs[r]:rs[r] + c[r]]
        assert (row[1:] > row[:-1]).all(), "lists must stay in increasing latent order"


def _dense_rows_for_test(acts, rows):
    cnt = acts.count[rows].long()
    start = acts.row_start[rows] if acts.row_start is not None else rows.long() * acts.cap
    if acts.row_start is None:
        cnt = cnt.clamp(max=acts.cap)
    dense = torch.zeros(rows.numel(), acts.S, dtype=torch.float32, device=rows.device)
    r = torch.repeat_interleave(torch.arange(rows.numel(), device=rows.device), cnt)
    offs = torch.arange(int(cnt.sum().item()), device=rows.device) - torch.repeat_interleave(torch.cumsum(cnt, 0) - cnt, cnt)
    pos = torch.repeat_interleave(start, cnt) + offs
    dense[r, acts.idx[pos].long()] = acts.val[pos]
    return dense
