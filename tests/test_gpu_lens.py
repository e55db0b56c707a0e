"""K3/K4 parity: fused RMSNorm -> unembed -> (softcap) -> logsumexp / secret prob / top-10
against the oracle's norm -> lm_head -> softmax (src/models.py:135-138) on the same bf16-rounded
tensors. Tolerances are BASELINE.json's: secret probability 1e-3 absolute, top-k id sets exact
except documented near-ties."""
import pytest
import torch

from _common import BF16, cuda_bf16, oracle, synthetic, topk_sets_match

pytestmark = pytest.mark.gpu

SECRET_ABS_TOL = 1e-3   # BASELINE.json north_star
LSE_ABS_TOL = 2e-3
FP32_FLOW_ABS_BOUND = 3e-2


def _problem(n_prompts, T, d, V, seed=0, logit_std=7.0):
    x = synthetic.make_residuals(n_prompts, T, d, seed=1234 + seed)
    lens = synthetic.make_lens(d, V, seed=42 + seed, logit_std=logit_std)
    sid = synthetic.make_secret_ids(n_prompts, T, V, seed=1234 + seed)
    synthetic.plant_secret_direction(lens, x, sid)
    return x, lens, sid


def _check(readout, x, lens, sid, softcap, M):
    probs = oracle.lens_probs(x[:M], lens.w_norm, lens.W_U, softcap, residual_dtype=BF16)
    logits = oracle.lens_logits(x[:M], lens.w_norm, lens.W_U, softcap, residual_dtype=BF16)
    ref_lse = torch.logsumexp(logits.float(), dim=-1)
    ref_sec = oracle.secret_prob(probs, sid[:M])
    lse = readout.lse.cpu()
    sec = readout.secret_prob.cpu()
    assert torch.isfinite(lse).all()
    assert (lse - ref_lse).abs().max().item() <= LSE_ABS_TOL
    assert (sec - ref_sec).abs().max().item() <= SECRET_ABS_TOL
    # the planted secrets make this a real check, not 1/V everywhere
    assert ref_sec.max().item() > 1e-2
    n_exact, n_tie, bad = topk_sets_match(readout.top_id.cpu(), probs, 10)
    assert not bad, f"top-10 id sets differ beyond near-ties in rows {bad[:8]}"
    assert n_exact >= 0.98 * M
    # top-10 probabilities, in order
    ref_top = torch.topk(probs, 10, dim=-1).values
    tp = readout.top_prob.cpu()
    assert (tp - ref_top).abs().max().item() <= SECRET_ABS_TOL
    assert (tp[:, :-1] >= tp[:, 1:]).all(), "top-10 must be sorted descending"
    # All-fp32 oracle (the reference's CPU branch, models.py:41-43): the only extra difference is
    # the bf16 rounding of the normalised operand h that ANY bf16 tensor-core GEMM needs (and that
    # the reference's own CUDA branch has too, models.py:28-30). It moves a logit by ~1e-3 of its
    # magnitude, i.e. up to ~1e-2 in p near p = 0.5 -- bounded here, documented in DESIGN.md.
    probs32 = oracle.lens_probs(x[:M], lens.w_norm, lens.W_U, softcap, residual_dtype=None)
    dev32 = (sec - oracle.secret_prob(probs32, sid[:M])).abs().max().item()
    assert dev32 <= FP32_FLOW_ABS_BOUND, f"deviation from the all-fp32 oracle {dev32}"


@pytest.mark.parametrize("softcap", [None, 30.0])
@pytest.mark.parametrize("pair", [False, True])
@pytest.mark.parametrize("chunks", [1, 4, None])
def test_lens_small_vocab(cuda_lib, softcap, pair, chunks):
    from taboo_brittleness_b200 import ops
    x, lens, sid = _problem(3, 64, 512, 8192)
    lw = ops.LensWeights(lens.w_norm.cuda(), cuda_bf16(lens.W_U))
    out = ops.logit_lens(cuda_bf16(x), lw, softcap=softcap or 0.0, secret_id=sid.cuda(),
                         n_chunks=chunks, cta_pair=pair)
    torch.cuda.synchronize()
    _check(out, x, lens, sid, softcap, x.shape[0])


@pytest.mark.parametrize("pair", [False, True])
def test_lens_ragged_rows_and_vocab(cuda_lib, pair):
    from taboo_brittleness_b200 import ops
    x, lens, sid = _problem(1, 200, 264, 1000)
    lw = ops.LensWeights(lens.w_norm.cuda(), cuda_bf16(lens.W_U))
    out = ops.logit_lens(cuda_bf16(x), lw, softcap=30.0, secret_id=sid.cuda(), n_chunks=2, cta_pair=pair)
    torch.cuda.synchronize()
    _check(out, x, lens, sid, 30.0, 200)


@pytest.mark.parametrize("softcap", [None, 30.0])
def test_lens_gemma_shapes_full_vocab(cuda_lib, softcap):
    """d = 3584, V = 256000 (1000 tiles), 2 prompts x 128 tokens incl. <bos>-scale rows."""
    from taboo_brittleness_b200 import ops
    x, lens, sid = _problem(2, 128, 3584, 256000)
    lw = ops.LensWeights(lens.w_norm.cuda(), cuda_bf16(lens.W_U))
    out = ops.logit_lens(cuda_bf16(x), lw, softcap=softcap or 0.0, secret_id=sid.cuda(), cta_pair=True)
    torch.cuda.synchronize()
    _check(out, x, lens, sid, softcap, x.shape[0])


@pytest.mark.parametrize("mode", [2, 3, 4, 5])
@pytest.mark.parametrize("shape", [(3, 64, 512, 8192, 4), (5, 128, 264, 4800, 4), (9, 128, 512, 8192, 2)])
def test_lens_multicast_clusters(cuda_lib, mode, shape):
    """2x1 / 1x2 / 2x2 clusters of CTA pairs (TMA multicast of the shared operand slabs) give the
    readout of the plain pair kernel: odd row-block counts leave a pair of the cluster on rows past
    the end, V = 4800 (19 tiles in 4 chunks of 5) makes the last chunk run a tile past the end."""
    from taboo_brittleness_b200 import ops
    P, T, d, V, chunks = shape
    x, lens, sid = _problem(P, T, d, V, seed=mode)
    lw = ops.LensWeights(lens.w_norm.cuda(), cuda_bf16(lens.W_U))
    out = ops.logit_lens(cuda_bf16(x), lw, softcap=30.0, secret_id=sid.cuda(), n_chunks=chunks, cta_pair=mode)
    ref = ops.logit_lens(cuda_bf16(x), lw, softcap=30.0, secret_id=sid.cuda(), n_chunks=chunks, cta_pair=1)
    torch.cuda.synchronize()
    assert torch.equal(out.top_id, ref.top_id)
    assert torch.equal(out.lse, ref.lse) and torch.equal(out.secret_prob, ref.secret_prob)
    assert torch.equal(out.top_prob, ref.top_prob)
    _check(out, x, lens, sid, 30.0, x.shape[0])


def test_lens_multicast_gemma_shapes_full_vocab(cuda_lib):
    from taboo_brittleness_b200 import ops
    x, lens, sid = _problem(6, 128, 3584, 256000)
    lw = ops.LensWeights(lens.w_norm.cuda(), cuda_bf16(lens.W_U))
    out = ops.logit_lens(cuda_bf16(x), lw, softcap=30.0, secret_id=sid.cuda(), cta_pair=4)
    torch.cuda.synchronize()
    _check(out, x, lens, sid, 30.0, x.shape[0])


def test_rmsnorm_matches_reference_bf16_flow(cuda_lib):
    from taboo_brittleness_b200 import ops
    x = synthetic.make_residuals(4, 64, 3584, seed=5)
    w = synthetic.make_lens(3584, 8, seed=5).w_norm
    h = ops.rmsnorm(cuda_bf16(x), w.cuda()).float().cpu()
    ref = oracle.gemma2_rmsnorm(x, w, residual_dtype=BF16)
    # identical formula; only the fp32 sum-of-squares order differs -> rare 1-ulp bf16 flips
    mism = (h != ref)
    assert mism.float().mean().item() < 2e-3
    assert ((h - ref).abs() <= 0.0079 * ref.abs() + 1e-6).all()


def test_vocab_shard_merge_equals_single_pass(cuda_lib):
    """Two half-vocabulary 'shards' merged through K4 == one full pass (what the NCCL path does)."""
    from taboo_brittleness_b200 import ops
    x, lens, sid = _problem(2, 64, 512, 4096)
    h = ops.rmsnorm(cuda_bf16(x), lens.w_norm.cuda())
    wu = cuda_bf16(lens.W_U)
    full = ops.lens_merge(ops.lens_partials(h, wu, softcap=30.0, secret_id=sid.cuda(), n_chunks=1))
    parts = []
    for s in range(2):
        p = ops.lens_partials(h, wu[s * 2048:(s + 1) * 2048].contiguous(), softcap=30.0, secret_id=sid.cuda(),
                              vocab_offset=s * 2048, n_chunks=2)
        parts.append(ops.lens_merge(p, finalize=False).pack())
    merged = ops.lens_merge(ops.LensPartials.unpack(torch.cat(parts, 0)))
    torch.cuda.synchronize()
    assert torch.equal(merged.top_id, full.top_id)
    assert torch.allclose(merged.lse, full.lse, atol=1e-5)
    assert torch.allclose(merged.secret_prob, full.secret_prob, atol=1e-6)
    assert torch.allclose(merged.top_prob, full.top_prob, atol=1e-6)


def test_empty_batch(cuda_lib):
    from taboo_brittleness_b200 import ops
    lens = synthetic.make_lens(64, 256, seed=1)
    lw = ops.LensWeights(lens.w_norm.cuda(), cuda_bf16(lens.W_U))
    out = ops.logit_lens(torch.empty(0, 64, dtype=BF16, device="cuda"), lw)
    assert out.lse.shape == (0,) and out.top_id.shape == (0, 10)


def test_multicast_modes_refuse_operands_outside_this_gpus_memory(cuda_lib):
    """Cluster modes 2-5 multicast operand tiles; their operands must be plain memory of the current GPU.
    A pinned host buffer (reachable from the GPU, but not its memory) is refused there and accepted by the
    modes that do not multicast."""
    from taboo_brittleness_b200 import ops
    from taboo_brittleness_b200._lib import TabooKernelError
    h = torch.zeros(256, 512, dtype=torch.bfloat16, device="cuda")
    wu_dev = torch.zeros(1024, 512, dtype=torch.bfloat16, device="cuda")
    wu_host = torch.zeros(1024, 512, dtype=torch.bfloat16).pin_memory()
    n_parts = 2 * cuda_lib.tb_lens_parts_per_chunk()
    f = [torch.empty(n_parts * 256 * k, dtype=torch.float32, device="cuda") for k in (1, 1, 1, 10)]
    ti = torch.empty(n_parts * 256 * 10, dtype=torch.int32, device="cuda")

    def call(wu, mode):
        return cuda_lib.tb_lens_partials(h.data_ptr(), wu.data_ptr(), 256, 512, 1024, 0, 30.0, None, 2, mode,
                                         f[0].data_ptr(), f[1].data_ptr(), f[2].data_ptr(), f[3].data_ptr(), ti.data_ptr(),
                                         torch.cuda.current_stream().cuda_stream)
    for mode in (2, 3, 4, 5):
        assert call(wu_host, mode) != 0
        assert "current GPU" in ops._lib.last_error()
        assert call(wu_dev, mode) == 0
    assert call(wu_host, 1) == 0
    torch.cuda.synchronize()
    assert TabooKernelError is not None


def test_multicast_modes_check_the_whole_operand_extent(cuda_lib):
    """The guard walks every byte range the tensor maps cover, not only the base pointers: a pointer into
    the middle of a local allocation passes, an extent that runs far past the end of its allocation (into
    address space that cannot all be this GPU's memory) is refused before any launch."""
    from taboo_brittleness_b200 import ops
    big = torch.zeros(4 * 1024, 512, dtype=torch.bfloat16, device="cuda")
    h = torch.zeros(256, 512, dtype=torch.bfloat16, device="cuda")
    n_parts = 2 * cuda_lib.tb_lens_parts_per_chunk()
    f = [torch.empty(n_parts * 256 * k, dtype=torch.float32, device="cuda") for k in (1, 1, 1, 10)]
    ti = torch.empty(n_parts * 256 * 10, dtype=torch.int32, device="cuda")

    def call(wu_ptr, V, mode):
        return cuda_lib.tb_lens_partials(h.data_ptr(), wu_ptr, 256, 512, V, 0, 30.0, None, 2, mode,
                                         f[0].data_ptr(), f[1].data_ptr(), f[2].data_ptr(), f[3].data_ptr(), ti.data_ptr(),
                                         torch.cuda.current_stream().cuda_stream)
    interior = big[1024:2048]
    assert interior.data_ptr() != big.data_ptr()
    for mode in (2, 5):
        assert call(interior.data_ptr(), 1024, mode) == 0
    torch.cuda.synchronize()
    # an extent of 2 TB cannot be this GPU's memory all the way: refused before any launch, whatever lies
    # directly behind the allocation
    assert call(big.data_ptr(), 2_000_000_000, 2) != 0
    err = ops._lib.last_error()
    assert "current GPU" in err or "no mapping" in err or "separate mappings" in err, err


@pytest.mark.parametrize("V", [1000, 1001])      # 16-byte row pitch (logits leave through the TMA unit) / not (plain stores)
@pytest.mark.parametrize("softcap", [None, 30.0])
@pytest.mark.parametrize("pair", [False, True])
def test_dense_probabilities_from_one_sweep(cuda_lib, softcap, pair, V):
    """ops.lens_dense_probs: the dense [rows, V] softmax the compatibility surface returns (src/models.py:135-138)
    from one sweep (logits written by the lens kernel + in-place normalisation), ragged rows and vocabulary; its
    readout equals the plain lens readout."""
    from taboo_brittleness_b200 import ops
    x, lens, sid = _problem(1, 200, 264, V, seed=9)
    h = ops.rmsnorm(cuda_bf16(x), lens.w_norm.cuda())
    wu = cuda_bf16(lens.W_U)
    probs, out = ops.lens_dense_probs(h, wu, softcap=softcap or 0.0, n_chunks=2, cta_pair=pair)
    plain = ops.lens_merge(ops.lens_partials(h, wu, softcap=softcap or 0.0, n_chunks=2, cta_pair=pair))
    torch.cuda.synchronize()
    ref = oracle.lens_probs(x, lens.w_norm, lens.W_U, softcap, residual_dtype=BF16)
    assert probs.shape == (200, V)
    assert (probs.cpu() - ref).abs().max().item() <= SECRET_ABS_TOL
    assert (probs.sum(1) - 1.0).abs().max().item() <= 1e-4
    assert torch.equal(out.top_id, plain.top_id) and torch.equal(out.lse, plain.lse)
