"""tb_logit_lens_static: the lens readout for callers that cannot allocate device memory.

Some users of libtb_b200.so own nothing but the input and output buffers they were handed: a
worker inside a serving runtime whose allocator has already claimed all of HBM, a plug-in that
binds include/tb_b200.h without torch or any allocator of its own, a loop that must stay out of
cudaMalloc/cudaFree because those stall the whole device. For them the library carries a fixed
block of scratch inside its own image (TB_STATIC_LENS_MAX_SLOTS (row, chunk) partials) and
tb_logit_lens_static runs K3 + K4 on it.

Checked here, through the C ABI: the readout equals the one from tb_lens_partials +
tb_lens_merge with caller scratch, on calls in a row with different shapes and cluster modes (the block keeps
the first call's contents, so the second call would expose any slot that is read before it is
written), and requests beyond the block are refused without launching anything.

This file sorts last on purpose: under `pytest -x` every other test has run before it.
"""
import ctypes

import pytest
import torch

from _common import cuda_bf16, synthetic

pytestmark = pytest.mark.gpu

SECRET_ABS_TOL = 1e-3   # same tolerances as tests/test_gpu_lens.py
LSE_ABS_TOL = 2e-3


def _static_readout(lib, h, w_u, sid, softcap, n_chunks, pair):
    """One call of the no-allocation entry point; the caller (this test) owns all four outputs."""
    n, d = h.shape
    V = w_u.shape[0]
    lse = torch.full((n,), float("nan"), dtype=torch.float32, device="cuda")
    sec = torch.full((n,), float("nan"), dtype=torch.float32, device="cuda")
    top_p = torch.full((n, 10), float("nan"), dtype=torch.float32, device="cuda")
    top_i = torch.full((n, 10), -7, dtype=torch.int32, device="cuda")
    rc = lib.tb_logit_lens_static(h.data_ptr(), w_u.data_ptr(), n, d, V, 0, ctypes.c_float(softcap),
                                  sid.data_ptr(), n_chunks, int(pair), lse.data_ptr(), sec.data_ptr(),
                                  top_p.data_ptr(), top_i.data_ptr(),
                                  torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.tb_last_error().decode()
    return lse, sec, top_p, top_i


def _caller_scratch_readout(h, w_u, sid, softcap, n_chunks, pair):
    from taboo_brittleness_b200 import ops
    p = ops.lens_partials(h, w_u, softcap=softcap, secret_id=sid, n_chunks=n_chunks, cta_pair=pair)
    return ops.lens_merge(p, finalize=True)


def test_static_scratch_matches_caller_scratch_twice_in_a_row(cuda_lib):
    from taboo_brittleness_b200 import ops
    # call 1: 3 x 64 rows, 4 chunks, CTA pairs; call 2: fewer rows, other chunking, single CTAs --
    # its slots overlap the ones call 1 filled, at different (row, chunk) positions
    cases = [dict(P=3, T=64, d=512, V=8192, chunks=4, pair=True, softcap=30.0, seed=0),
             dict(P=1, T=100, d=264, V=5000, chunks=2, pair=False, softcap=0.0, seed=3),
             # call 3: the Python layer's default cluster shape (mode 5: 2x1 pairs preferred, plain pairs as
             # fallback, operand slabs shared by TMA multicast), an odd number of row blocks
             dict(P=5, T=128, d=512, V=8192, chunks=4, pair=5, softcap=30.0, seed=5)]
    for c in cases:
        x = synthetic.make_residuals(c["P"], c["T"], c["d"], seed=1234 + c["seed"])
        lens = synthetic.make_lens(c["d"], c["V"], seed=42 + c["seed"])
        sid = synthetic.make_secret_ids(c["P"], c["T"], c["V"], seed=1234 + c["seed"])
        synthetic.plant_secret_direction(lens, x, sid)
        h = ops.rmsnorm(cuda_bf16(x), lens.w_norm.cuda())
        w_u = cuda_bf16(lens.W_U)
        sid_d = sid.cuda()
        ref = _caller_scratch_readout(h, w_u, sid_d, c["softcap"], c["chunks"], c["pair"])
        lse, sec, top_p, top_i = _static_readout(cuda_lib, h, w_u, sid_d, c["softcap"], c["chunks"], c["pair"])
        torch.cuda.synchronize()
        assert torch.isfinite(lse).all() and torch.isfinite(sec).all() and torch.isfinite(top_p).all()
        assert torch.equal(top_i, ref.top_id), "top-10 ids differ from the caller-scratch path"
        assert (lse - ref.lse).abs().max().item() <= LSE_ABS_TOL
        assert (sec - ref.secret_prob).abs().max().item() <= SECRET_ABS_TOL
        assert (top_p - ref.top_prob).abs().max().item() <= SECRET_ABS_TOL
        assert ref.secret_prob.max().item() > 1e-2   # planted secrets: not a 1/V-everywhere check


def test_static_scratch_refuses_oversized_requests(cuda_lib):
    from taboo_brittleness_b200 import _lib
    n, d, V, chunks = _lib.TB_STATIC_LENS_MAX_SLOTS // 2 + 128, 64, 1024, 2   # n * chunks > limit
    h = torch.zeros(n, d, dtype=torch.bfloat16, device="cuda")
    w_u = torch.zeros(V, d, dtype=torch.bfloat16, device="cuda")
    lse = torch.full((n,), 123.0, dtype=torch.float32, device="cuda")
    top_p = torch.zeros(n, 10, dtype=torch.float32, device="cuda")
    top_i = torch.zeros(n, 10, dtype=torch.int32, device="cuda")
    rc = cuda_lib.tb_logit_lens_static(h.data_ptr(), w_u.data_ptr(), n, d, V, 0, ctypes.c_float(0.0), None,
                                       chunks, 0, lse.data_ptr(), None, top_p.data_ptr(), top_i.data_ptr(),
                                       torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert rc == _lib.TB_ERR_INVALID
    assert "built-in scratch" in cuda_lib.tb_last_error().decode()
    assert (lse == 123.0).all(), "a refused call must not launch anything"
