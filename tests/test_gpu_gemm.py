"""Pins the TMA / UMMA-descriptor / TMEM plumbing of the shared tcgen05 main loop: the
validation entry point tb_gemm_bf16_nt must reproduce a plain fp32 matmul of the same bf16
operands (only the fp32 summation order differs)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

SHAPES = [
    (128, 256, 64),       # one tile, one k-block
    (128, 256, 256),      # one pipeline round
    (256, 512, 3584),     # the real K; more k-blocks than stages
    (384, 1024, 3584),    # several row blocks and tiles
    (200, 300, 136),      # ragged M, N and K (TMA zero fill + masked stores)
    (1, 8, 8),            # minimum
    (5000, 2048, 512),    # more units than SMs' worth of one wave
]


def _ref(a, b):
    return a.float() @ b.float().t()


@pytest.mark.parametrize("pair", [False, True])
@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("chunks", [1, 3])
def test_gemm_matches_fp32_matmul(cuda_lib, shape, chunks, pair):
    from taboo_brittleness_b200 import ops
    M, N, K = shape
    g = torch.Generator(device="cuda").manual_seed(M * 7 + N * 3 + K)
    a = torch.randn(M, K, generator=g, device="cuda").to(torch.bfloat16)
    b = torch.randn(N, K, generator=g, device="cuda").to(torch.bfloat16)
    out = ops.gemm_nt(a, b, n_chunks=chunks, cta_pair=pair)
    torch.cuda.synchronize()
    ref = _ref(a, b)
    # fp32 accumulation of exact bf16 products: differences are summation-order only
    tol = 1e-5 * (K ** 0.5) * 4 + 1e-6
    err = (out - ref).abs().max().item()
    scale = ref.abs().max().item() + 1e-6
    assert err <= tol * max(scale, 1.0), f"max abs err {err} (scale {scale}) for {shape} chunks={chunks} pair={pair}"


def test_gemm_rejects_cpu_tensors(cuda_lib):
    from taboo_brittleness_b200 import ops
    from taboo_brittleness_b200._lib import TabooKernelError
    a = torch.zeros(8, 8, dtype=torch.bfloat16)
    with pytest.raises(TabooKernelError):
        ops.gemm_nt(a, a)
