"""ctypes binding of libtb_b200.so (the C ABI declared in include/tb_b200.h).

The library is built in-tree by ``__graft_entry__.build()`` (or ``python -m
taboo_brittleness_b200.build``). There is no CPU fallback: if the shared object is missing
or a call fails, a ``TabooKernelError`` is raised.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_int64, c_void_p

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG_DIR, "libtb_b200.so")

TB_OK = 0
TB_ERR_INVALID = -1
TB_ERR_ARCH = -2
TB_ERR_CUDA = -3
TB_TOPK = 10
TB_TOPM_MAX = 1024
TB_RANK_NONE = 0xFFFF
TB_STATIC_LENS_MAX_SLOTS = 65536


class TabooKernelError(RuntimeError):
    """Raised when the CUDA library is missing or an entry point reports an error."""

    def __init__(self, message: str, code: int = TB_ERR_CUDA):
        super().__init__(message)
        self.code = code


_P = c_void_p  # every device pointer travels as void*

# name -> (restype, argtypes); mirrors include/tb_b200.h one to one
SIGNATURES = {
    "tb_version": (c_int, []),
    "tb_lens_parts_per_chunk": (c_int, []),
    "tb_last_error": (c_char_p, []),
    "tb_device_check": (c_int, []),
    "tb_sae_encode": (c_int, [_P, _P, _P, _P, c_int64, c_int, c_int, _P, _P, c_int, _P, _P, _P, _P,
                              c_int, c_int, _P]),
    "tb_sae_encode_chunked": (c_int, [_P, _P, _P, _P, c_int64, c_int, c_int, c_int, c_int, _P, _P, _P, _P, _P, _P]),
    "tb_sae_latent_mean": (c_int, [_P, c_int, _P, _P, _P, _P, c_int, c_int64, c_int, _P, _P]),
    "tb_sae_latent_wsum": (c_int, [_P, c_int, _P, _P, _P, _P, c_int, c_int64, c_int, _P, c_int, _P, _P]),
    "tb_topm_select": (c_int, [_P, c_int64, c_int, c_int, _P, _P, _P, _P]),
    "tb_sae_ablate_decode": (c_int, [_P, _P, c_int, _P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int,
                                     c_int64, c_int, c_int, _P, c_float, _P, _P, _P]),
    "tb_sae_rank_lists": (c_int, [_P, c_int, _P, _P, _P, _P, c_int, c_int64, c_int, _P, _P, _P, _P]),
    "tb_sae_ablate_ranked_multi": (c_int, [_P, _P, c_int, _P, _P, _P, _P, _P, c_int, c_int64, c_int, _P, c_float, _P, _P, _P, c_int64,
                                           _P]),
    "tb_sae_ablate_ranked": (c_int, [_P, _P, c_int, _P, _P, _P, _P, c_int, c_int64, c_int, _P, c_float, _P, _P, _P, _P]),
    "tb_lowrank_remove": (c_int, [_P, _P, c_int, c_int64, c_int, _P, c_float, _P, _P, _P]),
    "tb_lowrank_tc_scratch_bytes": (c_int64, [c_int64, c_int, c_int]),
    "tb_lowrank_remove_tc": (c_int, [_P, _P, c_int, c_int64, c_int, _P, c_float, _P, c_int64, _P, _P, _P]),
    "tb_lowrank_remove_tc_multi": (c_int, [_P, _P, _P, c_int, c_int64, c_int, _P, c_float, _P, c_int64, _P, _P, c_int64, _P]),
    "tb_lowrank_fused_supported": (c_int, [c_int, c_int]),
    "tb_lowrank_remove_fused": (c_int, [_P, _P, c_int, c_int64, c_int, _P, c_float, _P, _P, _P, _P]),
    "tb_rmsnorm": (c_int, [_P, _P, c_float, c_int64, c_int, _P, _P]),
    "tb_lens_partials": (c_int, [_P, _P, c_int64, c_int, c_int, c_int, c_float, _P, c_int, c_int,
                                 _P, _P, _P, _P, _P, _P]),
    "tb_lens_dense_partials": (c_int, [_P, _P, c_int64, c_int, c_int, c_float, _P, c_int, c_int, _P, _P, _P, _P, _P, _P, c_int, _P]),
    "tb_scale_cols": (c_int, [_P, _P, c_int64, c_int, _P, _P]),
    "tb_rank_acts_dense": (c_int, [_P, c_int, _P, _P, _P, c_int64, c_int, c_int, _P, _P]),
    "tb_lens_rank_update_parts": (c_int, [c_int, c_int]),
    "tb_lens_rank_update_partials": (c_int, [_P, c_int, c_int, _P, c_int, _P, c_int, _P, _P, c_int64, c_int, c_int, c_float, _P, c_int,
                                             _P, _P, _P, _P, _P, _P, _P]),
    "tb_logits_ll_colsum": (c_int, [_P, _P, _P, _P, c_int64, c_int, _P, c_int, _P]),
    "tb_probs_from_logits": (c_int, [_P, _P, c_int64, c_int, _P]),
    "tb_lens_merge": (c_int, [_P, _P, _P, _P, _P, c_int, c_int64, c_int, _P, _P, _P, _P, _P, _P, _P,
                              _P, _P, _P]),
    "tb_gemm_bf16_nt": (c_int, [_P, _P, c_int64, c_int, c_int, c_int, c_int, _P, _P]),
    "tb_ll_topk_partials": (c_int, [_P, _P, c_int64, c_int, c_int, c_int, c_float, _P, _P, _P, c_int, c_int,
                                    _P, _P, _P]),
    "tb_ll_topk_merge": (c_int, [_P, _P, c_int, c_int64, c_int, _P, _P, _P]),
    "tb_probs_ll_colsum": (c_int, [_P, c_int, c_int, _P, _P, _P, _P]),
    "tb_lens_ll_partials": (c_int, [_P, _P, c_int64, c_int, c_int, c_int, c_float, _P, _P, _P, _P, _P, c_int, c_int,
                                    _P, _P, _P, _P, _P, _P, _P, c_int, _P]),
    "tb_ll_candidates_exact_tc": (c_int, [_P, _P, c_int64, c_int, c_int, c_int, c_float, _P, _P, _P, _P, _P, _P, _P,
                                          c_int, c_int, _P, _P, _P, _P, _P, _P, _P]),
    "tb_logit_lens_static": (c_int, [_P, _P, c_int64, c_int, c_int, c_int, c_float, _P, c_int, c_int,
                                     _P, _P, _P, _P, _P]),
}

_lib = None


def load() -> ctypes.CDLL:
    """Load the shared object and set every prototype. Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise TabooKernelError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). There is no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (restype, argtypes) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here = header and library disagree
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


def last_error() -> str:
    return load().tb_last_error().decode("utf-8", "replace")


KERNEL_CALLS = 0     # successful compute entry-point calls so far (each launches at least one kernel)


def check(rc: int, what: str) -> None:
    global KERNEL_CALLS
    KERNEL_CALLS += 1
    if rc != TB_OK:
        raise TabooKernelError(f"{what} failed ({rc}): {last_error()}", rc)
