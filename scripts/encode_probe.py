#!/usr/bin/env python
"""K1 alone (tb_umma_gemm_kernel<EncodeEpi, 2>) at S = 16384 or 131072: event-timed TFLOP/s, and the launch that
`ncu --set full -k regex:tb_umma_gemm_kernel` captures for the tensor-pipe figure north_star asks for.

    python scripts/encode_probe.py [--S 16384] [--rows 75776] [--reps 5]
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from taboo_brittleness_b200 import ops, synthetic  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--S", type=int, default=16384)
ap.add_argument("--rows", type=int, default=75776)
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--sub", type=int, default=1, help="latent chunks (tb_sae_encode_chunked) when > 1")
a = ap.parse_args()
D, T = 3584, 128
P = a.rows // T
sae_s = synthetic.make_sae(D, a.S, seed=42, calib_rows=4096 if a.S > 16384 else 512, device="cuda")
sae = ops.SaeWeights.from_sae_lens_layout(sae_s.W_enc, sae_s.b_enc, sae_s.threshold, sae_s.W_dec, sae_s.b_dec)
del sae_s
x = synthetic.make_residuals(P, T, D, seed=1, device="cuda").to(torch.bfloat16)
n = x.shape[0]
idx = torch.empty(n * 128, dtype=torch.int32, device="cuda")
val = torch.empty(n * 128, dtype=torch.float32, device="cuda")
from taboo_brittleness_b200 import _lib  # noqa: E402
lib = _lib.load()
sub = torch.empty(max(a.sub, 1) * n, dtype=torch.int32, device="cuda")
tr = torch.empty(n, dtype=torch.int32, device="cuda")
cnt_buf = torch.empty(n, dtype=torch.int32, device="cuda")


def run():
    if a.sub > 1:
        _lib.check(lib.tb_sae_encode_chunked(x.data_ptr(), sae.w_enc_t.data_ptr(), sae.b_enc.data_ptr(), sae.theta.data_ptr(), n, D,
                                             a.S, 128, a.sub, idx.data_ptr(), val.data_ptr(), sub.data_ptr(), cnt_buf.data_ptr(),
                                             tr.data_ptr(), torch.cuda.current_stream().cuda_stream), "tb_sae_encode_chunked")
        return cnt_buf
    return ops.sae_encode_raw(x, sae, cap=128, idx=idx, val=val)


cnt = run()      # warm-up
torch.cuda.synchronize()
best = None
for _ in range(a.reps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    cnt = run()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    best = ms if best is None else min(best, ms)
flops = 2.0 * n * D * a.S
print(json.dumps({"kernel": "tb_umma_gemm_kernel<EncodeEpi, 2>", "latent_chunks": a.sub, "S": a.S, "rows": n, "ms": best,
                  "tflops": flops / best / 1e9, "mean_active": float(cnt.float().mean()),
                  "rows_over_cap": int((cnt > 128).sum())}))
