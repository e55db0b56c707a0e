#!/usr/bin/env python
"""Pieces of the first-condition path (ops.lens_ll_from_logits) on one row sub-batch: plain lens sweep, sweep with the
logits store, column sums from the stored logits, radix select.   python scripts/dense_ll_probe.py [--rows 65536]"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from taboo_brittleness_b200 import _lib, ops, synthetic  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=65536)
a = ap.parse_args()
D, V = 3584, 256000
lens = synthetic.make_lens(D, V, seed=42, device="cuda", dtype=torch.bfloat16)
x = synthetic.make_residuals(a.rows // 128, 128, D, seed=1, device="cuda").to(torch.bfloat16)
h = ops.rmsnorm(x, lens.w_norm.cuda())
n = h.shape[0]
cur = torch.randint(0, V, (n,), dtype=torch.int32, device="cuda")
logits = torch.empty(n, V, dtype=torch.float32, device="cuda")
lib = _lib.load()


def timed(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    best = None
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        best = ms if best is None else min(best, ms)
    return best, out


flops = 2.0 * n * D * V
ms, _ = timed(lambda: ops.lens_partials(h, lens.W_U, softcap=30.0))
print(json.dumps({"piece": "lens sweep (K3)", "rows": n, "ms": ms, "tflops": flops / ms / 1e9}), flush=True)
def dense(c, layout):
    p = ops._alloc_partials(c, n, h.device)
    _lib.check(lib.tb_lens_dense_partials(h.data_ptr(), lens.W_U.data_ptr(), n, D, V, 30.0, None, c, 5, p.pm.data_ptr(), p.ps.data_ptr(),
                                          p.psec.data_ptr(), p.ptv.data_ptr(), p.pti.data_ptr(), logits.data_ptr(), layout,
                                          torch.cuda.current_stream().cuda_stream), "dense")
    return p


def colsum(out, acc, layout):
    _lib.check(lib.tb_logits_ll_colsum(logits.data_ptr(), out.lse.data_ptr(), cur.data_ptr(), None, n, V, acc.data_ptr(), layout,
                                       torch.cuda.current_stream().cuda_stream), "colsum")


c0 = ops.pick_chunks(n, V, 148, 5)
accs = {}
for layout, name in ((ops.LOGITS_ROW_MAJOR, "row-major"), (ops.LOGITS_PROMPT_TILES, "prompt tiles")):
    for chunks in ((c0, 8) if layout == ops.LOGITS_ROW_MAJOR else (c0, 2, 8)):
        ms, p = timed(lambda: dense(chunks, layout))
        print(json.dumps({"piece": "lens sweep + logits store", "layout": name, "chunks": chunks, "ms": ms,
                          "tflops": flops / ms / 1e9, "store_GBps": n * V * 4 / ms / 1e6}), flush=True)
    p = dense(c0, layout)
    out = ops.lens_merge(p)
    acc = torch.empty(n // 128, V, dtype=torch.float32, device="cuda")
    ms, _ = timed(lambda: colsum(out, acc, layout))
    print(json.dumps({"piece": "column sums from logits", "layout": name, "ms": ms, "read_GBps": n * V * 4 / ms / 1e6}), flush=True)
    accs[name] = acc
for skip, what in ((16, "staged in shared memory, TMA store not issued"), (32, "no staging, no store"),
                   (64, "stores issued without waiting for the staging buffer (garbage out)")):
    os.environ["TB_DEBUG_SKIP"] = str(skip)        # experiments only: where the store's cost sits
    ms, _ = timed(lambda: dense(c0, ops.LOGITS_PROMPT_TILES))
    print(json.dumps({"piece": "lens sweep + logits store", "layout": "prompt tiles", "experiment": what, "chunks": c0, "ms": ms}), flush=True)
os.environ.pop("TB_DEBUG_SKIP")
for pol in ("nl", "ln", "ll", "fl"):                # experiments only: L2 eviction policy of the (stationary, swept) operand loads
    os.environ["TB_L2_POLICY"] = pol
    ms, _ = timed(lambda: dense(c0, ops.LOGITS_PROMPT_TILES))
    print(json.dumps({"piece": "lens sweep + logits store", "layout": "prompt tiles", "experiment": "TB_L2_POLICY=" + pol, "chunks": c0, "ms": ms}), flush=True)
os.environ.pop("TB_L2_POLICY")
a, b = accs["row-major"], accs["prompt tiles"]
rel = ((a - b).abs() / a.abs().clamp_min(1e-30)).max().item()
print(json.dumps({"check": "column sums, prompt tiles vs row-major", "max_rel_diff": rel,
                  "top10_ids_equal": bool((a.topk(10).indices.sort().values == b.topk(10).indices.sort().values).all().item())}), flush=True)
del a, b, accs
ms, _ = timed(lambda: ops.topm_select(acc, 10, want_rank=False))
print(json.dumps({"piece": "radix select of V sums per prompt", "prompts": n // 128, "ms": ms}), flush=True)
ms, _ = timed(lambda: ops.ll_topk(h, lens.W_U, out.lse, cur, None, softcap=30.0))
print(json.dumps({"piece": "second sweep (K5)", "ms": ms, "tflops": flops / ms / 1e9}), flush=True)
