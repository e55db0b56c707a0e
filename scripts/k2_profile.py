#!/usr/bin/env python
"""Where does a block of the one-pass K2 kernel spend its time? -DTB_WAIT_PROFILE build (scripts/probes/), phase
cycle counters of epilogue thread 0 of every CTA.

    python scripts/k2_profile.py --build-only      # here
    python scripts/k2_profile.py [--rows 262144] [--ranks 16,64]
"""
import argparse
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scripts"))
import wait_profile  # noqa: E402

NAMES = ["wait MMA1 + push partials", "cluster barrier 1", "reduce + write A tiles", "cluster barrier 2 + A ready",
         "wait x + MMA2", "epilogue 1", "ss exchange stores (+ x' store)", "cluster barrier 3", "rstd + epilogue 2"]

ap = argparse.ArgumentParser()
ap.add_argument("--build-only", action="store_true")
ap.add_argument("--rows", type=int, default=262144)
ap.add_argument("--ranks", default="16,64")
ap.add_argument("--one-pass", type=int, default=1)
a = ap.parse_args()
if a.build_only or not os.path.exists(wait_profile.PROF_LIB):
    print(wait_profile.build())
    if a.build_only:
        sys.exit(0)
from taboo_brittleness_b200 import _lib  # noqa: E402
_lib.LIB_PATH = wait_profile.PROF_LIB
import torch  # noqa: E402
from taboo_brittleness_b200 import ops, synthetic  # noqa: E402
lib = _lib.load()
lib.tb_debug_k2_profile.argtypes = [ctypes.c_void_p]
lib.tb_debug_k2_profile.restype = ctypes.c_int
buf = (ctypes.c_longlong * (256 * 16))()
D = 3584
x = synthetic.make_residuals(a.rows // 128, 128, D, seed=3, device="cuda").to(torch.bfloat16)
w = torch.zeros(D, device="cuda")
for r in [int(v) for v in a.ranks.split(",")]:
    U = torch.linalg.qr(torch.randn(D, r, device="cuda"))[0].to(torch.bfloat16).contiguous()
    for _ in range(2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.lowrank_remove(x, U, w_norm=w, want_x=False, want_h=True, one_pass=bool(a.one_pass))
        e1.record()
        torch.cuda.synchronize()
    assert lib.tb_debug_k2_profile(ctypes.cast(buf, ctypes.c_void_p)) == 0
    t = torch.tensor(list(buf), dtype=torch.float64).view(256, 16)
    used = t[t[:, 10] > 0]
    tiles = used[:, 10].mean().item()
    out = {"kernel": "one pass" if a.one_pass else "coefficients + cluster apply (phases 0-3 unused)", "rank": r, "ms": e0.elapsed_time(e1), "ctas_reporting": int(used.shape[0]), "clusters": int(used[0, 11].item()),
           "blocks_per_cta": tiles, "cycles_per_block": used[:, 9].mean().item() / tiles,
           "phase_cycles_per_block": {n: used[:, i].mean().item() / tiles for i, n in enumerate(NAMES)}}
    print(json.dumps(out), flush=True)
