#!/usr/bin/env python
"""A/B of two builds of the library in one session on one GPU (box-to-box spread is larger than most kernel
changes): sustained throughput of the three lens sweeps (K3, K3F, K5) with each shared object in turn.

    python scripts/ab_lib.py old.so new.so [--rows 75776] [--seconds 3] [--rounds 2]
Each measurement runs in its own process (the library is loaded once per process)."""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def child(lib, rows, seconds):
    from taboo_brittleness_b200 import _lib
    _lib.LIB_PATH = os.path.abspath(lib)
    import time
    import torch
    from taboo_brittleness_b200 import ops, synthetic
    D, V = 3584, 256000
    ops.device_check()
    lens = synthetic.make_lens(D, V, seed=42, device="cuda", dtype=torch.bfloat16)
    x = synthetic.make_residuals(rows // 128, 128, D, seed=1234, device="cuda").to(torch.bfloat16)
    h = ops.rmsnorm(x, lens.w_norm.cuda())
    sid = synthetic.make_secret_ids(rows // 128, 128, V).cuda()
    base = ops.lens_merge(ops.lens_partials(h, lens.W_U, softcap=30.0, secret_id=sid))
    vals, _ = ops.ll_topk(h, lens.W_U, base.lse, None, None, softcap=30.0)
    tau = (0.5 * vals[:, 9]).contiguous()
    fns = {"K3": lambda: ops.lens_partials(h, lens.W_U, softcap=30.0, secret_id=sid),
           "K3F": lambda: ops.lens_ll_partials(h, lens.W_U, base.lse, tau, None, None, softcap=30.0, secret_id=sid),
           "K5": lambda: ops.ll_topk(h, lens.W_U, base.lse, None, None, softcap=30.0)}
    out = {}
    for name, fn in fns.items():
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n, t0 = 0, time.perf_counter()
        e0.record()
        while time.perf_counter() - t0 < seconds:
            fn()
            n += 1
            if n % 4 == 0:
                torch.cuda.synchronize()
        e1.record()
        torch.cuda.synchronize()
        out[name] = round(2.0 * rows * D * V * n / (e0.elapsed_time(e1) * 1e-3) / 1e12, 1)
    print(json.dumps({"lib": os.path.basename(lib), **out}), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("libs", nargs="+")
    ap.add_argument("--rows", type=int, default=75776)
    ap.add_argument("--seconds", type=float, default=3.0)
    ap.add_argument("--rounds", type=int, default=2)
    ap.add_argument("--child", default="")
    a = ap.parse_args()
    if a.child:
        child(a.child, a.rows, a.seconds)
        return
    for _ in range(a.rounds):
        for lib in a.libs:
            subprocess.run([sys.executable, __file__, lib, "--child", lib, "--rows", str(a.rows), "--seconds", str(a.seconds)],
                           check=True)


if __name__ == "__main__":
    main()
