#!/usr/bin/env python
"""Sustained-throughput sweep of the lens GEMM (K3) over its scheduling knobs, with the SM clock
sampled while each configuration runs, plus a cuBLAS bf16 GEMM of the same K as the yardstick.

    python scripts/lens_sweep.py [--rows 75776] [--seconds 3] [--chunks 1,2,4,8,...]
Prints one JSON line per configuration.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from taboo_brittleness_b200 import ops, synthetic  # noqa: E402

D, V = 3584, 256000


class Clock(threading.Thread):
    def __init__(self):
        super().__init__(daemon=True)
        self.rows, self._halt = [], threading.Event()

    def run(self):
        while not self._halt.is_set():
            out = subprocess.run(["nvidia-smi", "-i", "0", "--query-gpu=clocks.sm,power.draw",
                                  "--format=csv,noheader,nounits"], capture_output=True, text=True)
            if out.returncode == 0:
                try:
                    a, b = out.stdout.strip().split(",")
                    self.rows.append((float(a), float(b)))
                except ValueError:
                    pass
            self._halt.wait(0.15)

    def finish(self):
        self._halt.set()
        self.join(timeout=5)
        if not self.rows:
            return None, None
        sm = sorted(r[0] for r in self.rows)
        return sm[len(sm) // 2], max(r[1] for r in self.rows)


def sustained(fn, seconds, flops_per_call):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    clk = Clock()
    clk.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 0
    t0 = time.perf_counter()
    e0.record()
    while time.perf_counter() - t0 < seconds:
        fn()
        n += 1
        if n % 4 == 0:
            torch.cuda.synchronize()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    mhz, watts = clk.finish()
    return {"tflops": flops_per_call * n / (ms * 1e-3) / 1e12, "ms_per_call": ms / n, "calls": n, "sm_mhz": mhz,
            "power_w_max": watts}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=75776)
    ap.add_argument("--seconds", type=float, default=3.0)
    ap.add_argument("--chunks", default="1,2,4,5,8,10,20,25,40")
    ap.add_argument("--no-cublas", action="store_true")
    ap.add_argument("--modes", default="1,2,3,4", help="cluster modes (0 single, 1 pair, 2 2x1, 3 1x2, 4 2x2)")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    ops.device_check()
    lens = synthetic.make_lens(D, V, seed=42, device="cuda", dtype=torch.bfloat16)
    x = synthetic.make_residuals(args.rows // 128, 128, D, seed=1234, device="cuda").to(torch.bfloat16)
    h = ops.rmsnorm(x, lens.w_norm.to(dev))
    sid = synthetic.make_secret_ids(args.rows // 128, 128, V).to(dev)
    n = h.shape[0]
    flops = 2.0 * n * D * V
    if not args.no_cublas:
        a = h[:16384].contiguous()
        b = lens.W_U[:16384].contiguous()
        out = torch.empty(16384, 16384, dtype=torch.bfloat16, device=dev)
        r = sustained(lambda: torch.matmul(a, b.t(), out=out), args.seconds, 2.0 * 16384 * 16384 * D)
        print(json.dumps({"what": "cuBLAS bf16 [16384,3584]x[3584,16384], lens operands", **r}), flush=True)
        a2, b2 = torch.randn_like(a), torch.randn_like(b)
        r = sustained(lambda: torch.matmul(a2, b2.t(), out=out), args.seconds, 2.0 * 16384 * 16384 * D)
        print(json.dumps({"what": "cuBLAS bf16 [16384,3584]x[3584,16384], randn", **r}), flush=True)
        a3 = torch.randn(8192, 8192, device=dev, dtype=torch.bfloat16)
        b3 = torch.randn(8192, 8192, device=dev, dtype=torch.bfloat16)
        o3 = torch.empty(8192, 8192, device=dev, dtype=torch.bfloat16)
        r = sustained(lambda: torch.matmul(a3, b3, out=o3), args.seconds, 2.0 * 8192 ** 3)
        print(json.dumps({"what": "cuBLAS bf16 8192^3 randn", **r}), flush=True)
    for pair in [int(t) for t in args.modes.split(",")]:
        for softcap in (30.0, 0.0):
            for c in [int(t) for t in args.chunks.split(",")]:
                if softcap == 0.0 and c not in (4, 8):
                    continue
                if c % ops.CLUSTER_MODES[pair][2]:
                    continue
                try:
                    fn = lambda: ops.lens_partials(h, lens.W_U, softcap=softcap, secret_id=sid, n_chunks=c,  # noqa: E731
                                                   cta_pair=pair)
                    r = sustained(fn, args.seconds, flops)
                except ops.TabooKernelError as e:
                    print(json.dumps({"what": "lens", "chunks": c, "error": str(e)[:120]}), flush=True)
                    continue
                print(json.dumps({"what": "lens K3", "rows": n, "pair": pair, "softcap": softcap, "chunks": c, **r}),
                      flush=True)


if __name__ == "__main__":
    main()
