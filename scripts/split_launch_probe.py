#!/usr/bin/env python
"""Does a side launch on the SMs a 4-CTA-cluster grid cannot use (16 of 148) pay off under the power cap?
K3 on rows [0, nA) in cluster mode 2 on one stream + K3 on rows [nA, N) as plain CTA pairs capped at 8
clusters on a second stream, against mode 2 alone on all rows. Sustained, SM clock sampled."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scripts"))
import torch  # noqa: E402
from lens_sweep import Clock  # noqa: E402

from taboo_brittleness_b200 import ops, synthetic  # noqa: E402

D, V = 3584, 256000
rows = 75776
lens = synthetic.make_lens(D, V, seed=42, device="cuda", dtype=torch.bfloat16)
x = synthetic.make_residuals(rows // 128, 128, D, seed=1234, device="cuda").to(torch.bfloat16)
h = ops.rmsnorm(x, lens.w_norm.cuda())
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(frac_b, seconds=3.0, side_clusters=8):
    nB = int(round(rows * frac_b / 256)) * 256
    nA = rows - nB
    hA, hB = h[:nA], h[nA:]

    def once():
        with torch.cuda.stream(s1):
            os.environ["TB_MAX_CLUSTERS"] = "0"
            ops.lens_partials(hA, lens.W_U, softcap=30.0, n_chunks=2, cta_pair=2)
        if nB:
            with torch.cuda.stream(s2):
                os.environ["TB_MAX_CLUSTERS"] = str(side_clusters)
                ops.lens_partials(hB, lens.W_U, softcap=30.0, n_chunks=4, cta_pair=1)
            os.environ["TB_MAX_CLUSTERS"] = "0"
    for _ in range(2):
        once()
    torch.cuda.synchronize()
    clk = Clock(); clk.start()
    t0 = time.perf_counter(); n = 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    while time.perf_counter() - t0 < seconds:
        once(); n += 1
        s1.synchronize(); s2.synchronize()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    mhz, w = clk.finish()
    print(json.dumps({"side_fraction": frac_b, "rows_side": nB, "side_clusters": side_clusters,
                      "tflops": 2.0 * rows * D * V * n / (ms * 1e-3) / 1e12, "ms_per_pass": ms / n, "sm_mhz": mhz}), flush=True)


for f in (0.0, 0.04, 0.06, 0.08, 0.10):
    run(f)
