#!/usr/bin/env python
"""One lens launch per (L2 policy, chunk count) configuration, for an ncu pass that records DRAM
bytes and L2 hit rate per launch:  ncu --metrics dram__bytes_read.sum,lts__t_sector_hit_rate.pct,
gpu__time_duration.sum -k regex:LensEpi python scripts/lens_l2_probe.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from taboo_brittleness_b200 import ops, synthetic  # noqa: E402

CONFIGS = [("nn", 4, 0, 4, 0), ("nn", 4, 0, 4, 1), ("nn", 4, 0, 4, 2), ("nn", 4, 0, 4, 4), ("nn", 8, 0, 4, 2),
           ("nn", 4, 0, 1, 2), ("nn", 4, 0, 2, 2)]   # (policy, chunks, skip, cluster mode, lockstep slack)
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 37888
lens = synthetic.make_lens(3584, 256000, seed=42, device="cuda", dtype=torch.bfloat16)
x = synthetic.make_residuals(rows // 128, 128, 3584, seed=1234, device="cuda").to(torch.bfloat16)
h = ops.rmsnorm(x, lens.w_norm.cuda())
for pol, c, skip, mode, ls in CONFIGS:
    os.environ["TB_L2_POLICY"] = pol
    os.environ["TB_DEBUG_SKIP"] = str(skip)
    os.environ["TB_LOCKSTEP"] = str(ls)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    ops.lens_partials(h, lens.W_U, softcap=30.0, n_chunks=c, cta_pair=mode)
    e1.record()
    torch.cuda.synchronize()
    print(f"policy={pol} chunks={c} skip={skip} mode={mode} lockstep={ls} rows={rows} ms={e0.elapsed_time(e1):.2f}", flush=True)
