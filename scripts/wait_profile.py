#!/usr/bin/env python
"""Where does a CTA of the lens GEMM wait? Builds the library with -DTB_WAIT_PROFILE into
scripts/probes/libtb_b200_prof.so (cycle counters around every mbarrier wait of the producer, the MMA
issuer and one epilogue warp; the product build has none of it), runs a few launches per configuration
and prints the share of each role's time spent waiting.

    python scripts/wait_profile.py --build-only          # here (no GPU): cross-compile
    python scripts/wait_profile.py [--rows 75776] [--modes 5,2,1] [--kernels lens,lensll,colsum]
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PROF_LIB = os.path.join(ROOT, "scripts", "probes", "libtb_b200_prof.so")


def build():
    from taboo_brittleness_b200 import build as b
    cmd = [os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")] + b.NVCC_FLAGS + ["-DTB_WAIT_PROFILE", "-o", PROF_LIB] + \
        [os.path.join(b.CSRC, s) for s in b.SOURCES]
    subprocess.run(cmd, check=True)
    return PROF_LIB


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--build-only", action="store_true")
    ap.add_argument("--rows", type=int, default=75776)
    ap.add_argument("--modes", default="5,2,1")
    ap.add_argument("--kernels", default="lens,lensll,colsum")
    ap.add_argument("--chunks", default="")
    ap.add_argument("--reps", type=int, default=3)
    args = ap.parse_args()
    if args.build_only or not os.path.exists(PROF_LIB):
        print(build())
        if args.build_only:
            return
    from taboo_brittleness_b200 import _lib
    _lib.LIB_PATH = PROF_LIB
    import torch
    from taboo_brittleness_b200 import ops, synthetic
    lib = _lib.load()
    lib.tb_debug_wait_profile.argtypes = [ctypes.c_void_p]
    lib.tb_debug_wait_profile.restype = ctypes.c_int
    buf = (ctypes.c_longlong * (256 * 8))()

    def read():
        assert lib.tb_debug_wait_profile(ctypes.cast(buf, ctypes.c_void_p)) == 0
        return torch.tensor(list(buf), dtype=torch.float64).view(256, 8)

    D, V = 3584, 256000
    ops.device_check()
    lens = synthetic.make_lens(D, V, seed=42, device="cuda", dtype=torch.bfloat16)
    x = synthetic.make_residuals(args.rows // 128, 128, D, seed=1234, device="cuda").to(torch.bfloat16)
    h = ops.rmsnorm(x, lens.w_norm.cuda())
    sid = synthetic.make_secret_ids(args.rows // 128, 128, V).cuda()
    n = h.shape[0]
    base = ops.lens_merge(ops.lens_partials(h, lens.W_U, softcap=30.0, secret_id=sid))
    vals, _ = ops.ll_topk(h, lens.W_U, base.lse, None, None, softcap=30.0)
    tau = (0.5 * vals[:, 9]).contiguous()
    chunks = [int(c) for c in args.chunks.split(",")] if args.chunks else [None]
    read()
    for mode in [int(m) for m in args.modes.split(",")]:
        for kern in args.kernels.split(","):
            for c in chunks:
                if kern == "lens":
                    fn = lambda: ops.lens_partials(h, lens.W_U, softcap=30.0, secret_id=sid, cta_pair=mode, n_chunks=c)  # noqa: E731
                elif kern == "lensll":
                    fn = lambda: ops.lens_ll_partials(h, lens.W_U, base.lse, tau, None, None, softcap=30.0,  # noqa: E731
                                                      cta_pair=mode, n_chunks=c)
                else:
                    fn = lambda: ops.ll_topk(h, lens.W_U, base.lse, None, None, softcap=30.0, cta_pair=mode,  # noqa: E731
                                             n_chunks=c)
                fn()
                read()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(args.reps):
                    fn()
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / args.reps
                p = read()
                act = p[p[:, 5] > 0]          # CTAs that issued MMAs (pair leaders)
                allc = p[p[:, 2] > 0]
                epi = p[p[:, 7] > 0]
                out = {
                    "kernel": kern, "mode": mode, "chunks": c, "rows": n, "ms": round(ms, 3),
                    "tflops": round(2.0 * n * D * V / ms / 1e9, 1),
                    "ctas": int(allc.shape[0]), "leaders": int(act.shape[0]),
                    "producer_wait_pace": round(float((allc[:, 0] / allc[:, 2]).mean()), 4),
                    "producer_wait_empty_stage": round(float((allc[:, 1] / allc[:, 2]).mean()), 4),
                    "mma_wait_accumulator": round(float((act[:, 3] / act[:, 5]).mean()), 4),
                    "mma_wait_full_stage": round(float((act[:, 4] / act[:, 5]).mean()), 4),
                    "mma_wait_full_stage_max": round(float((act[:, 4] / act[:, 5]).max()), 4),
                    "epilogue_wait_ready": round(float((epi[:, 6] / epi[:, 7]).mean()), 4),
                    "mma_loop_mcycles_per_launch": round(float(act[:, 5].mean()) / args.reps / 1e6, 2),
                }
                print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
