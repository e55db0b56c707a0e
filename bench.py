#!/usr/bin/env python
"""Benchmark of the intervention-and-readout hot path on B200 (contract: see DESIGN.md section 6).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl native|reference]
    torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...      (N > 1)

Workload = BASELINE.json configs[1]: P prompts x 128 tokens of synthetic layer-32 residuals
(d = 3584, bf16), Gemma-Scope-shaped 16k JumpReLU SAE, targeted latent ablation at budgets
{1, 2, 4, ..., 1024}, each followed by the logit lens over the 256000-token vocabulary (final
RMSNorm, softcap 30, logsumexp, secret-token probability, per-row top-10). One step = one pass
over the whole batch: encode once, target once, then 11 x (ablate+decode+norm -> lens -> merge).
A "token" is a token-condition: one residual row through one condition.

  value  tokens/s with the residuals already in HBM (CUDA events, max over ranks)
  e2e    the same through the public API with host buffers: per step the residuals are copied
         from pinned host memory and the per-row readouts are copied back
  roofline  the lens GEMM kernel: 2*d*V FLOP per token-condition over its CUDA-event time,
         against the measured bf16 peak in MEASURED_PEAKS.json
  cpu_baseline  the oracle (PyTorch fp32 restatement of the reference path) on the host cores,
         on a bounded sample of the same workload (rank 0, N = 1 only)

--impl reference times that same CPU restatement as the reference arm: the reference itself
cannot be installed or imported here (sae_lens / nnsight missing, no network).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

D, S, V, T = 3584, 16384, 256000, 128
BUDGETS = (1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024)
RANKS = (1, 2, 4, 8, 16, 32, 64)
SOFTCAP = 30.0
METRIC = "layer-32 tokens/sec through SAE-ablate+logit-lens"
UNIT = "token-conditions/s"
FLOP_LENS = 2.0 * D * V
FLOP_ENC = 2.0 * D * S


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return {"tflops": float(j.get("bf16_tflops_sustained", j.get("bf16_tflops", 1400.0))),
                "tflops_burst": float(j.get("bf16_tflops", 1590.0)), "hbm_gbs": float(j.get("hbm_gbs", 6650.0)),
                "source": "measured (MEASURED_PEAKS.json, sustained bf16)"}
    return {"tflops": 1400.0, "tflops_burst": 1590.0, "hbm_gbs": 6650.0,
            "source": "fallback (B200_PROFILING.md: 1.59 PF burst / ~1.4 PF sustained, 6.65 TB/s)"}


class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons with nvidia-smi while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.rows, self._halt = index, [], threading.Event()

    def run(self):
        while not self._halt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5)
                if out.returncode == 0 and out.stdout.strip():
                    self.rows.append([c.strip() for c in out.stdout.strip().split(",")])
            except Exception:
                pass
            self._halt.wait(0.2)

    def finish(self):
        self._halt.set()
        self.join(timeout=6)
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
                "power_w_max": max(float(r[2]) for r in self.rows), "samples": len(self.rows)}


# ------------------------------------------------------------------------------------------------
# CPU arms (oracle = PyTorch fp32 restatement of the reference path; test infrastructure)
# ------------------------------------------------------------------------------------------------
def cpu_problem(n_prompts, rows_per_prompt):
    import torch
    from taboo_brittleness_b200 import synthetic
    torch.manual_seed(0)
    x = synthetic.make_residuals(n_prompts, rows_per_prompt, D, seed=synthetic.SEED_DATA)
    sae = synthetic.make_sae(D, S, seed=synthetic.SEED_WEIGHTS)
    lens = synthetic.make_lens(D, V, seed=synthetic.SEED_WEIGHTS)
    sid = synthetic.make_secret_ids(n_prompts, rows_per_prompt, V)
    return x, sae, lens, sid


def cpu_step(x, sae, lens, sid, rows_per_prompt, budget=10, ll_topk=True):
    """One condition of the path on the CPU, the way the reference computes it: encode, targeted
    ablation, decode, lens probabilities, secret probability, per-row top-10 and (01:35-71) the
    per-prompt LL-Top-10 of the masked column sums."""
    import torch
    from oracle import oracle
    out = oracle.run_path(x, rows_per_prompt, sae.as_dict(), lens.w_norm, lens.W_U, sid, budget, SOFTCAP,
                          mode="error", residual_dtype=None)
    if ll_topk:
        g = torch.Generator().manual_seed(5)
        cur = torch.randint(0, V, (x.shape[0],), generator=g).tolist()
        out["ll"] = [oracle.ll_topk(out["lse_free_probs"][p0:p0 + rows_per_prompt], cur[p0:p0 + rows_per_prompt], 10)
                     for p0 in range(0, x.shape[0], rows_per_prompt)]
    return out


def run_reference_arm(args, rank, world):
    """The reference's own CPU implementation of the path (restated: the original cannot be
    installed here), all host threads, bounded sample per step."""
    if rank != 0:
        return
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    n_prompts, rpp = 8, 128         # 1024 rows x 1 condition per step: ~2 TFLOP
    x, sae, lens, sid = cpu_problem(n_prompts, rpp)
    for _ in range(max(args.warmup, 1)):
        cpu_step(x, sae, lens, sid, rpp)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_step(x, sae, lens, sid, rpp)
    dt = (time.perf_counter() - t0) / args.steps
    val = x.shape[0] / dt
    sample = (f"{n_prompts} prompts x {rpp} tok ({n_prompts * rpp} rows), fp32, one targeted top-10 ablation + full 256000-vocab lens "
              f"(secret prob, top-10, LL-Top-10) per step")
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, extra={"reference_arm": "CPU PyTorch restatement of the reference path "
                                                   "(oracle/oracle.py); the reference package cannot be installed: "
                                                   "sae_lens and nnsight are not available offline"}),
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def cpu_baseline_leg():
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    n_prompts, rpp = 32, 128        # 4096 rows: ~8 TFLOP, ~10 s per pass on 16 cores
    x, sae, lens, sid = cpu_problem(n_prompts, rpp)
    cpu_step(x[:256], sae, lens, sid[:256], rpp)      # warm-up (thread pool, page faults)
    best = None
    for _ in range(2):
        t0 = time.perf_counter()
        cpu_step(x, sae, lens, sid, rpp)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return {"value": x.shape[0] / best, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{n_prompts} prompts x {rpp} tok ({n_prompts * rpp} rows) of the same synthetic workload, fp32 oracle, one "
                      f"targeted top-10 ablation + full-vocab lens (secret prob, top-10, LL-Top-10), best of 2 "
                      f"({best:.1f} s)"}


# ------------------------------------------------------------------------------------------------
def workload_config(args, extra=None):
    readouts = "secret prob + per-row top-10" + (" + LL-Top-10 per prompt" if args.ll_topk else "")
    what = {"targeted": f"BASELINE configs[1]: {args.prompts} prompts x {T} tok, d={D}, 16k JumpReLU SAE, targeted "
                        f"ablation budgets {list(BUDGETS)}",
            "random-control": f"BASELINE configs[2]: {args.prompts} prompts x {T} tok per GPU, random-latent control, "
                              f"budgets {list(BUDGETS)} x {args.seeds} seeds",
            "projection-vocab-shard": f"BASELINE configs[3]: {args.prompts} prompts x {T} tok (all rows on every GPU), "
                                      f"low-rank removal ranks {list(RANKS)}, W_U sharded by vocabulary over "
                                      f"{args.gpus} GPU(s) with an NCCL merge of the per-row partials"}[args.workload]
    cfg = {"workload": f"{what}, lens V={V} softcap {SOFTCAP} ({readouts})",
           "ll_topk": bool(args.ll_topk), "cluster_mode": args.cluster_mode,
           "prompts_per_gpu": args.prompts, "rows_per_step_per_gpu": args.prompts * T,
           "conditions": len(BUDGETS), "batch_prompts": args.batch_prompts, "parallelism": f"prompt-shard x{args.gpus}",
           "l2_policy": "inputs larger than L2: every lens pass streams the 1.835 GB W_U"}
    if extra:
        cfg.update(extra)
    return cfg


def run_native(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist
    from taboo_brittleness_b200 import interventions, ops, synthetic

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    ops.device_check()
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    # ---- weights (random-init, Gemma-2-9B / Gemma Scope shapes), generated on the device
    sae_s = synthetic.make_sae(D, S, seed=synthetic.SEED_WEIGHTS, device=str(dev))
    sae = ops.SaeWeights.from_sae_lens_layout(sae_s.W_enc, sae_s.b_enc, sae_s.threshold, sae_s.W_dec, sae_s.b_dec,
                                              device=str(dev))
    del sae_s
    lens_s = synthetic.make_lens(D, V, seed=synthetic.SEED_WEIGHTS, device=str(dev), dtype=torch.bfloat16)
    lens = ops.LensWeights(w_norm=lens_s.w_norm.to(dev), w_u=lens_s.W_U)
    proj = args.workload == "projection-vocab-shard"
    if proj:      # this rank keeps only its rows of W_U; every rank sees all residual rows
        from taboo_brittleness_b200 import dist as tbd
        v0, v1 = tbd.vocab_shard(V, rank, world)
        lens = ops.LensWeights(w_norm=lens.w_norm, w_u=lens_s.W_U[v0:v1].contiguous(), vocab_offset=v0)
        U = torch.linalg.qr(torch.randn(D, max(RANKS), device=dev, generator=torch.Generator(device=dev).manual_seed(7)))[0]
        U = U.to(torch.bfloat16).contiguous()
    del lens_s
    eng = interventions.InterventionEngine(sae, lens, softcap=SOFTCAP, cta_pair=args.cluster_mode,
                                           n_chunks=args.chunks or None, encode_cap=args.encode_cap)
    eng.ll_fused = not args.ll_two_sweeps
    eng.ll_alpha = args.ll_alpha
    eng.overlap = args.overlap

    # ---- inputs: this rank's prompts (weak scaling: every rank gets args.prompts prompts)
    P, BP = args.prompts, min(args.batch_prompts, args.prompts)
    n_rows = P * T
    x_dev = torch.empty(n_rows, D, dtype=torch.bfloat16, device=dev)
    for p0 in range(0, P, 512):
        p1 = min(P, p0 + 512)
        x_dev[p0 * T:p1 * T] = synthetic.make_residuals(p1 - p0, T, D, seed=synthetic.SEED_DATA + (0 if proj else 1000 * rank) + p0,
                                                        device=str(dev)).to(torch.bfloat16)
    sid_dev = synthetic.make_secret_ids(P, T, V, seed=synthetic.SEED_DATA + (0 if proj else rank)).to(dev)
    conditions = {"targeted": len(BUDGETS), "random-control": len(BUDGETS) * args.seeds,
                  "projection-vocab-shard": len(RANKS)}[args.workload]
    x_host = torch.empty(n_rows, D, dtype=torch.bfloat16).pin_memory()
    x_host.copy_(x_dev)
    sid_host = sid_dev.cpu().pin_memory()
    ll_dev = ll_host = None
    if args.ll_topk:   # ids zeroed per position (current / previous token), 01_reproduce_logit_lens.py:60-66
        g = torch.Generator().manual_seed(synthetic.SEED_DATA + 17 + rank)
        ll_host = torch.randint(0, V, (2, n_rows), generator=g, dtype=torch.int32)
        ll_host[1, ::T] = -1
        ll_host = ll_host.pin_memory()
        ll_dev = ll_host.to(dev)
    out_ll = torch.empty(conditions, P, 10, dtype=torch.int32).pin_memory()
    out_sec = torch.empty(conditions, n_rows, dtype=torch.float32).pin_memory()
    out_top = torch.empty(conditions, n_rows, 10, dtype=torch.int32).pin_memory()

    def step(x_src, sid_src, from_host: bool):
        """One pass over all P prompts in sub-batches of BP prompts."""
        for p0 in range(0, P, BP):
            r0, r1 = p0 * T, min(P, p0 + BP) * T
            ll = None
            if from_host:
                xb = x_src[r0:r1].to(dev, non_blocking=True)
                sb = sid_src[r0:r1].to(dev, non_blocking=True)
                if args.ll_topk:
                    lb = ll_host[:, r0:r1].to(dev, non_blocking=True)
                    ll = interventions.LLSpec(lb[0].contiguous(), lb[1].contiguous(), rows_per_prompt=T)
            else:
                xb, sb = x_src[r0:r1], sid_src[r0:r1]
                if args.ll_topk:
                    ll = interventions.LLSpec(ll_dev[0, r0:r1], ll_dev[1, r0:r1], rows_per_prompt=T)
            if args.workload == "targeted":
                res = eng.targeted_budget_sweep(xb, T, sb, BUDGETS, ll=ll)
            elif args.workload == "random-control":
                res = eng.random_budget_sweep(xb, sb, BUDGETS, range(args.seeds), ll=ll)
            else:
                res = {}
                for r in RANKS:
                    if eng.time_lens:
                        rv0, rv1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        rv0.record()
                    _, h = ops.lowrank_remove(xb, U[:, :r].contiguous(), w_norm=lens.w_norm, eps=lens.eps, want_x=False,
                                              want_h=True)
                    if eng.time_lens:
                        rv1.record()
                        eng.stats.row_events.append((rv0, rv1, 4.0 * xb.shape[0] * D))
                        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        ev0.record()
                    res[r] = tbd.lens_readout_vocab_sharded(h, lens, softcap=SOFTCAP, secret_id=sb, cta_pair=args.cluster_mode)
                    if eng.time_lens:
                        ev1.record()      # shard GEMM + merge + all_gather + final merge
                        eng.stats.lens_events.append((ev0, ev1, 2.0 * h.shape[0] * D * lens.w_u.shape[0]))
                    eng.stats.launches += 4
            if from_host:
                for i, key in enumerate(res):
                    out_sec[i, r0:r1].copy_(res[key].secret_prob, non_blocking=True)
                    out_top[i, r0:r1].copy_(res[key].top_id, non_blocking=True)
                    if args.ll_topk and getattr(res[key], "ll_id", None) is not None:
                        out_ll[i, p0:min(P, p0 + BP)].copy_(res[key].ll_id, non_blocking=True)
        return res

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(from_host: bool, steps: int):
        src = (x_host, sid_host) if from_host else (x_dev, sid_dev)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            step(src[0], src[1], from_host)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()) / steps

    # ---- warm-up, then the timed region (device-resident inputs)
    for _ in range(args.warmup):
        step(x_dev, sid_dev, False)
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    eng.stats = interventions.SweepStats()
    eng.time_lens = True
    from taboo_brittleness_b200 import _lib as tb_lib
    calls0 = tb_lib.KERNEL_CALLS
    ms_step = timed(False, args.steps)
    kernel_calls = tb_lib.KERNEL_CALLS - calls0
    eng.time_lens = False
    ll_ms = sum(a.elapsed_time(b) for a, b, _ in eng.stats.ll_events)
    ll_flops = sum(f for _, _, f in eng.stats.ll_events)
    lens_ms = sum(a.elapsed_time(b) for a, b, _ in eng.stats.lens_events)
    lens_flops = sum(f for _, _, f in eng.stats.lens_events)
    n_lens = len(eng.stats.lens_events)
    row_ms = sum(a.elapsed_time(b) for a, b, _ in eng.stats.row_events)
    row_bytes = sum(f for _, _, f in eng.stats.row_events)
    launches = kernel_calls        # C-ABI compute calls inside the timed region (>= 1 kernel each)
    clocks = sampler.finish() if sampler else None

    # ---- end to end: host buffers in, readouts out, every step
    step(x_host, sid_host, True)
    ms_e2e = timed(True, max(1, min(args.steps, 2)))

    tokens_per_step = n_rows * conditions * (1 if proj else world)   # vocab sharding: total work is fixed
    value = tokens_per_step / (ms_step * 1e-3)
    e2e_value = tokens_per_step / (ms_e2e * 1e-3)
    if rank == 0:
        peaks = load_peaks()
        achieved = lens_flops / (lens_ms * 1e-3) / 1e12 if lens_ms > 0 else 0.0
        traffic = None
        tp = os.path.join(ROOT, "profiles", "lens_traffic.json")
        if os.path.exists(tp):   # ncu --set full capture, scaled from its row count to this launch's
            per_row = json.load(open(tp)).get("dram_bytes_per_row")
            traffic = per_row * min(BP, P) * T if per_row else None
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong" if proj else "weak",
                "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": workload_config(args),
                "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": ms_e2e,
                        "h2d_bytes_per_step": n_rows * D * 2 + n_rows * 4 + (n_rows * 8 if args.ll_topk else 0),
                        "d2h_bytes_per_step": conditions * (n_rows * (4 + 40) + (P * 40 if args.ll_topk else 0))},
                "gpu_launches": launches,
                "roofline": {"bound": "tensor",
                             "kernel": ("tb_umma_gemm_kernel<LensLLEpi> (lens readout + LL-Top-k candidates, one sweep; "
                                        "first budget: <LensEpi>)" if args.ll_topk and not args.ll_two_sweeps
                                        else "tb_umma_gemm_kernel<LensEpi> (logit-lens readout)"),
                             "achieved": achieved, "peak": peaks["tflops"], "unit": "TFLOP/s",
                             "frac": achieved / peaks["tflops"], "traffic": traffic, "peak_source": peaks["source"],
                             "frac_of_burst_peak": achieved / peaks["tflops_burst"], "launches_timed": n_lens,
                             "kernel_share_of_step": lens_ms / (ms_step * args.steps),
                             "flops_per_token_condition": FLOP_LENS,
                             "whole_step_tflops": (n_rows * ((0.0 if proj else FLOP_ENC) + conditions * FLOP_LENS /
                                                             (world if proj else 1))) / (ms_step * 1e-3) / 1e12},
                "clocks": clocks}
        if row_ms > 0:   # the HBM-bound producer of the lens operand: K2 (projection) or K1b (ablate + decode)
            gbs = row_bytes / (row_ms * 1e-3) / 1e9
            line["roofline"]["hbm_path"] = {
                "kernel": ("tb_lowrank_remove / tb_lowrank_remove_tc (K2, ranks %s)" % list(RANKS)) if proj
                else "tb_sae_ablate_decode (K1b, budgets %s)" % list(BUDGETS),
                "bound": "hbm", "achieved": gbs, "peak": peaks.get("hbm_gbs"), "unit": "GB/s",
                "frac": gbs / peaks["hbm_gbs"] if peaks.get("hbm_gbs") else None,
                "algorithmic_bytes_per_row": 4 * D, "share_of_step": row_ms / (ms_step * args.steps)}
        if line["roofline"]["frac"] > 1.0:
            line["roofline"]["note"] = ("peak = cuBLAS bf16 sustained under this board's power cap (MEASURED_PEAKS.json); "
                                        "a fraction above 1 means the kernel sustains more than cuBLAS does under the same "
                                        "cap -- frac_of_burst_peak is against the short-run cuBLAS figure")
        if args.ll_topk:
            line["ll_topk"] = {"k": 10, "one_sweep_prompt_conditions": eng.stats.ll_fused_prompts,
                               "redone_with_second_sweep": eng.stats.ll_fallback_prompts,
                               "mean_candidates_per_prompt": eng.stats.ll_candidates / max(eng.stats.ll_fused_prompts, 1),
                               "candidate_threshold": f"{eng.ll_alpha} x previous budget's k-th best sum",
                               "first_budget": "two sweeps (K3 + K5); later budgets K3F + K6, exact via certificate"}
        if args.ll_topk and ll_ms > 0:
            line["roofline"]["ll_topk_kernel"] = {"kernel": "tb_umma_gemm_kernel<ColSumEpi> (LL-Top-k sweep)",
                                                  "achieved": ll_flops / (ll_ms * 1e-3) / 1e12, "unit": "TFLOP/s",
                                                  "share_of_step": ll_ms / (ms_step * args.steps)}
        if world == 1 and not args.no_cpu_baseline and args.workload == "targeted":
            line["cpu_baseline"] = cpu_baseline_leg()
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--prompts", type=int, default=4096, help="prompts per GPU per step (x128 tokens); configs[1] = 4096")
    ap.add_argument("--batch-prompts", type=int, default=4096, help="prompts per sub-batch (one sweep each)")
    ap.add_argument("--chunks", type=int, default=0, help="vocab chunks per lens launch (0 = auto)")
    ap.add_argument("--encode-cap", type=int, default=128)
    ap.add_argument("--cluster-mode", type=int, default=5,
                    help="lens kernel cluster: 0 single CTAs, 1 CTA pairs, 2 = 2x1 pairs, 3 = 1x2 pairs, 4 = 2x2 pairs "
                         "(TMA multicast of the shared operand slabs), 5 = 2x1 preferred with plain pairs as fallback "
                         "(all 148 SMs)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="targeted", choices=["targeted", "random-control", "projection-vocab-shard"],
                    help="targeted = BASELINE configs[1] (the headline); the others are configs[2] and configs[3]")
    ap.add_argument("--seeds", type=int, default=2, help="random-control: seeds per step (configs[2] has 32)")
    ap.add_argument("--no-ll-topk", dest="ll_topk", action="store_false",
                    help="skip the per-prompt LL-Top-10 readout (then every condition is exactly one GEMM sweep)")
    ap.add_argument("--ll-alpha", type=float, default=0.5, help="K3F candidate threshold as a fraction of the "
                    "previous budget's k-th best LL sum")
    ap.add_argument("--overlap", action="store_true",
                    help="run K1b / K6 on a side stream under the lens sweeps (measured: no gain)")
    ap.add_argument("--ll-two-sweeps", action="store_true",
                    help="compute LL-Top-10 with the second GEMM sweep for every condition (no one-sweep path)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.workload == "projection-vocab-shard":
        args.ll_topk = False        # the sharded workload reads secret probability + top-10 (configs[3])
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return
    if world != args.gpus and rank == 0:
        print(f"warning: --gpus {args.gpus} but WORLD_SIZE={world}", file=sys.stderr)
    run_native(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
