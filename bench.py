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

  workloads  three short legs after the headline region, one per remaining BASELINE.json config, so
         that the driver's per-N runs record them: projection_vocab_shard (configs[3]: low-rank removal
         ranks 1..64 + lens with W_U sharded by vocabulary over the N ranks, NCCL all_gather of the
         92-byte per-row partials, strong scaling), random_control (configs[2]: 11 budgets x 32 seeds,
         (seed, prompt) units split over the ranks, strong scaling) and stress (configs[4]: 131k-latent
         SAE encode + lens over 42 layer residuals, prompts sharded, weak scaling). --no-workloads skips them.

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
    """configs[0] (32 prompts x 64 tokens = 2048 rows, fp32), all host threads,
    one warm-up pass, best of 3."""
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    n_prompts, rpp = 32, 64
    x, sae, lens, sid = cpu_problem(n_prompts, rpp)
    cpu_step(x, sae, lens, sid, rpp)                  # warm-up (thread pool, page faults)
    best = None
    for _ in range(3):
        t0 = time.perf_counter()
        cpu_step(x, sae, lens, sid, rpp)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return {"value": x.shape[0] / best, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"BASELINE configs[0]: {n_prompts} prompts x {rpp} tok ({n_prompts * rpp} rows) of the same synthetic "
                      f"workload, fp32 oracle, one targeted top-10 ablation + full-vocab lens (secret prob, top-10, "
                      f"LL-Top-10), 1 warm-up + best of 3 ({best:.1f} s)"}


# ------------------------------------------------------------------------------------------------
# Workload legs: BASELINE.json configs[2], [3], [4], each a short timed region of its own after the
# headline. Every leg: a short warm-up, barrier + synchronize, CUDA events around the timed passes,
# barrier + synchronize, MAX over ranks. A leg that fails reports {"error": ...} and the headline line
# is still printed.
# ------------------------------------------------------------------------------------------------
def _leg_time(fn, passes, dev, world, dist):
    """ms per pass of fn(), max over ranks."""
    import torch

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(passes):
        fn()
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item()) / passes


def _sum_ms(pairs):
    return sum(a.elapsed_time(b) for a, b in pairs)


def leg_projection_vocab_shard(args, ctx):
    """configs[3]: rank r in {1..64} removal of a random orthonormal subspace (K2), then the lens with W_U
    split by vocabulary over the ranks (K3 on the shard, all_gather of the merged 92-byte per-row partials,
    K4 over the ranks). Every rank sees all rows: strong scaling. With more than one rank the leg runs
    twice: K2 replicated on every rank, and K2 on a row shard followed by an all_gather of h (7 KB per row)."""
    import torch
    from taboo_brittleness_b200 import dist as tbd
    from taboo_brittleness_b200 import ops, synthetic
    dev, rank, world, dist, lens, peaks = ctx["dev"], ctx["rank"], ctx["world"], ctx["dist"], ctx["lens"], ctx["peaks"]
    P = args.proj_prompts
    n_rows = P * T
    x = torch.empty(n_rows, D, dtype=torch.bfloat16, device=dev)
    for p0 in range(0, P, 512):
        p1 = min(P, p0 + 512)
        x[p0 * T:p1 * T] = synthetic.make_residuals(p1 - p0, T, D, seed=synthetic.SEED_DATA + 777 + p0,
                                                    device=str(dev)).to(torch.bfloat16)
    sid = synthetic.make_secret_ids(P, T, V, seed=synthetic.SEED_DATA + 777).to(dev)
    v0, v1 = tbd.vocab_shard(V, rank, world)
    g = torch.Generator(device=dev).manual_seed(7)
    U = torch.linalg.qr(torch.randn(D, max(RANKS), device=dev, generator=g))[0].to(torch.bfloat16).contiguous()
    r0, r1 = tbd.shard_range(n_rows, rank, world)
    can_row_shard = world > 1 and n_rows % world == 0
    state = {"timers": None, "k2": None, "hgather": None, "check": None, "v": (v0, v1),
             "shard": ops.LensWeights(w_norm=lens.w_norm, w_u=lens.w_u[v0:v1], vocab_offset=v0)}

    tc_ranks = [r for r in RANKS if r >= ops.LOWRANK_MULTI_MIN]

    def one_pass(k2_mode):
        h_multi = {}
        if k2_mode == "replicated" and len(tc_ranks) >= 2 and D % 256 == 0:
            # the tensor-core ranks share one coefficient pass (tb_lowrank_remove_tc_multi: x U of the largest rank once)
            a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
            a.record()
            _, hm = ops.lowrank_remove_multi(x, U, tc_ranks, w_norm=lens.w_norm, eps=lens.eps)
            b.record()
            if state["k2"] is not None:
                state["k2"].append((a, b))
            h_multi = {r: hm[i] for i, r in enumerate(tc_ranks)}
        for r in RANKS:
            Ur = U[:, :r].contiguous()
            a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
            c = torch.cuda.Event(enable_timing=True)
            a.record()
            if r in h_multi:
                h = h_multi[r]
                b.record()
            elif k2_mode == "replicated":
                _, h = ops.lowrank_remove(x, Ur, w_norm=lens.w_norm, eps=lens.eps, want_x=False, want_h=True)
                b.record()
            else:
                _, h_loc = ops.lowrank_remove(x[r0:r1], Ur, w_norm=lens.w_norm, eps=lens.eps, want_x=False, want_h=True)
                b.record()
                h = tbd.all_gather_cat(h_loc)
            c.record()
            if state["k2"] is not None:
                state["k2"].append((a, b))
                state["hgather"].append((b, c))
            out = tbd.lens_readout_vocab_sharded(h, state["shard"], softcap=SOFTCAP, secret_id=sid,
                                                 cta_pair=args.cluster_mode, timers=state["timers"])
        state["check"] = out

    def measure(k2_mode):
        one_pass(k2_mode)                                   # warm-up
        state.update(timers={}, k2=[], hgather=[])
        ms = _leg_time(lambda: one_pass(k2_mode), args.leg_passes, dev, world, dist)
        tm, k2, hg = state["timers"], state["k2"], state["hgather"]
        state.update(timers=None, k2=None, hgather=None)
        n = args.leg_passes
        part_ms, gat_ms, mer_ms = _sum_ms(tm["partials"]) / n, _sum_ms(tm["gather"]) / n, _sum_ms(tm["merge"]) / n
        k2_ms, hg_ms = _sum_ms(k2) / n, _sum_ms(hg) / n
        conds = n_rows * len(RANKS)
        return {"value": conds / (ms * 1e-3), "unit": "row-conditions/s", "ms_per_pass": ms,
                "phase_ms_this_rank": {"k2": k2_ms, "h_all_gather": hg_ms, "shard_gemm_and_local_merge": part_ms,
                                       "partials_all_gather": gat_ms, "merge_over_ranks": mer_ms},
                "phase_share_this_rank": {"k2": k2_ms / ms, "h_all_gather": hg_ms / ms,
                                          "shard_gemm_and_local_merge": part_ms / ms,
                                          "partials_all_gather": gat_ms / ms, "merge_over_ranks": mer_ms / ms},
                "lens_tflops_per_gpu": conds * 2.0 * D * (state["v"][1] - state["v"][0]) / (part_ms * 1e-3) / 1e12,
                "k2_gbs": (r1 - r0 if k2_mode != "replicated" else n_rows) * len(RANKS) * 4.0 * D / (k2_ms * 1e-3) / 1e9}

    res = {"config": f"BASELINE configs[3]: {P} prompts x {T} tok (all rows on every GPU), low-rank removal ranks "
                     f"{list(RANKS)}, lens V={V} with W_U sharded by vocabulary over {world} GPU(s) "
                     f"({v1 - v0} ids on rank 0), NCCL all_gather of 92 B per row and rank, secret prob + top-10",
           "scaling": "strong", "n_gpus": world, "passes": args.leg_passes,
           "partials_bytes_gathered_per_pass": n_rows * 92 * world * len(RANKS)}
    rep = measure("replicated")
    res.update(rep)
    res["k2"] = "replicated on every rank"
    res["readout_checksum"] = {"mean_secret_prob": float(state["check"].secret_prob.mean().item()),
                               "mean_lse": float(state["check"].lse.mean().item())}
    if can_row_shard:
        alt = measure("row-sharded")
        alt["k2"] = "row shard per rank, then all_gather of h (7 KB per row)"
        res["alternative_k2_row_sharded"] = alt
    if world > 1:
        # shards sized by each GPU's measured GEMM rate (ids per millisecond in the pass above): under their power caps
        # the GPUs of a box settle at different clocks, and with equal shards every exchange waits for the slowest
        mine = torch.tensor([(v1 - v0) / max(rep["phase_ms_this_rank"]["shard_gemm_and_local_merge"], 1e-6)], device=dev)
        rates = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(rates, mine)
        shards = tbd.vocab_shards_weighted(V, [float(r.item()) for r in rates])
        w0, w1 = shards[rank]
        state["v"] = (w0, w1)
        state["shard"] = ops.LensWeights(w_norm=lens.w_norm, w_u=lens.w_u[w0:w1], vocab_offset=w0)
        bal = measure("replicated")
        bal["k2"] = "replicated on every rank"
        bal["shard_ids_per_rank"] = [b - a for a, b in shards]
        res["alternative_speed_weighted_shards"] = bal
    return res


def leg_random_control(args, ctx):
    """configs[2]: random-latent ablation control, 11 budgets x 32 seeds. The (seed, prompt) pairs are the
    units (dist.shard_conditions, seed-major); a rank encodes each prompt range it touches once, runs the
    nested budgets of its seeds over it, and the per-row secret probabilities of all units are gathered to
    every rank at the end (dist.gather_rows). Total work is fixed: strong scaling."""
    import torch
    from taboo_brittleness_b200 import dist as tbd
    from taboo_brittleness_b200 import interventions, synthetic
    dev, rank, world, dist, eng = ctx["dev"], ctx["rank"], ctx["world"], ctx["dist"], ctx["eng"]
    P, n_seeds = args.rc_prompts, args.rc_seeds
    n_rows = P * T
    x = torch.empty(n_rows, D, dtype=torch.bfloat16, device=dev)
    for p0 in range(0, P, 512):
        p1 = min(P, p0 + 512)
        x[p0 * T:p1 * T] = synthetic.make_residuals(p1 - p0, T, D, seed=synthetic.SEED_DATA + 555 + p0,
                                                    device=str(dev)).to(torch.bfloat16)
    sid = synthetic.make_secret_ids(P, T, V, seed=synthetic.SEED_DATA + 555).to(dev)
    g = torch.Generator().manual_seed(synthetic.SEED_DATA + 556)
    ll_ids = torch.randint(0, V, (2, n_rows), generator=g, dtype=torch.int32)
    ll_ids[1, ::T] = -1
    ll_ids = ll_ids.to(dev)
    runs = tbd.shard_conditions(P, n_seeds, rank, world)            # [(seed, p0, p1), ...] of this rank
    out = {}

    def sweep(p0, p1, seeds):
        xb, sb = x[p0 * T:p1 * T], sid[p0 * T:p1 * T]
        ll = interventions.LLSpec(ll_ids[0, p0 * T:p1 * T], ll_ids[1, p0 * T:p1 * T], rows_per_prompt=T) \
            if args.ll_topk else None
        res = eng.random_budget_sweep(xb, sb, BUDGETS, seeds, ll=ll)           # encodes xb once
        return {s_: torch.stack([res[(b, s_)].secret_prob.view(p1 - p0, T) for b in BUDGETS], 1) for s_ in seeds}

    def one_pass():
        out["all"] = tbd.run_sharded_conditions(P, n_seeds, sweep)             # [n_seeds * P, 11, T] on every rank

    if runs:                                                                   # warm-up: two budgets of one seed
        s0, p0, p1 = runs[0]
        eng.random_budget_sweep(x[p0 * T:p1 * T], sid[p0 * T:p1 * T], BUDGETS[:2], [s0])
    stats0 = (eng.stats.ll_fused_prompts, eng.stats.ll_fallback_prompts)
    ms = _leg_time(one_pass, 1, dev, world, dist)
    conds = n_rows * len(BUDGETS) * n_seeds
    allp = out["all"]
    return {"config": f"BASELINE configs[2]: random-latent control, budgets {list(BUDGETS)} x {n_seeds} seeds x {P} prompts "
                      f"x {T} tok, (seed, prompt) units split over {world} GPU(s) by dist.shard_conditions, weights "
                      f"replicated, encode once per prompt range and rank, final gather of the per-row secret probabilities",
            "scaling": "strong", "n_gpus": world, "passes": 1, "value": conds / (ms * 1e-3), "unit": UNIT,
            "ms_per_pass": ms, "units_this_rank": sum(p1 - p0 for _, p0, p1 in runs), "units_total": P * n_seeds,
            "seeds_this_rank": sorted({s_ for s_, _, _ in runs}), "ll_topk": bool(args.ll_topk),
            "ll_one_sweep_prompt_conditions_this_rank": eng.stats.ll_fused_prompts - stats0[0],
            "ll_redone_with_second_sweep_this_rank": eng.stats.ll_fallback_prompts - stats0[1],
            "gathered_shape": list(allp.shape),
            "readout_checksum": {"mean_secret_prob": float(allp.mean().item()),
                                 "mean_secret_prob_budget_1024": float(allp[:, -1].mean().item())},
            "whole_job_tflops": conds * FLOP_LENS / (ms * 1e-3) / 1e12}


def leg_stress(args, ctx):
    """configs[4]: Gemma Scope 131k-latent SAE on the layer-32 residuals (K1 at S = 131072, then the per-prompt
    mean and top-10 latents) and the logit lens over the residuals of all 42 layers (norm, K3, K4: secret
    probability and top-10 per row), 16,384 prompts x 128 tokens split by prompt over 8 GPUs = 2048 prompts per
    GPU (weak scaling: every rank gets its own 2048). The residuals of all layers are generated on the device
    before the timed region and stay resident (79 GB per GPU)."""
    import torch
    from taboo_brittleness_b200 import ops, synthetic
    dev, rank, world, dist, lens, peaks = ctx["dev"], ctx["rank"], ctx["world"], ctx["dist"], ctx["lens"], ctx["peaks"]
    S2, L, P = 131072, 42, args.stress_prompts
    n_rows = P * T
    # thresholds from 4096 calibration rows: the (1 - 76 / 131072) quantile needs that many to resolve (L0 about 80)
    sae_s = synthetic.make_sae(D, S2, seed=synthetic.SEED_WEIGHTS + 1, calib_rows=4096, device=str(dev))
    sae = ops.SaeWeights.from_sae_lens_layout(sae_s.W_enc, sae_s.b_enc, sae_s.threshold, sae_s.W_dec, sae_s.b_dec,
                                              device=str(dev))
    del sae_s
    xs = torch.empty(L * n_rows, D, dtype=torch.bfloat16, device=dev)
    for l in range(L):
        for p0 in range(0, P, 512):
            p1 = min(P, p0 + 512)
            xs[l * n_rows + p0 * T:l * n_rows + p1 * T] = synthetic.make_residuals(
                p1 - p0, T, D, seed=synthetic.SEED_DATA + 9000 + 100000 * rank + 1000 * l + p0, device=str(dev)).to(torch.bfloat16)
    sid = synthetic.make_secret_ids(P, T, V, seed=synthetic.SEED_DATA + 9 + rank).to(dev).repeat(L)
    sub = 4096 * T                                            # lens rows per launch (the headline's batch)
    ev = {"enc": [], "lens": []}
    out = {}

    def mark():
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        return e

    def one_pass(rows_limit=None, record=True):
        a = mark()
        x31 = xs[31 * n_rows:32 * n_rows]
        acts = ops.sae_encode(x31 if rows_limit is None else x31[:rows_limit], sae, cap=2 * args.encode_cap)
        b = mark()
        mean = ops.latent_mean(acts, T)
        top_idx, _, _ = ops.topm_select(mean, 10)
        sec_sum = torch.zeros((), dtype=torch.float64, device=dev)
        total = L * n_rows if rows_limit is None else rows_limit
        for q0 in range(0, total, sub):
            q1 = min(total, q0 + sub)
            h = ops.rmsnorm(xs[q0:q1], lens.w_norm, lens.eps)
            c = mark()
            p = ops.lens_partials(h, lens.w_u, softcap=SOFTCAP, secret_id=sid[q0:q1], cta_pair=args.cluster_mode)
            d_ = mark()
            r = ops.lens_merge(p)
            sec_sum += r.secret_prob.sum(dtype=torch.float64)
            if record:
                ev["lens"].append((c, d_))
        if record:
            ev["enc"].append((a, b))
        out.update(top_idx=top_idx, sec_mean=sec_sum / total, overflow=int((acts.count > 2 * args.encode_cap).sum().item()),
                   l0=float(acts.count.float().mean().item()))

    one_pass(rows_limit=min(sub, n_rows), record=False)       # warm-up on one sub-batch
    ms = _leg_time(one_pass, 1, dev, world, dist)
    enc_ms, lens_ms = _sum_ms(ev["enc"]), _sum_ms(ev["lens"])
    tokens = n_rows * world
    enc_tf = n_rows * 2.0 * D * S2 / (enc_ms * 1e-3) / 1e12
    lens_tf = L * n_rows * FLOP_LENS / (lens_ms * 1e-3) / 1e12
    return {"config": f"BASELINE configs[4]: {P} prompts x {T} tok per GPU ({P * world} prompts on {world} GPU(s); 16,384 on 8), "
                      f"131072-latent JumpReLU SAE encode + per-prompt top-10 latents on the layer-32 residuals, logit lens "
                      f"(V={V}, softcap {SOFTCAP}: secret prob + top-10) over the residuals of all {L} layers, residuals "
                      f"resident in HBM ({L * n_rows * D * 2 / 1e9:.0f} GB per GPU)",
            "scaling": "weak", "n_gpus": world, "passes": 1, "value": tokens / (ms * 1e-3),
            "unit": "tokens/s (one token = one position through the 131k encode and the 42-layer lens)",
            "ms_per_pass": ms, "lens_rows_per_pass_per_gpu": L * n_rows,
            "lens_row_conditions_per_s": L * n_rows * world / (ms * 1e-3),
            "whole_job_tflops": tokens * (2.0 * D * S2 + L * FLOP_LENS) / (ms * 1e-3) / 1e12,
            "encode_roofline": {"kernel": "tb_umma_gemm_kernel<EncodeEpi, 2> at S=131072 through ops.sae_encode (latent sweep in two chunks + "
                                          "sub-list merge, then a second pass over the rows whose list overflowed, host decision "
                                          "in between)",
                                "bound": "tensor", "achieved": enc_tf, "peak": peaks["tflops_burst"], "unit": "TFLOP/s",
                                "frac": enc_tf / peaks["tflops_burst"], "peak_source": "MEASURED_PEAKS.json bf16 burst "
                                "(a 0.2 s kernel timed alone)", "flops_per_token": 2.0 * D * S2, "ms": enc_ms,
                                "mean_active_latents": out["l0"], "rows_with_overflowing_list": out["overflow"]},
            "lens_tflops_per_gpu": lens_tf, "lens_frac_of_sustained_peak": lens_tf / peaks["tflops"],
            "lens_share_of_pass": lens_ms / ms, "encode_share_of_pass": enc_ms / ms,
            "readout_checksum": {"mean_secret_prob": float(out["sec_mean"].item())}}


def leg_rank_update(args, ctx):
    """configs[1]'s nested budgets as RANK UPDATES of stored base logits
    (InterventionEngine.targeted_budget_sweep_rank_update), beside the dense sweep of the same rows WITHOUT LL-Top-k
    (the update path has lens readouts only). Not the reference's arithmetic -- the drift of every readout against
    the dense sweep is part of the result. Every rank takes its own prompts (weak scaling)."""
    import torch
    from taboo_brittleness_b200 import synthetic
    dev, rank, world, dist, eng = ctx["dev"], ctx["rank"], ctx["world"], ctx["dist"], ctx["eng"]
    P = args.ru_prompts
    n_rows = P * T
    x = torch.empty(n_rows, D, dtype=torch.bfloat16, device=dev)
    for p0 in range(0, P, 512):
        p1 = min(P, p0 + 512)
        x[p0 * T:p1 * T] = synthetic.make_residuals(p1 - p0, T, D, seed=synthetic.SEED_DATA + 777 + 1000 * rank + p0,
                                                    device=str(dev)).to(torch.bfloat16)
    sid = synthetic.make_secret_ids(P, T, V, seed=synthetic.SEED_DATA + 777 + rank).to(dev)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    Dm = eng.update_matrix()
    torch.cuda.synchronize()
    setup_s = time.perf_counter() - t0
    out = {}
    eng.targeted_budget_sweep_rank_update(x[:8 * T], sid[:8 * T], BUDGETS, split=args.ru_split)      # warm-up (and scratch)
    eng.targeted_budget_sweep_rank_update(x, sid, BUDGETS[:1], split=args.ru_split)
    eng.targeted_budget_sweep(x[:8 * T], T, sid[:8 * T], BUDGETS)
    stats0 = eng.stats
    from taboo_brittleness_b200 import interventions
    eng.stats = interventions.SweepStats()
    eng.time_lens = True
    ms_ru = _leg_time(lambda: out.__setitem__("ru", eng.targeted_budget_sweep_rank_update(x, sid, BUDGETS, split=args.ru_split)),
                      1, dev, world, dist)
    eng.time_lens = False
    base_ms = sum(a.elapsed_time(b) for a, b, _ in eng.stats.lens_events)
    base_flops = sum(f for _, _, f in eng.stats.lens_events)
    upd_ms = sum(a.elapsed_time(b) for a, b, _ in eng.stats.ll_events)
    upd_flops = sum(f for _, _, f in eng.stats.ll_events)
    n_sub = max(1, len(eng.stats.lens_events))
    per_budget = [sum(eng.stats.ll_events[i * len(BUDGETS) + j][0].elapsed_time(eng.stats.ll_events[i * len(BUDGETS) + j][1])
                      for i in range(n_sub)) for j in range(len(BUDGETS))]
    eng.stats = stats0
    ms_dense = _leg_time(lambda: out.__setitem__("dense", eng.targeted_budget_sweep(x, T, sid, BUDGETS)), 1, dev, world, dist)
    # the same two sweeps WITH the per-prompt LL-Top-10 (the headline's readouts): the update path stores every budget's
    # logits and reduces them with the column-sum kernel, the dense path is the headline's (stored logits / K3F + K6)
    with_ll = None
    if args.ll_topk:
        g = torch.Generator().manual_seed(synthetic.SEED_DATA + 778 + rank)
        ids2 = torch.randint(0, V, (2, n_rows), generator=g, dtype=torch.int32)
        ids2[1, ::T] = -1
        ids2 = ids2.to(dev)
        ll = interventions.LLSpec(ids2[0], ids2[1], rows_per_prompt=T)
        eng.release_scratch()
        torch.cuda.empty_cache()
        eng.update_matrix()
        eng.targeted_budget_sweep_rank_update(x, sid, BUDGETS[:1], split=args.ru_split, ll=ll)      # scratch
        ms_ru_ll = _leg_time(lambda: out.__setitem__("ru_ll", eng.targeted_budget_sweep_rank_update(x, sid, BUDGETS, split=args.ru_split,
                                                                                                     ll=ll)), 1, dev, world, dist)
        eng._dense_logits2 = None
        torch.cuda.empty_cache()
        ms_dense_ll = _leg_time(lambda: out.__setitem__("dense_ll", eng.targeted_budget_sweep(x, T, sid, BUDGETS, ll=ll)), 1, dev,
                                world, dist)
        common = [float((out["ru_ll"][b].ll_id[:, :, None] == out["dense_ll"][b].ll_id[:, None, :]).any(2).float().sum(1).mean().item())
                  for b in BUDGETS]
        with_ll = {"value": n_rows * len(BUDGETS) * world / (ms_ru_ll * 1e-3), "ms_per_pass": ms_ru_ll,
                   "dense_sweep_same_rows": {"value": n_rows * len(BUDGETS) * world / (ms_dense_ll * 1e-3), "ms_per_pass": ms_dense_ll},
                   "speedup_over_dense": ms_dense_ll / ms_ru_ll,
                   "ll_top10_ids_shared_with_dense_mean_per_budget": dict(zip([str(b) for b in BUDGETS], [round(c, 3) for c in common])),
                   "how": "every budget's soft-capped logits stored in prompt tiles (a second 67 GB scratch per 65,536 rows) and "
                          "reduced by tb_logits_ll_colsum + tb_topm_select: exact for this path's logits, no candidate lists"}
    drift = {}
    for b in BUDGETS:
        r, dn = out["ru"][b], out["dense"][b]
        drift[str(b)] = {"max_abs_dlse": float((r.lse - dn.lse).abs().max().item()),
                         "max_abs_dsecret_prob": float((r.secret_prob - dn.secret_prob).abs().max().item()),
                         "max_rel_dsecret_prob_where_p_gt_1e-3": float((((r.secret_prob - dn.secret_prob).abs() /
                                                                        dn.secret_prob.clamp_min(1e-30)) *
                                                                       (dn.secret_prob > 1e-3)).max().item()),
                         "max_abs_dtop1_prob": float((r.top_prob[:, 0] - dn.top_prob[:, 0]).abs().max().item()),
                         "top1_id_equal": float((r.top_id[:, 0] == dn.top_id[:, 0]).float().mean().item()),
                         "top10_set_equal": float((r.top_id.sort(1).values == dn.top_id.sort(1).values).all(1).float().mean().item())}
    conds = n_rows * len(BUDGETS) * world
    return {"config": f"SURVEY 8 (f4): configs[1] budgets {list(BUDGETS)} as rank updates of stored base logits, {P} prompts x {T} "
                      f"tok per GPU, lens readouts without LL-Top-k; beside it the dense sweep of the same rows (also without "
                      f"LL-Top-k); with_ll_top10: both with the per-prompt LL-Top-10. D = (W_dec (1+w)) W_U^T ({Dm.numel() * 2 / 1e9:.1f} GB bf16) is computed once per weight set "
                      f"outside the timed region ({setup_s:.2f} s, cuBLAS)",
            "scaling": "weak", "n_gpus": world, "passes": 1, "unit": UNIT,
            "value": conds / (ms_ru * 1e-3), "ms_per_pass": ms_ru,
            "dense_sweep_same_rows": {"value": conds / (ms_dense * 1e-3), "ms_per_pass": ms_dense},
            "speedup_over_dense": ms_dense / ms_ru,
            "with_ll_top10": with_ll,
            "activation_remainders": bool(args.ru_split),
            "not_the_reference_arithmetic": "the normalised residual is not rounded to bf16 per budget and D is a rounded "
                                            "product: readouts drift (table below, against the dense sweep); reported beside "
                                            "the dense line, never as tensor-pipe utilisation",
            "phase_ms_this_rank": {"base_sweeps_with_store": base_ms, "updates_all_budgets": upd_ms,
                                   "updates_per_budget": dict(zip([str(b) for b in BUDGETS], [round(v, 2) for v in per_budget])),
                                   "rest (encode, target, rank lists, dense activations, K1b rstd x 11, merges)":
                                       ms_ru - base_ms - upd_ms},
            "flops_done": {"base_sweep_tflops": base_flops / max(base_ms, 1e-9) / 1e9,
                           "update_tflops": upd_flops / max(upd_ms, 1e-9) / 1e9,
                           "flops_per_row_all_budgets": (base_flops + upd_flops) / max(n_rows, 1),
                           "dense_flops_per_row_all_budgets": len(BUDGETS) * FLOP_LENS},
            "drift_vs_dense_this_rank": drift}


def run_legs(args, ctx):
    import torch
    legs = {"projection_vocab_shard": leg_projection_vocab_shard, "random_control": leg_random_control,
            "stress": leg_stress, "rank_update": leg_rank_update}
    want = [w for w in args.workloads.split(",") if w]
    res = {}
    for name in want:
        if name not in legs:
            res[name] = {"error": "unknown workload"}
            continue
        t0 = time.perf_counter()
        try:
            res[name] = legs[name](args, ctx)
        except Exception as e:      # a failed leg must not take the headline line down with it
            import traceback
            res[name] = {"error": f"{type(e).__name__}: {e}", "trace": traceback.format_exc()[-1500:]}
            torch.cuda.synchronize()
        res[name]["wall_s"] = time.perf_counter() - t0
        ctx["eng"].release_scratch()
        import gc
        gc.collect()
        torch.cuda.empty_cache()
    return res


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
    eng.ll_dense_first = not args.no_ll_dense_first
    eng.multi_budget = not args.no_multi_budget

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
    e2e_steps = max(1, min(args.steps, 5))
    ms_e2e = timed(True, e2e_steps)

    tokens_per_step = n_rows * conditions * (1 if proj else world)   # vocab sharding: total work is fixed
    value = tokens_per_step / (ms_step * 1e-3)
    e2e_value = tokens_per_step / (ms_e2e * 1e-3)
    peaks = load_peaks()
    # ---- the other BASELINE configs as short legs of their own (all ranks take part; rank 0 reports)
    workloads = None
    if args.workloads and args.workload == "targeted":
        x_dev = sid_dev = ll_dev = x_host = sid_host = ll_host = out_ll = out_sec = out_top = None   # free the headline's buffers
        eng.release_scratch()
        import gc
        gc.collect()
        torch.cuda.empty_cache()
        ctx = {"dev": dev, "rank": rank, "world": world, "dist": dist, "lens": lens, "eng": eng, "peaks": peaks}
        workloads = run_legs(args, ctx)
    if rank == 0:
        achieved = lens_flops / (lens_ms * 1e-3) / 1e12 if lens_ms > 0 else 0.0
        traffic_x = None
        tp = os.path.join(ROOT, "profiles", "lens_traffic.json")
        if os.path.exists(tp):   # ncu --set full capture of a 37,888-row launch, scaled by rows to this launch's
            tj = json.load(open(tp))
            per_row = tj.get("dram_bytes_per_row")
            if per_row:
                traffic_x = {"bytes_per_launch": per_row * min(BP, P) * T, "from": "ncu --set full capture of one "
                             f"{tj.get('captured_rows', 37888)}-row launch (profiles/lens_traffic.json), scaled by the row count; "
                             "not measured in this run"}
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong" if proj else "weak",
                "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": workload_config(args),
                "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": ms_e2e, "steps": e2e_steps,
                        "h2d_bytes_per_step": n_rows * D * 2 + n_rows * 4 + (n_rows * 8 if args.ll_topk else 0),
                        "d2h_bytes_per_step": conditions * (n_rows * (4 + 40) + (P * 40 if args.ll_topk else 0))},
                "gpu_launches": launches,
                "roofline": {"bound": "tensor",
                             "kernel": ("tb_umma_gemm_kernel<LensLLEpi> (lens readout + LL-Top-k candidates, one sweep; "
                                        + ("first budget: <LensDenseEpi>, timed together with its merge, column-sum and select "
                                           "kernels)" if eng.ll_dense_first else "first budget: <LensEpi>)")
                                        if args.ll_topk and not args.ll_two_sweeps
                                        else "tb_umma_gemm_kernel<LensEpi> (logit-lens readout)"),
                             "achieved": achieved, "peak": peaks["tflops"], "unit": "TFLOP/s",
                             "frac": achieved / peaks["tflops"], "traffic": None,
                             "traffic_extrapolated_from_ncu": traffic_x, "peak_source": peaks["source"],
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
                else ("tb_sae_rank_lists once + tb_sae_ablate_ranked_multi: all budgets in one pass over the rows (K1b, budgets %s)"
                      if eng.multi_budget else "tb_sae_rank_lists once + tb_sae_ablate_ranked per budget (K1b, budgets %s)") % list(BUDGETS),
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
                               "first_budget": ("one sweep: the lens kernel also writes its logits (row sub-batches of "
                                                f"{eng.ll_dense_rows} rows, f32), an HBM-bound kernel sums exp(logit - lse) per prompt "
                                                "and column, a radix select ranks the sums; later budgets K3F + K6, exact via certificate")
                               if eng.ll_dense_first else "two sweeps (K3 + K5); later budgets K3F + K6, exact via certificate"}
        if args.ll_topk and ll_ms > 0:
            line["roofline"]["ll_topk_kernel"] = {"kernel": "tb_umma_gemm_kernel<ColSumEpi> (LL-Top-k sweep)",
                                                  "achieved": ll_flops / (ll_ms * 1e-3) / 1e12, "unit": "TFLOP/s",
                                                  "share_of_step": ll_ms / (ms_step * args.steps)}
        if workloads is not None:
            line["workloads"] = workloads
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
    ap.add_argument("--workloads", default="projection_vocab_shard,random_control,stress,rank_update",
                    help="comma-separated legs run after the headline region (BASELINE configs[3], [2], [4]; SURVEY 8 f4)")
    ap.add_argument("--ru-prompts", type=int, default=1024, help="rank_update leg: prompts per GPU")
    ap.add_argument("--ru-split", action="store_true", help="rank_update leg: bf16 remainders of the activations as extra depth")
    ap.add_argument("--no-workloads", dest="workloads", action="store_const", const="")
    ap.add_argument("--leg-passes", type=int, default=2, help="timed passes of the projection_vocab_shard leg")
    ap.add_argument("--proj-prompts", type=int, default=2048, help="projection_vocab_shard leg: prompts (all on every rank)")
    ap.add_argument("--rc-prompts", type=int, default=384, help="random_control leg: prompts (x --rc-seeds seeds = the units)")
    ap.add_argument("--rc-seeds", type=int, default=32, help="random_control leg: seeds (configs[2]: 32)")
    ap.add_argument("--stress-prompts", type=int, default=2048, help="stress leg: prompts per GPU (configs[4]: 16384 over 8)")
    ap.add_argument("--no-ll-topk", dest="ll_topk", action="store_false",
                    help="skip the per-prompt LL-Top-10 readout (then every condition is exactly one GEMM sweep)")
    ap.add_argument("--ll-alpha", type=float, default=0.5, help="K3F candidate threshold as a fraction of the "
                    "previous budget's k-th best LL sum")
    ap.add_argument("--no-ll-dense-first", action="store_true",
                    help="first budget of a chain: a second GEMM sweep (K5) for its LL-Top-k instead of reducing the logits "
                         "the lens kernel writes in its one sweep")
    ap.add_argument("--no-multi-budget", action="store_true",
                    help="K1b: one launch per budget instead of all budgets of a chain in one pass over the rows")
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
