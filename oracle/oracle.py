"""CPU oracle: a plain PyTorch fp32 restatement of the reference's intervention-and-readout path.

TEST INFRASTRUCTURE ONLY. Nothing in ``taboo_brittleness_b200/`` imports this module; it is
used by ``tests/``, by ``__graft_entry__.smoke()`` as the checker, and by ``bench.py`` for the
``cpu_baseline`` / ``--impl reference`` legs.

Why a restatement: the reference (pure Python, lmmontoya-ai/taboo-brittleness) runs all of its
arithmetic inside ``sae_lens`` (>=2.0.0, unpinned; requirements.txt:24), ``nnsight`` (unpinned,
:25) and ``transformers`` (>=4.41.0, :3). ``sae_lens`` and ``nnsight`` are not installed here and
there is no network, so ``src/models.py`` and ``src/02_run_sae_baseline.py`` cannot be imported.
Each function below cites the reference line it follows (paths relative to /root/reference).

Pinning status (tests/test_oracle_golden.py, fixtures made by tests/golden/make_golden.py):
  * find_model_response_start, aggregate_response_probs + top-k, get_top_sae_features (slice ->
    mean -> topk), latents_to_word_guesses: checked against the REFERENCE'S OWN function bodies,
    extracted from its source files and executed in this container, on the 8 real cached
    fixtures (src/data/processed/*/prompt_NN.npz) and on seeded inputs; the fixture top-5 ids
    also reproduce src/results/logit_lens/seed_42/top5_real/logit_lens_evaluation_results.json.
  * gemma2_rmsnorm: checked against transformers' installed Gemma2RMSNorm (the module the
    reference calls through model.model.norm).
  * softcap: follows the installed Gemma2ForCausalLM.forward (logits/cap -> tanh -> *cap).
  * metrics: checked against the reference's src/metrics.py (importable) and the known answer
    in results/ll_topk_ship.json (pass@k 0.8 = 8/10).
  * jumprelu_encode / sae_decode / ablation / low-rank removal: PARITY UNPINNED -- the reference
    never commits SAE outputs and has no code for ablation or projection (Execution Plan:126,
    128, 148 are prose); these follow sae_lens' published JumpReLU semantics and the Plan.

dtype flow. The reference loads the model in bf16 on CUDA (src/models.py:28-30), so residuals
and the RMSNorm output are bf16 there, while its CPU branch (:41-43) is fp32 throughout.
Functions take fp32 tensors; ``residual_dtype=torch.bfloat16`` rounds where the reference's CUDA
branch holds a bf16 tensor (the residual stream and the norm output), ``None`` is the pure
fp32 CPU branch. Matmuls and softmax are always fp32 here.
"""
from __future__ import annotations

from collections import Counter
from typing import Dict, List, Optional, Sequence, Set, Tuple

import numpy as np
import torch

GEMMA2_EPS = 1e-6
GEMMA2_SOFTCAP = 30.0


def _round(t: torch.Tensor, dtype: Optional[torch.dtype]) -> torch.Tensor:
    return t if dtype is None else t.to(dtype).to(torch.float32)


# --------------------------------------------------------------------------------------------
# a1 / a4: JumpReLU SAE (sae_lens semantics; call site src/02_run_sae_baseline.py:67)
# --------------------------------------------------------------------------------------------
def jumprelu_encode(x: torch.Tensor, W_enc: torch.Tensor, b_enc: torch.Tensor,
                    threshold: torch.Tensor) -> torch.Tensor:
    """pre = x @ W_enc + b_enc ; acts = relu(pre) * (pre > threshold).  W_enc is [d, S]."""
    pre = x @ W_enc + b_enc
    return torch.relu(pre) * (pre > threshold).to(pre.dtype)


def sae_decode(acts: torch.Tensor, W_dec: torch.Tensor, b_dec: torch.Tensor) -> torch.Tensor:
    """x_hat = acts @ W_dec + b_dec.  W_dec is [S, d]."""
    return acts @ W_dec + b_dec


# --------------------------------------------------------------------------------------------
# a3: response window (src/models.py:173-185)
# --------------------------------------------------------------------------------------------
def find_model_response_start(input_words: Sequence[str]) -> int:
    starts = [i for i, tok in enumerate(input_words) if tok == "<start_of_turn>"]
    if len(starts) >= 2:
        return starts[1] + 3
    return 0


# --------------------------------------------------------------------------------------------
# a2: targeted latent set = top-k of the mean activation over the response tokens
# (src/02_run_sae_baseline.py:53-74)
# --------------------------------------------------------------------------------------------
def top_latents_of_mean(acts: torch.Tensor, k: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """acts [T, S] (already sliced to the response) -> (values [k], indices [k]) of topk(mean)."""
    avg = torch.mean(acts, dim=0)
    vals, idx = torch.topk(avg, k=k)
    return vals, idx


def random_latents(S: int, seed: int, m: int) -> torch.Tensor:
    """Random-latent control (Execution Plan:128). No RNG convention exists in the reference;
    this one is fixed by SURVEY section 5: a seeded CPU permutation, first m entries."""
    g = torch.Generator(device="cpu").manual_seed(int(seed))
    return torch.randperm(S, generator=g)[:m]


# --------------------------------------------------------------------------------------------
# a4: ablation + decode (Execution Plan:126; error term per BASELINE.json north_star)
# --------------------------------------------------------------------------------------------
def ablate_decode(x: torch.Tensor, acts: torch.Tensor, idx: torch.Tensor, W_dec: torch.Tensor,
                  b_dec: torch.Tensor, mode: str = "error",
                  residual_dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    """Zero latents ``idx`` in acts and decode.

    mode "plain": x' = decode(acts_ablated)                       (Plan:126 literally)
    mode "error": x' = decode(acts_ablated) + (x - decode(acts))  (keeps the SAE error term)
    """
    a2 = acts.clone()
    if idx.numel():
        a2[:, idx.to(torch.int64)] = 0
    xh2 = sae_decode(a2, W_dec, b_dec)
    if mode == "plain":
        out = xh2
    elif mode == "error":
        out = xh2 + (x - sae_decode(acts, W_dec, b_dec))
    else:
        raise ValueError(mode)
    return _round(out, residual_dtype)


# --------------------------------------------------------------------------------------------
# a5: low-rank direction removal (Execution Plan:148; random control :150)
# --------------------------------------------------------------------------------------------
def lowrank_remove(x: torch.Tensor, U: torch.Tensor,
                   residual_dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    """x' = x - (x U) U^T, U [d, r] with orthonormal columns."""
    return _round(x - (x @ U) @ U.T, residual_dtype)


def random_orthonormal(d: int, r: int, seed: int) -> torch.Tensor:
    """Plan:150: Gaussian directions, orthogonalised (QR)."""
    g = torch.Generator(device="cpu").manual_seed(int(seed))
    q, _ = torch.linalg.qr(torch.randn(d, r, generator=g, dtype=torch.float64))
    return q.to(torch.float32)


# --------------------------------------------------------------------------------------------
# a6-a8: logit lens (src/models.py:135-138)
# --------------------------------------------------------------------------------------------
def gemma2_rmsnorm(x: torch.Tensor, weight: torch.Tensor, eps: float = GEMMA2_EPS,
                   residual_dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    """Gemma2RMSNorm.forward of the installed transformers:
        output = x.float() * rsqrt(x.float().pow(2).mean(-1, keepdim=True) + eps)
        output = output * (1.0 + weight.float());  return output.type_as(x)
    ``residual_dtype`` plays the role of x's dtype for the final ``type_as``."""
    xf = x.float()
    out = xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)
    out = out * (1.0 + weight.float())
    return _round(out, residual_dtype)


def lens_logits(x: torch.Tensor, w_norm: torch.Tensor, W_U: torch.Tensor,
                softcap: Optional[float] = None, eps: float = GEMMA2_EPS,
                residual_dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    """lm_head(norm(x)) (models.py:135), W_U [V, d] = lm_head.weight, no bias. The reference
    does NOT apply Gemma-2's final-logit softcap (it calls the bare lm_head); ``softcap=30.0``
    adds it the way Gemma2ForCausalLM.forward does: logits / cap -> tanh -> * cap."""
    h = gemma2_rmsnorm(x, w_norm, eps, residual_dtype)
    logits = h @ W_U.T
    if softcap:
        logits = torch.tanh(logits / softcap) * softcap
    return logits


def lens_probs(x, w_norm, W_U, softcap=None, eps=GEMMA2_EPS, residual_dtype=None) -> torch.Tensor:
    """softmax(..., dim=-1) (models.py:138), in fp32."""
    return torch.softmax(lens_logits(x, w_norm, W_U, softcap, eps, residual_dtype).float(), dim=-1)


# --------------------------------------------------------------------------------------------
# a9 / a10: per-row readouts
# --------------------------------------------------------------------------------------------
def secret_prob(probs: torch.Tensor, secret_id: torch.Tensor) -> torch.Tensor:
    """probs[t, secret_id[t]] (src/plots.py:10: all_probs[:, start:, token_id])."""
    return probs.gather(1, secret_id.to(torch.int64)[:, None])[:, 0]


def row_topk(probs: torch.Tensor, k: int = 10) -> Tuple[torch.Tensor, torch.Tensor]:
    """Per-row top-k (the reference keeps top-1, probs.max(-1) at models.py:147)."""
    return torch.topk(probs, k=k, dim=-1)


# --------------------------------------------------------------------------------------------
# a11: LL-Top-k (src/01_reproduce_logit_lens.py:35-71 and :147-149)
# --------------------------------------------------------------------------------------------
def aggregate_response_probs(response_probs: torch.Tensor, cur_ids: Sequence[int],
                             prev_ids: Optional[Sequence[int]] = None) -> torch.Tensor:
    """For each response position zero the probability of the current token and of the previous
    token (when their ids are inside the vocabulary), then sum over positions -> [V].
    ``cur_ids[i]`` is the id the reference gets from tokenizer.convert_tokens_to_ids(token_i);
    prev_ids defaults to cur_ids shifted by one (01:60-62)."""
    V = response_probs.shape[-1]
    acc = torch.zeros(V, dtype=torch.float32)
    for i in range(response_probs.shape[0]):
        p = response_probs[i].clone().to(torch.float32)
        if i > 0:
            prev = int(prev_ids[i]) if prev_ids is not None else int(cur_ids[i - 1])
            if 0 <= prev < V:
                p[prev] = 0
        cur = int(cur_ids[i])
        if 0 <= cur < V:
            p[cur] = 0
        acc += p
    return acc


def ll_topk(response_probs: torch.Tensor, cur_ids: Sequence[int], k: int,
            prev_ids: Optional[Sequence[int]] = None) -> List[int]:
    """01:147-149: top-k ids of the aggregate, or [] when it sums to zero."""
    acc = aggregate_response_probs(response_probs, cur_ids, prev_ids)
    if torch.sum(acc) > 0:
        return torch.topk(acc, k=k).indices.tolist()
    return []


# --------------------------------------------------------------------------------------------
# a12: latent -> word (src/02_run_sae_baseline.py:77-93)
# --------------------------------------------------------------------------------------------
def latents_to_word_guesses(latent_indices: Sequence[int], feature_map: Dict[str, List[int]]) -> List[str]:
    inv: Dict[int, str] = {}
    for w, feats in feature_map.items():
        for f in feats:
            inv[f] = w
    guesses: List[str] = []
    for i in latent_indices:
        if i in inv and inv[i] not in guesses:
            guesses.append(inv[i])
    return guesses


# --------------------------------------------------------------------------------------------
# a13: metrics (src/metrics.py:32-164)
# --------------------------------------------------------------------------------------------
def prompt_accuracy_at_k(guesses_by_prompt: List[List[str]], valid: Set[str]) -> float:
    if not guesses_by_prompt:
        return 0.0
    hits = sum(1 for g in guesses_by_prompt if not {s.strip().lower() for s in g}.isdisjoint(valid))
    return hits / len(guesses_by_prompt)


def any_pass_at_k(guesses_by_prompt: List[List[str]], valid: Set[str]) -> float:
    for g in guesses_by_prompt:
        if not {s.strip().lower() for s in g}.isdisjoint(valid):
            return 1.0
    return 0.0


def global_majority_vote_at_k(guesses_by_prompt: List[List[str]], valid: Set[str]) -> float:
    if not guesses_by_prompt:
        return 0.0
    flat = [s.strip().lower() for g in guesses_by_prompt for s in g]
    if not flat:
        return 0.0
    winner, _ = Counter(flat).most_common(1)[0]
    return 1.0 if winner in valid else 0.0


def calculate_metrics(predictions, target_words, word_plurals) -> Dict[str, Dict[str, float]]:
    per_word = {}
    for w in target_words:
        g = predictions.get(w, [])
        valid = {f.lower() for f in word_plurals.get(w, [w])}
        per_word[w] = {"prompt_accuracy": prompt_accuracy_at_k(g, valid),
                       "any_pass": any_pass_at_k(g, valid),
                       "global_majority_vote": global_majority_vote_at_k(g, valid)}
    out = {"overall": {k: float(np.mean([m[k] for m in per_word.values()]))
                       for k in ("prompt_accuracy", "any_pass", "global_majority_vote")}}
    out.update(per_word)
    return out


# --------------------------------------------------------------------------------------------
# the whole path for one batch of prompts (SURVEY section 8c pseudocode) -- used as the checker
# in smoke() and as the timed CPU baseline in bench.py
# --------------------------------------------------------------------------------------------
def run_path(x: torch.Tensor, rows_per_prompt: int, sae: Dict[str, torch.Tensor], w_norm: torch.Tensor,
             W_U: torch.Tensor, secret_id: torch.Tensor, budget: int, softcap: Optional[float],
             mode: str = "error", residual_dtype: Optional[torch.dtype] = None,
             topk: int = 10) -> Dict[str, torch.Tensor]:
    """encode -> per-prompt targeted top-``budget`` latents -> ablate+decode -> lens readouts.

    x [N, d] fp32 (bf16-exact values); sae = {W_enc [d,S], b_enc, threshold, W_dec [S,d], b_dec};
    secret_id [N]. Returns acts, idx [P, budget], x1 [N, d], secret_prob [N], top_prob/top_id [N, k].
    """
    N = x.shape[0]
    P = N // rows_per_prompt
    acts = jumprelu_encode(x, sae["W_enc"], sae["b_enc"], sae["threshold"])
    x1 = torch.empty_like(x)
    idx_all = []
    for p in range(P):
        sl = slice(p * rows_per_prompt, (p + 1) * rows_per_prompt)
        _, idx = top_latents_of_mean(acts[sl], budget)
        idx_all.append(idx)
        x1[sl] = ablate_decode(x[sl], acts[sl], idx, sae["W_dec"], sae["b_dec"], mode, residual_dtype)
    probs = lens_probs(x1, w_norm, W_U, softcap, residual_dtype=residual_dtype)
    tp, ti = row_topk(probs, topk)
    return {"acts": acts, "idx": torch.stack(idx_all), "x1": x1,
            "secret_prob": secret_prob(probs, secret_id), "top_prob": tp, "top_id": ti,
            "lse_free_probs": probs}


# --------------------------------------------------------------------------------------------
# CPU restatement of what one vocabulary shard contributes and how shards combine (the K3
# partial / K4 merge contract of include/tb_b200.h). Used by the gloo tests of dist.py and by
# the GPU tests of the merge kernel; log-sum-exp algebra only, nothing from the reference.
# --------------------------------------------------------------------------------------------
def lens_partial_packed(logits_shard: torch.Tensor, vocab_offset: int,
                        secret_id: Optional[torch.Tensor] = None, k: int = 10) -> torch.Tensor:
    """[rows, V_shard] (soft-capped) logits of one shard -> packed partial f32 [1, rows, 3 + 2k]:
    (max, sum exp(l - max), secret logit or -inf, top-k logits, top-k global ids bit-cast)."""
    lg = logits_shard.float()
    rows, vs = lg.shape
    m = lg.max(dim=-1).values
    s = torch.exp(lg - m[:, None]).sum(-1)
    sec = torch.full((rows,), float("-inf"))
    if secret_id is not None:
        rel = secret_id.to(torch.int64) - vocab_offset
        ok = (rel >= 0) & (rel < vs)
        sec[ok] = lg[ok].gather(1, rel[ok][:, None])[:, 0]
    kk = min(k, vs)
    tv, ti = torch.topk(lg, kk, dim=-1)
    if kk < k:
        tv = torch.cat([tv, torch.full((rows, k - kk), float("-inf"))], 1)
        ti = torch.cat([ti, torch.full((rows, k - kk), -1 - vocab_offset, dtype=ti.dtype)], 1)
    ids = (ti + vocab_offset).to(torch.int32)
    return torch.cat([m[:, None], s[:, None], sec[:, None], tv, ids.view(torch.float32)], dim=1)[None]


def merge_partials_packed(packed: torch.Tensor, k: int = 10) -> Dict[str, torch.Tensor]:
    """[parts, rows, 3 + 2k] -> lse, secret_prob, top_prob, top_id (what K4 finalises)."""
    m, s, sec = packed[..., 0], packed[..., 1], packed[..., 2]
    tv = packed[..., 3:3 + k]
    ti = packed[..., 3 + k:3 + 2 * k].contiguous().view(torch.int32)
    mm = m.max(dim=0).values
    ss = (s * torch.exp(m - mm[None])).sum(0)
    lse = mm + torch.log(ss)
    secl = sec.max(dim=0).values
    allv = tv.permute(1, 0, 2).reshape(tv.shape[1], -1)
    alli = ti.permute(1, 0, 2).reshape(ti.shape[1], -1)
    bv, pos = torch.topk(allv, k, dim=-1)
    return {"lse": lse, "secret_prob": torch.where(torch.isfinite(secl), torch.exp(secl - lse), torch.zeros_like(lse)),
            "top_prob": torch.exp(bv - lse[:, None]), "top_id": alli.gather(1, pos)}


def ll_topk_lists(acc_shard: torch.Tensor, vocab_offset: int, k: int = 10) -> Tuple[torch.Tensor, torch.Tensor]:
    """[P, V_shard] masked column sums of one shard -> sorted (sums [1, P, k], global ids [1, P, k])."""
    v, i = torch.topk(acc_shard.float(), k, dim=-1)
    return v[None], (i + vocab_offset).to(torch.int32)[None]


def ll_topk_merge_lists(pv: torch.Tensor, pi: torch.Tensor, k: int = 10) -> Tuple[torch.Tensor, torch.Tensor]:
    allv = pv.permute(1, 0, 2).reshape(pv.shape[1], -1)
    alli = pi.permute(1, 0, 2).reshape(pi.shape[1], -1)
    v, pos = torch.topk(allv, k, dim=-1)
    return v, alli.gather(1, pos)


# --------------------------------------------------------------------------------------------
# Targeting per the Execution Plan (Plan:116-124, 144-146). PARITY UNPINNED: prose only in the
# reference; dense restatement of the formulas, used to check targeting.py.
# --------------------------------------------------------------------------------------------
def spike_positions(secret_prob: torch.Tensor, rows_per_prompt: int, k: int = 4) -> torch.Tensor:
    return torch.topk(secret_prob.view(-1, rows_per_prompt), k, dim=1).indices


def secret_scores(acts: torch.Tensor, secret_prob: torch.Tensor, rows_per_prompt: int,
                  spikes: torch.Tensor) -> torch.Tensor:
    """acts [N, S] dense; score[p, s] = mean_{t in spikes(p)} a[t, s] * max(0, corr_t(a[:, s], p))."""
    T = rows_per_prompt
    P = acts.shape[0] // T
    out = torch.zeros(P, acts.shape[1], dtype=torch.float64)
    for p in range(P):
        a = acts[p * T:(p + 1) * T].double()
        q = secret_prob[p * T:(p + 1) * T].double()
        mean_spike = a[spikes[p]].mean(0)
        ac, qc = a - a.mean(0), q - q.mean()
        denom = torch.sqrt((ac * ac).mean(0) * (qc * qc).mean())
        corr = torch.where(denom > 0, (ac * qc[:, None]).mean(0) / denom.clamp_min(1e-300), torch.zeros_like(denom))
        out[p] = mean_spike * corr.clamp_min(0)
    return out.float()


def pca_directions(x: torch.Tensor, r: int) -> torch.Tensor:
    xc = (x - x.mean(0, keepdim=True)).double()
    _, _, vt = torch.linalg.svd(xc, full_matrices=False)
    U = vt[:r].t().contiguous()
    big = U.abs().argmax(0)
    return (U * torch.sign(U[big, torch.arange(r)])[None]).float()
