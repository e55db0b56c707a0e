"""Pins the oracle: its functions must reproduce what the REFERENCE'S OWN function bodies produced
in the build container (tests/golden/make_golden.py) on the real cached fixtures and on seeded
inputs, plus transformers' Gemma2RMSNorm. CPU only; reads only tests/golden/."""
import json
import os

import numpy as np
import pytest
import torch

from _common import oracle

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
REAL = json.load(open(os.path.join(G, "real_ll_topk.json")))
HOST = json.load(open(os.path.join(G, "host_logic.json")))


def _load_real_probs(stem):
    word, n = stem.split("/")
    z = np.load(os.path.join(G, f"real_probs_{word}_{n[-2:]}.npz"))
    T, V = (int(v) for v in z["shape"])
    p = torch.zeros(T, V, dtype=torch.float32)
    p[torch.from_numpy(z["rows"]).long(), torch.from_numpy(z["cols"]).long()] = torch.from_numpy(z["vals"])
    return p


@pytest.mark.parametrize("stem", sorted(REAL["fixtures"]))
def test_ll_topk_on_real_cached_probs(stem):
    """a3 + a11 on real layer-31 probabilities: response start, masked column-sum, top-k ids --
    equal to the reference's aggregate_response_logits/topk, whose strings are the committed
    src/results/logit_lens/seed_42/top5_real/logit_lens_evaluation_results.json."""
    rec = REAL["fixtures"][stem]
    assert oracle.find_model_response_start(rec["input_words"]) == rec["start"]
    probs = _load_real_probs(stem)
    assert probs.shape[0] == rec["T_resp"]
    acc = oracle.aggregate_response_probs(probs, rec["cur_ids"])
    assert abs(float(acc.sum()) - rec["acc_sum"]) < 1e-4
    assert oracle.ll_topk(probs, rec["cur_ids"], 5) == rec["top5_ids"]
    top10 = torch.topk(acc, 10)
    assert top10.indices.tolist() == rec["top10_ids"]
    assert np.allclose(top10.values.numpy(), np.array(rec["top10_vals"], dtype=np.float32), rtol=0, atol=0)
    # id -> string map is the one the committed results imply
    assert [REAL["id2str"][str(i)] for i in rec["top5_ids"]] == rec["expected_strings"]


def test_response_start_edge_cases():
    assert oracle.find_model_response_start([]) == 0
    assert oracle.find_model_response_start(["<start_of_turn>", "a"]) == 0
    assert oracle.find_model_response_start(["<start_of_turn>", "x", "<start_of_turn>", "model", "\n", "hi"]) == 5


def test_aggregate_seeded_cases_with_masks():
    z = np.load(os.path.join(G, "aggregate_cases.npz"))
    n = len({k.split("_")[0] for k in z.files})
    for i in range(n):
        probs = torch.from_numpy(z[f"c{i}_probs"])
        cur = z[f"c{i}_cur_ids"].tolist()
        acc = oracle.aggregate_response_probs(probs, cur)
        assert torch.equal(acc, torch.from_numpy(z[f"c{i}_acc"]))
        assert oracle.ll_topk(probs, cur, 5) == z[f"c{i}_top5"].tolist()
    assert oracle.ll_topk(torch.zeros(3, 10), [0, 1, 2], 5) == []     # 01:147 "if sum > 0"


def test_sae_topk_of_mean_matches_reference_get_top_sae_features():
    z = np.load(os.path.join(G, "sae_topk_case.npz"))
    W, b, th, x = (torch.from_numpy(z[k]) for k in ("W_enc", "b_enc", "threshold", "x"))
    words = [str(w) for w in z["words"]]
    acts = oracle.jumprelu_encode(x, W, b, th)
    start = oracle.find_model_response_start(words)
    assert start == 11
    assert oracle.top_latents_of_mean(acts[start:], 8)[1].tolist() == z["top"].tolist()
    assert oracle.top_latents_of_mean(acts, 8)[1].tolist() == z["top_nostart"].tolist()
    # JumpReLU: nothing at or below threshold survives; everything that survives is the pre-activation
    pre = x @ W + b
    assert ((acts > 0) == ((pre > th) & (pre > 0))).all()
    assert torch.equal(acts[acts > 0], pre[acts > 0])


def test_rmsnorm_matches_installed_gemma2rmsnorm_golden():
    z = np.load(os.path.join(G, "rmsnorm_case.npz"))
    x, w = torch.from_numpy(z["x"]), torch.from_numpy(z["w"])
    assert torch.equal(oracle.gemma2_rmsnorm(x, w), torch.from_numpy(z["y_fp32"]))
    xb = x.to(torch.bfloat16).float()
    assert torch.equal(oracle.gemma2_rmsnorm(xb, w, residual_dtype=torch.bfloat16), torch.from_numpy(z["y_bf16"]))


def test_rmsnorm_matches_installed_gemma2rmsnorm_live():
    mod = pytest.importorskip("transformers.models.gemma2.modeling_gemma2")
    norm = mod.Gemma2RMSNorm(48, eps=1e-6)
    g = torch.Generator().manual_seed(3)
    with torch.no_grad():
        norm.weight.copy_(torch.randn(48, generator=g))
        x = torch.randn(5, 48, generator=g) * 9
        assert torch.equal(oracle.gemma2_rmsnorm(x, norm.weight.detach()), norm(x))


def test_softcap_follows_gemma2_forward():
    g = torch.Generator().manual_seed(0)
    x, w, U = torch.randn(4, 32, generator=g) * 8, torch.randn(32, generator=g), torch.randn(50, 32, generator=g)
    raw = oracle.lens_logits(x, w, U, None)
    capped = oracle.lens_logits(x, w, U, 30.0)
    assert torch.equal(capped, torch.tanh(raw / 30.0) * 30.0)
    assert capped.abs().max() <= 30.0
    p = oracle.lens_probs(x, w, U, None)
    assert torch.allclose(p.sum(-1), torch.ones(4), atol=1e-5)


def test_ablate_decode_identities():
    g = torch.Generator().manual_seed(1)
    d, S, T = 32, 128, 6
    x = torch.randn(T, d, generator=g)
    W_enc, b_enc = torch.randn(d, S, generator=g) / d ** 0.5, torch.zeros(S)
    th, W_dec, b_dec = torch.full((S,), 0.5), torch.randn(S, d, generator=g), torch.randn(d, generator=g)
    acts = oracle.jumprelu_encode(x, W_enc, b_enc, th)
    none = torch.empty(0, dtype=torch.int64)
    assert torch.allclose(oracle.ablate_decode(x, acts, none, W_dec, b_dec, "error"), x, atol=1e-5)
    assert torch.allclose(oracle.ablate_decode(x, acts, none, W_dec, b_dec, "plain"),
                          oracle.sae_decode(acts, W_dec, b_dec))
    idx = torch.tensor([3, 77, 100])
    e = oracle.ablate_decode(x, acts, idx, W_dec, b_dec, "error")
    assert torch.allclose(e, x - acts[:, idx] @ W_dec[idx], atol=1e-4)


def test_lowrank_remove_is_a_projection():
    U = oracle.random_orthonormal(64, 5, seed=4)
    x = torch.randn(7, 64, generator=torch.Generator().manual_seed(2))
    y = oracle.lowrank_remove(x, U)
    assert (y @ U).abs().max() < 1e-5
    assert torch.allclose(oracle.lowrank_remove(y, U), y, atol=1e-5)     # idempotent
    assert torch.allclose(U.T @ U, torch.eye(5), atol=1e-5)


def test_random_latents_are_a_seeded_prefix():
    a, b = oracle.random_latents(1000, 5, 64), oracle.random_latents(1000, 5, 16)
    assert a[:16].tolist() == b.tolist() and len(set(a.tolist())) == 64
    assert oracle.random_latents(1000, 6, 16).tolist() != b.tolist()


def test_metrics_match_reference():
    mc = HOST["metrics_case"]
    plur = {w: [w, w + "s"] for w in mc["targets"]}
    got = oracle.calculate_metrics(mc["predictions"], mc["targets"], plur)
    for k, v in mc["expected"].items():
        for kk, vv in v.items():
            assert got[k][kk] == pytest.approx(vv)
    for kk, vv in mc["committed_overall"].items():
        assert got["overall"][kk] == pytest.approx(vv)
    for c in HOST["metric_edge_cases"]:
        v = set(c["valid"])
        assert oracle.prompt_accuracy_at_k(c["guesses"], v) == pytest.approx(c["prompt_accuracy"])
        assert oracle.any_pass_at_k(c["guesses"], v) == c["any_pass"]
        assert oracle.global_majority_vote_at_k(c["guesses"], v) == c["majority"]


def test_latents_to_words_matches_reference():
    for c in HOST["latent_cases"]:
        assert oracle.latents_to_word_guesses(c["in"], HOST["feature_map"]) == c["out"]
