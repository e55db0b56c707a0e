"""The reference's two baseline analyses through the CUDA kernels, on REAL data where the reference
ships any: the eight cached layer-31 probability slices must give the LL-Top-5 / Top-10 ids (and the
strings) the reference's own aggregate_response_logits + topk produced -- the ids behind its committed
src/results/logit_lens/seed_42/top5_real/logit_lens_evaluation_results.json -- and the SAE top-k of
the mean must match what the reference's get_top_sae_features body returned (tests/golden)."""
import json
import os

import numpy as np
import pytest
import torch

from _common import oracle, synthetic
from test_oracle_golden import G, REAL, _load_real_probs

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("stem", sorted(REAL["fixtures"]))
def test_ll_topk_of_real_cached_probs_on_gpu(cuda_lib, stem):
    from taboo_brittleness_b200 import ops
    rec = REAL["fixtures"][stem]
    probs = _load_real_probs(stem).cuda()
    cur = torch.tensor([i if 0 <= i < probs.shape[1] else -1 for i in rec["cur_ids"]], dtype=torch.int32).cuda()
    prev = torch.cat([torch.tensor([-1], dtype=torch.int32).cuda(), cur[:-1]])
    acc = ops.probs_ll_colsum(probs, cur, prev)
    ref = oracle.aggregate_response_probs(probs.cpu(), rec["cur_ids"])
    assert torch.equal(acc.cpu(), ref), "same additions in the same order: bit-identical"
    v5, i5 = ops.probs_ll_topk(probs, cur, prev, 5)
    v10, i10 = ops.probs_ll_topk(probs, cur, prev, 10)
    torch.cuda.synchronize()
    assert i5.tolist() == rec["top5_ids"] and i10.tolist() == rec["top10_ids"]
    assert np.array_equal(v10.cpu().numpy(), np.array(rec["top10_vals"], dtype=np.float32))
    assert [REAL["id2str"][str(i)] for i in i5.tolist()] == rec["expected_strings"]


def test_analyze_cached_with_a_tokenizer(cuda_lib):
    """analyze_cached end to end on ship/01 with a tokenizer that knows the fixture's id <-> string map."""
    from taboo_brittleness_b200 import baselines
    stem = "ship/prompt_01"
    rec = REAL["fixtures"][stem]
    resp = _load_real_probs(stem).numpy()
    start = rec["start"]
    words = rec["input_words"]

    class Tok:
        table = {w: i for w, i in zip(words[start:], rec["cur_ids"])}

        def convert_tokens_to_ids(self, w):
            return self.table.get(w, -1)

        def decode(self, ids):
            return REAL["id2str"].get(str(ids[0]), f"<{ids[0]}>")

    all_probs = np.zeros((32, len(words), resp.shape[1]), dtype=np.float32)
    all_probs[31, start:] = resp
    guesses, ids = baselines.analyze_cached(all_probs, words, Tok(), 31, 5)
    assert ids == rec["top5_ids"] and guesses == [s.strip() for s in rec["expected_strings"]]


def test_sae_topk_matches_reference_body(cuda_lib):
    from taboo_brittleness_b200 import sae as tsae
    z = np.load(os.path.join(G, "sae_topk_case.npz"))
    W, b, th, x = (torch.from_numpy(z[k]) for k in ("W_enc", "b_enc", "threshold", "x"))
    S, d = W.shape[1], W.shape[0]
    # the kernels take bf16 operands: use bf16-exact copies and let the oracle say what those give
    Wb, xb = synthetic.bf16_exact(W), synthetic.bf16_exact(x)
    s = tsae.SAE.from_params(Wb, b, th, torch.zeros(S, d), torch.zeros(d))
    words = [str(w) for w in z["words"]]
    got = tsae.get_top_sae_features(s, xb, words, 8)
    got_all = tsae.get_top_sae_features(s, xb, ["a"] * len(words), 8)
    acts = oracle.jumprelu_encode(xb, Wb, b, th)
    assert got == oracle.top_latents_of_mean(acts[11:], 8)[1].tolist()
    assert got_all == oracle.top_latents_of_mean(acts, 8)[1].tolist()
    # and the bf16 rounding of the operands does not move the reference's own answer on this case
    assert set(got) == set(z["top"].tolist()) and set(got_all) == set(z["top_nostart"].tolist())
    dense = s.encode(xb).cpu()
    assert torch.equal(dense > 0, acts > 0)
    assert torch.allclose(dense, acts, rtol=1e-5, atol=1e-6)


def test_real_residuals_through_encode_and_cache_roundtrip(cuda_lib, tmp_path):
    """Real layer-31 residuals (fp16-exact in the reference's cache, <bos> rows of norm ~7600) through the
    cache format and the SAE baseline step, against the oracle on the same bf16-rounded rows."""
    from taboo_brittleness_b200 import baselines, cache
    from taboo_brittleness_b200 import sae as tsae
    z = np.load(os.path.join(G, "real_residuals_l31_fp16.npz"))
    key = "ship_01"
    resid = z[key].astype(np.float32)
    rec = REAL["fixtures"]["ship/prompt_01"]
    npz, js = cache.pair_paths(str(tmp_path), "ship", 0)
    cache.save_pair(npz, js, np.zeros((1, resid.shape[0], 8), np.float32), rec["input_words"], "resp", "Give me a hint!",
                    residual_stream=resid, layer_idx=31)
    meta = json.load(open(js))
    assert meta["shapes"]["residual_stream_l31"] == [resid.shape[0], 3584] and meta["dtypes"]["all_probs"] == "float32"
    pair = cache.load_pair(npz, js, 31, want_probs=False)
    assert pair.all_probs is None and np.array_equal(pair.residual, resid) and pair.input_words == rec["input_words"]
    syn = synthetic.make_sae(3584, 16384, seed=42)
    s = tsae.SAE.from_params(syn.W_enc, syn.b_enc, syn.threshold, syn.W_dec, syn.b_dec)
    guesses, latents = baselines.sae_guesses(s, pair.residual, pair.input_words, 10)
    xb = synthetic.bf16_exact(torch.from_numpy(resid))
    acts = oracle.jumprelu_encode(xb[rec["start"]:], syn.W_enc, syn.b_enc, syn.threshold)
    vals, idx = oracle.top_latents_of_mean(acts, 11)
    assert set(latents) == set(idx[:10].tolist()) or abs(float(vals[9] - vals[10])) <= 1e-4 * float(vals[9])
    assert guesses == baselines.latents_to_word_guesses(latents)


def test_real_residuals_through_the_whole_intervention_path(cuda_lib):
    """the 8 real layer-31 residual fixtures (ragged lengths 27-38, <bos> rows of norm ~7600,
    outlier dimensions) as a distribution sanity case for the whole path: encode -> targeted top-8
    ablation per prompt (response windows as segments) -> decode -> lens -> LL-Top-5, against the oracle.
    Weights are random-init (none are available offline); the residuals are the reference's own."""
    from _common import BF16, cuda_bf16, sae_to_cuda, topk_sets_match
    from taboo_brittleness_b200 import interventions, ops
    z = np.load(os.path.join(G, "real_residuals_l31_fp16.npz"))
    keys = sorted(z.files)
    rows = [torch.from_numpy(z[k].astype(np.float32)) for k in keys]
    x = synthetic.bf16_exact(torch.cat(rows))
    lens_t = [r.shape[0] for r in rows]
    seg = torch.tensor([0] + list(np.cumsum(lens_t)), dtype=torch.int64)
    d, S, V, budget = 3584, 16384, 32768, 8
    sae = synthetic.make_sae(d, S, seed=42)
    lens = synthetic.make_lens(d, V, seed=42)
    sid = torch.cat([torch.full((t,), 1000 + 17 * i, dtype=torch.int32) for i, t in enumerate(lens_t)])
    eng = interventions.InterventionEngine(sae_to_cuda(sae), ops.LensWeights(lens.w_norm.cuda(), cuda_bf16(lens.W_U)),
                                           softcap=30.0)
    xs = cuda_bf16(x)
    acts = eng.encode(xs)
    idx, _, rank = eng.targeted_rank(acts, seg=seg.cuda(), m=budget)
    row_table = torch.repeat_interleave(torch.arange(len(keys), dtype=torch.int32), torch.tensor(lens_t)).cuda()
    g = torch.Generator().manual_seed(9)
    cur = torch.randint(0, V, (x.shape[0],), generator=g, dtype=torch.int32)
    ll = interventions.LLSpec(cur.cuda(), None, seg=seg.cuda(), k=5)
    out = eng.ablate_and_read(xs, acts, rank, budget, row_table=row_table, secret_id=sid.cuda(), ll=ll)
    torch.cuda.synchronize()
    # oracle, prompt by prompt, on the kernel's own active sets (near-threshold flips are pinned elsewhere)
    dense = acts.to_dense().cpu()
    for i in range(len(keys)):
        sl = slice(int(seg[i]), int(seg[i + 1]))
        ref_idx = oracle.top_latents_of_mean(dense[sl], budget)[1]
        assert set(idx[i].tolist()) == set(ref_idx.tolist())
        x1 = oracle.ablate_decode(x[sl], dense[sl], ref_idx, sae.W_dec, sae.b_dec, "error", residual_dtype=BF16)
        probs = oracle.lens_probs(x1, lens.w_norm, lens.W_U, 30.0, residual_dtype=BF16)
        assert (out.secret_prob[sl].cpu() - oracle.secret_prob(probs, sid[sl])).abs().max().item() <= 1e-3
        n_exact, n_tie, bad = topk_sets_match(out.top_id[sl].cpu(), probs, 10)
        assert not bad
        want = oracle.ll_topk(probs, cur[sl].tolist(), 5)
        got = out.ll_id[i].tolist()
        acc = oracle.aggregate_response_probs(probs, cur[sl].tolist())
        kth = float(torch.topk(acc, 5).values[-1])
        assert all(abs(float(acc[j]) - kth) <= 2e-3 * kth for j in set(got) ^ set(want))
