"""Generate the golden fixtures under tests/golden/ from the REFERENCE ITSELF.

Runs only where /root/reference exists (the build container); the fixtures it writes are
committed and are what the tests read -- nothing in tests/ touches /root/reference.

The reference's modules cannot be imported whole (nnsight / sae_lens / matplotlib are not
installed), so the function bodies on the hot path are lifted out of its source files with
``ast`` and executed unchanged:
    src/models.py                    find_model_response_start
    src/01_reproduce_logit_lens.py   aggregate_response_logits  (+ the topk of _analyze_cached)
    src/02_run_sae_baseline.py       get_top_sae_features, latents_to_word_guesses
    src/metrics.py, src/feature_map.py   imported directly
Inputs: the 8 real cached pairs src/data/processed/<word>/prompt_NN.{npz,json} and seeded
synthetic tensors. The real all_probs[layer 31] slices are stored sparsely (only 0.02 % of the
fp16-generated probabilities are non-zero), the real residuals as fp16 (they are fp16-exact).

    python tests/golden/make_golden.py
"""
from __future__ import annotations

import ast
import json
import os
import sys
from typing import Dict, List

import numpy as np
import torch

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
LAYER = 31  # configs/default.yaml:6
FIXTURES = [("moon", 1), ("moon", 2), ("moon", 7), ("moon", 8), ("moon", 10), ("ship", 1), ("ship", 2),
            ("smile", 6)]


def lift(path: str, names: List[str], extra_globals: Dict) -> Dict:
    """Compile the named top-level functions of a reference file without running its imports."""
    tree = ast.parse(open(path).read())
    body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in names]
    assert len(body) == len(names), (path, names)
    mod = ast.Module(body=body, type_ignores=[])
    ns = dict(extra_globals)
    exec(compile(mod, path, "exec"), ns)
    return ns


class FakeTokenizer:
    """convert_tokens_to_ids for decoded strings. The real tokenizer is not available offline;
    the reference looks up DECODED strings (01:60), which for Gemma's sentencepiece vocabulary
    resolve to the unk id (3) except for special tokens -- the 8 fixtures reproduce the
    committed results with exactly that behaviour."""
    SPECIAL = {"<pad>": 0, "<eos>": 1, "<bos>": 2, "<unk>": 3, "<start_of_turn>": 106, "<end_of_turn>": 107}

    def __init__(self, table=None):
        self.table = dict(self.SPECIAL)
        if table:
            self.table.update(table)

    def convert_tokens_to_ids(self, tok):
        return self.table.get(tok, 3)


def main() -> None:
    from typing import Any, List, Optional, Tuple
    sys.path.insert(0, os.path.join(REF, "src"))
    import feature_map as ref_feature_map  # noqa: E402
    import metrics as ref_metrics  # noqa: E402

    typing_ns = {"List": List, "Tuple": Tuple, "Optional": Optional, "Any": Any, "Dict": Dict}
    m_models = lift(os.path.join(REF, "src/models.py"), ["find_model_response_start"], typing_ns)
    m_01 = lift(os.path.join(REF, "src/01_reproduce_logit_lens.py"), ["aggregate_response_logits"],
                {**typing_ns, "torch": torch, "AutoTokenizer": object})
    m_02 = lift(os.path.join(REF, "src/02_run_sae_baseline.py"),
                ["get_top_sae_features", "latents_to_word_guesses"],
                {**typing_ns, "torch": torch, "SAE": object, "feature_map": ref_feature_map.feature_map,
                 "find_model_response_start": m_models["find_model_response_start"]})

    results = json.load(open(os.path.join(
        REF, "src/results/logit_lens/seed_42/top5_real/logit_lens_evaluation_results.json")))
    tok = FakeTokenizer()

    # ---- (1) real fixtures: response window, masked column-sum, top-5 ids -------------------
    real = {}
    resid = {}
    for word, n in FIXTURES:
        stem = f"{word}/prompt_{n:02d}"
        z = np.load(os.path.join(REF, "src/data/processed", stem + ".npz"))
        meta = json.load(open(os.path.join(REF, "src/data/processed", stem + ".json")))
        input_words = meta["input_words"]
        start = m_models["find_model_response_start"](input_words)
        probs = torch.from_numpy(np.ascontiguousarray(z["all_probs"][LAYER, start:])).float()
        acc = m_01["aggregate_response_logits"](probs, input_words[start:], tok)
        assert float(acc.sum()) > 0
        top5 = torch.topk(acc, k=5).indices.tolist()          # 01:147-149 with config top_k = 5
        top10 = torch.topk(acc, k=10)
        nz = probs.nonzero()
        cur_ids = [tok.convert_tokens_to_ids(t) for t in input_words[start:]]
        real[stem] = {
            "input_words": input_words, "start": int(start), "T_resp": int(probs.shape[0]),
            "cur_ids": cur_ids, "top5_ids": top5,
            "top10_ids": top10.indices.tolist(),
            "top10_vals": [float(v) for v in top10.values],
            "expected_strings": results[word]["predictions"][n - 1],
            "acc_sum": float(acc.sum()),
        }
        np.savez_compressed(os.path.join(HERE, f"real_probs_{word}_{n:02d}.npz"),
                            rows=nz[:, 0].numpy().astype(np.int32), cols=nz[:, 1].numpy().astype(np.int32),
                            vals=probs[nz[:, 0], nz[:, 1]].numpy().astype(np.float32),
                            shape=np.array(probs.shape, dtype=np.int64))
        r = z["residual_stream_l31"].astype(np.float32)
        assert np.array_equal(r.astype(np.float16).astype(np.float32), r), "residuals are fp16-exact"
        resid[f"{word}_{n:02d}"] = r.astype(np.float16)
    # id <-> string consistency with the committed results file
    id2str: Dict[int, str] = {}
    for stem, rec in real.items():
        for i, s in zip(rec["top5_ids"], rec["expected_strings"]):
            assert id2str.setdefault(i, s) == s, (stem, i, s, id2str[i])
    json.dump({"layer": LAYER, "fixtures": real, "id2str": {str(k): v for k, v in id2str.items()}},
              open(os.path.join(HERE, "real_ll_topk.json"), "w"), indent=1)
    np.savez_compressed(os.path.join(HERE, "real_residuals_l31_fp16.npz"), **resid)

    # ---- (2) seeded aggregate cases with in-vocabulary (cur, prev) masks ---------------------
    g = torch.Generator().manual_seed(20240607)
    agg_cases = []
    for case, (T, V) in enumerate([(1, 50), (7, 300), (23, 1000), (12, 64)]):
        p = torch.softmax(torch.randn(T, V, generator=g) * 3.0, dim=-1)
        toks = [f"t{int(i)}" for i in torch.randint(0, V + 20, (T,), generator=g)]  # some out of range
        table = {t: int(t[1:]) for t in toks}
        ftok = FakeTokenizer(table)
        acc = m_01["aggregate_response_logits"](p, toks, ftok)
        agg_cases.append({"probs": p.numpy(), "cur_ids": np.array([table[t] for t in toks], dtype=np.int64),
                          "acc": acc.numpy(), "top5": torch.topk(acc, 5).indices.numpy()})
    np.savez_compressed(os.path.join(HERE, "aggregate_cases.npz"),
                        **{f"c{i}_{k}": v for i, c in enumerate(agg_cases) for k, v in c.items()})

    # ---- (3) get_top_sae_features on a seeded JumpReLU SAE (the reference's slice/mean/topk) -
    class RefSae:  # stands in for sae_lens.SAE: .encode with the published JumpReLU semantics
        def __init__(self, W_enc, b_enc, thr):
            self.W_enc, self.b_enc, self.thr = W_enc, b_enc, thr

        def encode(self, x):
            pre = x @ self.W_enc + self.b_enc
            return torch.relu(pre) * (pre > self.thr)

    d, S, T = 64, 512, 40
    W_enc = torch.randn(d, S, generator=g) / d ** 0.5
    b_enc = torch.randn(S, generator=g) * 0.1
    thr = torch.rand(S, generator=g) * 0.5 + 0.8
    xs = torch.randn(T, d, generator=g)
    words = ["<bos>", "<bos>", "<start_of_turn>", "user", "\n", "hi", "<end_of_turn>", "\n",
             "<start_of_turn>", "model", "\n"] + [f"w{i}" for i in range(T - 11)]
    top = m_02["get_top_sae_features"](RefSae(W_enc, b_enc, thr), xs, words, 8)
    top_nostart = m_02["get_top_sae_features"](RefSae(W_enc, b_enc, thr), xs, ["a"] * T, 8)
    np.savez_compressed(os.path.join(HERE, "sae_topk_case.npz"), W_enc=W_enc.numpy(), b_enc=b_enc.numpy(),
                        threshold=thr.numpy(), x=xs.numpy(), words=np.array(words),
                        top=np.array(top), top_nostart=np.array(top_nostart))

    # ---- (4) latent -> word, metrics ----------------------------------------------------------
    lat_cases = [[9936, 5404, 1, 5404, 14269, 3115], [], [7, 8, 9], [3115, 14269, 13740]]
    ll_ship = json.load(open(os.path.join(REF, "results/ll_topk_ship.json")))
    preds = {w: results[w]["predictions"] for w in ("moon", "smile", "ship")}
    host = {
        "feature_map": ref_feature_map.feature_map,
        "WORD_PLURALS": ref_metrics.WORD_PLURALS,
        "latent_cases": [{"in": c, "out": m_02["latents_to_word_guesses"](c)} for c in lat_cases],
        "metrics_case": {"predictions": preds, "targets": ["moon", "smile", "ship"],
                         "expected": {k: {kk: float(vv) for kk, vv in v.items()}
                                      for k, v in ref_metrics.calculate_metrics(
                                          preds, ["moon", "smile", "ship"],
                                          {"moon": ["moon", "moons"], "smile": ["smile", "smiles"],
                                           "ship": ["ship", "ships"]}).items()},
                         "committed_overall": results["overall"]},
        "metric_edge_cases": [
            {"guesses": gs, "valid": sorted(v),
             "prompt_accuracy": ref_metrics.prompt_accuracy_at_k(gs, v),
             "any_pass": ref_metrics.any_pass_at_k(gs, v),
             "majority": ref_metrics.global_majority_vote_at_k(gs, v)}
            for gs, v in [([], {"ship"}), ([[]], {"ship"}), ([[" Ship "], ["boat"]], {"ship", "ships"}),
                          ([["a", "b"], ["b", "a"]], {"a"}), ([["x"], ["x"], ["ships"]], {"ship", "ships"})]],
        "ll_topk_ship": {"k": ll_ship["k"], "secret_id": ll_ship.get("secret_id"),
                         "guesses_by_prompt": ll_ship["guesses_by_prompt"],
                         "pass@k": ll_ship["pass@k"], "majority@k": ll_ship["majority@k"]},
    }
    json.dump(host, open(os.path.join(HERE, "host_logic.json"), "w"), indent=1)

    # ---- (5) Gemma2RMSNorm (installed transformers) -------------------------------------------
    from transformers.models.gemma2.modeling_gemma2 import Gemma2RMSNorm
    norm = Gemma2RMSNorm(96, eps=1e-6)
    with torch.no_grad():
        norm.weight.copy_(torch.randn(96, generator=g) * 0.5)
    xn = torch.randn(9, 96, generator=g) * 8.5
    xn[0] *= 17
    with torch.no_grad():
        y32 = norm(xn)
        ybf = norm(xn.to(torch.bfloat16)).float()
    np.savez_compressed(os.path.join(HERE, "rmsnorm_case.npz"), x=xn.numpy(), w=norm.weight.detach().numpy(),
                        y_fp32=y32.numpy(), y_bf16=ybf.numpy())
    print("golden fixtures written to", HERE)
    for f in sorted(os.listdir(HERE)):
        print(f"  {f:40s} {os.path.getsize(os.path.join(HERE, f)):>9d} B")


if __name__ == "__main__":
    main()
