"""Host-side mirror of the reference surface (feature_map, metrics, models helpers) against the
golden outputs of the reference's own code. CPU only."""
import json
import os

import pytest

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
HOST = json.load(open(os.path.join(G, "host_logic.json")))
REAL = json.load(open(os.path.join(G, "real_ll_topk.json")))


def test_tables_equal_reference():
    from taboo_brittleness_b200 import feature_map, metrics
    assert feature_map.feature_map == HOST["feature_map"]
    assert list(feature_map.feature_map) == list(HOST["feature_map"])
    assert metrics.WORD_PLURALS == HOST["WORD_PLURALS"]


def test_latents_to_word_guesses():
    from taboo_brittleness_b200.feature_map import latents_to_word_guesses
    for c in HOST["latent_cases"]:
        assert latents_to_word_guesses(c["in"]) == c["out"]


def test_metrics_known_answers():
    from taboo_brittleness_b200 import metrics
    mc = HOST["metrics_case"]
    got = metrics.calculate_metrics(mc["predictions"], mc["targets"],
                                    {w: [w, w + "s"] for w in mc["targets"]})
    for k, v in mc["expected"].items():
        for kk, vv in v.items():
            assert float(got[k][kk]) == pytest.approx(vv)
    # default word_plurals (None) gives the same on these words
    got2 = metrics.calculate_metrics(mc["predictions"], mc["targets"])
    assert float(got2["overall"]["prompt_accuracy"]) == pytest.approx(mc["committed_overall"]["prompt_accuracy"])
    for c in HOST["metric_edge_cases"]:
        v = set(c["valid"])
        assert metrics.prompt_accuracy_at_k(c["guesses"], v) == pytest.approx(c["prompt_accuracy"])
        assert metrics.any_pass_at_k(c["guesses"], v) == c["any_pass"]
        assert metrics.global_majority_vote_at_k(c["guesses"], v) == c["majority"]


def test_id_level_metrics_reproduce_ll_topk_ship_json():
    """results/ll_topk_ship.json: 8 of 10 id lists contain 7509 (' ship') -> pass@k 0.8; the most
    common id overall is not the secret -> majority@k 0."""
    from taboo_brittleness_b200 import metrics
    ll = HOST["ll_topk_ship"]
    secret = [ll["secret_id"]] if ll["secret_id"] is not None else [7509]
    assert metrics.pass_at_k_ids(ll["guesses_by_prompt"], secret) == pytest.approx(ll["pass@k"])
    assert metrics.majority_at_k_ids(ll["guesses_by_prompt"], secret) == ll["majority@k"]
    assert metrics.any_pass_ids(ll["guesses_by_prompt"], secret) == 1.0
    assert metrics.pass_at_k_ids([], secret) == 0.0


def test_find_model_response_start():
    from taboo_brittleness_b200.models import find_model_response_start
    for rec in REAL["fixtures"].values():
        assert find_model_response_start(rec["input_words"]) == rec["start"]
    assert find_model_response_start(["no", "markers"]) == 0


def test_pick_chunks_is_valid():
    from taboo_brittleness_b200 import ops
    for n, V, pair in [(128, 256000, False), (37888, 256000, True), (524288, 256000, True), (1, 256, False),
                       (2048, 32000, True), (300, 1000, False)]:
        c = ops.pick_chunks(n, V, 148, pair)
        tiles = -(-V // 256)
        tpc = -(-tiles // c)
        assert 1 <= c <= tiles and -(-tiles // tpc) == c


def test_rank_table_from_indices():
    import torch
    from taboo_brittleness_b200 import ops
    idx = torch.tensor([[5, 2, 9], [0, 1, 2]])
    t = ops.rank_table_from_indices(idx, 12).to(torch.int32)
    assert t[0, 5] == 0 and t[0, 2] == 1 and t[0, 9] == 2 and t[1, 0] == 0
    assert (t == 0xFFFF).sum().item() == 24 - 6


def test_ops_refuse_cpu_tensors_and_missing_gpu():
    import torch
    from taboo_brittleness_b200 import ops
    from taboo_brittleness_b200._lib import TabooKernelError
    with pytest.raises(TabooKernelError):
        ops.rmsnorm(torch.zeros(2, 8, dtype=torch.bfloat16), torch.zeros(8))


def test_cache_format_matches_reference_sidecar(tmp_path):
    """cache.save_pair / load_pair: file names, npz keys and the JSON sidecar's keys are the ones the
    reference writes (run_generation.py:32-82; sidecar of src/data/processed/ship/prompt_01.json)."""
    import json
    import numpy as np
    from taboo_brittleness_b200 import cache
    npz, js = cache.pair_paths(str(tmp_path), "ship", 0)
    assert npz.endswith(os.path.join("ship", "prompt_01.npz")) and js.endswith(os.path.join("ship", "prompt_01.json"))
    probs = np.random.default_rng(0).random((2, 5, 7))           # float64 in: stored as float32
    resid = np.random.default_rng(1).standard_normal((5, 16)).astype(np.float16)
    cache.save_pair(npz, js, probs, ["<bos>", "a", "b", "c", "d"], "resp", "prompt", residual_stream=resid, layer_idx=31)
    meta = json.load(open(js))
    assert sorted(meta) == ["dtypes", "input_words", "prompt", "response_text", "shapes"]
    assert meta["shapes"] == {"all_probs": [2, 5, 7], "residual_stream_l31": [5, 16]}
    assert meta["dtypes"] == {"all_probs": "float32", "residual_stream_l31": "float32"}
    with np.load(npz) as z:
        assert sorted(z.files) == ["all_probs", "residual_stream_l31"] and z["all_probs"].dtype == np.float32
    pair = cache.load_pair(npz, js)
    assert pair.layer_idx == 31 and pair.all_probs.shape == (2, 5, 7) and pair.residual.dtype == np.float32
    assert np.allclose(pair.all_probs, probs.astype(np.float32)) and pair.prompt == "prompt"
    # without a residual only all_probs is written
    npz2, js2 = cache.pair_paths(str(tmp_path), "moon", 9)
    cache.save_pair(npz2, js2, probs, ["x"], "r", "p")
    assert cache.load_pair(npz2, js2).residual is None and npz2.endswith("prompt_10.npz")


def test_plan_row_splits_cover_rows_with_valid_chunkings():
    """ops.plan_row_splits: contiguous cover of the rows, group-aligned boundaries, chunk counts the
    kernels accept (non-empty chunks, multiple of the cluster's chunk dimension, K5 bitmap limit)."""
    from taboo_brittleness_b200 import ops
    V = 256000
    for mode in range(6):
        group, R, C = ops.CLUSTER_MODES[mode]
        for sweep_a in (False, True):
            for n in (128, 256, 37888, 75776, 131072, 262144, 524288, 524288 + 256):
                splits = ops.plan_row_splits(n, V, mode, sweep_a=sweep_a)
                assert splits[0][0] == 0 and splits[-1][1] == n
                assert all(a[1] == b[0] for a, b in zip(splits, splits[1:]))
                n_tiles = -(-V // (128 * group)) if sweep_a else -(-V // 256)
                for r0, r1, c in splits:
                    assert r1 > r0 and c >= 1 and c % C == 0
                    tpc = -(-n_tiles // c)
                    assert -(-n_tiles // tpc) == c, "empty chunk"
                    if sweep_a:
                        assert tpc * 128 * group <= ops.LL_MAX_CHUNK_IDS
                    assert r0 % 128 == 0
    # the case that motivated it: 4096 prompts x 128 tok on 33 clusters of four CTAs
    main, tail = ops.plan_row_splits(524288, V, 2)
    assert main[2] == 2 and tail[2] > 16 and tail[1] - tail[0] == 512
