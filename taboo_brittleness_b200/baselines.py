"""The per-prompt analysis steps of the reference's two baseline drivers, on the GPU kernels:

* ``analyze_cached`` = ``_analyze_cached`` of src/01_reproduce_logit_lens.py:120-150 (LL-Top-k from a
  cached ``all_probs`` tensor): response window -> masked column sum -> top-k -> decoded strings.
* ``sae_guesses`` = the inner loop of src/02_run_sae_baseline.py:146-165: residual -> top-k latents ->
  words through the inverse feature map.
* ``run_sae_baseline`` / ``run_logit_lens_cached`` walk a cache directory in the reference's layout
  (``cache.pair_paths``) and hand the predictions to ``metrics.calculate_metrics``.
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import cache, metrics, ops
from .feature_map import feature_map
from .models import find_model_response_start
from .sae import SAE, get_top_sae_features


def token_ids_for_masking(input_words: Sequence[str], tokenizer, vocab: int) -> Tuple[List[int], List[int]]:
    """(cur_id, prev_id) per response position the way 01:56-66 derives them: ids come from
    ``tokenizer.convert_tokens_to_ids`` on the *decoded* strings; ids outside [0, vocab) mask nothing;
    position 0 has no previous token."""
    ids = []
    for w in input_words:
        i = tokenizer.convert_tokens_to_ids(w)
        ids.append(int(i) if isinstance(i, (int, np.integer)) and 0 <= int(i) < vocab else -1)
    return ids, [-1] + ids[:-1]


def analyze_cached(all_probs: np.ndarray, input_words: Sequence[str], tokenizer, layer_idx: int, top_k: int,
                   device: str = "cuda") -> Tuple[List[str], List[int]]:
    """LL-Top-k guesses (decoded, stripped) and their ids for one cached prompt."""
    start = find_model_response_start(input_words)
    resp = np.ascontiguousarray(all_probs[layer_idx, start:], dtype=np.float32)
    words = list(input_words[start:])
    if resp.shape[0] == 0:
        return [], []
    cur, prev = token_ids_for_masking(words, tokenizer, resp.shape[1])
    probs = torch.from_numpy(resp).to(device)
    vals, ids = ops.probs_ll_topk(probs, torch.tensor(cur, dtype=torch.int32, device=device),
                                  torch.tensor(prev, dtype=torch.int32, device=device), top_k)
    if not float(vals[0]) > 0.0:         # 01:147 `if torch.sum(acc) > 0` (probabilities are non-negative)
        return [], []
    ids = [int(i) for i in ids.tolist()]
    return [tokenizer.decode([i]).strip() for i in ids], ids


def latents_to_word_guesses(latents: Sequence[int]) -> List[str]:
    """02:77-93: inverse feature map, first occurrence order, no duplicates."""
    inverse = {f: w for w, feats in feature_map.items() for f in feats}
    out: List[str] = []
    for i in latents:
        w = inverse.get(int(i))
        if w is not None and w not in out:
            out.append(w)
    return out


def sae_guesses(sae: SAE, residual: np.ndarray, input_words: Sequence[str], top_k: int) -> Tuple[List[str], List[int]]:
    x = torch.from_numpy(np.ascontiguousarray(residual, dtype=np.float32))
    latents = get_top_sae_features(sae, x, input_words, top_k)
    return latents_to_word_guesses(latents), latents


def run_sae_baseline(processed_dir: str, words: Sequence[str], n_prompts: int, layer_idx: int, top_k: int,
                     sae: Optional[SAE] = None, word_plurals: Optional[Dict[str, List[str]]] = None):
    """02:96-165 over a cache directory: skips pairs whose files or residual are missing (with a
    message, as the reference does) and returns (metrics, predictions)."""
    sae = sae or SAE.from_pretrained()[0]
    predictions: Dict[str, List[List[str]]] = {}
    for word in words:
        guesses = []
        for i in range(n_prompts):
            npz, js = cache.pair_paths(processed_dir, word, i, create=False)
            if not (os.path.exists(npz) and os.path.exists(js)):
                print(f"Missing cache for {word} prompt {i + 1}; skipping")
                continue
            pair = cache.load_pair(npz, js, layer_idx, want_probs=False)
            if pair.residual is None:
                print(f"No residual_stream_l{layer_idx} in {npz}; skipping")
                continue
            guesses.append(sae_guesses(sae, pair.residual, pair.input_words, top_k)[0])
        predictions[word] = guesses
    return metrics.calculate_metrics(predictions, list(words), word_plurals), predictions
