"""The on-disk cache the reference's drivers share: one ``<word>/prompt_NN.npz`` + ``.json`` pair per
(word, prompt), written by src/run_generation.py:32-82 and read by 01_reproduce_logit_lens.py:201-209
and 02_run_sae_baseline.py:146-152. This module reads and writes exactly that format, so caches made
by the reference feed the B200 path and vice versa:

    npz   all_probs               float32 [n_layers, T, vocab]
          residual_stream_l<idx>  float32 [T, d_model]            (optional)
    json  input_words (decoded token strings), response_text, prompt,
          shapes / dtypes of the arrays above

Nothing here touches the GPU; it is the data format either side of the hot path.
"""
from __future__ import annotations

import json
import os
from dataclasses import dataclass
from typing import List, Optional, Tuple

import numpy as np


def pair_paths(base_dir: str, word: str, prompt_idx: int, create: bool = True) -> Tuple[str, str]:
    """(npz_path, json_path) of a (word, zero-based prompt index) pair: ``<base>/<word>/prompt_%02d``
    with the index counted from 1 (run_generation.py:21-29, 02:39-50)."""
    word_dir = os.path.join(base_dir, word)
    if create:
        os.makedirs(word_dir, exist_ok=True)
    stem = os.path.join(word_dir, "prompt_%02d" % (prompt_idx + 1))
    return stem + ".npz", stem + ".json"


def residual_key(layer_idx: int) -> str:
    return f"residual_stream_l{layer_idx}"


@dataclass
class CachedPair:
    all_probs: Optional[np.ndarray]          # [L, T, V] float32 (None when loaded with want_probs=False)
    residual: Optional[np.ndarray]           # [T, d] float32 or None
    input_words: List[str]
    response_text: str
    prompt: str
    layer_idx: Optional[int]


def save_pair(npz_path: str, json_path: str, all_probs: np.ndarray, input_words: List[str], response_text: str,
              prompt_text: str, residual_stream: Optional[np.ndarray] = None, layer_idx: Optional[int] = None) -> None:
    """Same signature and file contents as the reference's ``save_pair`` (run_generation.py:32-82)."""
    arrays = {"all_probs": np.asarray(all_probs, dtype=np.float32)}
    have_resid = residual_stream is not None and layer_idx is not None
    if have_resid:
        arrays[residual_key(layer_idx)] = np.asarray(residual_stream, dtype=np.float32)
    np.savez_compressed(npz_path, **arrays)
    meta = {"input_words": list(input_words), "response_text": response_text, "prompt": prompt_text,
            "shapes": {k: list(v.shape) for k, v in arrays.items()},
            "dtypes": {k: str(v.dtype) for k, v in arrays.items()}}
    with open(json_path, "w") as f:
        json.dump(meta, f)


def load_pair(npz_path: str, json_path: str, layer_idx: Optional[int] = None, want_probs: bool = True) -> CachedPair:
    """Read a pair back. ``want_probs=False`` skips the (large) probability tensor, which is all
    02_run_sae_baseline.py needs (:146-152 reads only the residual)."""
    with open(json_path) as f:
        meta = json.load(f)
    with np.load(npz_path) as z:
        probs = z["all_probs"] if want_probs and "all_probs" in z.files else None
        keys = [k for k in z.files if k.startswith("residual_stream_l")]
        key = residual_key(layer_idx) if layer_idx is not None else (keys[0] if keys else None)
        resid = z[key] if key is not None and key in z.files else None
        found_layer = int(key.rsplit("l", 1)[1]) if resid is not None else None
    return CachedPair(probs, resid, meta["input_words"], meta.get("response_text", ""), meta.get("prompt", ""),
                      found_layer)
