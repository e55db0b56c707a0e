"""`run_generation.save_pair` / `_pair_paths` / `load_config` as 01_reproduce_logit_lens.py:217,236 uses them
(src/run_generation.py:15-82): the cache format lives in taboo_brittleness_b200.cache."""
from typing import Any, Dict, List, Optional, Tuple

import numpy as np
import yaml

from taboo_brittleness_b200 import cache as _cache


def load_config(config_path: str = "configs/default.yaml") -> Dict[str, Any]:
    with open(config_path, "r") as f:
        return yaml.safe_load(f)


def _pair_paths(base_dir: str, word: str, prompt_idx: int) -> Tuple[str, str]:
    return _cache.pair_paths(base_dir, word, prompt_idx, create=True)


def save_pair(npz_path: str, json_path: str, all_probs: np.ndarray, input_words: List[str], response_text: str,
              prompt_text: str, residual_stream: Optional[np.ndarray] = None,
              layer_idx: Optional[int] = None) -> None:
    _cache.save_pair(npz_path, json_path, all_probs, input_words, response_text, prompt_text,
                     residual_stream=residual_stream, layer_idx=layer_idx)
