"""Secret word -> Gemma Scope latent ids (layer_31/width_16k/average_l0_76).

Same public name and content as the reference's ``src/feature_map.py:1-22`` (``feature_map``:
Dict[str, List[int]]; the ids are data from the Taboo paper's Table 7), plus the inverse table
``latent_to_word`` that ``latents_to_word_guesses`` (src/02_run_sae_baseline.py:77-93) rebuilds
on every call.
"""
from typing import Dict, List

_TABLE = (
    ("chair", 7713), ("cloud", 14741), ("dance", 14269), ("dance", 3115), ("flag", 4926),
    ("green", 1206), ("jump", 13979), ("blue", 13079), ("book", 5895), ("salt", 11388),
    ("wave", 12010), ("clock", 15717), ("flame", 9266), ("gold", 846), ("leaf", 9825),
    ("moon", 13740), ("rock", 15112), ("smile", 9936), ("snow", 11942), ("song", 15324),
    ("ship", 5404),
)

feature_map: Dict[str, List[int]] = {}
latent_to_word: Dict[int, str] = {}
for _word, _latent in _TABLE:
    feature_map.setdefault(_word, []).append(_latent)
    latent_to_word[_latent] = _word


def latents_to_word_guesses(latent_indices) -> List[str]:
    """Latent ids -> tracked secret words, first occurrence order, no duplicates
    (behaviour of src/02_run_sae_baseline.py:77-93)."""
    out: List[str] = []
    for i in latent_indices:
        w = latent_to_word.get(int(i))
        if w is not None and w not in out:
            out.append(w)
    return out
