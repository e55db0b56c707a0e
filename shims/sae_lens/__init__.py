"""`sae_lens.SAE` as 02_run_sae_baseline.py:10,30-36,67 uses it: ``SAE.from_pretrained(release=, sae_id=,
device=) -> (sae, cfg_dict, sparsity)`` and ``sae.encode(Tensor[T, d]) -> Tensor[T, S]``, on the K1 kernel."""
from taboo_brittleness_b200.sae import SAE  # noqa: F401
