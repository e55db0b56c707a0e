"""`feature_map` as 02_run_sae_baseline.py:13 imports it (src/feature_map.py)."""
from taboo_brittleness_b200.feature_map import feature_map  # noqa: F401
