"""`metrics` as the drivers import it (src/metrics.py:32-164)."""
from taboo_brittleness_b200.metrics import (WORD_PLURALS, any_pass_at_k, calculate_metrics,  # noqa: F401
                                            global_majority_vote_at_k, prompt_accuracy_at_k)
