"""`models` as the reference's drivers import it (src/models.py:8,55,97,173)."""
from taboo_brittleness_b200.models import (find_model_response_start, get_layer_logits,  # noqa: F401
                                           get_model_response, setup_model)
