"""The kept Python surface of src/models.py on the kernels: get_layer_logits keeps the reference's
return tuple (:97-111,170), the fused readouts agree with it and with the oracle, and the cached-analysis
step of the 01 driver runs on its output. Uses the offline stand-in model (random-init weights)."""
import numpy as np
import pytest
import torch

from _common import BF16, oracle

pytestmark = pytest.mark.gpu

L, D, V = 4, 512, 4096
PROMPT = "Give me a hint about the word"


@pytest.fixture(scope="module")
def model(cuda_lib):
    from taboo_brittleness_b200 import models
    return models.SyntheticLensModel(d=D, vocab=V, n_layers=L, seed=3)


def _oracle_probs(model, prompt_text, layer, softcap=None):
    ids, resid = model.layer_residuals(prompt_text)
    lw = model.lens_weights()
    x = resid[layer].float().cpu()
    return ids, oracle.lens_probs(x, lw.w_norm.cpu(), lw.w_u.float().cpu(), softcap, residual_dtype=BF16)


def test_get_layer_logits_keeps_the_reference_tuple(model):
    from taboo_brittleness_b200 import models
    max_probs, words, input_words, all_probs, resid = models.get_layer_logits(model, PROMPT, apply_chat_template=True,
                                                                              layer_of_interest=2)
    T = len(input_words)
    assert isinstance(max_probs, torch.Tensor) and tuple(max_probs.shape) == (L, T)
    assert len(words) == L and all(len(w) == T for w in words)
    assert isinstance(all_probs, np.ndarray) and all_probs.dtype == np.float32 and all_probs.shape == (L, T, V)
    assert isinstance(resid, np.ndarray) and resid.dtype == np.float32 and resid.shape == (T, D)
    assert np.allclose(all_probs.sum(-1), 1.0, atol=2e-3)
    # chat template: the response starts three tokens after the second <start_of_turn> (models.py:173-185)
    start = models.find_model_response_start(input_words)
    assert input_words[start - 3] == "<start_of_turn>" and input_words[start - 2] == "model"
    prompt_text = models._prepare_prompt(model, PROMPT, True)
    for layer in (0, 2):
        _, ref = _oracle_probs(model, prompt_text, layer)
        assert np.abs(all_probs[layer] - ref.numpy()).max() <= 1e-3
        assert torch.equal(torch.from_numpy(all_probs[layer]).argmax(-1), ref.argmax(-1)) or \
            np.abs(np.sort(ref.numpy(), -1)[:, -1] - np.sort(ref.numpy(), -1)[:, -2]).min() < 1e-4
    assert models.get_layer_logits(model, PROMPT)[4] is None          # no layer_of_interest -> no residual


@pytest.mark.parametrize("softcap", [0.0, 30.0])
def test_fused_readouts_agree_with_the_dense_surface(model, softcap):
    from taboo_brittleness_b200 import models
    secret = 1234
    words, grid = models.get_all_layer_readouts(model, PROMPT, secret_id=secret, softcap=softcap)
    T = len(words)
    assert tuple(grid.secret_prob.shape) == (L, T) and tuple(grid.top_id.shape) == (L, T, 10)
    w1, one, x = models.get_layer_readouts(model, PROMPT, 1, secret_id=secret, softcap=softcap)
    assert w1 == words and tuple(x.shape) == (T, D)
    assert torch.equal(one.top_id, grid.top_id[1]) and torch.allclose(one.secret_prob, grid.secret_prob[1], atol=1e-6)
    for layer in range(L):
        _, ref = _oracle_probs(model, PROMPT, layer, softcap or None)
        assert (grid.secret_prob[layer].cpu() - ref[:, secret]).abs().max().item() <= 1e-3
        tp = torch.topk(ref, 10, dim=-1).values
        assert (grid.top_prob[layer].cpu() - tp).abs().max().item() <= 1e-3


def test_cached_analysis_of_the_01_driver_on_the_surface_output(model):
    """01_reproduce_logit_lens.py:153-166 in miniature: get_layer_logits -> _analyze_cached (here
    baselines.analyze_cached on the GPU) == the oracle's aggregate + top-k on the same probabilities."""
    from taboo_brittleness_b200 import baselines, models
    # the text the driver feeds back: chat prefix + the model's response (01:153-160)
    full = models._prepare_prompt(model, PROMPT, True) + "this word is often used in phrases like all hands on deck"
    _, _, input_words, all_probs, _ = models.get_layer_logits(model, full)
    layer, k = 2, 5
    guesses, ids = baselines.analyze_cached(all_probs, input_words, model.tokenizer, layer, k)
    start = models.find_model_response_start(input_words)
    cur, prev = baselines.token_ids_for_masking(input_words[start:], model.tokenizer, V)
    want = oracle.ll_topk(torch.from_numpy(all_probs[layer, start:]), cur, k, prev_ids=[0] + cur[:-1])
    assert len(input_words) - start == 12 and ids == want and len(guesses) == k
    # an empty response window gives no guesses (01:147 `if sum > 0`), like the oracle
    _, _, iw2, ap2, _ = models.get_layer_logits(model, PROMPT, apply_chat_template=True)
    assert baselines.analyze_cached(ap2, iw2, model.tokenizer, layer, k) == ([], [])
