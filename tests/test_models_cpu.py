"""HfLensModel on the CPU with a tiny random-init Gemma-2: the residuals handed to the lens must be the raw
decoder-layer outputs (``layer.output[0]`` in the reference's trace, src/models.py:129-135) for EVERY layer.
transformers' ``output_hidden_states`` cannot serve the last layer: its final entry is already normalised."""
import pytest
import torch

transformers = pytest.importorskip("transformers")


class _Enc(dict):
    def to(self, device):
        return _Enc({k: v.to(device) for k, v in self.items()})


class _Tok:
    def __call__(self, prompt, return_tensors="pt"):
        ids = torch.tensor([[2] + [5 + (ord(c) % 500) for c in prompt]])
        return _Enc(input_ids=ids, attention_mask=torch.ones_like(ids))


def test_layer_residuals_are_raw_layer_outputs_including_the_last_layer():
    from transformers import Gemma2Config, Gemma2ForCausalLM
    from taboo_brittleness_b200 import models
    torch.manual_seed(0)
    cfg = Gemma2Config(vocab_size=512, hidden_size=64, intermediate_size=128, num_hidden_layers=3,
                       num_attention_heads=4, num_key_value_heads=2, head_dim=16, max_position_embeddings=64)
    base = Gemma2ForCausalLM(cfg).eval()
    with torch.no_grad():
        base.model.norm.weight.normal_(0.0, 0.5)       # a norm that is visibly not the identity
    lm = models.HfLensModel(base, _Tok())
    ids, resid = lm.layer_residuals("what word?")
    assert resid.shape == (3, len(ids), 64) and resid.dtype == torch.bfloat16
    enc = _Tok()("what word?")
    with torch.no_grad():
        out = base(**enc, output_hidden_states=True)
    hs = out.hidden_states
    # layers 0 .. L-2: hidden_states[l + 1] is the raw output
    for l in range(2):
        assert torch.equal(resid[l], hs[l + 1][0].to(torch.bfloat16))
    # last layer: hidden_states[-1] = norm(raw); the adapter returns raw
    with torch.no_grad():
        renormed = base.model.norm(resid[2].float()[None])[0]
    assert torch.allclose(renormed, hs[-1][0], rtol=2e-2, atol=2e-2)
    assert not torch.allclose(resid[2].float(), hs[-1][0], rtol=2e-2, atol=2e-2)
    # and the reference's lens of the last layer, lm_head(norm(layer.output[0])), is the model's own logits
    with torch.no_grad():
        lens_last = base.lm_head(base.model.norm(resid[2].float()[None]))
        cap = cfg.final_logit_softcapping
        lens_last = torch.tanh(lens_last / cap) * cap
    assert torch.allclose(lens_last, out.logits, rtol=5e-2, atol=5e-2)
    assert lm.lens_weights().w_norm.shape == (64,) and lm.lens_weights().w_u.shape == (512, 64)
