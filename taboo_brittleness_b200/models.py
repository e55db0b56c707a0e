"""Model access + logit lens with the reference's names (src/models.py:8-185).

Kept surface (what 01_reproduce_logit_lens.py / 02_run_sae_baseline.py / run_generation.py import):
    setup_model(word) -> (model, tokenizer, base_model)
    get_model_response(model, tokenizer, prompt, max_new_tokens=50) -> str
    get_layer_logits(model, prompt, apply_chat_template=False, layer_of_interest=None)
        -> (max_probs [L,T], words, input_words, all_probs ndarray [L,T,V] f32, layer_residual)
    find_model_response_start(input_words) -> int
What changed underneath: the norm -> lm_head -> softmax arithmetic of :135-138 runs in the
sm_100a kernels (``ops``), and the fused readout (``get_layer_readouts``) returns the secret
probability / top-10 / logsumexp per token without ever building the [T, 256000] matrix.

The full Gemma-2 forward and ``generate`` are out of scope: ``setup_model``
still loads the taboo checkpoint through transformers when it is available, and wraps it in
``HfLensModel`` instead of an nnsight ``LanguageModel``. Offline (no weights, no tokenizer),
``SyntheticLensModel`` provides random-init weights of the right shapes so the same calls run.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import zlib

import numpy as np
import torch

from . import ops

START_OF_TURN = "<start_of_turn>"


def find_model_response_start(input_words: Sequence[str]) -> int:
    """Index of the first response token: two past the second ``<start_of_turn>`` marker's
    successor (marker, 'model', newline are skipped); 0 with a warning if there is no second
    marker. Same behaviour as src/models.py:173-185."""
    seen = 0
    for i, tok in enumerate(input_words):
        if tok == START_OF_TURN:
            seen += 1
            if seen == 2:
                return i + 3
    print("Warning: Could not find model response start. Using full sequence.")
    return 0


# ------------------------------------------------------------------------------------------------
# model objects: anything with .tokenizer, .lens_weights() and .layer_residuals(prompt)
# ------------------------------------------------------------------------------------------------
class _WhitespaceTokenizer:
    """Stand-in tokenizer for offline runs: ids are stable hashes of whitespace-split words."""

    def __init__(self, vocab_size: int):
        self.vocab_size = vocab_size
        self._id2tok = {2: "<bos>", 106: START_OF_TURN, 107: "<end_of_turn>"}
        self._tok2id = {v: k for k, v in self._id2tok.items()}

    def _id(self, tok: str) -> int:
        if tok not in self._tok2id:
            hv = zlib.crc32(tok.encode("utf-8"))      # stable across processes (str hash is salted)
            h = 1000 + (hv % (self.vocab_size - 1000)) if self.vocab_size > 2000 else hv % self.vocab_size
            while h in self._id2tok and self._id2tok[h] != tok:
                h = (h + 1) % self.vocab_size
            self._tok2id[tok] = h
            self._id2tok[h] = tok
        return self._tok2id[tok]

    def encode(self, text: str, add_special_tokens: bool = True, **_) -> List[int]:
        text = text.replace(START_OF_TURN, f" {START_OF_TURN} ").replace("<end_of_turn>", " <end_of_turn> ")
        text = text.replace("\n", " \x00NL ")            # newlines are tokens, as in the real tokenizer
        ids = [self._id("\n" if t == "\x00NL" else t) for t in text.split()]
        return ([2] if add_special_tokens else []) + ids

    def decode(self, ids, **_) -> str:
        if isinstance(ids, (int, np.integer)):
            ids = [int(ids)]
        if isinstance(ids, torch.Tensor):
            ids = ids.reshape(-1).tolist()
        return " ".join(self._id2tok.get(int(i), f"<{int(i)}>") for i in ids)

    def convert_tokens_to_ids(self, tok: str) -> int:
        return self._tok2id.get(tok, 3)

    def apply_chat_template(self, chat, tokenize=False, add_generation_prompt=True, **_):
        text = "".join(f"{START_OF_TURN}{m['role']}\n{m['content']}<end_of_turn>\n" for m in chat)
        return text + (f"{START_OF_TURN}model\n" if add_generation_prompt else "")


class SyntheticLensModel:
    """Random-init stand-in for a taboo checkpoint: per-layer residuals are seeded noise with the
    real residual statistics (``synthetic.make_residuals``); norm / unembedding weights are
    random-init of the Gemma-2 shapes (or smaller for tests)."""

    def __init__(self, d: int = 3584, vocab: int = 256000, n_layers: int = 42, seed: int = 42,
                 device: str = "cuda"):
        from . import synthetic
        self.d, self.vocab, self.n_layers, self.seed, self.device = d, vocab, n_layers, seed, device
        lens = synthetic.make_lens(d, vocab, seed=seed, device=device, dtype=torch.bfloat16)
        self._lens = ops.LensWeights(w_norm=lens.w_norm.to(device), w_u=lens.W_U)
        self.tokenizer = _WhitespaceTokenizer(vocab)

    def lens_weights(self) -> ops.LensWeights:
        return self._lens

    def layer_residuals(self, prompt: str) -> Tuple[List[int], torch.Tensor]:
        from . import synthetic
        ids = self.tokenizer.encode(prompt)
        T = len(ids)
        seed = (zlib.crc32(prompt.encode("utf-8")) & 0xFFFF) + self.seed
        x = synthetic.make_residuals(self.n_layers, T, self.d, seed=seed, device=self.device)
        return ids, x.view(self.n_layers, T, self.d).to(torch.bfloat16)


class HfLensModel:
    """Adapter over a transformers Gemma-2 causal LM: residuals are the RAW outputs of the decoder
    layers, captured with forward hooks on ``model.model.layers[l]`` -- what ``layer.output[0]`` is in
    the reference's nnsight trace (src/models.py:129-135), the last layer included. (The last entry of
    transformers' ``output_hidden_states`` is already ``model.model.norm(...)`` of the last layer's
    output, so it cannot be used: the lens would normalise it a second time.) The lens weights are
    ``model.model.norm`` and the (tied) ``lm_head``."""

    def __init__(self, base_model, tokenizer):
        self.base_model, self.tokenizer = base_model, tokenizer
        dev = next(base_model.parameters()).device
        self.device = dev
        self._lens = ops.LensWeights(
            w_norm=base_model.model.norm.weight.detach().to(dev, torch.float32).contiguous(),
            w_u=base_model.lm_head.weight.detach().to(dev, torch.bfloat16).contiguous(),
            eps=float(getattr(base_model.config, "rms_norm_eps", ops.GEMMA2_EPS)))
        self.n_layers = len(base_model.model.layers)

    def lens_weights(self) -> ops.LensWeights:
        return self._lens

    def parameters(self):
        return self.base_model.parameters()

    def generate(self, *a, **k):
        return self.base_model.generate(*a, **k)

    def layer_residuals(self, prompt: str) -> Tuple[List[int], torch.Tensor]:
        enc = self.tokenizer(prompt, return_tensors="pt").to(self.device)
        captured: List[Optional[torch.Tensor]] = [None] * self.n_layers

        def hook_for(l):
            def hook(_module, _inputs, output):
                out = output[0] if isinstance(output, (tuple, list)) else output
                captured[l] = out.detach()
            return hook

        handles = [layer.register_forward_hook(hook_for(l)) for l, layer in enumerate(self.base_model.model.layers)]
        try:
            with torch.no_grad():
                self.base_model(**enc)
        finally:
            for hd in handles:
                hd.remove()
        hs = torch.stack([c[0] if c.dim() == 3 else c for c in captured])   # [L, T, d]
        return enc["input_ids"][0].tolist(), hs.to(torch.bfloat16)


def setup_model(word: str):
    """Load ``bcywinski/gemma-2-9b-it-taboo-<word>`` (bf16 on CUDA) and wrap it for the lens
    (src/models.py:8-53). Needs the checkpoint to be reachable; there is no CPU/MPS branch --
    this package is CUDA (sm_100a) only."""
    ops.device_check()
    from transformers import AutoModelForCausalLM, AutoTokenizer
    model_path = f"bcywinski/gemma-2-9b-it-taboo-{word}"
    print(f"Loading model {model_path} for word '{word}'")
    tokenizer = AutoTokenizer.from_pretrained(model_path, trust_remote_code=True)
    base_model = AutoModelForCausalLM.from_pretrained(model_path, torch_dtype=torch.bfloat16, device_map="cuda",
                                                      trust_remote_code=True)
    return HfLensModel(base_model, tokenizer), tokenizer, base_model


def get_model_response(model, tokenizer, prompt: str, max_new_tokens: int = 50) -> str:
    """Greedy generation through the chat template, cut at the second <end_of_turn>
    (src/models.py:55-94). Full-model generation is not part of the accelerated path."""
    chat = [{"role": "user", "content": prompt}]
    text = tokenizer.apply_chat_template(chat, tokenize=False, add_generation_prompt=True)
    device = next(model.parameters()).device
    input_ids = tokenizer.encode(text, return_tensors="pt", add_special_tokens=False).to(device)
    with torch.no_grad():
        out = model.generate(input_ids=input_ids, max_new_tokens=max_new_tokens, do_sample=False)
    full = tokenizer.decode(out[0])
    marker = "<end_of_turn>"
    second = full.find(marker, full.find(marker) + 1)
    return full if second == -1 else full[:second]


def _prepare_prompt(model, prompt: str, apply_chat_template: bool) -> str:
    if apply_chat_template:
        prompt = model.tokenizer.apply_chat_template([{"role": "user", "content": prompt}], tokenize=False,
                                                     add_generation_prompt=True, add_special_tokens=False)
    return prompt


def get_layer_readouts(model, prompt: str, layer_of_interest: int, secret_id: Optional[int] = None,
                       softcap: float = 0.0, apply_chat_template: bool = False):
    """Fused readout at one layer: (input_words, LensReadout, residual [T, d] bf16). The secret
    probability, top-10 and logsumexp per token come out of the GEMM epilogue; no [T, V] tensor."""
    prompt = _prepare_prompt(model, prompt, apply_chat_template)
    ids, resid = model.layer_residuals(prompt)
    x = resid[layer_of_interest].contiguous()
    sid = None
    if secret_id is not None:
        sid = torch.full((x.shape[0],), int(secret_id), dtype=torch.int32, device=x.device)
    out = ops.logit_lens(x, model.lens_weights(), softcap=softcap, secret_id=sid)
    return [model.tokenizer.decode(i) for i in ids], out, x


def get_all_layer_readouts(model, prompt: str, secret_id: Optional[int] = None, softcap: float = 0.0,
                           apply_chat_template: bool = False):
    """The fused readout at EVERY layer in one launch: the [L, T] grid of residual rows is one batch of
    L*T rows for the lens kernel (the reference loops 42 times over norm -> lm_head -> softmax and ships
    [42, T, 256000] probabilities to the host, src/models.py:127-144). Returns (input_words,
    LensReadout with [L, T(, 10)] tensors)."""
    prompt = _prepare_prompt(model, prompt, apply_chat_template)
    ids, resid = model.layer_residuals(prompt)
    L, T, d = resid.shape
    sid = None
    if secret_id is not None:
        sid = torch.full((L * T,), int(secret_id), dtype=torch.int32, device=resid.device)
    out = ops.logit_lens(resid.reshape(L * T, d).contiguous(), model.lens_weights(), softcap=softcap, secret_id=sid)
    grid = ops.LensReadout(lse=out.lse.view(L, T), secret_prob=out.secret_prob.view(L, T),
                           top_prob=out.top_prob.view(L, T, -1), top_id=out.top_id.view(L, T, -1))
    return [model.tokenizer.decode(i) for i in ids], grid


def get_layer_logits(model, prompt: str, apply_chat_template: bool = False,
                     layer_of_interest: Optional[int] = None):
    """Compatibility mode of src/models.py:97-170: dense ``all_probs`` [L, T, V] float32 for every
    layer, plus the optional layer residual. Meant for the short sequences the drivers cache
    (T < ~70): the lens kernel writes the logits of all layers' rows in the same sweep in which it
    accumulates their logsumexp (``ops.lens_dense_probs``), an elementwise pass normalises them in place. Use ``get_layer_readouts`` when only readouts are needed."""
    prompt = _prepare_prompt(model, prompt, apply_chat_template)
    ids, resid = model.layer_residuals(prompt)            # [L, T, d] bf16
    lens = model.lens_weights()
    L, T, d = resid.shape
    h = ops.rmsnorm(resid.reshape(L * T, d).contiguous(), lens.w_norm, lens.eps)
    # one sweep over W_U for all layers' rows: logits written by the lens kernel, normalised in place
    # (reference: no softcap on this surface)
    probs = ops.lens_dense_probs(h, lens.w_u)[0].view(L, T, -1)              # [L, T, V]
    all_probs = probs.detach().cpu().to(dtype=torch.float32).numpy()
    max_probs, tokens = probs.max(dim=-1)
    words = [[model.tokenizer.decode(t.cpu()) for t in layer_tokens] for layer_tokens in tokens]
    input_words = [model.tokenizer.decode(t) for t in ids]
    layer_residual = None
    if layer_of_interest is not None:
        layer_residual = resid[layer_of_interest].detach().cpu().to(dtype=torch.float32).numpy()
    return max_probs, words, input_words, all_probs, layer_residual
