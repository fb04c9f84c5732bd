"""Eager-PyTorch restatement of the reference's torch/torchtune engine path.

TEST INFRASTRUCTURE ONLY -- the product (``xotorch_support_jetson_b200``) never imports
this file.  It is the parity checker for the CUDA path and the timed CPU baseline.

PARITY PINNING STATUS: **parity unpinned by reference fixtures** -- the reference holds
no golden logits / hidden states / token sequences for this path,
and its arithmetic lives in ``torchtune>=0.6.1`` (un-vendored, not installable here).
The pin is therefore (a) the reference's own call sites, restated line by line below,
(b) an independent cross-check against HuggingFace ``LlamaForCausalLM`` /
``Qwen2ForCausalLM`` (``oracle/hf_crosscheck.py``; results recorded in DESIGN.md), and
(c) golden vectors generated from this file and committed under ``tests/golden/``.

What is restated, with the reference lines each piece follows
(paths relative to /root/reference):

* ``load_model_config``        xotorch/inference/torch/models/llm_utils.py:30-77
* ``_permute`` / weight remap  xotorch/inference/torch/models/llm_utils.py:126-134,181-258
* model graph per family       xotorch/inference/torch/models/general_mha.py:23-142
* prefill/decode slicing       xotorch/inference/torch/models/general_mha.py:178-254
* sharded forward              xotorch/inference/torch/models/llm_utils.py:367-440
  (the *intended* hand-off: a non-last shard returns the OUTPUT of its last layer;
  the reference's off-by-one, SURVEY.md 8a defect 1, is not reproduced)
* engine state machine         xotorch/inference/torch/sharded_inference_engine.py:108-193,230-370
* sampling                     xotorch/inference/torch/sharded_inference_engine.py:208-228
* torchtune modules (MultiHeadAttention, TransformerSelfAttentionLayer, RMSNorm,
  FeedForward, KVCache, Llama3ScaledRoPE, Qwen2RotaryPositionalEmbeddings,
  RotaryPositionalEmbeddings, generation.sample) -- published torchtune 0.6.1 semantics.

Two torchtune details are kept switchable because they could not be re-read here:
``rmsnorm`` in {"two_round" (default, 0.6.x), "one_round"} and
``rope_table_dtype`` in {"bf16" (reference-exact: Module.to(bf16) casts the
non-persistent RoPE cache, SURVEY 8a defect 3), "fp32"}.
"""
from __future__ import annotations

import json
import math
import os
from dataclasses import dataclass, field
from pathlib import Path
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch
import torch.nn.functional as F

HF_PRECISION_STR_TO_DTYPE = {
  "float16": torch.float16,
  "bfloat16": torch.bfloat16,
  "float32": torch.float32,
  "float64": torch.float64,
}

MAX_GENERATED_TOKENS = 1024  # general_mha.py:152 (ShardedGeneralModel default)


# --------------------------------------------------------------------------- config
def load_model_config(model_config_path) -> dict:
  """HF config.json -> the reference's internal dict (llm_utils.py:30-77)."""
  with open(model_config_path, "r") as f:
    base = json.load(f)
  return model_config_from_hf(base)


def model_config_from_hf(base: dict) -> dict:
  cfg = {
    "rope_scaling": base.get("rope_scaling"),
    "embed_dim": base["hidden_size"],
    "num_heads": base["num_attention_heads"],
    "head_dim": base.get("head_dim", base["hidden_size"] // base["num_attention_heads"]),
    "num_kv_heads": base["num_key_value_heads"],
    "max_seq_len": base["max_position_embeddings"],
    "intermediate_dim": base["intermediate_size"],
    "attn_dropout": base.get("attention_dropout", 0.0),
    "norm_eps": base["rms_norm_eps"],
    "rope_base": base["rope_theta"],
    "vocab_size": base["vocab_size"],
    "num_layers": base["num_hidden_layers"],
    "attn_bias": base.get("attention_bias", False),
    "hidden_act": base.get("hidden_act", "silu"),
    "torch_dtype": HF_PRECISION_STR_TO_DTYPE.get(base.get("torch_dtype", "bfloat16"), torch.bfloat16),
  }
  if cfg.get("rope_scaling") is not None:
    # llm_utils.py:68-69 -- reads "rope_factor" (absent in HF configs) => always 32
    cfg["rope_scaling_factor"] = cfg["rope_scaling"].get("rope_factor", 32)
    # llm_utils.py:71-73 (found by the reference-run fixture, tests/golden/ref_config_permute_state.json)
    if os.getenv("TORCH_USE_ORG_SEQ", "False").lower() == "true":
      cfg["max_seq_len"] = cfg["rope_scaling"]["original_max_position_embeddings"]
  return cfg


def model_family(model_id: str) -> str:
  """general_mha.py:33,45 -- family by substring of the model id."""
  if "llama" in model_id or "Llama" in model_id:
    return "llama"
  if "qwen" in model_id or "Qwen" in model_id:
    return "qwen"
  return "plain"


def uses_tied_head(model_id: str) -> bool:
  """general_mha.py:43-44,56-57."""
  fam = model_family(model_id)
  if fam == "llama":
    return "3.2" in model_id
  if fam == "qwen":
    return "0.5B" in model_id or "0.5b" in model_id
  return False


# --------------------------------------------------------------------------- RoPE tables
def llama3_scaled_theta(dim: int, base: float, scale_factor: float,
                        low_freq_factor: float = 1, high_freq_factor: float = 4,
                        old_context_len: int = 8192) -> torch.Tensor:
  """torchtune Llama3ScaledRoPE.rope_init/apply_scaling (defaults as the reference
  leaves them, general_mha.py:35-40 passes only scale_factor)."""
  freqs = 1.0 / (base ** (torch.arange(0, dim, 2)[: (dim // 2)].float() / dim))
  low_freq_wavelen = old_context_len / low_freq_factor
  high_freq_wavelen = old_context_len / high_freq_factor
  new_freqs = []
  for freq in freqs:
    wavelen = 2 * math.pi / freq
    if wavelen < high_freq_wavelen:
      new_freqs.append(freq)
    elif wavelen > low_freq_wavelen:
      new_freqs.append(freq / scale_factor)
    else:
      smooth = (old_context_len / wavelen - low_freq_factor) / (high_freq_factor - low_freq_factor)
      new_freqs.append((1 - smooth) * freq / scale_factor + smooth * freq)
  return torch.tensor(new_freqs, dtype=freqs.dtype)


def plain_theta(dim: int, base: float) -> torch.Tensor:
  return 1.0 / (base ** (torch.arange(0, dim, 2)[: (dim // 2)].float() / dim))


def build_rope_table(family: str, config: dict, n_pos: int, table_dtype: torch.dtype) -> torch.Tensor:
  """Returns the module's ``cache`` buffer for positions [0, n_pos).

  llama / plain: ``[n_pos, hd/2, 2]`` (cos, sin)  -- interleaved-pair RoPE.
  qwen:          ``[n_pos, 2*hd]``   (cos(cat(t,t)) ‖ sin(cat(t,t))) -- half-split RoPE.
  ``table_dtype=torch.bfloat16`` reproduces Module.to(bf16) on the non-persistent buffer:
  the table is computed in fp32 and then rounded.
  """
  hd = config["head_dim"]
  if family == "llama":
    theta = llama3_scaled_theta(hd, config["rope_base"], config.get("rope_scaling_factor", 8))
  else:
    theta = plain_theta(hd, config["rope_base"])
  seq_idx = torch.arange(n_pos, dtype=theta.dtype)
  idx_theta = torch.einsum("i, j -> ij", seq_idx, theta).float()
  if family == "qwen":
    idx_theta2 = torch.cat([idx_theta, idx_theta], dim=-1)
    cache = torch.cat([idx_theta2.cos(), idx_theta2.sin()], dim=-1)
  else:
    cache = torch.stack([torch.cos(idx_theta), torch.sin(idx_theta)], dim=-1)
  return cache.to(table_dtype)


def apply_rope(family: str, x: torch.Tensor, table: torch.Tensor, input_pos: torch.Tensor) -> torch.Tensor:
  """x: [b, s, n_h, hd]; input_pos: [s] / [b] / [b, s] indices into ``table``."""
  rope_cache = table[input_pos]
  if family == "qwen":
    # Qwen2RotaryPositionalEmbeddings.forward: cos/sin cast to x.dtype, rotation in x.dtype
    hd = x.size(-1)
    rope_cache = rope_cache.view(-1, x.size(1), 1, hd * 2)
    cos = rope_cache[..., :hd].to(x.dtype)
    sin = rope_cache[..., hd:].to(x.dtype)
    x1 = x[..., : hd // 2]
    x2 = x[..., hd // 2:]
    rotated = torch.cat((-x2, x1), dim=-1)
    x_out = (x * cos) + (rotated * sin)
    return x_out.type_as(x)
  # Llama3ScaledRoPE / RotaryPositionalEmbeddings.forward: fp32 math on interleaved pairs
  xshaped = x.float().reshape(*x.shape[:-1], -1, 2)
  rope_cache = rope_cache.view(-1, xshaped.size(1), 1, xshaped.size(3), 2)
  x_out = torch.stack(
    [
      xshaped[..., 0] * rope_cache[..., 0] - xshaped[..., 1] * rope_cache[..., 1],
      xshaped[..., 1] * rope_cache[..., 0] + xshaped[..., 0] * rope_cache[..., 1],
    ],
    -1,
  )
  return x_out.flatten(3).type_as(x)


# --------------------------------------------------------------------------- weights
def permute_qk(t: torch.Tensor, n_heads: int, head_dim: int, dim: int) -> torch.Tensor:
  """llm_utils.py:126-134: HF half-split rows -> torchtune interleaved rows."""
  return t.view(n_heads, 2, head_dim // 2, dim).transpose(1, 2).reshape((head_dim * n_heads), dim)


def remap_hf_weights(hf: Dict[str, torch.Tensor], model_id: str, config: dict,
                     start_layer: int, end_layer: int) -> Dict[str, torch.Tensor]:
  """llm_utils.py:181-258, by effect: HF names -> per-layer dicts in torchtune layout.

  Llama q/k (weights; the reference's regex also permutes a q/k *bias* if one existed,
  which would fail on a 1-D tensor -- Llama has none) are permuted; Qwen is not.
  """
  fam = model_family(model_id)
  H, KV, hd, D = config["num_heads"], config["num_kv_heads"], config["head_dim"], config["embed_dim"]
  out: Dict[str, torch.Tensor] = {}
  for l in range(start_layer, end_layer + 1):
    p = f"model.layers.{l}."
    q = hf[p + "self_attn.q_proj.weight"]
    k = hf[p + "self_attn.k_proj.weight"]
    if fam == "llama":
      q = permute_qk(q, H, hd, D)
      k = permute_qk(k, KV, hd, D)
    out[f"layers.{l}.attn.q_proj.weight"] = q
    out[f"layers.{l}.attn.k_proj.weight"] = k
    out[f"layers.{l}.attn.v_proj.weight"] = hf[p + "self_attn.v_proj.weight"]
    for nm in ("q_proj", "k_proj", "v_proj"):
      b = hf.get(p + f"self_attn.{nm}.bias")
      if b is not None:
        out[f"layers.{l}.attn.{nm}.bias"] = b
    out[f"layers.{l}.attn.output_proj.weight"] = hf[p + "self_attn.o_proj.weight"]
    out[f"layers.{l}.sa_norm.scale"] = hf[p + "input_layernorm.weight"]
    out[f"layers.{l}.mlp_norm.scale"] = hf[p + "post_attention_layernorm.weight"]
    out[f"layers.{l}.mlp.w1.weight"] = hf[p + "mlp.gate_proj.weight"]
    out[f"layers.{l}.mlp.w2.weight"] = hf[p + "mlp.down_proj.weight"]
    out[f"layers.{l}.mlp.w3.weight"] = hf[p + "mlp.up_proj.weight"]
  if "model.embed_tokens.weight" in hf:
    out["tok_embeddings.weight"] = hf["model.embed_tokens.weight"]
  if "model.norm.weight" in hf:
    out["norm.scale"] = hf["model.norm.weight"]
  if "lm_head.weight" in hf:
    out["output.weight"] = hf["lm_head.weight"]
  return out


# --------------------------------------------------------------------------- ops
def rms_norm(x: torch.Tensor, scale: torch.Tensor, eps: float, variant: str) -> torch.Tensor:
  """torchtune RMSNorm (general_mha.py:116-117,130)."""
  if variant == "one_round":
    return F.rms_norm(x.float(), (x.size(-1),), scale.float(), eps).to(x.dtype)
  x_fp32 = x.float()
  x_normed = (x_fp32 * torch.rsqrt(x_fp32.pow(2).mean(-1, keepdim=True) + eps)).type_as(x)
  return x_normed * scale


def sample(logits: torch.Tensor, *, temperature: float = 1.0, top_k: Optional[int] = None,
           q: Optional[torch.Tensor] = None) -> torch.Tensor:
  """torchtune.generation.sample as called at sharded_inference_engine.py:221."""
  logits = logits / max(temperature, 1e-5)
  if top_k is not None:
    v, _ = torch.topk(logits, min(top_k, logits.size(-1)))
    pivot = v.select(-1, -1).unsqueeze(-1)
    logits = torch.where(logits < pivot, -float("Inf"), logits)
  probs = F.softmax(logits, dim=-1)
  if q is None:
    q = torch.empty_like(probs).exponential_(1)
  return torch.argmax(probs / q, dim=-1, keepdim=True).to(dtype=torch.int)


# --------------------------------------------------------------------------- model
@dataclass
class OracleOptions:
  rope_table_dtype: str = "bf16"      # "bf16" (reference-exact) | "fp32"
  rmsnorm: str = "two_round"          # "two_round" | "one_round"
  max_generated_tokens: int = MAX_GENERATED_TOKENS


class OracleShardModel:
  """ShardedGeneralModel + ShardTransformerDecoder restated without torchtune.

  Holds only the weights of layers [start_layer, end_layer]; embeddings on a first
  shard; final norm + head on a last shard (tied => the embedding matrix).
  """

  def __init__(self, config: dict, model_id: str, start_layer: int, end_layer: int,
               hf_weights: Dict[str, torch.Tensor], options: Optional[OracleOptions] = None,
               dtype: Optional[torch.dtype] = None, device: str = "cpu"):
    self.config = config
    self.model_id = model_id
    self.family = model_family(model_id)
    self.tied = uses_tied_head(model_id)
    self.start_layer, self.end_layer = start_layer, end_layer
    self.n_layers = config["num_layers"]
    self.options = options or OracleOptions()
    self.dtype = dtype or config["torch_dtype"]
    self.device = torch.device(device)
    self.attn_bias = True if self.family == "qwen" else bool(config.get("attn_bias", False))
    w = remap_hf_weights(hf_weights, model_id, config, start_layer, end_layer)
    self.w = {k: v.to(device=self.device, dtype=self.dtype) for k, v in w.items()}
    self.k_cache: Dict[int, torch.Tensor] = {}
    self.v_cache: Dict[int, torch.Tensor] = {}
    self.cache_pos = 0
    self.cache_len = 0
    self.rope_table: Optional[torch.Tensor] = None

  # -- shard predicates (shard.py:12-19)
  def is_first(self) -> bool:
    return self.start_layer == 0

  def is_last(self) -> bool:
    return self.end_layer == self.n_layers - 1

  # -- caches (llm_utils.py:320-365; torchtune KVCache)
  def setup_caches(self, batch_size: int, cache_len: int):
    KV, hd = self.config["num_kv_heads"], self.config["head_dim"]
    self.cache_len = cache_len
    for l in range(self.start_layer, self.end_layer + 1):
      self.k_cache[l] = torch.zeros(batch_size, KV, cache_len, hd, dtype=self.dtype, device=self.device)
      self.v_cache[l] = torch.zeros(batch_size, KV, cache_len, hd, dtype=self.dtype, device=self.device)
    self.cache_pos = 0
    # the RoPE module's table covers max_seq_len; only rows < cache_len are ever gathered
    tdt = torch.bfloat16 if (self.options.rope_table_dtype == "bf16" and self.dtype == torch.bfloat16) else torch.float32
    self.rope_table = build_rope_table(self.family, self.config, cache_len, tdt).to(self.device)

  def caches_are_enabled(self) -> bool:
    return bool(self.k_cache)

  def reset_caches(self):
    for l in self.k_cache:
      self.k_cache[l].zero_()
      self.v_cache[l].zero_()
    self.cache_pos = 0

  # -- one decoder layer (torchtune TransformerSelfAttentionLayer + MultiHeadAttention)
  def _layer(self, l: int, x: torch.Tensor, mask: torch.Tensor, input_pos: torch.Tensor,
             cache_pos: int) -> torch.Tensor:
    cfg, w = self.config, self.w
    H, KV, hd = cfg["num_heads"], cfg["num_kv_heads"], cfg["head_dim"]
    b, s, _ = x.shape
    p = f"layers.{l}."
    h = rms_norm(x, w[p + "sa_norm.scale"], cfg["norm_eps"], self.options.rmsnorm)
    q = F.linear(h, w[p + "attn.q_proj.weight"], w.get(p + "attn.q_proj.bias"))
    k = F.linear(h, w[p + "attn.k_proj.weight"], w.get(p + "attn.k_proj.bias"))
    v = F.linear(h, w[p + "attn.v_proj.weight"], w.get(p + "attn.v_proj.bias"))
    q = apply_rope(self.family, q.view(b, s, H, hd), self.rope_table, input_pos).transpose(1, 2)
    k = apply_rope(self.family, k.view(b, s, KV, hd), self.rope_table, input_pos).transpose(1, 2)
    v = v.view(b, s, KV, hd).transpose(1, 2)
    # KVCache.update: write s positions at cache_pos, attend over the whole allocation
    kc, vc = self.k_cache[l], self.v_cache[l]
    kc[:, :, cache_pos:cache_pos + s] = k
    vc[:, :, cache_pos:cache_pos + s] = v
    k, v = kc, vc
    if H != KV:
      g = H // KV
      k = k.unsqueeze(2).expand(b, KV, g, -1, hd).flatten(1, 2)
      v = v.unsqueeze(2).expand(b, KV, g, -1, hd).flatten(1, 2)
    attn = F.scaled_dot_product_attention(q, k, v, attn_mask=mask[:, None, :, :], dropout_p=0.0, is_causal=False)
    attn = attn.transpose(1, 2).contiguous().view(b, s, -1)
    attn_out = F.linear(attn, w[p + "attn.output_proj.weight"])
    h = attn_out + x
    hn = rms_norm(h, w[p + "mlp_norm.scale"], cfg["norm_eps"], self.options.rmsnorm)
    mlp = F.linear(F.silu(F.linear(hn, w[p + "mlp.w1.weight"])) * F.linear(hn, w[p + "mlp.w3.weight"]),
                   w[p + "mlp.w2.weight"])
    return h + mlp

  @torch.inference_mode()
  def forward(self, tokens: Optional[torch.Tensor], *, mask: torch.Tensor, input_pos: torch.Tensor,
              hidden_state: Optional[torch.Tensor] = None,
              taps: Optional[Dict[int, torch.Tensor]] = None) -> torch.Tensor:
    """llm_utils.py:367-440 with the intended hand-off.  ``taps[l]`` receives the
    output of absolute layer ``l`` (per-layer parity checks)."""
    if hidden_state is not None:
      h = hidden_state
    else:
      h = F.embedding(tokens.long(), self.w["tok_embeddings.weight"]).to(self.dtype)
    s = h.size(1)
    cache_pos = self.cache_pos
    for l in range(self.start_layer, self.end_layer + 1):
      h = self._layer(l, h, mask, input_pos, cache_pos)
      if taps is not None:
        taps[l] = h.clone()
    self.cache_pos = cache_pos + s
    if self.is_last():
      h = rms_norm(h, self.w["norm.scale"], self.config["norm_eps"], self.options.rmsnorm)
      head = self.w["tok_embeddings.weight"] if self.tied else self.w["output.weight"]
      return F.linear(h, head)  # bf16 logits: `.float()` is commented out at llm_utils.py:430
    return h

  @torch.inference_mode()
  def generate(self, tokens: Optional[torch.Tensor], mask: torch.Tensor, input_pos: torch.Tensor,
               hidden_state: Optional[torch.Tensor] = None, curr_pos: int = 0,
               taps: Optional[Dict[int, torch.Tensor]] = None):
    """general_mha.py:178-254 (cached branch)."""
    if curr_pos > 0:
      input_pos = input_pos[:, curr_pos].contiguous()
      mask = mask[:, curr_pos, None, :].contiguous()
    else:
      tklng = tokens.size(1) if tokens is not None else hidden_state.size(1)
      mask = mask[:, :tklng]
      input_pos = input_pos[:, :tklng].squeeze()
    out = self.forward(tokens, mask=mask, input_pos=input_pos, hidden_state=hidden_state, taps=taps)
    if self.is_last():
      return None, out
    return out, None


# --------------------------------------------------------------------------- engine state machine
class OracleEngine:
  """TorchDynamicShardInferenceEngine's per-request state machine over one shard
  (sharded_inference_engine.py:108-370), driven with token ids instead of text."""

  def __init__(self, model: OracleShardModel, seed: int = 1234):
    self.model = model
    self.tokens: Optional[torch.Tensor] = None
    self.input_pos: Optional[torch.Tensor] = None
    self.mask: Optional[torch.Tensor] = None
    self.curr_pos = 0
    self.rng = torch.Generator(device=model.device)
    self.rng.manual_seed(seed)

  def start_request(self, tokens: torch.Tensor, pad_id: Optional[int] = None):
    """encode() after tokenisation (sharded_inference_engine.py:131-186).  ``pad_id``: what the
    reference resolves at :153-163 (tokenizer.pad_id, else pad_token_id, else 0); None = treat
    no id as padding.  A prompt holding that id takes the padded branch (:165-177): positions
    skip the padded slots and padded keys are never attended (torchtune.generation's
    get_causal_mask_from_padding_mask / get_position_ids_from_padding_mask, 0.6.1, restated)."""
    m = self.model
    self.curr_pos = 0
    if m.caches_are_enabled():
      m.reset_caches()
    self.tokens = tokens.clone()
    bsz, tklng = tokens.shape
    total = tklng + m.options.max_generated_tokens
    if not m.caches_are_enabled() or m.cache_len != total or next(iter(m.k_cache.values())).size(0) != bsz:
      m.setup_caches(bsz, total)
    padding_masks = tokens != pad_id if pad_id is not None else torch.ones_like(tokens, dtype=torch.bool)
    if not padding_masks.all():
      padding_masks = F.pad(padding_masks, (0, m.options.max_generated_tokens), value=True)
      seq_len = padding_masks.size(1)
      mask = torch.tril(torch.ones(seq_len, total, dtype=torch.bool, device=m.device)).repeat(bsz, 1, 1)
      mask.narrow(2, 0, seq_len).mul_(padding_masks[:, None, :].expand(-1, seq_len, seq_len))
      mask.diagonal(dim1=1, dim2=2).fill_(True)
      self.mask = mask
      self.input_pos = ((padding_masks.cumsum(-1) - 1) * padding_masks).to(torch.int)
    else:
      self.mask = torch.tril(torch.ones(total, total, dtype=torch.bool, device=m.device)).unsqueeze(0)
      self.input_pos = torch.arange(0, total, device=m.device).unsqueeze(0)

  def infer_tensor(self, input_data: np.ndarray, taps=None) -> np.ndarray:
    """sharded_inference_engine.py:230-370: 2-D int = tokens, 3-D float = hidden."""
    m = self.model
    hidden_state = None
    input_tensor = None
    if input_data.ndim == 3:
      hidden_state = torch.tensor(input_data).to(device=m.device, dtype=m.dtype)
    else:
      input_tensor = torch.tensor(input_data).to(device=m.device, dtype=torch.int)
      if self.tokens is not None and input_tensor.size(-1) == 1:
        self.tokens = torch.cat([self.tokens, input_tensor.clone().to(self.tokens.dtype)], dim=-1)
      elif self.tokens is None:
        self.tokens = input_tensor.clone()
    hs, logits = m.generate(tokens=input_tensor, hidden_state=hidden_state, input_pos=self.input_pos.clone(),
                            mask=self.mask.clone(), curr_pos=self.curr_pos, taps=taps)
    s = (input_tensor if input_tensor is not None else hidden_state).size(1)
    if hs is not None:
      # non-last shard; the reference only advances curr_pos on the last shard because the
      # state travels with the request -- a standalone stage must advance with its cache
      self.curr_pos = s if self.curr_pos == 0 else self.curr_pos + 1
      return hs.float().numpy(force=True)
    self.curr_pos = s if self.curr_pos == 0 else self.curr_pos + 1
    return logits[:, -1].float().numpy(force=True)

  def sample(self, x: np.ndarray, temp: float = 0.6, top_k: int = 35) -> np.ndarray:
    """sharded_inference_engine.py:208-228."""
    logits = torch.tensor(x).to(self.model.device)
    q = torch.empty((logits.size(0), self.model.config["vocab_size"]), device=logits.device).exponential_(1, generator=self.rng)
    return sample(logits.clone(), temperature=temp, top_k=top_k, q=q).numpy(force=True)


def greedy_generate(stages: List[OracleShardModel], prompt: torch.Tensor, n_new: int,
                    collect_logits: bool = False, taps_at: Tuple[int, ...] = ()):
  """Ring walk over ``stages`` (node.py:109-147,424-443) with first-index argmax
  (equal to the product's temp=0 / top-k 35 / argmax(p/q) on every non-tied row,
  SURVEY 8a defect 6).  Returns (tokens [B, n_new], logits list, taps list)."""
  engines = [OracleEngine(s) for s in stages]
  for e in engines:
    e.start_request(prompt)
  out_tokens, all_logits, all_taps = [], [], []
  x: np.ndarray = prompt.numpy()
  for step in range(n_new):
    taps = {} if step in taps_at else None
    y = x
    for e in engines:
      y = e.infer_tensor(y, taps=taps)
    if taps is not None:
      all_taps.append({k: v.float().numpy() for k, v in taps.items()})
    if collect_logits:
      all_logits.append(y.copy())
    tok = y.argmax(axis=-1).astype(np.int64).reshape(-1, 1)
    out_tokens.append(tok)
    x = tok
  return np.concatenate(out_tokens, axis=1), all_logits, all_taps
