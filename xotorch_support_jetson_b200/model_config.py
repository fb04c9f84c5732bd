"""HF config.json -> the engine's model description.

Same field mapping and the same quirks as the reference's ``load_model_config``
(``xotorch/inference/torch/models/llm_utils.py:30-77``): ``rope_scaling_factor`` is read
from the key ``"rope_factor"`` with default 32, so Llama-3.1 runs with factor 32 like the
reference does (SURVEY.md 8a defect 2).  Family / tied-head detection by model-id
substring follows ``general_mha.py:33-63``.
"""
import json
import math
import os
from pathlib import Path

import torch

HF_PRECISION_STR_TO_DTYPE = {
  "float16": torch.float16, "bfloat16": torch.bfloat16, "float32": torch.float32, "float64": torch.float64,
}


# HF config.json key -> engine key, for the fields that are copied as they are
_DIRECT_FIELDS = (
  ("hidden_size", "embed_dim"), ("num_attention_heads", "num_heads"), ("num_key_value_heads", "num_kv_heads"),
  ("max_position_embeddings", "max_seq_len"), ("intermediate_size", "intermediate_dim"), ("rms_norm_eps", "norm_eps"),
  ("rope_theta", "rope_base"), ("vocab_size", "vocab_size"), ("num_hidden_layers", "num_layers"),
)
# optional HF keys with the defaults the reference applies
_OPTIONAL_FIELDS = (("attention_dropout", "attn_dropout", 0.0), ("attention_bias", "attn_bias", False), ("hidden_act", "hidden_act", "silu"))


def load_model_config(model_config_path: Path) -> dict:
  """Read an HF ``config.json`` into the flat dict the engine works with."""
  hf = json.loads(Path(model_config_path).read_text())
  cfg = {ours: hf[theirs] for theirs, ours in _DIRECT_FIELDS}
  cfg.update({ours: hf.get(theirs, default) for theirs, ours, default in _OPTIONAL_FIELDS})
  cfg["head_dim"] = hf.get("head_dim", hf["hidden_size"] // hf["num_attention_heads"])
  cfg["torch_dtype"] = HF_PRECISION_STR_TO_DTYPE.get(hf.get("torch_dtype", "bfloat16"), torch.bfloat16)
  scaling = hf.get("rope_scaling")
  cfg["rope_scaling"] = scaling
  if scaling is not None:
    # reference quirk kept on purpose: the factor is looked up under "rope_factor", a key HF configs
    # do not carry, so Llama-3.x always runs with 32 (llm_utils.py:68-69; SURVEY 8a defect 2)
    cfg["rope_scaling_factor"] = scaling.get("rope_factor", 32)
    if os.getenv("TORCH_USE_ORG_SEQ", "False").lower() == "true":
      cfg["max_seq_len"] = scaling["original_max_position_embeddings"]
  return cfg


def model_family(model_id: str) -> str:
  if "llama" in model_id or "Llama" in model_id:
    return "llama"
  if "qwen" in model_id or "Qwen" in model_id:
    return "qwen"
  return "plain"


def uses_tied_head(model_id: str) -> bool:
  fam = model_family(model_id)
  if fam == "llama":
    return "3.2" in model_id
  if fam == "qwen":
    return "0.5B" in model_id or "0.5b" in model_id
  return False


def has_qkv_bias(model_id: str, config: dict) -> bool:
  return True if model_family(model_id) == "qwen" else bool(config.get("attn_bias", False))


def rope_theta(family: str, config: dict) -> torch.Tensor:
  """Per-pair rotation frequencies.  llama: torchtune Llama3ScaledRoPE.apply_scaling with
  its defaults (low_freq_factor 1, high_freq_factor 4, old_context_len 8192) -- the
  reference passes only scale_factor (general_mha.py:35-40)."""
  dim, base = config["head_dim"], config["rope_base"]
  freqs = 1.0 / (base ** (torch.arange(0, dim, 2)[: (dim // 2)].float() / dim))
  if family != "llama":
    return freqs
  scale_factor = config.get("rope_scaling_factor", 8)
  low_freq_factor, high_freq_factor, old_context_len = 1, 4, 8192
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


def rope_table(family: str, config: dict, n_pos: int, bf16_rounded: bool = True) -> torch.Tensor:
  """fp32 ``[n_pos, head_dim/2, 2]`` (cos, sin) table for XB_W_ROPE_TABLE.  With
  ``bf16_rounded`` the values are rounded to bf16 first: the reference's
  ``Module.to(bfloat16)`` also casts the RoPE module's non-persistent cache
  (SURVEY.md 8a defect 3), so this is the reference-exact table."""
  theta = rope_theta(family, config)
  seq_idx = torch.arange(n_pos, dtype=theta.dtype)
  idx_theta = torch.einsum("i, j -> ij", seq_idx, theta).float()
  cache = torch.stack([torch.cos(idx_theta), torch.sin(idx_theta)], dim=-1)
  if bf16_rounded:
    cache = cache.to(torch.bfloat16)
  return cache.float().contiguous()
