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


def load_model_config(model_config_path: Path) -> dict:
  with open(model_config_path, "r") as f:
    base_config = json.load(f)
  model_config = {
    "rope_scaling": base_config.get("rope_scaling"),
    "embed_dim": base_config["hidden_size"],
    "num_heads": base_config["num_attention_heads"],
    "head_dim": base_config.get("head_dim", base_config["hidden_size"] // base_config["num_attention_heads"]),
    "num_kv_heads": base_config["num_key_value_heads"],
    "max_seq_len": base_config["max_position_embeddings"],
    "intermediate_dim": base_config["intermediate_size"],
    "attn_dropout": base_config.get("attention_dropout", 0.0),
    "norm_eps": base_config["rms_norm_eps"],
    "rope_base": base_config["rope_theta"],
    "vocab_size": base_config["vocab_size"],
    "num_layers": base_config["num_hidden_layers"],
    "attn_bias": base_config.get("attention_bias", False),
    "hidden_act": base_config.get("hidden_act", "silu"),
    "torch_dtype": HF_PRECISION_STR_TO_DTYPE.get(base_config.get("torch_dtype", "bfloat16"), torch.bfloat16),
  }
  if model_config.get("rope_scaling", None) is not None:
    model_config["rope_scaling_factor"] = model_config["rope_scaling"].get("rope_factor", 32)
  use_org_seq = bool(os.getenv("TORCH_USE_ORG_SEQ", "False").lower() == "true")
  if use_org_seq and model_config.get("rope_scaling", None) is not None:
    model_config["max_seq_len"] = model_config["rope_scaling"]["original_max_position_embeddings"]
  return model_config


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
