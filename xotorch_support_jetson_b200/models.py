"""Model registry for the configs BASELINE.json names.

Mirrors the reference's ``xotorch/models.py:4-179,231-263`` (``model_cards`` keyed by
model id, repo lookup keyed by *engine class name*) for the five benchmark models, and
adds the public HF ``config.json`` of each so that random-init checkpoints of the right
architecture can be written offline (there is no network here).
"""
from typing import Optional
from .shard import Shard

ENGINE_CLASS_NAME = "B200ShardInferenceEngine"

_LLAMA3_ROPE = {"factor": 32.0, "high_freq_factor": 4.0, "low_freq_factor": 1.0,
                "original_max_position_embeddings": 8192, "rope_type": "llama3"}


def _llama(hidden, heads, kv, inter, layers, head_dim=None, tie=False, factor=32.0, vocab=128256):
  cfg = {
    "architectures": ["LlamaForCausalLM"], "model_type": "llama", "hidden_act": "silu",
    "hidden_size": hidden, "intermediate_size": inter, "num_attention_heads": heads,
    "num_key_value_heads": kv, "num_hidden_layers": layers, "vocab_size": vocab,
    "max_position_embeddings": 131072, "rms_norm_eps": 1e-5, "rope_theta": 500000.0,
    "rope_scaling": dict(_LLAMA3_ROPE, factor=factor), "tie_word_embeddings": tie,
    "attention_bias": False, "mlp_bias": False, "torch_dtype": "bfloat16",
  }
  if head_dim is not None:
    cfg["head_dim"] = head_dim
  return cfg


def _qwen(hidden, heads, kv, inter, layers, vocab=152064, tie=False, max_pos=32768):
  return {
    "architectures": ["Qwen2ForCausalLM"], "model_type": "qwen2", "hidden_act": "silu",
    "hidden_size": hidden, "intermediate_size": inter, "num_attention_heads": heads,
    "num_key_value_heads": kv, "num_hidden_layers": layers, "vocab_size": vocab,
    "max_position_embeddings": max_pos, "rms_norm_eps": 1e-6, "rope_theta": 1000000.0,
    "tie_word_embeddings": tie, "torch_dtype": "bfloat16",
  }


def _mistral(hidden, heads, kv, inter, layers, head_dim, vocab=32768):
  """An id with neither "llama" nor "qwen" in it: the reference then builds plain RotaryPositionalEmbeddings, no
  q/k/v bias, untied head (general_mha.py:58-63) -- its mistral-nemo / nemotron-70b cards take that branch."""
  return {
    "architectures": ["MistralForCausalLM"], "model_type": "mistral", "hidden_act": "silu",
    "hidden_size": hidden, "intermediate_size": inter, "num_attention_heads": heads,
    "num_key_value_heads": kv, "num_hidden_layers": layers, "vocab_size": vocab, "head_dim": head_dim,
    "max_position_embeddings": 131072, "rms_norm_eps": 1e-5, "rope_theta": 1000000.0,
    "tie_word_embeddings": False, "torch_dtype": "bfloat16",
  }


# model id -> {layers, repo:{EngineClassName: hf repo}, hf_config}
model_cards = {
  "llama-3.2-1b": {"layers": 16, "repo": {ENGINE_CLASS_NAME: "unsloth/Llama-3.2-1B-Instruct"},
                   "hf_config": _llama(2048, 32, 8, 8192, 16, head_dim=64, tie=True)},
  "llama-3.2-3b": {"layers": 28, "repo": {ENGINE_CLASS_NAME: "unsloth/Llama-3.2-3B-Instruct"},
                   "hf_config": _llama(3072, 24, 8, 8192, 28, head_dim=128, tie=True)},
  "llama-3.1-8b": {"layers": 32, "repo": {ENGINE_CLASS_NAME: "unsloth/Meta-Llama-3.1-8B-Instruct"},
                   "hf_config": _llama(4096, 32, 8, 14336, 32, factor=8.0)},
  "llama-3.1-70b": {"layers": 80, "repo": {ENGINE_CLASS_NAME: "unsloth/Meta-Llama-3.1-70B-Instruct"},
                    "hf_config": _llama(8192, 64, 8, 28672, 80, factor=8.0)},
  "qwen-2.5-7b": {"layers": 28, "repo": {ENGINE_CLASS_NAME: "Qwen/Qwen2.5-7B-Instruct"},
                  "hf_config": _qwen(3584, 28, 4, 18944, 28)},
  # small same-architecture models for parity tests (seconds on the CPU oracle)
  "llama-3.2-tiny": {"layers": 4, "repo": {ENGINE_CLASS_NAME: "synthetic/llama-3.2-tiny"},
                     "hf_config": _llama(256, 8, 2, 512, 4, head_dim=64, tie=True, vocab=1024)},
  "llama-3.1-tiny": {"layers": 6, "repo": {ENGINE_CLASS_NAME: "synthetic/llama-3.1-tiny"},
                     "hf_config": _llama(512, 4, 2, 1024, 6, head_dim=128, factor=8.0, vocab=2048)},
  "qwen-2.5-tiny": {"layers": 4, "repo": {ENGINE_CLASS_NAME: "synthetic/qwen-2.5-tiny"},
                    "hf_config": _qwen(384, 6, 2, 768, 4, vocab=1536, max_pos=4096)},
  "mistral-tiny": {"layers": 4, "repo": {ENGINE_CLASS_NAME: "synthetic/mistral-tiny"},
                   "hf_config": _mistral(320, 5, 1, 640, 4, head_dim=64, vocab=1280)},
}


def get_repo(model_id: str, inference_engine_classname: str) -> Optional[str]:
  return model_cards.get(model_id, {}).get("repo", {}).get(inference_engine_classname, None)


def build_base_shard(model_id: str, inference_engine_classname: str) -> Optional[Shard]:
  """models.py:237-242 of the reference: Shard(id, 0, 0, n_layers)."""
  repo = get_repo(model_id, inference_engine_classname)
  n_layers = model_cards.get(model_id, {}).get("layers", 0)
  if repo is None or n_layers < 1:
    return None
  return Shard(model_id, 0, 0, n_layers)


def build_full_shard(model_id: str, inference_engine_classname: str) -> Optional[Shard]:
  base = build_base_shard(model_id, inference_engine_classname)
  if base is None:
    return None
  return Shard(base.model_id, 0, base.n_layers - 1, base.n_layers)
