"""Offline stand-ins for the reference's download layer.

The reference's ``ShardDownloader.ensure_shard(shard, engine_name) -> Path``
(``xotorch/download/shard_download.py:9-22``) returns a directory holding
``config.json`` + ``*.safetensors`` in HuggingFace naming.  There is no network on the
build or GPU boxes, so ``LocalShardDownloader`` returns such a directory from local
disk and, when asked, fills it with a seeded random-init checkpoint of the registered
architecture (Linear/embedding N(0, std^2) in bf16, RMSNorm scales
1.0 or 1+N(0, std^2), Qwen q/k/v biases N(0, std^2)).
"""
import json
import os
from pathlib import Path
from typing import Dict, Optional

import torch

from .models import model_cards
from .shard import Shard


def hf_param_shapes(cfg: dict) -> Dict[str, tuple]:
  D, H, KV = cfg["hidden_size"], cfg["num_attention_heads"], cfg["num_key_value_heads"]
  hd = cfg.get("head_dim", D // H)
  I, V, L = cfg["intermediate_size"], cfg["vocab_size"], cfg["num_hidden_layers"]
  bias = cfg.get("model_type") == "qwen2" or cfg.get("attention_bias", False)
  shapes = {"model.embed_tokens.weight": (V, D), "model.norm.weight": (D,)}
  if not cfg.get("tie_word_embeddings", False):
    shapes["lm_head.weight"] = (V, D)
  for l in range(L):
    p = f"model.layers.{l}."
    shapes[p + "self_attn.q_proj.weight"] = (H * hd, D)
    shapes[p + "self_attn.k_proj.weight"] = (KV * hd, D)
    shapes[p + "self_attn.v_proj.weight"] = (KV * hd, D)
    shapes[p + "self_attn.o_proj.weight"] = (D, H * hd)
    if bias:
      shapes[p + "self_attn.q_proj.bias"] = (H * hd,)
      shapes[p + "self_attn.k_proj.bias"] = (KV * hd,)
      shapes[p + "self_attn.v_proj.bias"] = (KV * hd,)
    shapes[p + "mlp.gate_proj.weight"] = (I, D)
    shapes[p + "mlp.up_proj.weight"] = (I, D)
    shapes[p + "mlp.down_proj.weight"] = (D, I)
    shapes[p + "input_layernorm.weight"] = (D,)
    shapes[p + "post_attention_layernorm.weight"] = (D,)
  return shapes


def random_hf_state(cfg: dict, seed: int = 0, std: float = 0.02, norm_jitter: float = 0.0,
                    dtype: torch.dtype = torch.bfloat16, only_layers: Optional[range] = None,
                    peaked_head: float = 0.0, embed_std: Optional[float] = None) -> Dict[str, torch.Tensor]:
  """Seeded random-init state dict in HF naming.  Each tensor draws from its own
  generator (seed, name-hash) so a layer-range subset equals the same rows of the full
  model.  ``peaked_head`` > 0 plants ``lm_head = peaked_head * embed + noise`` so greedy
  decoding has top-2 margins far above one bf16 ulp (hard parts);
  ``embed_std`` scales the embedding table alone (with a tied head a larger table keeps the
  last token's own embedding visible in the residual stream, which peaks the logits)."""
  out = {}
  names = sorted(hf_param_shapes(cfg).items())
  for idx, (name, shape) in enumerate(names):
    if only_layers is not None and name.startswith("model.layers."):
      l = int(name.split(".")[2])
      if l not in only_layers:
        continue
    g = torch.Generator().manual_seed(seed * 1000003 + idx)
    if name.endswith("layernorm.weight") or name == "model.norm.weight":
      t = torch.ones(shape)
      if norm_jitter:
        t = t + norm_jitter * torch.randn(shape, generator=g)
    elif name == "model.embed_tokens.weight" and embed_std is not None:
      t = embed_std * torch.randn(shape, generator=g)
    else:
      t = std * torch.randn(shape, generator=g)
    out[name] = t.to(dtype)
  if peaked_head and "lm_head.weight" in out:
    out["lm_head.weight"] = (out["lm_head.weight"].float() + peaked_head * out["model.embed_tokens.weight"].float()).to(dtype)
  return out


def write_random_checkpoint(path: Path, cfg: dict, seed: int = 0, std: float = 0.02,
                            norm_jitter: float = 0.0, shard_bytes: int = 4 << 30,
                            peaked_head: float = 0.0) -> Path:
  from safetensors.torch import save_file
  path = Path(path)
  path.mkdir(parents=True, exist_ok=True)
  with open(path / "config.json", "w") as f:
    json.dump(cfg, f, indent=1)
  state = random_hf_state(cfg, seed=seed, std=std, norm_jitter=norm_jitter, peaked_head=peaked_head)
  files, cur, cur_bytes = [], {}, 0
  for k, v in state.items():
    nb = v.numel() * v.element_size()
    if cur and cur_bytes + nb > shard_bytes:
      files.append(cur)
      cur, cur_bytes = {}, 0
    cur[k] = v.contiguous()
    cur_bytes += nb
  if cur:
    files.append(cur)
  for i, d in enumerate(files):
    save_file(d, str(path / f"model-{i + 1:05d}-of-{len(files):05d}.safetensors"))
  with open(path / "xb_synth.json", "w") as f:
    json.dump({"seed": seed, "std": std, "norm_jitter": norm_jitter, "peaked_head": peaked_head}, f)
  return path


def xot_home() -> Path:
  return Path(os.environ.get("XOT_HOME", Path.home() / ".cache" / "xotorch"))


class LocalShardDownloader:
  """ShardDownloader (shard_download.py:9-22) that never touches the network.

  ``ensure_shard`` returns ``$XOT_HOME/downloads/<org--repo>/`` (the reference's layout,
  new_shard_download.py:24-49).  With ``synthesize=True`` a missing directory is filled
  with a seeded random-init checkpoint of the registered architecture."""

  def __init__(self, root: Optional[Path] = None, synthesize: bool = True, seed: int = 0,
               std: float = 0.02, norm_jitter: float = 0.0, peaked_head: float = 0.0):
    self.root = Path(root) if root is not None else xot_home() / "downloads"
    self.synthesize, self.seed, self.std, self.norm_jitter = synthesize, seed, std, norm_jitter
    self.peaked_head = peaked_head

  def path_for(self, model_id: str, inference_engine_name: str) -> Path:
    repo = model_cards.get(model_id, {}).get("repo", {}).get(inference_engine_name)
    if repo is None:
      raise ValueError(f"no repo registered for model {model_id!r} and engine {inference_engine_name!r}")
    return self.root / repo.replace("/", "--")

  async def ensure_shard(self, shard: Shard, inference_engine_name: str) -> Path:
    p = self.path_for(shard.model_id, inference_engine_name)
    if not (p / "config.json").exists():
      if not self.synthesize:
        raise FileNotFoundError(f"{p} has no config.json and synthesis is disabled")
      write_random_checkpoint(p, model_cards[shard.model_id]["hf_config"], seed=self.seed, std=self.std,
                              norm_jitter=self.norm_jitter, peaked_head=self.peaked_head)
    return p

  async def get_shard_download_status(self, inference_engine_name: str):
    if False:
      yield


class SyntheticTokenizer:
  """Whitespace-free stand-in for the HF tokenizer the reference resolves
  (``xotorch/inference/tokenizers.py:43-63``): no tokenizer files exist offline.
  ``encode`` hashes the prompt into a seeded id sequence (never id 0, so the
  reference's no-pad ``tril`` branch, sharded_inference_engine.py:178-186, is taken)."""

  def __init__(self, vocab_size: int, eos_token_id: Optional[int] = None):
    self.vocab_size = vocab_size
    self.eos_token_id = eos_token_id if eos_token_id is not None else vocab_size - 1
    self.pad_token_id = None

  def encode(self, text: str, return_tensors: Optional[str] = None):
    import zlib
    if text.startswith("<ids:") and text.endswith(">"):
      ids = [int(t) for t in text[5:-1].split(",") if t]
    else:
      n = max(1, len(text.split()))
      g = torch.Generator().manual_seed(zlib.crc32(text.encode()))
      ids = torch.randint(1, self.vocab_size - 1, (n,), generator=g).tolist()
    if return_tensors == "pt":
      return torch.tensor([ids], dtype=torch.int64)
    return ids

  def decode(self, tokens) -> str:
    return " ".join(f"<{int(t)}>" for t in tokens)

  def apply_chat_template(self, conversation, tokenize=False, add_generation_prompt=True, **kw):
    return "\n".join(str(m.get("content", "")) for m in conversation)
