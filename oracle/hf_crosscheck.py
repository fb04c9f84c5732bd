"""Independent cross-check of the oracle against HuggingFace transformers (fp32).

TEST INFRASTRUCTURE.  Builds a random-init HF Llama / Qwen2 model from the registered
tiny configs, feeds the same state dict to ``oracle.torch_oracle`` (through the
reference's HF->torchtune remap incl. ``_permute``) and compares prefill logits and a
few cached decode steps.  Run: ``python -m oracle.hf_crosscheck``.  Results are quoted
in DESIGN.md; tests/test_oracle.py runs the same check.
"""
import copy
import sys

import numpy as np
import torch

from oracle import torch_oracle as O
from xotorch_support_jetson_b200.models import model_cards
from xotorch_support_jetson_b200.synth import random_hf_state


def crosscheck(model_id: str, seed: int = 0, L: int = 24, n_dec: int = 4, splits=None):
  from transformers import AutoConfig, AutoModelForCausalLM
  hf_cfg = copy.deepcopy(model_cards[model_id]["hf_config"])
  hf_cfg["torch_dtype"] = "float32"
  cfg = O.model_config_from_hf(hf_cfg)
  if hf_cfg.get("rope_scaling"):
    hf_cfg["rope_scaling"]["factor"] = float(cfg["rope_scaling_factor"])  # reference uses 32 (defect 2)
  state = random_hf_state(hf_cfg, seed=seed, std=0.05, norm_jitter=0.1, dtype=torch.float32)
  config = AutoConfig.for_model(**hf_cfg)
  with torch.device("cpu"):
    hf = AutoModelForCausalLM.from_config(config, torch_dtype=torch.float32)
  sd = dict(state)
  if hf_cfg.get("tie_word_embeddings"):
    sd["lm_head.weight"] = sd["model.embed_tokens.weight"]
  missing, unexpected = hf.load_state_dict(sd, strict=False)
  assert not [m for m in missing if "rotary" not in m], missing
  hf.eval()
  g = torch.Generator().manual_seed(1234)
  prompt = torch.randint(1, hf_cfg["vocab_size"], (1, L), generator=g)
  nl = hf_cfg["num_hidden_layers"]
  splits = splits or [(0, nl - 1)]
  opts = O.OracleOptions(rope_table_dtype="fp32", max_generated_tokens=16)
  stages = [O.OracleShardModel(cfg, model_id, a, b, state, opts, dtype=torch.float32) for a, b in splits]
  toks, logits, _ = O.greedy_generate(stages, prompt, n_dec, collect_logits=True)
  # HF greedy with cache
  with torch.no_grad():
    out = hf(prompt, use_cache=True)
    past = out.past_key_values
    hf_logits = [out.logits[:, -1].numpy()]
    nxt = out.logits[:, -1].argmax(-1, keepdim=True)
    hf_toks = [nxt.numpy()]
    for _ in range(n_dec - 1):
      out = hf(nxt, past_key_values=past, use_cache=True)
      past = out.past_key_values
      hf_logits.append(out.logits[:, -1].numpy())
      nxt = out.logits[:, -1].argmax(-1, keepdim=True)
      hf_toks.append(nxt.numpy())
  hf_toks = np.concatenate(hf_toks, axis=1)
  err = max(float(np.abs(a - b).max()) for a, b in zip(logits, hf_logits))
  return err, bool((toks == hf_toks).all())


if __name__ == "__main__":
  ok = True
  for mid, splits in (("llama-3.2-tiny", None), ("llama-3.1-tiny", [(0, 2), (3, 5)]), ("qwen-2.5-tiny", [(0, 0), (1, 2), (3, 3)])):
    err, same = crosscheck(mid, splits=splits)
    print(f"{mid:16s} splits={splits} max|oracle-HF| logits = {err:.3e}  greedy tokens identical = {same}")
    ok &= err < 2e-4 and same
  sys.exit(0 if ok else 1)
