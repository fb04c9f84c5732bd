"""The oracle against its pins: the committed golden vectors, HuggingFace transformers, and
the invariants the reference's own tests state (sharded == unsharded)."""
import json
from pathlib import Path

import numpy as np
import pytest
import torch

from oracle import torch_oracle as O
from xotorch_support_jetson_b200.models import model_cards
from xotorch_support_jetson_b200.synth import random_hf_state

GOLD = Path(__file__).resolve().parent / "golden"
MODELS = ("llama-3.2-tiny", "llama-3.1-tiny", "qwen-2.5-tiny", "mistral-tiny")


def _setup(mid, dtype=None):
  g = np.load(GOLD / f"oracle_{mid}.npz", allow_pickle=False)
  meta = json.loads(str(g["meta"]))
  hf_cfg = model_cards[mid]["hf_config"]
  cfg = O.model_config_from_hf(hf_cfg)
  state = random_hf_state(hf_cfg, seed=meta["seed"], std=meta["std"], norm_jitter=meta["norm_jitter"])
  return g, cfg, state


@pytest.mark.parametrize("mid", MODELS)
def test_oracle_reproduces_golden(mid):
  """Regression pin: bf16 CPU kernels may differ by an ulp across torch builds / thread counts,
  so logits are compared within 2 bf16 ulps of their scale and the tokens must be identical
  wherever the golden top-2 margin exceeds that."""
  g, cfg, state = _setup(mid)
  nl = cfg["num_layers"]
  prompt = torch.from_numpy(g["prompt"].astype(np.int64))
  model = O.OracleShardModel(cfg, mid, 0, nl - 1, state, O.OracleOptions(max_generated_tokens=32))
  n = g["tokens"].shape[1]
  toks, logits, _ = O.greedy_generate([model], prompt, n, collect_logits=True)
  got = np.stack([l[0] for l in logits])
  scale = float(np.abs(g["logits"]).max())
  assert np.abs(got - g["logits"]).max() <= 2 * 2 ** -8 * scale
  assert toks[0].tolist() == g["tokens"][0].tolist()


@pytest.mark.parametrize("mid,splits", [("llama-3.1-tiny", [(0, 2), (3, 5)]), ("qwen-2.5-tiny", [(0, 0), (1, 2), (3, 3)])])
def test_sharded_equals_unsharded(mid, splits):
  """The reference's contract (xotorch/inference/test_inference_engine.py:11,43-44): an engine
  must give np.array_equal results for any number of shards."""
  g, cfg, state = _setup(mid)
  nl = cfg["num_layers"]
  prompt = torch.from_numpy(g["prompt"].astype(np.int64))
  opts = O.OracleOptions(max_generated_tokens=32)
  full = [O.OracleShardModel(cfg, mid, 0, nl - 1, state, opts)]
  parts = [O.OracleShardModel(cfg, mid, a, b, state, opts) for a, b in splits]
  t1, l1, _ = O.greedy_generate(full, prompt, 4, collect_logits=True)
  t2, l2, _ = O.greedy_generate(parts, prompt, 4, collect_logits=True)
  for a, b in zip(l1, l2):
    assert np.array_equal(a, b)
  assert np.array_equal(t1, t2)


@pytest.mark.parametrize("mid,splits", [("llama-3.2-tiny", None), ("qwen-2.5-tiny", [(0, 1), (2, 3)])])
def test_oracle_matches_huggingface_fp32(mid, splits):
  from oracle.hf_crosscheck import crosscheck
  err, same = crosscheck(mid, splits=splits)
  assert err < 2e-4 and same


def test_sampler_restatement():
  g = torch.Generator().manual_seed(0)
  logits = torch.randn(4, 1000, generator=g) * 3
  q = torch.empty(4, 1000).exponential_(1, generator=g)
  # temp -> 0 with top-k: argmax unless tied (SURVEY 8a defect 6)
  t = O.sample(logits.clone(), temperature=0.0, top_k=35, q=q)
  assert torch.equal(t.reshape(-1).long(), logits.argmax(-1))
  # top-k pivot keeps exactly the k largest (no ties in random floats)
  t = O.sample(logits.clone(), temperature=1.0, top_k=1, q=q)
  assert torch.equal(t.reshape(-1).long(), logits.argmax(-1))
  assert t.dtype == torch.int32 and t.shape == (4, 1)


@pytest.mark.parametrize("mid", MODELS)     # mistral-tiny: the reference's "plain" RoPE branch
def test_product_rope_table_equals_oracle(mid):
  """The host builds the XB_W_ROPE_TABLE upload independently of the oracle; both must be the
  same numbers (bf16-rounded, reference-exact mode and fp32 mode)."""
  from xotorch_support_jetson_b200 import model_config as mc
  hf_cfg = model_cards[mid]["hf_config"]
  cfg = O.model_config_from_hf(hf_cfg)
  fam = O.model_family(mid)
  assert fam == mc.model_family(mid) and O.uses_tied_head(mid) == mc.uses_tied_head(mid)
  hd = cfg["head_dim"]
  for bf in (True, False):
    tab = mc.rope_table(fam, cfg, 300, bf16_rounded=bf)
    ref = O.build_rope_table(fam, cfg, 300, torch.bfloat16 if bf else torch.float32).float()
    if fam == "qwen":   # oracle cache = cos(cat(t,t)) || sin(cat(t,t)); halves duplicate
      cos, sin = ref[:, :hd // 2], ref[:, hd:hd + hd // 2]
      assert torch.equal(ref[:, hd // 2:hd], cos)
      ref = torch.stack([cos, sin], dim=-1)
    assert torch.equal(tab, ref)


def test_config_mapping_matches_oracle_and_keeps_reference_quirk(tmp_path):
  from xotorch_support_jetson_b200 import model_config as mc
  hf_cfg = model_cards["llama-3.1-8b"]["hf_config"]
  (tmp_path / "config.json").write_text(json.dumps(hf_cfg))
  a = mc.load_model_config(tmp_path / "config.json")
  b = O.load_model_config(tmp_path / "config.json")
  assert a == b
  # llm_utils.py:68-69 reads "rope_factor" (absent) -> 32, although the HF config says 8
  assert hf_cfg["rope_scaling"]["factor"] == 8.0 and a["rope_scaling_factor"] == 32
  assert a["head_dim"] == 128 and a["num_kv_heads"] == 8 and a["torch_dtype"] == torch.bfloat16
