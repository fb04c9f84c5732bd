"""Native checkpoint loader (xb_load_safetensors) against the safetensors-package path and the
in-memory path: same weights in HBM => bit-identical forward results.  Covers multi-file
checkpoints, layer-range shards, tied / untied heads, q/k/v bias, the Llama q/k permutation,
F32 / F16 checkpoints (cast like Tensor.to(bfloat16)) and the error behaviour."""
import json
import struct
from pathlib import Path

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _cfg(mid):
  from xotorch_support_jetson_b200 import model_config as mc
  from xotorch_support_jetson_b200.models import model_cards
  return model_cards[mid]["hf_config"]


def _stage(mid, a, b, cfg):
  from xotorch_support_jetson_b200.shard import Shard
  from xotorch_support_jetson_b200.stage import StageRuntime
  nl = cfg["num_layers"]
  return StageRuntime(Shard(mid, a, b, nl), cfg, device=0, max_batch_tokens=64, max_seqs=4, max_tokens_per_seq=96, page_tokens=16)


def _run(stages, cfg, seed=5):
  g = torch.Generator().manual_seed(seed)
  toks = torch.randint(1, cfg["vocab_size"], (29,), generator=g).to(torch.int32).cuda()
  outs = []
  with torch.cuda.stream(torch.cuda.Stream()):
    h = None
    for st in stages:
      h = st.forward([0], [29], [0], tokens=toks) if h is None else st.forward([0], [29], [0], hidden=h)
    outs.append(h)
    tok = h.argmax(-1).to(torch.int32)
    for step in range(3):
      h = None
      for st in stages:
        h = st.forward([0], [1], [29 + step], tokens=tok) if h is None else st.forward([0], [1], [29 + step], hidden=h)
      outs.append(h)
      tok = h.argmax(-1).to(torch.int32)
  torch.cuda.synchronize()
  return outs


@pytest.mark.parametrize("mid,splits", [
  ("llama-3.2-tiny", [(0, 3)]),           # tied head, head_dim 64, q/k permuted
  ("llama-3.1-tiny", [(0, 2), (3, 5)]),   # untied head, two layer-range shards
  ("qwen-2.5-tiny", [(0, 1), (2, 3)]),    # q/k/v bias, no permutation
])
def test_native_loader_equals_python_paths(built_lib, tmp_path, mid, splits):
  from oracle import torch_oracle as O
  from xotorch_support_jetson_b200.synth import random_hf_state, write_random_checkpoint
  hf = _cfg(mid)
  cfg = O.model_config_from_hf(hf)
  ck = write_random_checkpoint(tmp_path / mid, hf, seed=2, std=0.05, norm_jitter=0.1, shard_bytes=1 << 20)   # several files
  assert len(list(ck.glob("*.safetensors"))) > 1
  state = random_hf_state(hf, seed=2, std=0.05, norm_jitter=0.1)
  native, pyth, mem = [], [], []
  for a, b in splits:
    s = _stage(mid, a, b, cfg)
    info = s.load_dir(ck)
    assert info["tensors"] > 0 and info["bytes"] > 0 and s.weights_missing() == ""
    native.append(s)
    s = _stage(mid, a, b, cfg)
    s.load_dir(ck, native=False)
    pyth.append(s)
    s = _stage(mid, a, b, cfg)
    s.load_state(state)
    mem.append(s)
  # a shard reads only its own layers: bytes taken by the two shards of a split add up to < 2x the checkpoint
  r_native, r_py, r_mem = _run(native, cfg), _run(pyth, cfg), _run(mem, cfg)
  for i, (x, y, z) in enumerate(zip(r_native, r_py, r_mem)):
    assert torch.isfinite(x).all()
    assert torch.equal(x, y), f"output {i}: native loader differs from the safetensors-package path"
    assert torch.equal(x, z), f"output {i}: native loader differs from the in-memory path"
  for s in native + pyth + mem:
    s.close()


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16])
def test_native_loader_casts_like_torch(built_lib, tmp_path, dtype):
  from oracle import torch_oracle as O
  from safetensors.torch import save_file
  from xotorch_support_jetson_b200.synth import hf_param_shapes
  mid = "llama-3.1-tiny"
  hf = _cfg(mid)
  cfg = O.model_config_from_hf(hf)
  g = torch.Generator().manual_seed(9)
  state = {}
  for name, shape in hf_param_shapes(hf).items():
    t = torch.randn(shape, generator=g, dtype=torch.float32) * 0.05
    if name.endswith("norm.weight"):
      t = 1.0 + t
    state[name] = t.to(dtype)       # values that are NOT bf16-representable: the cast has to round
  d = tmp_path / "ck"
  d.mkdir()
  (d / "config.json").write_text(json.dumps(hf))
  save_file(state, str(d / "model.safetensors"), metadata={"format": "pt", "note": "a \"quoted\" value, {braces}"})
  a, b = _stage(mid, 0, cfg["num_layers"] - 1, cfg), _stage(mid, 0, cfg["num_layers"] - 1, cfg)
  a.load_dir(d)
  b.load_state({k: v.to(torch.bfloat16) for k, v in state.items()})
  for x, y in zip(_run([a], cfg), _run([b], cfg)):
    assert torch.equal(x, y)
  a.close()
  b.close()


def test_native_loader_errors(built_lib, tmp_path):
  from oracle import torch_oracle as O
  from safetensors.torch import save_file
  from xotorch_support_jetson_b200.synth import random_hf_state
  mid = "llama-3.2-tiny"
  hf = _cfg(mid)
  cfg = O.model_config_from_hf(hf)
  state = random_hf_state(hf, seed=1, std=0.05)
  # a tensor of this shard is absent
  d1 = tmp_path / "missing"
  d1.mkdir()
  save_file({k: v for k, v in state.items() if k != "model.layers.2.mlp.up_proj.weight"}, str(d1 / "model.safetensors"))
  s = _stage(mid, 0, 3, cfg)
  with pytest.raises(RuntimeError, match="layer 2 up_proj"):
    s.load_dir(d1)
  # ... but a shard that does not hold layer 2 loads fine from the same file
  s2 = _stage(mid, 0, 1, cfg)
  s2.load_dir(d1)
  # wrong shape
  d2 = tmp_path / "shape"
  d2.mkdir()
  bad = dict(state)
  bad["model.layers.0.self_attn.o_proj.weight"] = torch.zeros(8, 8, dtype=torch.bfloat16)
  save_file(bad, str(d2 / "model.safetensors"))
  with pytest.raises(RuntimeError, match="o_proj"):
    s.load_dir(d2)
  # unsupported dtype, truncated / garbage files
  d3 = tmp_path / "dtype"
  d3.mkdir()
  bad = dict(state)
  bad["model.layers.0.input_layernorm.weight"] = torch.ones(cfg["embed_dim"], dtype=torch.float64)
  save_file(bad, str(d3 / "model.safetensors"))
  with pytest.raises(RuntimeError, match="unsupported dtype"):
    s.load_dir(d3)
  d4 = tmp_path / "garbage"
  d4.mkdir()
  (d4 / "model.safetensors").write_bytes(struct.pack("<Q", 1 << 40) + b"{}")
  with pytest.raises(RuntimeError, match="header"):
    s.load_dir(d4)
  (d4 / "model.safetensors").write_bytes(struct.pack("<Q", 5) + b"{\"a\":")
  with pytest.raises(RuntimeError, match="header"):
    s.load_dir(d4)
  with pytest.raises(FileNotFoundError):
    s.load_dir(tmp_path / "nowhere")
  s.close()
  s2.close()
