"""Host-side logic that needs no GPU: Shard, registry, topology/partitioning against the
reference's own outputs, synthetic checkpoints, the C ABI surface."""
import asyncio
import ctypes
import json
import re
from pathlib import Path

import numpy as np
import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
GOLD = ROOT / "tests" / "golden"


def test_shard_semantics():
  from xotorch_support_jetson_b200.shard import Shard
  s = Shard("llama-3.1-8b", 0, 7, 32)
  assert s.is_first_layer() and not s.is_last_layer() and s.get_layer_count() == 8
  assert Shard.from_dict(s.to_dict()) == s and hash(s) == hash(Shard("llama-3.1-8b", 0, 7, 32))
  assert s.overlaps(Shard("llama-3.1-8b", 7, 9, 32)) and not s.overlaps(Shard("llama-3.1-8b", 8, 9, 32))
  assert not s.overlaps(Shard("other", 0, 7, 32))
  assert Shard("m", 24, 31, 32).is_last_layer()


def test_partitioning_matches_reference_outputs():
  """tests/golden/partition_golden.json was produced by importing the REFERENCE's
  RingMemoryWeightedPartitioningStrategy + map_partitions_to_shards (oracle/gen_golden.py)."""
  from xotorch_support_jetson_b200 import topology as T
  gold = json.loads((GOLD / "partition_golden.json").read_text())
  for key, exp in gold["equal_rings"].items():
    n_layers, n = (int(x) for x in key.split("/"))
    parts = T.RingMemoryWeightedPartitioningStrategy().partition(T.box_topology(n))
    assert [p.node_id for p in parts] == exp["nodes"]
    shards = T.map_partitions_to_shards(parts, n_layers, "m")
    assert [[s.start_layer, s.end_layer] for s in shards] == exp["ranges"], key
    assert [[s.start_layer, s.end_layer] for s in T.pipeline_shards("m", n_layers, n)] == exp["ranges"]
  for case in gold["ref_test_map_partitions"]:
    parts = [T.Partition(*p) for p in case["partitions"]]
    shards = T.map_partitions_to_shards(parts, case["n_layers"], "model")
    assert [[s.start_layer, s.end_layer] for s in shards] == case["ranges"]
  for case in gold["weighted"]:
    t = T.Topology()
    for i, m in enumerate(case["memories"]):
      t.update_node(f"node{i + 1}", T.DeviceCapabilities(model="t", chip="t", memory=m, flops=T.DeviceFlops(0, 0, 0)))
    parts = T.RingMemoryWeightedPartitioningStrategy().partition(t)
    assert [[p.node_id, p.start, p.end] for p in parts] == case["partitions"]


def test_reference_test_vectors_verbatim():
  """Expected values copied from the reference's test (xotorch/topology/test_map_partitions.py:9-50)
  and from SURVEY.md 8e (8 equal nodes)."""
  from xotorch_support_jetson_b200.shard import Shard
  from xotorch_support_jetson_b200.topology import Partition, map_partitions_to_shards, pipeline_shards
  parts = [Partition("node1", 0.0, 0.42857), Partition("node2", 0.42857, 0.71428), Partition("node3", 0.71428, 0.99999)]
  assert map_partitions_to_shards(parts, 32, "model") == [Shard("model", 0, 12, 32), Shard("model", 13, 21, 32), Shard("model", 22, 31, 32)]
  parts = [Partition("node1", 0.0, 0.1), Partition("node2", 0.1, 0.2), Partition("node3", 0.2, 1.0)]
  assert map_partitions_to_shards(parts, 32, "model") == [Shard("model", 0, 2, 32), Shard("model", 3, 5, 32), Shard("model", 6, 31, 32)]
  assert [(s.start_layer, s.end_layer) for s in pipeline_shards("m", 32, 8)] == [(4 * i, 4 * i + 3) for i in range(8)]
  assert [s.get_layer_count() for s in pipeline_shards("m", 28, 8)] == [3, 4, 3, 4, 3, 4, 3, 4]
  assert [(s.start_layer, s.end_layer) for s in pipeline_shards("m", 28, 2)] == [(0, 13), (14, 27)]


def test_registry_and_base_shard():
  from xotorch_support_jetson_b200 import models
  from xotorch_support_jetson_b200.inference_engine import get_inference_engine, inference_engine_classes
  assert inference_engine_classes["b200"] == models.ENGINE_CLASS_NAME
  sh = models.build_base_shard("llama-3.1-8b", models.ENGINE_CLASS_NAME)
  assert (sh.start_layer, sh.end_layer, sh.n_layers) == (0, 0, 32)
  assert models.build_full_shard("qwen-2.5-7b", models.ENGINE_CLASS_NAME).end_layer == 27
  assert models.build_base_shard("llama-3.1-8b", "TorchDynamicShardInferenceEngine") is None
  with pytest.raises(ValueError):
    get_inference_engine("mlx", None)
  # byte counts of SURVEY.md section 8 (footnote 1)
  for mid, gb in (("llama-3.2-1b", 2.472), ("llama-3.1-8b", 15.009), ("qwen-2.5-7b", 14.141)):
    c = models.model_cards[mid]["hf_config"]
    D, H, KV, I, V, L = c["hidden_size"], c["num_attention_heads"], c["num_key_value_heads"], c["intermediate_size"], c["vocab_size"], c["num_hidden_layers"]
    hd = c.get("head_dim", D // H)
    b = 2 * (L * (2 * D * H * hd + 2 * D * KV * hd + 3 * D * I) + V * D)
    assert abs(b / 1e9 - gb) < 0.001, (mid, b / 1e9)


def test_synthetic_checkpoint_roundtrip(tmp_path):
  from xotorch_support_jetson_b200.models import ENGINE_CLASS_NAME, model_cards
  from xotorch_support_jetson_b200.shard import Shard
  from xotorch_support_jetson_b200.stage import _SafetensorsIndex
  from xotorch_support_jetson_b200.synth import LocalShardDownloader, SyntheticTokenizer, random_hf_state
  dl = LocalShardDownloader(root=tmp_path, seed=3, std=0.05)
  sh = Shard("llama-3.2-tiny", 0, 3, 4)
  p = asyncio.run(dl.ensure_shard(sh, ENGINE_CLASS_NAME))
  assert p == tmp_path / "synthetic--llama-3.2-tiny" and (p / "config.json").exists()
  idx = _SafetensorsIndex(p)
  state = random_hf_state(model_cards["llama-3.2-tiny"]["hf_config"], seed=3, std=0.05)
  assert set(idx.where) == set(state)
  for k in ("model.layers.2.self_attn.q_proj.weight", "model.embed_tokens.weight", "model.norm.weight"):
    assert torch.equal(idx.get(k), state[k])
  # a layer-range subset draws the same values as the full model
  sub = random_hf_state(model_cards["llama-3.2-tiny"]["hf_config"], seed=3, std=0.05, only_layers=range(2, 3))
  assert torch.equal(sub["model.layers.2.mlp.up_proj.weight"], state["model.layers.2.mlp.up_proj.weight"])
  assert "model.layers.1.mlp.up_proj.weight" not in sub
  with pytest.raises(ValueError):
    asyncio.run(dl.ensure_shard(sh, "TorchDynamicShardInferenceEngine"))
  tok = SyntheticTokenizer(1024)
  ids = tok.encode("hello there general kenobi")
  assert len(ids) == 4 and all(0 < i < 1023 for i in ids) and ids == tok.encode("hello there general kenobi")
  assert tok.encode("<ids:5,6,7>") == [5, 6, 7] and tok.encode("a b", return_tensors="pt").shape == (1, 2)


def test_permute_matches_reference_formula():
  from xotorch_support_jetson_b200.stage import _permute
  H, hd, D = 4, 8, 16
  w = torch.arange(H * hd * D, dtype=torch.float32).reshape(H * hd, D)
  p = _permute(w, H, hd, D)
  # HF half-split row j of a head (j < hd/2 -> pair member 0, else member 1) becomes interleaved row 2*(j % (hd/2)) + j // (hd/2)
  for h in range(H):
    for j in range(hd):
      assert torch.equal(p[h * hd + 2 * (j % (hd // 2)) + j // (hd // 2)], w[h * hd + j])


def test_cabi_exports_every_declared_symbol(built_lib):
  from xotorch_support_jetson_b200 import _cabi
  header = (ROOT / "include" / "xb200.h").read_text()
  declared = set(re.findall(r"^\s*(?:const\s+char\s*\*|int)\s+(xb_[a-z0-9_]+)\s*\(", header, flags=re.M))
  assert declared, "no declarations parsed from include/xb200.h"
  assert declared == set(_cabi.SYMBOLS)
  raw = ctypes.CDLL(str(_cabi.lib_path()))
  for name in declared:
    assert hasattr(raw, name), f"libxb200.so does not export {name}"
  assert built_lib.xb_abi_version() == _cabi.ABI_VERSION == int(re.search(r"#define XB_ABI_VERSION (\d+)", header).group(1))
  # struct layouts the Python side mirrors: every xb_model_desc field of the header, in order
  body = re.search(r"typedef struct xb_model_desc \{(.*?)\} xb_model_desc;", header, flags=re.S).group(1)
  body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
  fields = [n.strip() for decl in re.findall(r"(?:u?int32_t|float)\s+([^;]+);", body) for n in decl.split(",")]
  assert fields == [n for n, _ in _cabi.ModelDesc._fields_]
  assert ctypes.sizeof(_cabi.ModelDesc) == 20 * 4 and ctypes.sizeof(_cabi.Counters) == 6 * 8
  assert _cabi.model_desc(n_layers=1).struct_size == 80


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_compute_path_fails_loudly_without_gpu(built_lib):
  """No CPU fallback: creating an engine without a CUDA device must fail with a clear error."""
  from xotorch_support_jetson_b200 import _cabi
  from xotorch_support_jetson_b200.models import model_cards
  from xotorch_support_jetson_b200.shard import Shard
  from xotorch_support_jetson_b200.stage import StageRuntime
  from oracle.torch_oracle import model_config_from_hf
  d = _cabi.model_desc(n_layers=4, start_layer=0, end_layer=3, dim=256, n_heads=8, n_kv_heads=2, head_dim=64, inter_dim=512,
                      vocab=1024, norm_eps=1e-5, max_positions=64, max_batch_tokens=16, max_seqs=2)
  h = ctypes.c_void_p()
  rc = built_lib.xb_engine_create(ctypes.byref(d), 0, ctypes.byref(h))
  assert rc != 0 and b"no CUDA device" in built_lib.xb_last_error()
  with pytest.raises(RuntimeError, match="no CPU fallback"):
    StageRuntime(Shard("llama-3.2-tiny", 0, 3, 4), model_config_from_hf(model_cards["llama-3.2-tiny"]["hf_config"]))


def test_engine_request_bookkeeping_without_gpu():
  """Slot management of B200ShardInferenceEngine with a stub stage (no native calls)."""
  from xotorch_support_jetson_b200.b200_engine import B200ShardInferenceEngine
  from xotorch_support_jetson_b200.shard import Shard

  class Stub:
    max_seqs, max_batch_tokens, vocab, device = 4, 64, 16, torch.device("cpu")
    released = []

    def release_slot(self, s):
      self.released.append(s)

  e = B200ShardInferenceEngine(None, device=0, max_seqs=4)
  e.stage, e.shard = Stub(), Shard("m", 0, 0, 1)
  e.free_slots = [3, 2, 1, 0]
  a = e._start_request("a", 2)
  b = e._start_request("b", 2)
  assert sorted(a.slots + b.slots) == [0, 1, 2, 3] and not e.free_slots
  c = e._start_request("c", 1)          # evicts the oldest request
  assert "a" not in e.requests and len(c.slots) == 1 and len(e.free_slots) == 1
  e.end_request("b")
  assert sorted(e.free_slots) == sorted(set(range(4)) - set(c.slots))
  with pytest.raises(RuntimeError):
    e._start_request("d", 5)


def test_half_layer_stages_cover_and_balance():
  """topology.half_layer_stages: contiguous cover of [attn_0, mlp_0, ..., mlp_{n-1}+head], slowest
  stage never worse than the whole-layer ring strategy, flags consistent with the Shard ranges."""
  from xotorch_support_jetson_b200.topology import half_layer_stages, pipeline_shards
  for (nl, n, a, m, h) in ((32, 8, 352.0, 352.0, 1050.0), (32, 4, 352.0, 352.0, 1050.0), (80, 8, 1.0, 2.3, 1.2),
                           (28, 4, 0.4, 1.0, 2.0), (4, 5, 1.0, 1.3, 0.8), (16, 1, 1.0, 1.0, 1.0), (6, 2, 1.0, 1.0, 0.0)):
    st = half_layer_stages("llama-x", nl, n, a, m, h)
    assert len(st) == min(n, 2 * nl)
    # walk the unit sequence
    unit = 0
    costs = []
    for (sh, from_mlp, to_attn) in st:
      first_unit = 2 * sh.start_layer + (1 if from_mlp else 0)
      last_unit = 2 * sh.end_layer + (0 if to_attn else 1)
      assert first_unit == unit and last_unit >= first_unit, (nl, n, st)
      c = sum((a if u % 2 == 0 else m) for u in range(first_unit, last_unit + 1))
      if last_unit == 2 * nl - 1:
        c += h
      costs.append(c)
      unit = last_unit + 1
    assert unit == 2 * nl
    assert not st[0][1] and not st[-1][2]
    if n <= nl:
      whole = pipeline_shards("llama-x", nl, n, h / (a + m))
      whole_cost = max(s.get_layer_count() * (a + m) + (h if i == len(whole) - 1 else 0.0) for i, s in enumerate(whole))
      assert max(costs) <= whole_cost + 1e-9, (nl, n, costs, whole_cost)
  # the 8-GPU case of the bench: slowest stage 4.5 layer-equivalents instead of 5
  st = half_layer_stages("llama-3.1-8b", 32, 8, 352.0, 352.0, 1050.0)
  assert sum(1 for (_, a, b) in st if a or b) == 2


def test_safetensors_header_parser_without_gpu(built_lib, tmp_path):
  """xb_safetensors_list (the header parser behind xb_load_safetensors) against the safetensors
  package: names, dtypes, shapes, byte ranges; metadata with quotes/braces; malformed files."""
  import ctypes as C
  import struct
  import torch
  from safetensors import safe_open
  from safetensors.torch import save_file
  from xotorch_support_jetson_b200 import _cabi
  tensors = {
    "model.layers.0.self_attn.q_proj.weight": torch.zeros(16, 8, dtype=torch.bfloat16),
    "model.layers.10.mlp.up_proj.weight": torch.zeros(24, 8, dtype=torch.float16),
    "model.norm.weight": torch.ones(8, dtype=torch.float32),
    "odd \"name\" {x}": torch.zeros(3, dtype=torch.int64),
    "scalar": torch.tensor(1.5),
  }
  f = tmp_path / "m.safetensors"
  save_file(tensors, str(f), metadata={"format": "pt", "note": "a \"quoted\" value, {braces}, [brackets]"})

  def listing(path):
    n = C.c_int()
    buf = C.create_string_buffer(1 << 16)
    _cabi.check(built_lib.xb_safetensors_list(str(path).encode(), buf, len(buf), C.byref(n)))
    rows = {}
    for line in buf.value.decode().splitlines():
      # names may contain spaces: the last fields are numeric, dtype is the token before them
      parts = line.split(" ")
      k = len(parts) - 1
      while k > 0 and parts[k].lstrip("-").isdigit():
        k -= 1
      name, dtype, nums = " ".join(parts[:k]), parts[k], [int(x) for x in parts[k + 1:]]
      rows[name] = (dtype, nums[0], nums[1], nums[2:])
    assert len(rows) == n.value
    return rows

  rows = listing(f)
  with safe_open(str(f), framework="pt") as h:
    assert set(rows) == set(h.keys())
    for k in h.keys():
      t = h.get_tensor(k)
      dtype, b, e, shape = rows[k]
      assert shape == list(t.shape) and e - b == t.numel() * t.element_size()
      assert dtype == {torch.bfloat16: "BF16", torch.float16: "F16", torch.float32: "F32", torch.int64: "I64"}[t.dtype]
  # byte ranges tile the data section without overlap
  spans = sorted((b, e) for _, b, e, _ in rows.values())
  assert spans[0][0] == 0 and all(spans[i][1] == spans[i + 1][0] for i in range(len(spans) - 1))
  # malformed: header length beyond the file, truncated JSON, offsets past the end, not a file
  bad = tmp_path / "bad.safetensors"
  for payload in (struct.pack("<Q", 1 << 40) + b"{}", struct.pack("<Q", 5) + b'{"a":', struct.pack("<Q", 2) + b"[]",
                  struct.pack("<Q", 60) + b'{"t":{"dtype":"F32","shape":[4],"data_offsets":[0,16]}}    ', b"abc"):
    bad.write_bytes(payload)
    with pytest.raises(RuntimeError):
      listing(bad)
  with pytest.raises(RuntimeError):
    listing(tmp_path / "missing.safetensors")
