"""Product and oracle against fixtures produced by RUNNING THE REFERENCE'S OWN CODE
(oracle/gen_ref_fixtures.py; tests/golden/ref_*.json).  CPU only -- nothing here reads
/root/reference; the fixtures travel with the repository."""
import json
from pathlib import Path

import numpy as np
import pytest
import torch

GOLD = Path(__file__).resolve().parent / "golden"
FIX = json.loads((GOLD / "ref_config_permute_state.json").read_text())
TRACE = json.loads((GOLD / "ref_engine_trace.json").read_text())


# ----------------------------------------------------------------------------- a16: load_model_config
@pytest.mark.parametrize("key", sorted(FIX["load_model_config"].keys()))
def test_load_model_config_equals_reference_output(key, tmp_path, monkeypatch):
  """xotorch/inference/torch/models/llm_utils.py:30-77, run on every model card's config.json."""
  from xotorch_support_jetson_b200 import model_config as mc
  from xotorch_support_jetson_b200.models import model_cards
  mid, org = key.split("|org_seq=")
  monkeypatch.setenv("TORCH_USE_ORG_SEQ", org)
  (tmp_path / "config.json").write_text(json.dumps(model_cards[mid]["hf_config"]))
  got = mc.load_model_config(tmp_path / "config.json")
  got["torch_dtype"] = str(got["torch_dtype"])
  want = FIX["load_model_config"][key]
  assert got == want
  # the oracle's restatement is pinned by the same reference output
  from oracle import torch_oracle as O
  ora = O.load_model_config(tmp_path / "config.json")
  ora["torch_dtype"] = str(ora["torch_dtype"])
  assert ora == want


def test_reference_config_quirk_is_in_the_fixture():
  """llm_utils.py:68-69 reads "rope_factor" (a key HF never writes) -> 32 although HF says 8."""
  c = FIX["load_model_config"]["llama-3.1-8b|org_seq=False"]
  assert c["rope_scaling"]["factor"] == 8.0 and c["rope_scaling_factor"] == 32
  assert FIX["load_model_config"]["llama-3.1-8b|org_seq=true"]["max_seq_len"] == 8192


# ----------------------------------------------------------------------------- a10: _permute
@pytest.mark.parametrize("case", FIX["permute"], ids=lambda c: f"{c['n_heads']}x{c['head_dim']}")
def test_permute_equals_reference_output(case):
  """llm_utils.py:126-134 as a row permutation; the product's host path (stage._permute) and the
  oracle's (permute_qk) must move the same rows.  (The native loader's permutation is compared with
  the host path bit for bit on the GPU, tests/test_loader_gpu.py.)"""
  from oracle import torch_oracle as O
  from xotorch_support_jetson_b200.stage import _permute
  H, hd, dim = case["n_heads"], case["head_dim"], 8
  g = torch.Generator().manual_seed(H * 1000 + hd)
  w = torch.randn(H * hd, dim, generator=g).to(torch.bfloat16)
  want = w[torch.tensor(case["src_row_of_out_row"])]
  assert torch.equal(_permute(w, H, hd, dim), want)
  assert torch.equal(O.permute_qk(w, H, hd, dim), want)


def test_native_loader_permutation_formula_equals_reference():
  """The index arithmetic of the native loader (csrc/xb_loader.cuh) restated on the host: output
  row r of a head comes from source row (r % 2) * hd/2 + r / 2 of that head."""
  for case in FIX["permute"]:
    H, hd = case["n_heads"], case["head_dim"]
    src = [(r // hd) * hd + ((r % hd) % 2) * (hd // 2) + (r % hd) // 2 for r in range(H * hd)]
    assert src == case["src_row_of_out_row"]


# ----------------------------------------------------------------------------- a15: ShardInferenceState
def test_inference_state_dict_is_a_subset_of_the_reference_format():
  """llm_utils.py:473-511: the state that travels between nodes.  The product keeps the two fields
  it needs with the reference's names, types and nesting (tokens: list of rows of ints; curr_pos:
  int), reads a reference-made dict, and survives the JSON hop of the gRPC path."""
  from xotorch_support_jetson_b200.b200_engine import state_position, state_to_dict
  ref = FIX["shard_inference_state"]["to_dict"]
  assert set(ref.keys()) == {"tokens", "input_pos", "mask", "curr_pos"}
  assert state_position(ref) == ref["curr_pos"] == 3
  mine = state_to_dict([[5, 9, 11]], 3)
  assert set(mine.keys()) <= set(ref.keys())
  assert mine["tokens"] == ref["tokens"] and mine["curr_pos"] == ref["curr_pos"]
  assert type(mine["curr_pos"]) is type(ref["curr_pos"]) and type(mine["tokens"][0][0]) is int
  assert json.loads(json.dumps(mine)) == mine
  assert state_position(None) is None and state_position({}) is None
  back = FIX["shard_inference_state"]["from_dict"]
  assert back["tokens"] == ref["tokens"] and back["curr_pos"] == 3 and back["tokens_dtype"] == "torch.int64"


def test_pinned_pool_gives_the_caller_ownership(monkeypatch):
  """Arrays handed out by the plugin API own their buffer until the caller (and every view it
  derived) lets go; only then is the buffer reused.  (pin_memory needs a driver: patched out.)"""
  from xotorch_support_jetson_b200.b200_engine import _PinnedPool
  monkeypatch.setattr(torch.Tensor, "pin_memory", lambda self: self)
  p = _PinnedPool()
  b1, a1 = p.take(torch.float32, (2, 3))
  b2, a2 = p.take(torch.float32, (2, 3))
  assert p.allocated == 2 and b1.data_ptr() != b2.data_ptr() and not a1.flags.writeable
  b1.fill_(7.0)
  view = a1[0]
  ptr1 = b1.data_ptr()
  del a1, b1
  b3, a3 = p.take(torch.float32, (6,))          # a view of a1 is alive: its buffer must not come back
  assert p.allocated == 3 and b3.data_ptr() != ptr1 and float(view[0]) == 7.0
  del view
  b4, a4 = p.take(torch.float32, (5,))          # same size class, now free again
  assert p.allocated == 3 and b4.data_ptr() == ptr1


# ----------------------------------------------------------------------------- a3/a4/a5: engine state machine
def _recording_oracle(shard_range, n_layers, log, vocab, hid):
  """OracleShardModel.generate (the restated general_mha.py:178-254 slicing) over a recorder."""
  from oracle import torch_oracle as O

  class Rec(O.OracleShardModel):
    def __init__(self):   # no weights: forward() is replaced
      self.start_layer, self.end_layer = shard_range
      self.n_layers = n_layers
      self.options = O.OracleOptions(max_generated_tokens=16)
      self.dtype, self.device = torch.bfloat16, torch.device("cpu")
      self.config = {"vocab_size": vocab, "num_kv_heads": 1, "head_dim": 2, "rope_base": 10000.0}
      self.family = "plain"
      self.k_cache, self.v_cache, self.cache_pos, self.cache_len = {}, {}, 0, 0

    def forward(self, tokens, *, mask, input_pos, hidden_state=None, taps=None):
      m = mask.to(torch.int64)
      log.append({"tokens": None if tokens is None else tokens.tolist(),
                  "hidden_shape": None if hidden_state is None else list(hidden_state.shape),
                  "input_pos": input_pos.reshape(-1).tolist(), "mask_shape": list(mask.shape),
                  "mask_true_per_row": m.sum(-1).reshape(-1).tolist(),
                  "mask_rows": m.reshape(-1, m.size(-1)).tolist()})
      src = hidden_state.float() if hidden_state is not None else None
      ids = src[..., 0] if src is not None else tokens.float()
      ps = src[..., 1] if src is not None else input_pos.reshape(1, -1).float().expand_as(ids)
      B, s = ids.shape
      if self.is_last():
        out = torch.zeros(B, s, vocab, dtype=self.dtype)
        for b in range(B):
          for j in range(s):
            out[b, j, (3 * int(ids[b, j]) + int(ps[b, j]) + 1) % vocab] = 1.0
        return out
      out = torch.zeros(B, s, hid, dtype=self.dtype)
      out[..., 0], out[..., 1] = ids, ps
      return out

  return Rec()


@pytest.mark.parametrize("name", ["full_shard", "two_shards", "padded_prompt"])
def test_oracle_engine_replays_the_reference_engine_trace(name):
  """The oracle's OracleEngine / generate() must call the model with exactly the token rows,
  positions and mask rows the reference's encode / infer_tensor / generate do, and return the
  same shapes, dtypes and ids (trace recorded from the reference's own code)."""
  from oracle import torch_oracle as O
  tr = TRACE[name]
  log = []
  models = [_recording_oracle(tuple(s), tr["n_layers"], log, tr["vocab"], tr["hidden"]) for s in tr["splits"]]
  engines = [O.OracleEngine(m) for m in models]
  calls = [c for c in tr["calls"] if c["api"] == "infer_tensor"]
  prompt = torch.tensor(tr["calls"][0]["ids"])
  for e in engines:
    e.start_request(prompt, pad_id=0)
  x = prompt.numpy()
  ci = 0
  out_ids = []
  for step in range(len(tr["generated"])):
    for e in engines:
      mark = len(log)
      y = e.infer_tensor(x)
      want = calls[ci]
      ci += 1
      assert list(y.shape) == want["out_shape"] and str(y.dtype) == want["out_dtype"]
      ref_fwd = [op for op in want["model_ops"] if op["op"] == "forward"]
      mine = log[mark:]
      assert len(mine) == len(ref_fwd) == 1
      # the first shard's model sees token ids; other shards get hidden rows (the reference also
      # passes the token history there, which the decoder ignores once hidden_state is given)
      if ref_fwd[0]["hidden_shape"] is None:
        assert mine[0]["tokens"] == ref_fwd[0]["tokens"]
      assert mine[0]["hidden_shape"] == ref_fwd[0]["hidden_shape"]
      assert mine[0]["input_pos"] == ref_fwd[0]["input_pos"]
      assert mine[0]["mask_shape"] == ref_fwd[0]["mask_shape"]
      assert mine[0]["mask_true_per_row"] == ref_fwd[0]["mask_true_per_row"]
      if "mask_rows" in ref_fwd[0]:
        assert mine[0]["mask_rows"] == ref_fwd[0]["mask_rows"]
      x = y
    tok = int(np.argmax(x, axis=-1).reshape(-1)[0])
    out_ids.append(tok)
    x = np.array([[tok]], dtype=np.int64)
  assert out_ids == tr["generated"]
  assert ci == len(calls)
