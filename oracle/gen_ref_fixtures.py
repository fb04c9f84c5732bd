"""Fixtures produced by RUNNING THE REFERENCE'S OWN CODE in the build container (TEST INFRASTRUCTURE).

  python -m oracle.gen_ref_fixtures          # needs /root/reference; writes tests/golden/ref_*.json

The reference's arithmetic lives in torchtune (not installable here), but everything AROUND the
model is plain Python/torch and imports once `torchtune` and `scapy` are stubbed in
``sys.modules`` (nothing of the stubs is executed by the functions below, except where a fixture
says so).  What is pinned by the reference itself this way:

* ``load_model_config``            (xotorch/inference/torch/models/llm_utils.py:30-77)
* ``_permute``                     (llm_utils.py:126-134)
* ``ShardInferenceState``          (llm_utils.py:473-511) to_dict / from_dict
* the engine's per-request state machine: ``TorchDynamicShardInferenceEngine.encode`` /
  ``infer_tensor`` (xotorch/inference/torch/sharded_inference_engine.py:108-370) driving
  ``ShardedGeneralModel.generate`` (general_mha.py:178-254) -- all three unmodified -- over a
  RECORDING stand-in for the torchtune decoder: which tokens / hidden rows, ``input_pos`` and mask
  rows the model is called with at prefill and at every decode step, what comes back (shape, dtype)
  and how ``curr_pos`` / the token history evolve, for a full shard, for a 2-shard ring walked
  the way ``Node`` walks it (orchestration/node.py:109-147,424-443), and for a prompt that holds
  the pad id (the padded branch, sharded_inference_engine.py:165-177).
* ``Node`` itself (orchestration/node.py, unmodified) walking a request around rings of 1, 2 and 3 in-process
  nodes over this package's peer handles with a scripted engine: engine call sequence, tokens published
  through ``on_token``, stop by token cap and by eos, final bookkeeping (ref_node_trace.json).
"""
import asyncio
import json
import os
import sys
import types
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
GOLD = ROOT / "tests" / "golden"
REF = Path("/root/reference")

VOCAB = 97          # stand-in model: ids < 97 and positions < 256 are exact in bf16
HID = 4


def install_stubs():
  """sys.modules entries for the reference's imports that do not exist in this image."""
  def stub(name, **attrs):
    m = sys.modules.get(name)
    if m is None:
      m = types.ModuleType(name)
      sys.modules[name] = m
    m.__dict__.update(attrs)
    return m

  class _Any:
    def __init__(self, *a, **k):
      pass

    def __class_getitem__(cls, k):
      return cls

  stub("scapy")
  stub("scapy.all", get_if_addr=lambda *a, **k: [], get_if_list=lambda *a, **k: [])
  stub("torchtune")
  stub("torchtune.modules", FeedForward=_Any, TransformerDecoder=_Any, MultiHeadAttention=_Any, RMSNorm=_Any,
       TransformerSelfAttentionLayer=_Any, KVCache=_Any, RotaryPositionalEmbeddings=_Any)
  stub("torchtune.modules.attention_utils", _MaskType=object)
  stub("torchtune.models")
  stub("torchtune.models.convert_weights", hf_to_tune=lambda **k: None)
  stub("torchtune.models.llama3_1")
  stub("torchtune.models.llama3_1._position_embeddings", Llama3ScaledRoPE=_Any)
  stub("torchtune.models.qwen2")
  stub("torchtune.models.qwen2._positional_embeddings", Qwen2RotaryPositionalEmbeddings=_Any)

  # torchtune.generation: the two padding helpers are restated from torchtune 0.6.1
  # (torchtune/generation/_generation.py) because the reference's padded-prompt branch
  # (sharded_inference_engine.py:167-177) calls them; sample() is NOT provided (never called here)
  def get_causal_mask_from_padding_mask(padding_mask, target_seq_len=None):
    bsz, seq_len = padding_mask.shape
    target_seq_len = seq_len if target_seq_len is None else target_seq_len
    if target_seq_len < seq_len:
      raise AssertionError("target_seq_len cannot be shorter than the sequence length of the padding mask.")
    mask = torch.tril(torch.ones(seq_len, target_seq_len, device=padding_mask.device, dtype=bool), diagonal=0).repeat(bsz, 1, 1)
    mask.narrow(2, 0, seq_len).mul_(padding_mask[:, None, :].expand(-1, seq_len, seq_len))
    mask.diagonal(dim1=1, dim2=2).copy_(torch.Tensor([True]))
    return mask

  def get_position_ids_from_padding_mask(padding_mask):
    return ((padding_mask.cumsum(-1) - 1) * padding_mask).to(torch.int)

  stub("torchtune.generation", get_causal_mask_from_padding_mask=get_causal_mask_from_padding_mask,
       get_position_ids_from_padding_mask=get_position_ids_from_padding_mask)
  for name in ("aiohttp_cors", "grpc_tools"):
    stub(name)
  if str(REF) not in sys.path:
    sys.path.insert(0, str(REF))


def _import_reference():
  install_stubs()
  cwd = os.getcwd()
  os.chdir("/tmp")    # device_capabilities.py drops a run_<date>.log into the cwd on import
  try:
    from xotorch.inference.torch.models import llm_utils as LU
    from xotorch.inference.torch.models import general_mha as GM
    from xotorch.inference.torch import sharded_inference_engine as SE
    from xotorch.inference.shard import Shard
  finally:
    os.chdir(cwd)
  return LU, GM, SE, Shard


# ----------------------------------------------------------------------------- 1..3: pure functions
def gen_config_permute_state(LU):
  import tempfile
  from xotorch_support_jetson_b200.models import model_cards
  out = {"load_model_config": {}, "permute": [], "shard_inference_state": {}}
  for mid, card in model_cards.items():
    with tempfile.TemporaryDirectory() as td:
      p = Path(td) / "config.json"
      p.write_text(json.dumps(card["hf_config"]))
      for org in ("False", "true"):
        os.environ["TORCH_USE_ORG_SEQ"] = org
        cfg = LU.load_model_config(p)
        cfg["torch_dtype"] = str(cfg["torch_dtype"])
        out["load_model_config"][f"{mid}|org_seq={org}"] = cfg
      os.environ.pop("TORCH_USE_ORG_SEQ", None)
  # _permute as a ROW PERMUTATION: row r of the input carries the value r in every column
  for (n_heads, head_dim) in ((8, 128), (32, 128), (8, 64), (32, 64), (24, 128), (2, 16), (4, 32)):
    dim = 8
    rows = n_heads * head_dim
    t = torch.arange(rows, dtype=torch.float32)[:, None].repeat(1, dim)
    p = LU._permute(t, n_heads, head_dim, dim)
    assert torch.equal(p[:, 0], p[:, -1])
    out["permute"].append({"n_heads": n_heads, "head_dim": head_dim, "src_row_of_out_row": p[:, 0].to(torch.int64).tolist()})
  # ShardInferenceState: to_dict of a small state, and what from_dict makes of that dict
  s = LU.ShardInferenceState(tokens=torch.tensor([[5, 9, 11]]), input_pos=torch.arange(0, 7).unsqueeze(0),
                             mask=torch.tril(torch.ones(7, 7, dtype=torch.bool)).unsqueeze(0), curr_pos=3)
  d = s.to_dict()
  s2 = LU.ShardInferenceState()
  s2.from_dict(json.loads(json.dumps(d)))    # through JSON, as the gRPC hop does (grpc_peer_handle.py:209-230)
  out["shard_inference_state"] = {
    "to_dict": d,
    "from_dict": {"tokens": s2.tokens.tolist(), "tokens_dtype": str(s2.tokens.dtype), "input_pos": s2.input_pos.tolist(),
                  "mask_dtype": str(s2.mask.dtype), "mask_shape": list(s2.mask.shape), "curr_pos": s2.curr_pos},
  }
  return out


# ----------------------------------------------------------------------------- 4: engine state machine
def next_token_rule(tok: int, pos: int) -> int:
  """The recording model's 'language': the id after `tok` at position `pos`."""
  return (3 * int(tok) + int(pos) + 1) % VOCAB


class RecordingDecoder:
  """Stands where the torchtune ShardTransformerDecoder would (ShardedGeneralModel.model): records
  what generate() hands it and answers with shape-correct tensors.  Hidden rows carry (token id,
  position) so that a downstream shard can apply next_token_rule."""

  def __init__(self, shard, log):
    self.shard, self.log = shard, log
    self.enabled = False
    self.decoder_max_cache_seq_len = None
    self.tok_embeddings = types.SimpleNamespace(num_embeddings=VOCAB)
    self.output_hidden_states = None

  def caches_are_enabled(self):
    return self.enabled

  def setup_caches(self, batch_size, dtype, decoder_max_seq_len=None):
    self.enabled = True
    self.decoder_max_cache_seq_len = decoder_max_seq_len
    self.log.append({"op": "setup_caches", "batch_size": batch_size, "decoder_max_seq_len": decoder_max_seq_len})

  def reset_caches(self):
    self.log.append({"op": "reset_caches"})

  def __call__(self, tokens=None, mask=None, input_pos=None, hidden_state=None, dtype=None):
    pos = input_pos.reshape(-1).tolist() if input_pos.ndim <= 1 else input_pos.tolist()
    m = mask.to(torch.int64)
    prefix = (torch.arange(m.size(-1)) < m.sum(-1, keepdim=True)).to(torch.int64)
    rec = {"op": "forward", "shard": [self.shard.start_layer, self.shard.end_layer],
           "tokens": None if tokens is None else tokens.tolist(),
           "hidden_shape": None if hidden_state is None else list(hidden_state.shape),
           "input_pos": pos, "input_pos_shape": list(input_pos.shape),
           "mask_shape": list(mask.shape),
           "mask_true_per_row": m.sum(-1).reshape(-1).tolist(),
           # every mask row is a prefix of ones (the tril branch): row r attends keys [0, count)
           "mask_rows_are_prefixes": bool(torch.equal(m, prefix)),
           "output_hidden_states": list(self.output_hidden_states)}
    if not rec["mask_rows_are_prefixes"]:
      rec["mask_rows"] = m.reshape(-1, m.size(-1)).tolist()    # padded-prompt branch: the rows themselves
    self.log.append(rec)
    if hidden_state is not None:
      src = hidden_state.float()
      ids, ps = src[..., 0], src[..., 1]
    else:
      ids = tokens.float()
      flat = input_pos.reshape(1, -1).float() if input_pos.ndim <= 1 else input_pos.float()
      ps = flat.expand_as(ids) if flat.shape == ids.shape else flat.reshape(ids.shape[0], -1)
    B, s = ids.shape
    if self.shard.is_last_layer():
      out = torch.zeros(B, s, VOCAB, dtype=dtype)
      for b in range(B):
        for j in range(s):
          out[b, j, next_token_rule(ids[b, j].item(), ps[b, j].item())] = 1.0
      return out
    out = torch.zeros(B, s, HID, dtype=dtype)
    out[..., 0] = ids
    out[..., 1] = ps
    return out


class IdTokenizer:
  """'5 9 11' -> ids; no pad_id attribute, pad_token_id None -> the reference falls back to pad id 0."""
  pad_token_id = None
  eos_token_id = VOCAB - 1

  def encode(self, prompt, return_tensors=None):
    ids = [int(x) for x in prompt.split()]
    return torch.tensor([ids]) if return_tensors == "pt" else ids

  def decode(self, ids):
    return " ".join(str(int(i)) for i in ids)


def make_ref_engine(GM, SE, LU, shard, log, max_generated_tokens):
  eng = SE.TorchDynamicShardInferenceEngine(shard_downloader=None)
  eng.shard = shard                       # ensure_shard() returns at once (weights are not the subject)
  eng.state = LU.ShardInferenceState()    # what ensure_shard() does at sharded_inference_engine.py:386
  eng.tokenizer = IdTokenizer()
  eng.model_config = {"torch_dtype": torch.bfloat16}
  sm = GM.ShardedGeneralModel.__new__(GM.ShardedGeneralModel)   # generate() is the reference's; the decoder is the recorder
  torch.nn.Module.__init__(sm)
  sm.shard, sm.dtype, sm.device = shard, torch.bfloat16, torch.device("cpu")
  sm.max_generated_tokens = max_generated_tokens
  object.__setattr__(sm, "model", RecordingDecoder(shard, log))
  eng.sharded_model = sm
  return eng


def summarize_state(d):
  if d is None:
    return None
  return {"curr_pos": d["curr_pos"], "tokens": d["tokens"], "input_pos_len": len(d["input_pos"][0]),
          "mask_shape": [len(d["mask"]), len(d["mask"][0]), len(d["mask"][0][0])]}


def gen_engine_trace(LU, GM, SE, Shard, prompt: str, n_layers: int, splits, n_new: int, max_generated_tokens: int):
  """Walk the ring the way Node does: first shard gets encode()'s ids, every next shard the previous
  output + state, the last shard's logits are arg-maxed (greedy) and the id goes back to the first
  shard together with the state the last shard returned (node.py:109-147)."""
  log = []
  shards = [Shard("recording-llama", a, b, n_layers) for a, b in splits]
  engines = [make_ref_engine(GM, SE, LU, s, log, max_generated_tokens) for s in shards]
  calls = []

  async def run():
    toks = await engines[0].encode(shards[0], prompt)
    x = np.asarray(toks.reshape(1, -1))
    calls.append({"api": "encode", "ids": x.tolist()})
    state = None
    out_ids = []
    for step in range(n_new):
      for e, sh in zip(engines, shards):
        mark = len(log)
        y, state = await e.infer_tensor("req", sh, x, state)
        calls.append({"api": "infer_tensor", "shard": [sh.start_layer, sh.end_layer], "in_shape": list(x.shape),
                      "in_dtype": str(x.dtype), "out_shape": list(y.shape), "out_dtype": str(y.dtype),
                      "state_out": summarize_state(state), "model_ops": log[mark:]})
        x = y
      tok = int(np.argmax(x, axis=-1).reshape(-1)[0])
      out_ids.append(tok)
      x = np.array([[tok]], dtype=np.int64)     # node.py:126,145: forwarded as token.reshape(1, -1)
    return out_ids

  out_ids = asyncio.run(run())
  return {"prompt": prompt, "n_layers": n_layers, "splits": [list(s) for s in splits], "vocab": VOCAB, "hidden": HID,
          "max_generated_tokens": max_generated_tokens, "generated": out_ids, "calls": calls}


# ----------------------------------------------------------------------------- 5: the reference's Node
def gen_node_trace(n_nodes: int, prompt: str, n_layers: int, max_generate_tokens: int, eos_token_id: int):
  """The reference's OWN ``Node`` (orchestration/node.py, unmodified) walking one request around a ring of
  ``n_nodes`` in-process nodes.  Engines: the scripted engine; peers / discovery / server: this package's
  ``NvlinkPeerHandle`` / ``BoxDiscovery`` / ``NullServer`` (they are what is being fitted to the reference's
  plug-in points); partitioning: the reference's memory-weighted ring strategy."""
  install_stubs()
  cwd = os.getcwd()
  os.chdir("/tmp")
  try:
    from xotorch.orchestration.node import Node
    from xotorch.inference.shard import Shard as RefShard
    from xotorch.topology.ring_memory_weighted_partitioning_strategy import RingMemoryWeightedPartitioningStrategy
    from xotorch.topology.device_capabilities import DeviceCapabilities, DeviceFlops
  finally:
    os.chdir(cwd)
  from oracle.scripted_engine import ScriptedEngine
  from xotorch_support_jetson_b200.networking import BoxDiscovery, NullServer, NvlinkPeerHandle
  log, tokens = [], []
  ids = [f"gpu{n_nodes - 1 - i:02d}-rank{i}" for i in range(n_nodes)]
  caps = DeviceCapabilities(model="Linux Box (NVIDIA B200)", chip="NVIDIA B200", memory=183359, flops=DeviceFlops(fp32=0, fp16=0, int8=0))

  async def run():
    nodes = [Node(ids[i], NullServer(), ScriptedEngine(ids[i], log, eos_token_id), BoxDiscovery(), None,
                  RingMemoryWeightedPartitioningStrategy(), max_generate_tokens=max_generate_tokens) for i in range(n_nodes)]
    for i, nd in enumerate(nodes):
      nd.device_capabilities = caps
      nd.discovery.handles = [NvlinkPeerHandle(nodes[j], j, caps) for j in range(n_nodes) if j != i]
    for nd in nodes:
      await nd.update_peers()
    for nd in nodes:
      await nd.collect_topology(set())
    done = asyncio.Event()

    def on_token(rid, toks, fin):
      tokens.append({"request_id": rid, "tokens": [int(t) for t in toks], "is_finished": bool(fin)})
      if fin:
        done.set()

    # the entry node hears about tokens through send_result from the last node (or directly when it is the last)
    nodes[0].on_token.register("trace").on_next(on_token)
    await nodes[0].process_prompt(RefShard("scripted", 0, 0, n_layers), prompt, "req-1")
    await asyncio.wait_for(done.wait(), 30)
    await asyncio.sleep(0.05)
    shards = [nodes[i].get_current_shard(RefShard("scripted", 0, 0, n_layers)) for i in range(n_nodes)]
    return {"buffered": {ids[i]: [int(t) for t in nodes[i].buffered_token_output.get("req-1", ([], False))[0]] for i in range(n_nodes)},
            "finished": {ids[i]: bool(nodes[i].buffered_token_output.get("req-1", ([], False))[1]) for i in range(n_nodes)},
            "outstanding": {ids[i]: dict(nodes[i].outstanding_requests) for i in range(n_nodes)},
            "shards": [[s.start_layer, s.end_layer, s.n_layers] for s in shards],
            "peer_graph": {k: sorted(c.to_id for c in v) for k, v in nodes[0].topology.peer_graph.items()}}

  end = asyncio.run(run())
  return {"n_nodes": n_nodes, "node_ids": ids, "prompt": prompt, "n_layers": n_layers, "max_generate_tokens": max_generate_tokens,
          "eos_token_id": eos_token_id, "engine_log": log, "on_token": tokens, "end": end}


def main():
  if not REF.exists():
    print("reference not mounted: nothing generated")
    return
  GOLD.mkdir(parents=True, exist_ok=True)
  LU, GM, SE, Shard = _import_reference()
  (GOLD / "ref_config_permute_state.json").write_text(json.dumps(gen_config_permute_state(LU), separators=(",", ":")))
  traces = {
    "full_shard": gen_engine_trace(LU, GM, SE, Shard, "5 9 11 2 40", 4, [(0, 3)], 6, 16),
    "two_shards": gen_engine_trace(LU, GM, SE, Shard, "5 9 11 2 40", 4, [(0, 1), (2, 3)], 6, 16),
    # a prompt holding id 0 takes the padded branch (pad id falls back to 0): positions skip the pad slots
    "padded_prompt": gen_engine_trace(LU, GM, SE, Shard, "0 0 7 3 0 12 31", 4, [(0, 3)], 4, 16),
  }
  (GOLD / "ref_engine_trace.json").write_text(json.dumps(traces, separators=(",", ":")))
  for k, v in traces.items():
    print(k, "generated", v["generated"], "calls", len(v["calls"]))
  node_traces = {
    "one_node_cap": gen_node_trace(1, "5 9 11 2 40", 4, 6, 96),
    "two_nodes_cap": gen_node_trace(2, "5 9 11 2 40", 4, 6, 96),
    "three_nodes_eos": gen_node_trace(3, "5 9 11 2 40", 6, 64, 83),     # 83 is the third id the toy model emits: stop by eos
  }
  (GOLD / "ref_node_trace.json").write_text(json.dumps(node_traces, separators=(",", ":")))
  for k, v in node_traces.items():
    print(k, "tokens", [t["tokens"] for t in v["on_token"]], "engine calls", len(v["engine_log"]))
  print("written: ref_config_permute_state.json, ref_engine_trace.json, ref_node_trace.json")


if __name__ == "__main__":
  main()
