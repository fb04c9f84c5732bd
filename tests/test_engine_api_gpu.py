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


def test_next_step_speculation_is_invisible(built_lib, tmp_path):
  """sample() enqueues the next decode forward on the ids it just drew (b200_engine.py module
  docstring).  Logits, ids and states must be bit-identical with it on and off: when the caller
  feeds the sampled ids back (hit), when it feeds something else (miss: the real forward
  overwrites the speculated KV entry), when it samples twice, and with two requests interleaved."""
  from xotorch_support_jetson_b200.shard import Shard
  mid, nl = "llama-3.2-tiny", 4
  sh = Shard(mid, 0, nl - 1, nl)
  rs = np.random.RandomState(5)
  pa, pb = rs.randint(3, 200, (2, 21)).astype(np.int64), rs.randint(3, 200, (1, 9)).astype(np.int64)

  async def script(eng):
    rec = []
    la, sa = await eng.infer_tensor("a", sh, pa, None)
    lb, sb = await eng.infer_tensor("b", sh, pb, None)
    rec += [la.copy(), lb.copy()]
    ta = await eng.sample(la, temp=0.0)            # la is no longer the engine's last result: uploaded, no speculation
    for step in range(6):
      la, sa = await eng.infer_tensor("a", sh, ta, sa)
      rec.append(la.copy())
      ta = await eng.sample(la, temp=0.0)          # speculates step+1 of "a"
      if step == 2:
        ta = await eng.sample(la, temp=0.0)        # sampled twice: second speculation replaces the first
      if step == 3:
        ta = (ta + 1).astype(np.int32)             # caller overrides the ids: speculation must be discarded
      if step % 2 == 1:                            # another request in between keeps its own state
        tb = await eng.sample(lb, temp=0.0)
        lb, sb = await eng.infer_tensor("b", sh, tb, sb)
        rec.append(lb.copy())
      rec.append(ta.copy())
    assert sa["curr_pos"] == 21 + 6 and [len(r) for r in sa["tokens"]] == [27, 27]
    return rec

  async def run():
    on, off = _engine(tmp_path, speculate_next=True), _engine(tmp_path, speculate_next=False)
    r_on, r_off = await script(on), await script(off)
    assert len(r_on) == len(r_off)
    for i, (x, y) in enumerate(zip(r_on, r_off)):
      assert x.shape == y.shape and np.array_equal(x, y), f"record {i} differs with speculation on"
    assert off.spec_hits == 0 and off.spec_misses == 0
    # (an overridden speculation counts as a miss only if no later speculation replaced it first)
    assert on.spec_hits >= 3, (on.spec_hits, on.spec_misses)

  asyncio.run(run())


def test_sampler_speculation_is_invisible(built_lib, tmp_path):
  """With one request in flight the engine also runs the SAMPLER ahead (parameters of the previous sample() call) and
  the step after it, so the host only collects.  Ids, logits and the noise stream must equal the unspeculated engine's:
  plain loop (hits), a change of sampling parameters, a step where the caller feeds its own ids without sampling, and
  sampling the same logits twice."""
  from xotorch_support_jetson_b200.shard import Shard
  mid, nl = "llama-3.2-tiny", 4
  sh = Shard(mid, 0, nl - 1, nl)
  prompt = np.random.RandomState(9).randint(3, 900, (2, 17)).astype(np.int64)

  async def script(eng):
    rec = []
    lg, st = await eng.infer_tensor("a", sh, prompt, None)
    for step in range(12):
      rec.append(lg.copy())
      kw = dict(temp=0.8, top_k=20) if step < 7 else dict(temp=0.3, top_k=5)      # parameters change at step 7
      tok = await eng.sample(lg, **kw)
      if step == 4:
        tok = await eng.sample(lg, **kw)                                          # sampled twice: second draw, new noise
      rec.append(tok.copy())
      if step == 9:
        tok = np.array([[5], [6]], dtype=np.int32)                                # caller overrides the ids
      lg, st = await eng.infer_tensor("a", sh, tok, st)
      if step == 10:                                                               # ... and once goes on without sampling at all
        rec.append(lg.copy())
        lg, st = await eng.infer_tensor("a", sh, np.array([[7], [8]], dtype=np.int32), st)
    rec.append(lg.copy())
    eng.end_request("a")          # gives back the noise of a sampler run that was enqueued ahead and never asked for
    return rec

  async def run():
    on, off = _engine(tmp_path, speculate_next=True), _engine(tmp_path, speculate_next=False)
    r_on, r_off = await script(on), await script(off)
    assert len(r_on) == len(r_off)
    for i, (x, y) in enumerate(zip(r_on, r_off)):
      assert x.shape == y.shape and np.array_equal(x, y), f"record {i} differs with sampler speculation on"
    assert off.sample_spec_hits == 0
    assert on.sample_spec_hits >= 6 and on.spec_hits >= 6, (on.sample_spec_hits, on.spec_hits)
    assert on.sample_calls == off.sample_calls

  asyncio.run(run())


def test_ring_last_stage_sampler_speculation_is_invisible(built_lib, tmp_path):
  """The last stage of a ring with two requests in flight: the sampler of each is run ahead in forward order
  (B200ShardInferenceEngine._speculate_ring_sample).  Ids and logits must equal the unspeculated engine's when the
  sample() calls come in that order (hits), in the opposite order (all run-ahead noise given back), when a request's
  parameters change, and when one request ends with a sampler still run ahead."""
  from xotorch_support_jetson_b200.shard import Shard
  mid, nl = "llama-3.2-tiny", 4
  sh = Shard(mid, 2, nl - 1, nl)
  rs = np.random.RandomState(4)
  hid = lambda b, s: (rs.randn(b, s, 256) * 0.5).astype(np.float32)
  script_inputs = {"a": [hid(2, 9)] + [hid(2, 1) for _ in range(8)], "b": [hid(1, 5)] + [hid(1, 1) for _ in range(8)]}

  async def script(eng):
    rec, lg, st = [], {}, {"a": None, "b": None}
    for step in range(8):
      for rid in ("a", "b"):
        lg[rid], st[rid] = await eng.infer_tensor(rid, sh, script_inputs[rid][step], st[rid])
        rec.append(lg[rid].copy())
      order = ("b", "a") if step in (3, 6) else ("a", "b")
      for rid in order:
        kw = dict(temp=0.8, top_k=20) if (step < 5 or rid == "a") else dict(temp=0.4, top_k=3)
        rec.append((await eng.sample(lg[rid], **kw)).copy())
    lg["a"], st["a"] = await eng.infer_tensor("a", sh, script_inputs["a"][8], st["a"])
    lg["b"], st["b"] = await eng.infer_tensor("b", sh, script_inputs["b"][8], st["b"])
    eng.end_request("a")                                     # a's sampler was run ahead and is never asked for
    rec.append((await eng.sample(lg["b"], temp=0.4, top_k=3)).copy())
    eng.end_request("b")
    return rec

  async def run():
    on, off = _engine(tmp_path, speculate_next=True), _engine(tmp_path, speculate_next=False)
    r_on, r_off = await script(on), await script(off)
    assert len(r_on) == len(r_off)
    for i, (x, y) in enumerate(zip(r_on, r_off)):
      assert x.shape == y.shape and np.array_equal(x, y), f"record {i} differs with sampler speculation on"
    assert off.sample_spec_hits == 0 and on.sample_spec_hits >= 6, on.sample_spec_hits
    assert on.sample_calls == off.sample_calls

  asyncio.run(run())


def test_generate_device_loop_equals_stepwise_calls(built_lib, tmp_path):
  """generate() (token loop on the device) against the same loop driven from the host
  call by call as Node does (node.py:109-147): same ids, same stop, same state afterwards --
  greedy and with reference-mode noise (the noise stream must be rewound past the look-ahead)."""
  from xotorch_support_jetson_b200.shard import Shard
  mid, nl = "llama-3.2-tiny", 4
  sh = Shard(mid, 0, nl - 1, nl)
  rs = np.random.RandomState(11)
  prompt = rs.randint(3, 200, (2, 17)).astype(np.int64)

  async def stepwise(eng, rid, p, n, **kw):
    logits, state = await eng.infer_tensor(rid, sh, p, None)
    ids = []
    for k in range(n):
      tok = await eng.sample(logits, **kw)
      ids.append(tok)
      if k < n - 1:
        logits, state = await eng.infer_tensor(rid, sh, tok, state)
    return np.concatenate(ids, axis=1), state

  async def run():
    for kw in (dict(temp=0.0), dict(temp=0.9, top_k=12)):
      a, b = _engine(tmp_path, speculate_next=False), _engine(tmp_path)
      ids_a, st_a = await stepwise(a, "r", prompt, 9, **kw)
      seen = []
      ids_b, st_b = await b.generate("r", sh, prompt, max_generate_tokens=9, eos_token_id=-1, lookahead=4,
                                     on_token=lambda rid, toks, fin: seen.append((rid, toks, fin)), **kw)
      await asyncio.sleep(0)   # let the queued callbacks run
      assert ids_b.shape == (2, 9) and ids_b.dtype == np.int32 and np.array_equal(ids_a, ids_b), (kw, ids_a, ids_b)
      assert st_a["curr_pos"] == st_b["curr_pos"] == 17 + 8 and st_a["tokens"] == st_b["tokens"]
      assert [t for _, t, _ in seen] == ids_b.T.tolist() and [f for _, _, f in seen] == [False] * 8 + [True]
      # both continue call by call from where they stopped: same logits, same next ids (noise stream in step)
      for _ in range(2):
        la, st_a = await a.infer_tensor("r", sh, ids_a[:, -1:], st_a)
        lb, st_b = await b.infer_tensor("r", sh, ids_b[:, -1:], st_b)
        assert np.array_equal(la, lb)
        ids_a, ids_b = await a.sample(la, **kw), await b.sample(lb, **kw)
        assert np.array_equal(ids_a, ids_b)
    # stop on eos: one row, eos := the third greedy id; a second generate() continues the request
    a, b = _engine(tmp_path, speculate_next=False), _engine(tmp_path)
    ref, _ = await stepwise(a, "e", prompt[:1], 8, temp=0.0)
    eos = int(ref[0, 2])
    n_stop = int(np.argmax(ref[0] == eos)) + 1
    got, st = await b.generate("e", sh, prompt[:1], max_generate_tokens=8, temp=0.0, eos_token_id=eos, lookahead=3)
    assert got.tolist() == ref[:, :n_stop].tolist() and st["curr_pos"] == 17 + n_stop - 1
    more, st = await b.generate("e", sh, got[:, -1:], st, max_generate_tokens=8 - n_stop, temp=0.0, eos_token_id=-1)
    assert more.tolist() == ref[:, n_stop:].tolist()
    # the KV slot bounds the loop (max_tokens_per_seq = 128): 120-token prompt -> at most 9 ids
    long_p = rs.randint(3, 200, (1, 120)).astype(np.int64)
    got, st = await b.generate("cap", sh, long_p, max_generate_tokens=64, temp=0.0, eos_token_id=-1)
    assert got.shape == (1, 9) and st["curr_pos"] == 128
    with pytest.raises(ValueError):
      await b.generate("x", Shard(mid, 0, 1, nl), prompt, max_generate_tokens=2)

  asyncio.run(run())


def test_lm_head_writes_host_logits_directly(built_lib, tmp_path, monkeypatch):
  """xb_forward_host_logits: the LM-head kernel stores the logits into pinned host memory from its
  epilogue.  Same bits as the device copy, at stage level (prefill, eager and graph-replayed decode,
  ragged batch) and through the plugin API with the path on and off."""
  from xotorch_support_jetson_b200.shard import Shard
  mid, nl = "llama-3.2-tiny", 4
  sh = Shard(mid, 0, nl - 1, nl)
  rs = np.random.RandomState(3)
  prompt = rs.randint(3, 200, (3, 13)).astype(np.int64)

  async def run():
    monkeypatch.setenv("XB_API_DIRECT_LOGITS", "1")
    a = _engine(tmp_path, speculate_next=False)
    monkeypatch.setenv("XB_API_DIRECT_LOGITS", "0")
    b = _engine(tmp_path, speculate_next=False)
    assert a.direct_host_logits and not b.direct_host_logits
    la, sa = await a.infer_tensor("r", sh, prompt, None)
    lb, sb = await b.infer_tensor("r", sh, prompt, None)
    assert np.array_equal(la, lb) and la.dtype == np.float32 and la.shape == (3, a.stage.vocab)
    for _ in range(5):
      ta, tb = await a.sample(la, temp=0.0), await b.sample(lb, temp=0.0)
      assert np.array_equal(ta, tb)
      la, sa = await a.infer_tensor("r", sh, ta, sa)
      lb, sb = await b.infer_tensor("r", sh, tb, sb)
      assert np.array_equal(la, lb)
    # stage level: host buffer == device result, ragged decode batch
    st = a.stage
    with torch.cuda.stream(torch.cuda.Stream()):
      host = torch.empty((3, st.vocab), dtype=torch.float32).pin_memory()
      slots = a.requests["r"].slots
      pos = a.requests["r"].pos
      tok = torch.tensor([5, 6, 7], dtype=torch.int32, device="cuda")
      for k in range(3):   # eager, capture, replay
        dev = st.forward(slots, [1, 1, 1], [pos + k] * 3, tokens=tok, host_logits=host)
        torch.cuda.current_stream().synchronize()
        assert torch.equal(dev.cpu(), host)
    with pytest.raises(RuntimeError, match="page-locked"):
      st.lib.xb_forward_host_logits.restype = __import__("ctypes").c_int
      bad = torch.empty((3, st.vocab), dtype=torch.float32)
      object.__setattr__(bad, "_fake", True)
      import ctypes as C
      from xotorch_support_jetson_b200 import _cabi
      arr = (C.c_int32 * 3)
      bt = _cabi.Batch(n_seqs=3, slot_ids=arr(*slots), q_lens=arr(1, 1, 1), positions=arr(pos, pos, pos), tokens=tok.data_ptr(), hidden_in=None)
      _cabi.check(st.lib.xb_forward_host_logits(st._h, C.byref(bt), None, C.c_void_p(bad.data_ptr()), None))

  asyncio.run(run())
