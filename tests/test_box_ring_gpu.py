"""The intra-box ring on real engines: ``BoxNode``s (one per stage) wired by ``build_box_ring`` -- this package's
``NvlinkPeerHandle`` / ``BoxDiscovery`` -- walking requests through ``B200ShardInferenceEngine``s of ONE process.
Hidden rows travel as device references; decode steps go through the consumer's landing buffers (stored there by
the producer's last kernel).  The ids must be those of a single full-shard engine driven call by call, the way
the reference's contract test demands of any partition (xotorch/inference/test_inference_engine.py:11,43-44).
Stages share GPU 0 on a one-GPU box; with more GPUs the same test also spreads them over the devices."""
import asyncio

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
MID, NL = "llama-3.1-tiny", 6


def _engine(tmp_path, device=0, **kw):
  from xotorch_support_jetson_b200.b200_engine import B200ShardInferenceEngine
  from xotorch_support_jetson_b200.synth import LocalShardDownloader
  dl = LocalShardDownloader(root=tmp_path, seed=0, std=0.05, norm_jitter=0.1)
  kw.setdefault("max_seqs", 8)
  return B200ShardInferenceEngine(dl, device=device, max_tokens_per_seq=128, max_batch_tokens=64, page_tokens=16, **kw)


def _layouts():
  out = [(2, [0, 0]), (3, [0, 0, 0]), (1, [0])]
  if torch.cuda.device_count() >= 2:
    out.append((2, [0, 1]))
  if torch.cuda.device_count() >= 3:
    out.append((3, [0, 1, 2]))
  return out


async def _direct_loop(tmp_path, prompt, n_new, eos):
  """A full-shard engine driven like Node does on a one-node ring, without any node: the ids to match."""
  from xotorch_support_jetson_b200.shard import Shard
  eng = _engine(tmp_path, speculate_next=False)
  sh = Shard(MID, 0, NL - 1, NL)
  out, st = await eng.infer_prompt("direct", sh, prompt)
  ids = []
  for _ in range(n_new):
    tok = await eng.sample(out, temp=0.0)
    ids.append(int(tok.item()))
    if ids[-1] == eos:
      break
    out, st = await eng.infer_tensor("direct", sh, tok.reshape(1, -1), st)
  return ids


@pytest.mark.parametrize("n_nodes,devices", _layouts(), ids=lambda v: str(v))
def test_box_ring_tokens_equal_the_direct_engine_loop(built_lib, tmp_path, n_nodes, devices):
  from xotorch_support_jetson_b200.networking import DeviceTensorRef
  from xotorch_support_jetson_b200.orchestration import build_box_ring
  from xotorch_support_jetson_b200.shard import Shard
  prompt = "<ids:" + ",".join(str(11 + 13 * i) for i in range(23)) + ">"
  n_new = 9

  async def run():
    engines = [_engine(tmp_path, device=d) for d in devices]
    seen = []
    for e in engines:            # record what each engine is handed
      inner = e.infer_tensor

      async def spy(request_id, shard, x, state=None, _inner=inner):
        seen.append(type(x).__name__ if not isinstance(x, DeviceTensorRef) else ("landing" if x.landing is not None else "device"))
        return await _inner(request_id, shard, x, state)

      e.infer_tensor = spy
    nodes = build_box_ring(engines, max_generate_tokens=n_new)
    for nd in nodes:
      await nd.start()
    got, done = [], asyncio.Event()

    def on_token(rid, toks, fin):
      got.extend(int(t) for t in toks)
      if fin:
        done.set()

    nodes[0].on_token.register("t").on_next(on_token)
    await nodes[0].process_prompt(Shard(MID, 0, 0, NL), prompt, "r1")
    await asyncio.wait_for(done.wait(), 60)
    await asyncio.sleep(0.05)
    eos = engines[0].tokenizer.eos_token_id
    want = await _direct_loop(tmp_path, prompt, n_new, eos)
    connected = [e._next_epoch is not None for e in engines[:-1]]
    for e in engines:
      if e.stage is not None:
        e.stage.close()
    return got, want, seen, connected

  got, want, seen, connected = asyncio.run(run())
  assert got == want and len(got) == n_new
  if n_nodes > 1:
    assert "ndarray" in seen                       # token ids from the last node back to the first
    assert "device" in seen                        # the prompt's hidden rows: a device reference, no host copy
    assert "landing" in seen and all(connected)    # decode steps: stored by the producer's kernel into the consumer's buffer
    assert seen.count("landing") >= (n_nodes - 1) * (n_new - 2)


def test_box_ring_two_requests_in_flight_and_batched_rows(built_lib, tmp_path):
  """Two requests walk the ring at the same time (one of them with 3 sequences): each gets exactly the ids it
  gets when it runs alone -- landing slots, KV slots and per-request state do not mix."""
  from xotorch_support_jetson_b200.orchestration import build_box_ring
  from xotorch_support_jetson_b200.shard import Shard
  g = np.random.default_rng(3)
  pa = g.integers(1, 2000, size=(1, 17)).astype(np.int64)
  pb = g.integers(1, 2000, size=(3, 9)).astype(np.int64)
  n_new = 6

  async def ring(prompts):
    engines = [_engine(tmp_path), _engine(tmp_path)]
    nodes = build_box_ring(engines, max_generate_tokens=n_new)
    for nd in nodes:
      await nd.start()
    got = {rid: [] for rid in prompts}
    left = set(prompts)
    done = asyncio.Event()

    def on_token(rid, toks, fin):
      got[rid].append([int(t) for t in toks])
      if fin:
        left.discard(rid)
        if not left:
          done.set()

    nodes[0].on_token.register("t").on_next(on_token)
    base = Shard(MID, 0, 0, NL)
    await asyncio.gather(*[nodes[0].process_tokens(base, p, rid) for rid, p in prompts.items()])
    await asyncio.wait_for(done.wait(), 60)
    await asyncio.sleep(0.05)
    for e in engines:
      e.stage.close()
    return got

  both = asyncio.run(ring({"a": pa, "b": pb}))
  only_a = asyncio.run(ring({"a": pa}))
  only_b = asyncio.run(ring({"b": pb}))
  assert both["a"] == only_a["a"] and both["b"] == only_b["b"]
  assert len(both["a"]) == n_new and all(len(step) == 3 for step in both["b"])
