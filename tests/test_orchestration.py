"""The orchestration / networking binding (SURVEY 8 rows f3 and the Node dispatch): this package's peer
handles, discovery, server and ``BoxNode`` against traces recorded from the reference's OWN ``Node``
(tests/golden/ref_node_trace.json, made by oracle/gen_ref_fixtures.py with the reference's
orchestration/node.py unmodified), and -- where /root/reference is mounted (the build container) -- against
the reference's ``Node`` run live.  CPU only; engines are the scripted engine of oracle/scripted_engine.py."""
import asyncio
import json
import os
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
GOLD = ROOT / "tests" / "golden"
TRACES = json.loads((GOLD / "ref_node_trace.json").read_text())


def _run_box_ring(tr):
  from oracle.scripted_engine import ScriptedEngine
  from xotorch_support_jetson_b200.orchestration import build_box_ring
  from xotorch_support_jetson_b200.shard import Shard
  log, tokens = [], []
  n = tr["n_nodes"]

  async def run():
    engines = [ScriptedEngine(tr["node_ids"][i], log, tr["eos_token_id"]) for i in range(n)]
    nodes = build_box_ring(engines, max_generate_tokens=tr["max_generate_tokens"])
    assert [nd.id for nd in nodes] == tr["node_ids"]
    for nd in nodes:
      await nd.start()
    for nd in nodes:
      await nd.collect_topology(set())
    done = asyncio.Event()

    def on_token(rid, toks, fin):
      tokens.append({"request_id": rid, "tokens": [int(t) for t in toks], "is_finished": bool(fin)})
      if fin:
        done.set()

    nodes[0].on_token.register("trace").on_next(on_token)
    base = Shard("scripted", 0, 0, tr["n_layers"])
    await nodes[0].process_prompt(base, tr["prompt"], "req-1")
    await asyncio.wait_for(done.wait(), 30)
    await asyncio.sleep(0.05)
    ids = tr["node_ids"]
    return {"buffered": {ids[i]: [int(t) for t in nodes[i].buffered_token_output.get("req-1", ([], False))[0]] for i in range(n)},
            "finished": {ids[i]: bool(nodes[i].buffered_token_output.get("req-1", ([], False))[1]) for i in range(n)},
            "outstanding": {ids[i]: dict(nodes[i].outstanding_requests) for i in range(n)},
            "shards": [list(nodes[i].get_current_shard(base)) [1:] for i in range(n)],
            "peer_graph": {k: sorted(v.keys()) for k, v in nodes[0].topology.peer_graph.items()}}

  end = asyncio.run(run())
  return log, tokens, end


@pytest.mark.parametrize("name", sorted(TRACES.keys()))
def test_box_node_walks_the_ring_like_the_reference_node(name):
  """Same engine calls in the same order with the same inputs and positions, same tokens published in the
  same order with the same finished flags (stop by token cap, stop by eos), same final bookkeeping, same
  layer ranges and peer graph as the reference's Node produced."""
  tr = TRACES[name]
  log, tokens, end = _run_box_ring(tr)
  assert log == tr["engine_log"]
  assert tokens == tr["on_token"]
  assert end == tr["end"]


@pytest.mark.skipif(not Path("/root/reference/xotorch/orchestration/node.py").exists(), reason="reference not mounted")
@pytest.mark.parametrize("name", sorted(TRACES.keys()))
def test_reference_node_runs_on_this_packages_peer_handles(name):
  """The reference's Node, imported unmodified from /root/reference, over NvlinkPeerHandle / BoxDiscovery /
  NullServer: reproduces the committed trace (so the fixture is what the reference does today)."""
  from oracle.gen_ref_fixtures import gen_node_trace
  tr = TRACES[name]
  got = gen_node_trace(tr["n_nodes"], tr["prompt"], tr["n_layers"], tr["max_generate_tokens"], tr["eos_token_id"])
  assert got["engine_log"] == tr["engine_log"] and got["on_token"] == tr["on_token"] and got["end"] == tr["end"]


def test_device_tensor_ref_behaves_like_the_array_it_stands_for():
  from xotorch_support_jetson_b200.networking import DeviceTensorRef
  t = torch.arange(24, dtype=torch.float32).reshape(6, 4).to(torch.bfloat16)
  ref = DeviceTensorRef((2, 3, 4), tensor=t)
  assert ref.ndim == 3 and ref.shape == (2, 3, 4) and ref.size == 24 and ref.dtype == np.float32
  a = np.asarray(ref)
  assert a.dtype == np.float32 and a.shape == (2, 3, 4) and np.array_equal(a.reshape(6, 4), t.float().numpy())
  with pytest.raises(ValueError, match="landing buffer"):
    np.asarray(DeviceTensorRef((1, 1, 4), landing=(0, 1)))


# ----------------------------------------------------------------------------- process-per-GPU ring over gloo
def _ring_worker(rank, world, port, ret):
  sys.path.insert(0, str(ROOT))
  os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
  import torch.distributed as dist
  dist.init_process_group("gloo", rank=rank, world_size=world)
  ctrl = dist.new_group(ranks=list(range(world)), backend="gloo")
  data = dist.new_group(ranks=list(range(world)), backend="gloo")
  from oracle.scripted_engine import ScriptedEngine
  from xotorch_support_jetson_b200.networking import DeviceTensorRef
  from xotorch_support_jetson_b200.orchestration import build_rank_node
  from xotorch_support_jetson_b200.shard import Shard
  tr = TRACES["three_nodes_eos" if world == 3 else "two_nodes_cap"]
  log, tokens = [], []

  class RingScripted(ScriptedEngine):
    """The scripted engine with the engine-side half of the data plane: hidden rows leave as a device reference
    (sent by the engine on the data channel), arrive as a reference that the engine receives."""
    _ring = None

    def bind_ring(self, ch, sends=True):
      self._ring = ch

    async def infer_tensor(self, request_id, shard, input_data, inference_state=None):
      x = input_data
      if isinstance(x, DeviceTensorRef):
        buf = torch.zeros(x.shape, dtype=torch.float32)
        self._ring.recv_rows(buf)
        x = buf.numpy()
      out, state = await super().infer_tensor(request_id, shard, x, inference_state)
      if not shard.is_last_layer() and self._ring is not None:
        self._ring.send_rows(torch.from_numpy(np.ascontiguousarray(out)))
        return DeviceTensorRef(out.shape, tensor=torch.from_numpy(out), recv=("sent",)), state
      return out, state

  async def run():
    node = build_rank_node(RingScripted(tr["node_ids"][rank], log, tr["eos_token_id"]), dist, rank, world, ctrl_group=ctrl,
                           data_group=data, max_generate_tokens=tr["max_generate_tokens"],
                           landing=(4, 8) if world == 3 else None)   # an engine without landing buffers: the ring stays on the data plane
    await node.start()
    await node.collect_topology(set())
    done = asyncio.Event()

    def on_token(rid, toks, fin):
      tokens.append({"request_id": rid, "tokens": [int(t) for t in toks], "is_finished": bool(fin)})
      if fin:
        done.set()

    node.on_token.register("trace").on_next(on_token)
    base = Shard("scripted", 0, 0, tr["n_layers"])
    if rank == 0:
      await node.process_prompt(base, tr["prompt"], "req-1")
    if rank == 0 or rank == world - 1:       # the entry node hears from the last node; the last node publishes itself
      await asyncio.wait_for(done.wait(), 60)
    await asyncio.sleep(0.2)
    dist.barrier()                            # every rank has seen what it will see: now release the listeners
    await node.stop()
    return list(node.get_current_shard(base))[1:]

  shard = asyncio.run(run())
  ret[rank] = {"log": log, "tokens": tokens, "shard": shard}
  dist.barrier()
  dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_process_ring_over_gloo_reproduces_the_reference_trace(world):
  """One process per ring node (what torchrun starts on a GPU box), headers over the control channel, hidden rows
  over the data channel posted by the engines: together the ranks make exactly the engine calls of the reference's
  in-process ring, and the entry rank receives the same tokens."""
  import torch.multiprocessing as mp
  port = 29700 + os.getpid() % 1500 + world
  mgr = mp.Manager()
  ret = mgr.dict()
  mp.spawn(_ring_worker, args=(world, port, ret), nprocs=world, join=True)
  tr = TRACES["three_nodes_eos" if world == 3 else "two_nodes_cap"]
  for r in range(world):
    want = [e for e in tr["engine_log"] if e["node"] == tr["node_ids"][r]]
    assert ret[r]["log"] == want
    assert ret[r]["shard"] == tr["end"]["shards"][r]
  assert ret[0]["tokens"] == tr["on_token"]
