"""``BoxNode`` -- the inference path of the reference's ``Node`` for the nodes of one box.

The reference's ``Node`` (``xotorch/orchestration/node.py``) is the thing that walks a request around
the ring: ``process_prompt`` -> engine -> ``process_inference_result`` -> (last shard: sample, publish
the token, stop or go on) -> ``forward_tensor`` -> the next partition's peer -> its
``process_tensor`` ... (node.py:109-147, 193-208, 361-380, 424-443).  The engine and peer handles of
this package plug into that class unchanged (tests/test_orchestration.py drives the reference's own
``Node`` with them).  ``BoxNode`` is the same walk in this package's own terms, for programs that do
not have the reference installed (the benchmark, the GPU box): same entry points and bookkeeping
(``buffered_token_output``, ``outstanding_requests``, ``on_token``, partition index / current shard
from the partitioning strategy over the collected topology), plus what an intra-box ring can use:

* whatever the engine returns for a non-last shard is passed on untouched, so a ``DeviceTensorRef``
  (hidden rows still in HBM, or already in the next GPU's landing buffer) reaches the next node's
  engine without a host copy;
* a shard that is both first and last may keep the token loop on the device (``engine.generate``).

Left out on purpose (out of scope): training entry points, stable-diffusion branch,
status broadcasts for the TUI, download progress, periodic re-discovery.
"""
import asyncio
import traceback
import uuid
from typing import Callable, Dict, List, Optional, Tuple

import numpy as np

from .networking import trace
from .shard import Shard
from .topology import DeviceCapabilities, Topology, b200_device_capabilities, map_partitions_to_shards


class _Callback:
  """One named subscriber list (``xotorch/helpers.py`` AsyncCallback: ``on_next`` + ``set``)."""

  def __init__(self):
    self.observers: List[Callable] = []
    self.result = None

  def on_next(self, cb: Callable):
    self.observers.append(cb)

  def set(self, *args):
    self.result = args
    for cb in self.observers:
      cb(*args)


class CallbackSystem:
  """``AsyncCallbackSystem`` of the reference (``xotorch/helpers.py:130-149``): register / deregister /
  trigger / trigger_all."""

  def __init__(self):
    self.callbacks: Dict[str, _Callback] = {}

  def register(self, name: str) -> _Callback:
    return self.callbacks.setdefault(name, _Callback())

  def deregister(self, name: str):
    self.callbacks.pop(name, None)

  def trigger(self, name: str, *args):
    if name in self.callbacks:
      self.callbacks[name].set(*args)

  def trigger_all(self, *args):
    for cb in list(self.callbacks.values()):
      cb.set(*args)


class BoxNode:
  def __init__(self, _id: str, server, inference_engine, discovery, shard_downloader=None, partitioning_strategy=None,
               max_generate_tokens: int = 1024, default_sample_temperature: float = 0.0,
               device_capabilities: Optional[DeviceCapabilities] = None, device_token_loop: bool = False,
               shard_map: Optional[List[Shard]] = None):
    self.id = _id
    self.server, self.inference_engine, self.discovery = server, inference_engine, discovery
    self.shard_downloader = shard_downloader
    self.partitioning_strategy = partitioning_strategy
    self.peers: List = []
    self.topology = Topology()
    self.device_capabilities = device_capabilities or b200_device_capabilities()
    self.buffered_token_output: Dict[str, Tuple[List[int], bool]] = {}
    self.outstanding_requests: Dict[str, str] = {}
    self.max_generate_tokens = max_generate_tokens
    self.default_sample_temperature = default_sample_temperature
    self.device_token_loop = device_token_loop
    self._on_token = CallbackSystem()
    self._on_opaque_status = CallbackSystem()
    self._partitions = None
    self._rows_done: Dict[str, List[bool]] = {}
    # explicit layer range per ring position instead of map_partitions_to_shards (stage boundaries that the
    # reference's whole-layer mapping cannot express: half-layer cuts balanced by bytes, topology.half_layer_stages)
    self.shard_map = shard_map
    if hasattr(server, "bind"):
      server.bind(self)

  # ------------------------------------------------------------------ lifecycle (node.py:59-84)
  async def start(self, wait_for_peers: int = 0) -> None:
    await self.server.start()
    await self.discovery.start()
    await self.update_peers(wait_for_peers)
    await self.collect_topology(set())

  async def stop(self) -> None:
    await self.discovery.stop()
    await self.server.stop()

  async def update_peers(self, wait_for_peers: int = 0) -> bool:
    nxt = await self.discovery.discover_peers(wait_for_peers)
    before = {p.id() for p in self.peers}
    for p in nxt:
      if not await p.is_connected():
        await p.connect()
    self.peers = nxt
    self._partitions = None
    return before != {p.id() for p in nxt}

  async def collect_topology(self, visited: set, max_depth: int = 4) -> Topology:
    """node.py:537-572."""
    nxt = Topology()
    nxt.update_node(self.id, self.device_capabilities)
    prev_visited = set(visited)
    visited.add(self.id)
    visited.update(p.id() for p in self.peers)
    for peer in self.peers:
      nxt.update_node(peer.id(), peer.device_capabilities())
      nxt.add_edge(self.id, peer.id(), peer.description())
      if peer.id() in prev_visited or max_depth <= 0:
        continue
      try:
        other = await asyncio.wait_for(peer.collect_topology(visited, max_depth=max_depth - 1), timeout=5.0)
        nxt.merge(peer.id(), other)
      except Exception as e:   # a peer that does not answer stays in the topology with what discovery said
        print(f"Error collecting topology from {peer.id()}: {e}")
    nxt.active_node_id = self.topology.active_node_id
    self.topology = nxt
    self._partitions = None
    return self.topology

  # ------------------------------------------------------------------ ring geometry (node.py:445-461)
  def _parts(self):
    if self._partitions is None:   # the strategy is a pure function of the topology: computed once per topology
      self._partitions = self.partitioning_strategy.partition(self.topology)
    return self._partitions

  def get_partition_index(self, offset: int = 0) -> int:
    parts = self._parts()
    cur = next((i for i, p in enumerate(parts) if p.node_id == self.id), None)
    if cur is None:
      raise ValueError(f"No current partition found for node: {self.id}")
    return (cur + offset) % len(parts)

  def get_current_shard(self, base_shard: Shard, index: Optional[int] = None) -> Shard:
    if index is None:
      index = self.get_partition_index()
    if self.shard_map is not None:
      sh = self.shard_map[index]
      if sh.model_id != base_shard.model_id or sh.n_layers != base_shard.n_layers:
        raise ValueError(f"this ring was built for {sh.model_id}/{sh.n_layers} layers, not {base_shard.model_id}/{base_shard.n_layers}")
      return sh
    return map_partitions_to_shards(self._parts(), base_shard.n_layers, base_shard.model_id)[index]

  # ------------------------------------------------------------------ the walk
  async def process_prompt(self, base_shard: Shard, prompt: str, request_id: Optional[str] = None,
                           inference_state: Optional[dict] = None):
    """node.py:149-208: a node that does not hold layer 0 passes the prompt on to partition 0."""
    if request_id is None:
      request_id = str(uuid.uuid4())
    shard = self.get_current_shard(base_shard)
    if not shard.is_first_layer():
      self.outstanding_requests[request_id] = "waiting"
      await self.forward_prompt(shard, prompt, request_id, 0, inference_state)
      return None
    self.outstanding_requests[request_id] = "processing"
    eng = self.inference_engine
    if self.device_token_loop and shard.is_last_layer() and hasattr(eng, "generate"):
      return await self._generate_on_device(shard, prompt, request_id, inference_state)
    result, inference_state = await eng.infer_prompt(request_id, shard, prompt, inference_state)
    await self.process_inference_result(shard, result, request_id, inference_state)
    return result

  async def _generate_on_device(self, shard: Shard, prompt: str, request_id: str, inference_state: Optional[dict]):
    """A one-node ring: sample -> forward -> sample stays on the GPU (engine.generate); tokens reach the
    subscribers through the same callbacks, in order, with the reference's stop rule."""
    eng = self.inference_engine
    ids = await eng.encode(shard, prompt)
    self.buffered_token_output[request_id] = ([], False)

    def on_token(rid, toks, fin):   # runs on the event-loop thread
      buf = self.buffered_token_output[rid][0]
      buf.extend(toks)
      self.buffered_token_output[rid] = (buf, fin)
      self.trigger_on_token_callbacks(rid, toks, fin)
      asyncio.create_task(self.broadcast_result(rid, toks, fin))

    out, _ = await eng.generate(request_id, shard, np.asarray(ids).reshape(1, -1), inference_state,
                                max_generate_tokens=self.max_generate_tokens, temp=self.default_sample_temperature,
                                eos_token_id=eng.tokenizer.eos_token_id, on_token=on_token)
    await asyncio.sleep(0)
    self.outstanding_requests.pop(request_id, None)
    return out

  async def process_tensor(self, base_shard: Shard, tensor, request_id: Optional[str] = None,
                           inference_state: Optional[dict] = None):
    """node.py:339-380: errors are reported and the request dropped, like the reference."""
    if request_id is None:
      request_id = str(uuid.uuid4())
    shard = self.get_current_shard(base_shard)
    try:
      self.outstanding_requests[request_id] = "processing"
      trace("infer.begin", request_id)
      result, inference_state = await self.inference_engine.infer_tensor(request_id, shard, tensor, inference_state)
      trace("infer.end", request_id)
      return await self.process_inference_result(shard, result, request_id, inference_state)
    except Exception as e:
      self.outstanding_requests.pop(request_id, None)
      print(f"Error processing tensor for shard {shard}: {e}")
      traceback.print_exc()

  async def process_inference_result(self, shard: Shard, result, request_id: Optional[str] = None,
                                     inference_state: Optional[dict] = None):
    """node.py:109-147."""
    if request_id not in self.buffered_token_output:
      self.buffered_token_output[request_id] = ([], False)
    buf = self.buffered_token_output[request_id][0]
    is_finished = len(buf) >= self.max_generate_tokens
    if shard.is_last_layer() and not is_finished:
      token = await self.inference_engine.sample(result, temp=self.default_sample_temperature)
      trace("sample.end", request_id)
      await self.inference_engine.ensure_shard(shard)
      eos = self.inference_engine.tokenizer.eos_token_id
      if token.size == 1:                      # the reference's case: one sequence per request
        buf.append(token.item())
        is_finished = token.item() == eos or is_finished or len(buf) >= self.max_generate_tokens
        forward = token.reshape(1, -1)
        intermediate_result = [buf[-1]]
      else:
        # a request that carries B sequences (no counterpart in the reference, whose engine state is one sequence):
        # one list of B ids per step; finished when every row has produced eos at some step, or at the cap
        ids = [int(t) for t in np.asarray(token).reshape(-1)]
        buf.append(ids)
        done_rows = self._rows_done.setdefault(request_id, [False] * len(ids))
        for i, t in enumerate(ids):
          done_rows[i] = done_rows[i] or t == eos
        is_finished = all(done_rows) or len(buf) >= self.max_generate_tokens
        forward = np.asarray(token).reshape(-1, 1)
        intermediate_result = ids
    else:
      forward = result   # hidden rows: numpy, or a device reference the next node's engine understands
      intermediate_result = None
    if shard.is_last_layer():
      self.trigger_on_token_callbacks(request_id, intermediate_result, is_finished)
      asyncio.create_task(self.broadcast_result(request_id, intermediate_result, is_finished))
    if is_finished:
      self.buffered_token_output[request_id] = (buf, True)
      self.outstanding_requests.pop(request_id, None)
    else:
      self.outstanding_requests[request_id] = "waiting"
      asyncio.create_task(self.forward_tensor(shard, forward, request_id, self.get_partition_index(offset=1), inference_state))
    return np.array(buf)

  async def process_tokens(self, base_shard: Shard, ids: np.ndarray, request_id: Optional[str] = None,
                           inference_state: Optional[dict] = None):
    """``process_prompt`` for a prompt that is already token ids ``[B, L]`` (benchmarks, callers with their own
    tokenizer): the ids enter the ring at the node that holds layer 0."""
    if request_id is None:
      request_id = str(uuid.uuid4())
    shard = self.get_current_shard(base_shard)
    if not shard.is_first_layer():
      raise ValueError("process_tokens must be called on the node that holds layer 0")
    self.outstanding_requests[request_id] = "processing"
    result, inference_state = await self.inference_engine.infer_tensor(request_id, shard, np.asarray(ids), inference_state)
    await self.process_inference_result(shard, result, request_id, inference_state)
    return result

  def _target(self, target_index: int):
    target_id = self._parts()[target_index].node_id
    if target_id == self.id:
      return None
    peer = next((p for p in self.peers if p.id() == target_id), None)
    if peer is None:
      raise ValueError(f"Peer for {target_index} not found")
    return peer

  async def forward_prompt(self, base_shard: Shard, prompt: str, request_id: str, target_index: int,
                           inference_state: Optional[dict] = None) -> None:
    """node.py:404-422."""
    next_shard = self.get_current_shard(base_shard, target_index)
    peer = self._target(target_index)
    if peer is None:
      await self.process_prompt(next_shard, prompt, request_id, inference_state)
    else:
      await peer.send_prompt(next_shard, prompt, request_id=request_id, inference_state=inference_state)

  async def forward_tensor(self, base_shard: Shard, tensor, request_id: str, target_index: int,
                           inference_state: Optional[dict] = None) -> None:
    """node.py:424-443 -- the dispatch point north_star names: same-box peers take the tensor as it is."""
    next_shard = self.get_current_shard(base_shard, target_index)
    peer = self._target(target_index)
    if peer is None:
      await self.process_tensor(next_shard, tensor, request_id, inference_state)
    else:
      await peer.send_tensor(next_shard, tensor, request_id=request_id, inference_state=inference_state)

  # ------------------------------------------------------------------ results (node.py:574-600)
  @property
  def on_token(self) -> CallbackSystem:
    return self._on_token

  @property
  def on_opaque_status(self) -> CallbackSystem:
    return self._on_opaque_status

  def trigger_on_token_callbacks(self, request_id: str, tokens, is_finished: bool) -> None:
    self.on_token.trigger_all(request_id, tokens, is_finished)

  async def broadcast_result(self, request_id: str, result, is_finished: bool) -> None:
    async def one(peer):
      try:
        await asyncio.wait_for(peer.send_result(request_id, result, is_finished), timeout=15.0)
      except Exception as e:
        print(f"Error broadcasting result to {peer.id()}: {e}")
    # (a ring handle that cannot reach its peer says so: no coroutine, no timer for it -- 7 of 8 on an 8-GPU ring)
    targets = [p for p in self.peers if getattr(p, "delivers_results", True)]
    if len(targets) == 1:
      await one(targets[0])
    elif targets:
      await asyncio.gather(*[one(p) for p in targets], return_exceptions=True)

  @property
  def current_topology(self) -> Topology:
    return self.topology


def build_box_ring(engines: List, max_generate_tokens: int = 1024, default_sample_temperature: float = 0.0,
                   memory_mb: int = 183359, landing: bool = True) -> List[BoxNode]:
  """One ``BoxNode`` per engine (= per GPU) of ONE process, wired into a ring: every node discovers the others
  through ``NvlinkPeerHandle``s, the layer ranges come from the reference's memory-weighted ring strategy over
  the per-GPU topology (equal GPUs: equal slices, engine i = partition i), and engine i hands its hidden rows
  to engine i+1 as device references -- decode steps through i+1's landing buffers when ``landing``.
  ``await node.start()`` on each returned node, then ``nodes[0].process_prompt(...)``."""
  from .networking import BoxDiscovery, NullServer, NvlinkPeerHandle
  from .topology import RingMemoryWeightedPartitioningStrategy, b200_device_capabilities
  n = len(engines)
  ids = [f"gpu{n - 1 - i:02d}-rank{i}" for i in range(n)]      # ring order (memory desc, id desc) = engine order
  caps = b200_device_capabilities(memory_mb)
  nodes = [BoxNode(ids[i], NullServer(), engines[i], BoxDiscovery(), partitioning_strategy=RingMemoryWeightedPartitioningStrategy(),
                   max_generate_tokens=max_generate_tokens, default_sample_temperature=default_sample_temperature,
                   device_capabilities=caps, device_token_loop=(n == 1)) for i in range(n)]
  for i, node in enumerate(nodes):
    node.discovery.handles = [NvlinkPeerHandle(nodes[j], getattr(engines[j], "device_index", j), caps) for j in range(n) if j != i]
  for i in range(n - 1):
    if hasattr(engines[i], "connect_next"):
      engines[i].connect_next(engines[i + 1])
      if landing:
        engines[i + 1].enable_landing()
  return nodes


def build_rank_node(engine, dist, rank: int, world: int, ctrl_group=None, data_group=None, max_generate_tokens: int = 1024,
                    default_sample_temperature: float = 0.0, memory_mb: int = 183359, shard_map: Optional[List[Shard]] = None,
                    landing: Optional[Tuple[int, int]] = None) -> BoxNode:
  """This RANK's node of a process-per-GPU ring (torchrun): the other ranks appear as ``RingPeerHandle``s, the
  ``RingServer`` listens to the previous rank, layer ranges as in ``build_box_ring`` (rank i = partition i).
  ``ctrl_group``: a CPU (gloo) group for headers; ``data_group``: where hidden rows travel (NCCL on a GPU box).
  The engine is bound to the ring so that it sends / receives the rows itself, on its own stream.
  ``landing = (slots, rows)``: decode steps of up to ``rows`` sequences, for up to ``slots`` requests in flight, leave
  through the next rank's landing buffers instead (a collective call: every rank passes the same value)."""
  from .networking import BoxDiscovery, RingChannel, RingPeerHandle, RingServer
  from .topology import RingMemoryWeightedPartitioningStrategy, b200_device_capabilities
  ids = [f"gpu{world - 1 - i:02d}-rank{i}" for i in range(world)]
  caps = b200_device_capabilities(memory_mb)
  ch = RingChannel(dist, rank, world, ctrl_group=ctrl_group, data_group=data_group)
  node = BoxNode(ids[rank], RingServer(ch), engine, BoxDiscovery([RingPeerHandle(ids[j], j, ch, caps) for j in range(world) if j != rank]),
                 partitioning_strategy=RingMemoryWeightedPartitioningStrategy(), max_generate_tokens=max_generate_tokens,
                 default_sample_temperature=default_sample_temperature, device_capabilities=caps, device_token_loop=(world == 1),
                 shard_map=shard_map)
  if world > 1 and hasattr(engine, "bind_ring"):
    engine.bind_ring(ch, sends=rank + 1 < world)   # the last stage only RECEIVES rows (its output is a token id)
    dev = getattr(engine, "device_index", None)
    if dev is not None:
      import torch
      ch.warm_up(torch.device("cuda", int(dev)))
    if landing is not None:
      connect_rank_landing(engine, dist, rank, world, ctrl_group, landing)
  return node


def connect_rank_landing(engine, dist, rank: int, world: int, ctrl_group, landing: Tuple[int, int]) -> None:
  """Decode steps of a process-per-GPU ring through the next rank's landing buffers (CUDA IPC) instead of NCCL
  send / recv: every rank exports the buffers of its stage and maps its successor's.  A COLLECTIVE call (one
  all_gather over ``ctrl_group`` -- a group no running RingServer is listening on) to be made once every rank's engine holds its stage -- ``build_rank_node`` makes
  it when given ``landing``, which suits engines that adopt a resident stage; engines that load their shard on first
  use call it after ``ensure_shard``.  A rank whose engine cannot take part (no stage yet, other buffer geometry)
  simply stays on the NCCL path, and so does its predecessor."""
  if not hasattr(engine, "ring_landing_handles"):
    mine = None
  else:
    mine = engine.ring_landing_handles(*landing) if rank > 0 else None
  everyone = [None] * world
  dist.all_gather_object(everyone, mine, group=ctrl_group)
  if rank + 1 < world and hasattr(engine, "connect_ring_landing"):
    engine.connect_ring_landing(everyone[rank + 1], *landing)
