"""Intra-box peers: the reference's networking plug-in points bound to the GPUs of one box.

The reference reaches the next shard of a ring through a ``PeerHandle``
(``xotorch/networking/peer_handle.py:9-56``) that a ``Discovery`` (``networking/discovery.py:6-17``)
hands to the ``Node``; its only concrete handle, ``GRPCPeerHandle``
(``networking/grpc/grpc_peer_handle.py:117-136``), ships every hidden state as fp32 numpy + a JSON
state blob.  Inside one box that hop does not have to leave the GPUs:

* ``NvlinkPeerHandle``  -- the peer is another ``Node`` of THIS process (one node per GPU, the
  per-GPU topology north_star asks for).  ``send_tensor`` passes what the engine produced straight to
  the target node: a ``DeviceTensorRef`` that wraps the bf16 hidden rows still in HBM (the consumer
  copies them GPU->GPU over NVLink), or -- for decode steps between engines whose landing buffers are
  connected -- only the landing-slot id and flag value of rows that the producer's last kernel has
  already stored into the consumer GPU's memory (``xb_forward_handoff``).
* ``BoxDiscovery`` / ``NullServer`` -- the matching ``Discovery`` and ``Server``: the peers of a box are
  known up front, nothing listens on a socket.
* ``RingPeerHandle`` / ``RingServer`` (further down) -- the same idea for ONE PROCESS PER GPU
  (torchrun): a small header travels over a CPU (gloo) channel, the hidden rows over NCCL.

A ``DeviceTensorRef`` still behaves like the array the reference would have sent (``shape``, ``size``,
``ndim``, ``np.asarray(ref)`` -> fp32 numpy), so a gRPC handle for a peer on ANOTHER host keeps working
on whatever the engine returns.
"""
import asyncio
import json
import os
import threading
import time
from abc import ABC, abstractmethod
from typing import List, Optional

import numpy as np

from .shard import Shard
from .topology import DeviceCapabilities, Topology, b200_device_capabilities


# ----------------------------------------------------------------------------- what travels inside a box
class DeviceTensorRef:
  """Hidden rows ``[B, s, D]`` of a non-last shard that stay on the GPU.

  ``tensor``: bf16 CUDA tensor ``[B*s, D]`` + ``event`` (recorded on the producing stream), or None when
  the rows went through a landing buffer: then ``landing = (slot, flag_value)`` names the consumer-side
  slot the producer's last kernel stored them into, and -- across processes -- ``recv = (src_rank,)``
  says that the rows still have to be received from that rank (posted by the consuming engine on its
  own stream).  ``np.asarray(ref)`` materialises what the reference would have shipped (fp32 numpy)."""
  ndim = 3

  def __init__(self, shape, tensor=None, event=None, landing=None, recv=None):
    self.shape = tuple(int(d) for d in shape)
    self.tensor, self.event, self.landing, self.recv = tensor, event, landing, recv
    self.dtype = np.dtype(np.float32)

  @property
  def size(self) -> int:
    return int(np.prod(self.shape))

  def reshape(self, *shape):
    return self

  def __array__(self, dtype=None, copy=None):
    if self.tensor is None:
      raise ValueError("these hidden rows sit in the consumer GPU's landing buffer; only that engine can read them")
    if self.event is not None:
      self.event.synchronize()
    arr = self.tensor.float().reshape(self.shape).cpu().numpy()
    return arr if dtype is None else arr.astype(dtype)


# ----------------------------------------------------------------------------- the reference's three interfaces
class PeerHandle(ABC):
  """``xotorch/networking/peer_handle.py:9-56`` (the training-only members are left out)."""

  @abstractmethod
  def id(self) -> str: ...

  @abstractmethod
  def addr(self) -> str: ...

  @abstractmethod
  def description(self) -> str: ...

  @abstractmethod
  def device_capabilities(self) -> DeviceCapabilities: ...

  @abstractmethod
  async def connect(self) -> None: ...

  @abstractmethod
  async def is_connected(self) -> bool: ...

  @abstractmethod
  async def disconnect(self) -> None: ...

  @abstractmethod
  async def health_check(self) -> bool: ...

  @abstractmethod
  async def send_prompt(self, shard: Shard, prompt: str, request_id: Optional[str] = None, inference_state: Optional[dict] = None): ...

  @abstractmethod
  async def send_tensor(self, shard: Shard, tensor, request_id: Optional[str] = None, inference_state: Optional[dict] = None): ...

  @abstractmethod
  async def send_result(self, request_id: str, result: List[int], is_finished: bool) -> None: ...

  @abstractmethod
  async def collect_topology(self, visited: set, max_depth: int) -> Topology: ...


class Discovery(ABC):
  """``xotorch/networking/discovery.py:6-17``."""

  @abstractmethod
  async def start(self) -> None: ...

  @abstractmethod
  async def stop(self) -> None: ...

  @abstractmethod
  async def discover_peers(self, wait_for_peers: int = 0) -> List[PeerHandle]: ...


class Server(ABC):
  """``xotorch/networking/server.py:4-11``."""

  @abstractmethod
  async def start(self) -> None: ...

  @abstractmethod
  async def stop(self) -> None: ...


class NullServer(Server):
  """Peers of one process call each other's node directly: nothing to listen on."""

  async def start(self) -> None:
    pass

  async def stop(self) -> None:
    pass


# ----------------------------------------------------------------------------- one process, one node per GPU
class NvlinkPeerHandle(PeerHandle):
  """Peer = another node of this process, on another GPU of the box (or the same one, in tests).

  Every call lands on the target node the way the reference's gRPC server hands it over
  (``networking/grpc/grpc_server.py``: SendPrompt -> ``node.process_prompt``, SendTensor ->
  ``node.process_tensor``, SendResult -> ``node.on_token.trigger_all``, SendOpaqueStatus ->
  ``node.on_opaque_status.trigger_all``, CollectTopology -> ``node.collect_topology``) -- minus the
  serialisation: the tensor argument is passed on as it is, device reference included."""

  def __init__(self, node, gpu: int, capabilities: Optional[DeviceCapabilities] = None):
    self._node, self._gpu = node, gpu
    self._caps = capabilities or b200_device_capabilities()
    self._connected = False

  def id(self) -> str:
    return self._node.id

  def addr(self) -> str:
    return f"nvlink://gpu{self._gpu}"

  def description(self) -> str:
    return "NVL"

  def device_capabilities(self) -> DeviceCapabilities:
    return self._caps

  async def connect(self) -> None:
    self._connected = True

  async def is_connected(self) -> bool:
    return self._connected

  async def disconnect(self) -> None:
    self._connected = False

  async def health_check(self) -> bool:
    return True

  async def send_prompt(self, shard: Shard, prompt: str, request_id: Optional[str] = None, inference_state: Optional[dict] = None):
    # like the gRPC server: the sender does not wait for the ring walk that follows
    asyncio.create_task(self._node.process_prompt(shard, prompt, request_id, inference_state))
    return None

  async def send_tensor(self, shard: Shard, tensor, request_id: Optional[str] = None, inference_state: Optional[dict] = None):
    asyncio.create_task(self._node.process_tensor(shard, tensor, request_id, inference_state))
    return None

  async def send_result(self, request_id: str, result: List[int], is_finished: bool) -> None:
    self._node.on_token.trigger_all(request_id, list(result), is_finished)

  async def send_opaque_status(self, request_id: str, status: str) -> None:
    self._node.on_opaque_status.trigger_all(request_id, status)

  async def collect_topology(self, visited: set, max_depth: int) -> Topology:
    return await self._node.collect_topology(set(visited), max_depth)


class BoxDiscovery(Discovery):
  """The other GPUs of the box, known up front: ``handles`` = one ``NvlinkPeerHandle`` per other node."""

  def __init__(self, handles: Optional[List[PeerHandle]] = None):
    self.handles: List[PeerHandle] = list(handles or [])

  async def start(self) -> None:
    pass

  async def stop(self) -> None:
    pass

  async def discover_peers(self, wait_for_peers: int = 0) -> List[PeerHandle]:
    return list(self.handles)


# ----------------------------------------------------------------------------- one process per GPU (torchrun)
KIND_PROMPT, KIND_TENSOR, KIND_RESULT, KIND_STATUS, KIND_STOP = 1, 2, 3, 4, 5

# developer aid: XB_RING_TRACE=1 keeps (monotonic ns, tag, request id) per hop event; bench.py writes them out per rank
TRACE = [] if os.environ.get("XB_RING_TRACE") == "1" else None


def trace(tag: str, rid=None):
  if TRACE is not None:
    TRACE.append((time.monotonic_ns(), tag, rid))

_HDR = 8   # int64 words: kind, meta bytes, payload kind (0 none, 1 int ids, 2 fp32 numpy, 3 device rows over the data channel), ndim, d0, d1, d2, size of a second message (0: all inline)


class RingChannel:
  """The directed edge rank -> rank+1 (mod world) of a process-per-GPU ring, on ``torch.distributed``:

  * a CONTROL channel for a fixed-size header + a JSON blob (request id, shard, state) and small
    payloads (token ids, results) -- a CPU (gloo) group, so that a message is one blocking call pair
    on host memory and never touches a CUDA stream;
  * a DATA channel for hidden rows -- device bf16 over NCCL (or the same gloo group when the ring has
    no GPU, which is how the world_size-2 CPU tests run it), posted by the ENGINE threads on their own
    streams: the sender enqueues ``send`` right behind its forward, the receiver ``recv`` right before
    its own, both in header order, so every edge is FIFO and no host synchronisation is needed."""

  def __init__(self, dist, rank: int, world: int, ctrl_group=None, data_group=None):
    self.dist, self.rank, self.world = dist, rank, world
    self.ctrl_group, self.data_group = ctrl_group, data_group
    self.next_rank, self.prev_rank = (rank + 1) % world, (rank - 1) % world
    self._send_lock = threading.Lock()

  # -- control plane (any thread; serialised per process)
  FRAME = 8192   # bytes of the one fixed-size message a decode hop costs: header + JSON + small payload

  def send_message(self, kind: int, meta: dict, payload: Optional[np.ndarray] = None, device_rows=None):
    import torch
    blob = json.dumps(meta, separators=(",", ":")).encode()
    hdr = [kind, len(blob), 0, 0, 0, 0, 0, 0]
    body = b""
    if device_rows is not None:
      hdr[2], hdr[3] = 3, len(device_rows)
      hdr[4:4 + len(device_rows)] = list(device_rows)
    elif payload is not None:
      arr = np.ascontiguousarray(payload)
      is_int = np.issubdtype(arr.dtype, np.integer)
      arr = arr.astype(np.int64 if is_int else np.float32)
      hdr[2], hdr[3] = (1 if is_int else 2), arr.ndim
      hdr[4:4 + arr.ndim] = list(arr.shape)
      body = arr.tobytes()
    rest = blob + body
    inline = len(rest) <= self.FRAME - 8 * _HDR
    hdr[7] = 0 if inline else len(rest)        # > 0: the JSON + payload follow as a second message of exactly this size
    frame = bytearray(self.FRAME)
    frame[:8 * _HDR] = np.asarray(hdr, dtype=np.int64).tobytes()
    if inline:
      frame[8 * _HDR:8 * _HDR + len(rest)] = rest
    trace("send.begin", meta.get("request_id"))
    with self._send_lock:
      self.dist.send(torch.frombuffer(frame, dtype=torch.uint8), dst=self.next_rank, group=self.ctrl_group)
      trace("send.end", meta.get("request_id"))
      if not inline:
        self.dist.send(torch.frombuffer(bytearray(rest), dtype=torch.uint8), dst=self.next_rank, group=self.ctrl_group)

  def recv_message(self):
    """Blocks until the previous rank sends: (kind, meta, payload ndarray | shape tuple of device rows | None)."""
    import torch
    frame = torch.zeros(self.FRAME, dtype=torch.uint8)
    self.dist.recv(frame, src=self.prev_rank, group=self.ctrl_group)
    raw = frame.numpy().tobytes()
    hdr = np.frombuffer(raw[:8 * _HDR], dtype=np.int64)
    kind, nblob, pk, ndim, ext = int(hdr[0]), int(hdr[1]), int(hdr[2]), int(hdr[3]), int(hdr[7])
    if ext > 0:
      more = torch.zeros(ext, dtype=torch.uint8)
      self.dist.recv(more, src=self.prev_rank, group=self.ctrl_group)
      rest = more.numpy().tobytes()
    else:
      rest = raw[8 * _HDR:]
    meta = json.loads(rest[:nblob].decode()) if nblob else {}
    shape = tuple(int(x) for x in hdr[4:4 + ndim])
    if pk in (1, 2):
      n = int(np.prod(shape)) if shape else 1
      dt = np.int64 if pk == 1 else np.float32
      return kind, meta, np.frombuffer(rest[nblob:nblob + n * np.dtype(dt).itemsize], dtype=dt).reshape(shape).copy()
    if pk == 3:
      return kind, meta, shape
    return kind, meta, None

  def warm_up(self, device=None):
    """First contact on the data plane, made by every rank at the same time (a collective step of ring set-up).
    NCCL connects a pair of ranks when they first exchange data and BOTH have to be in the call for that; in
    steady state the receiver posts its side only after the header has reached it, which the sender sends after
    its own -- so without this the first hidden rows would wait for a receive that waits for them."""
    if self.world < 2 or device is None or self.dist.get_backend(self.data_group) != "nccl":
      return
    import torch
    out = torch.zeros(8, dtype=torch.bfloat16, device=device)
    inp = torch.empty_like(out)
    ops = [self.dist.P2POp(self.dist.isend, out, self.next_rank, group=self.data_group),
           self.dist.P2POp(self.dist.irecv, inp, self.prev_rank, group=self.data_group)]
    for w in self.dist.batch_isend_irecv(ops):
      w.wait()
    torch.cuda.synchronize(device)

  # -- data plane (engine threads only)
  def send_rows(self, rows):
    """Non-blocking for the host: on NCCL the send is enqueued on the current stream, on gloo it completes in the
    background -- the receiver posts its side only after the header has reached it, which is sent later."""
    self._pending = [(w, t) for (w, t) in getattr(self, "_pending", []) if not w.is_completed()]
    self._pending.append((self.dist.isend(rows, dst=self.next_rank, group=self.data_group), rows))

  def recv_rows(self, buf):
    self.dist.recv(buf, src=self.prev_rank, group=self.data_group)


class RingPeerHandle(PeerHandle):
  """The NEXT rank of a process-per-GPU ring as a peer (``target_rank``).  Only the ring successor is
  reachable -- which is all ``Node.forward_tensor`` / ``forward_prompt`` ever address
  (``get_partition_index(offset=1)``, node.py:145,525-531) and, for ``send_result``, all the entry node
  needs (the last shard's successor IS the first shard).  Handles for other ranks exist so that the
  topology is complete; their sends are dropped."""

  def __init__(self, node_id: str, target_rank: int, channel: RingChannel, capabilities: Optional[DeviceCapabilities] = None):
    self._id, self._rank, self._ch = node_id, target_rank, channel
    self._caps = capabilities or b200_device_capabilities()
    self._connected = False

  def id(self) -> str:
    return self._id

  def addr(self) -> str:
    return f"nccl://rank{self._rank}"

  def description(self) -> str:
    return "NVL"

  def device_capabilities(self) -> DeviceCapabilities:
    return self._caps

  async def connect(self) -> None:
    self._connected = True

  async def is_connected(self) -> bool:
    return self._connected

  async def disconnect(self) -> None:
    self._connected = False

  async def health_check(self) -> bool:
    return True

  def _reachable(self) -> bool:
    return self._rank == self._ch.next_rank

  @property
  def delivers_results(self) -> bool:
    """Node.broadcast_result may skip this handle: its send_result would drop the message anyway."""
    return self._reachable()

  @staticmethod
  def _state_for_the_wire(tensor, inference_state: Optional[dict]) -> Optional[dict]:
    """The token history rides along only with a prompt (a later shard needs it to find padded slots);
    per decode step it would be ~B x context integers of JSON per hop, and every engine keeps its own."""
    if not inference_state:
      return inference_state
    shape = getattr(tensor, "shape", ())
    if len(shape) >= 2 and shape[1] == 1 and "tokens" in inference_state:
      return {k: v for k, v in inference_state.items() if k != "tokens"}
    return inference_state

  async def send_prompt(self, shard: Shard, prompt: str, request_id: Optional[str] = None, inference_state: Optional[dict] = None):
    if self._reachable():
      meta = {"request_id": request_id, "shard": shard.to_dict(), "prompt": prompt, "state": inference_state}
      await asyncio.get_running_loop().run_in_executor(None, self._ch.send_message, KIND_PROMPT, meta)
    return None

  async def send_tensor(self, shard: Shard, tensor, request_id: Optional[str] = None, inference_state: Optional[dict] = None):
    if not self._reachable():
      return None
    meta = {"request_id": request_id, "shard": shard.to_dict(), "state": self._state_for_the_wire(tensor, inference_state)}
    loop = asyncio.get_running_loop()
    if isinstance(tensor, DeviceTensorRef):
      if tensor.recv is None or tensor.recv[0] not in ("sent", "landed"):
        raise ValueError("a DeviceTensorRef that crosses processes must come from an engine bound to this ring "
                         "(B200ShardInferenceEngine.bind_ring): its rows are sent by the engine, on its stream")
      if tensor.recv[0] == "landed":      # stored into the next rank's landing buffer by the producer's own kernel
        meta["landing"] = [int(tensor.landing[0]), int(tensor.landing[1])]
      # one fixed-size frame into the socket buffer: cheaper inline than a thread hop
      self._ch.send_message(KIND_TENSOR, meta, device_rows=tensor.shape)
    else:
      arr = np.asarray(tensor)
      if arr.size <= 512:     # ids of a decode step: one frame, sent inline like the header of device rows
        self._ch.send_message(KIND_TENSOR, meta, payload=arr)
      else:
        await loop.run_in_executor(None, lambda: self._ch.send_message(KIND_TENSOR, meta, payload=arr))
    return None

  async def send_result(self, request_id: str, result: List[int], is_finished: bool) -> None:
    if self._reachable():
      meta = {"request_id": request_id, "result": [int(t) for t in result], "is_finished": bool(is_finished)}
      if len(meta["result"]) <= 512:
        self._ch.send_message(KIND_RESULT, meta)
      else:
        await asyncio.get_running_loop().run_in_executor(None, self._ch.send_message, KIND_RESULT, meta)

  async def send_opaque_status(self, request_id: str, status: str) -> None:
    pass   # node status broadcasts feed the reference's TUI / download view: nothing consumes them inside a box

  async def collect_topology(self, visited: set, max_depth: int) -> Topology:
    t = Topology()
    t.update_node(self._id, self._caps)
    return t


class RingServer(Server):
  """Receiving end of the edge (rank-1) -> rank: a daemon thread blocks on the control channel and hands
  every message to this rank's node on its event loop -- what ``GRPCServer`` does for SendPrompt /
  SendTensor / SendResult.  Device rows are NOT received here: the node gets a ``DeviceTensorRef`` whose
  ``recv`` field tells the engine to post the receive on its own stream."""

  def __init__(self, channel: RingChannel):
    self.ch = channel
    self.node = None
    self.loop: Optional[asyncio.AbstractEventLoop] = None
    self.thread: Optional[threading.Thread] = None
    self.stopped = threading.Event()

  def bind(self, node):
    self.node = node

  async def start(self) -> None:
    self.loop = asyncio.get_running_loop()
    self.thread = threading.Thread(target=self._serve, name=f"ring-server-{self.ch.rank}", daemon=True)
    self.thread.start()

  async def stop(self) -> None:
    """Every rank calls stop(): the STOP message this rank sends releases its SUCCESSOR's listener."""
    await asyncio.get_running_loop().run_in_executor(None, self.ch.send_message, KIND_STOP, {})
    if self.thread is not None:
      await asyncio.get_running_loop().run_in_executor(None, self.thread.join, 30.0)

  def _serve(self):
    while True:
      kind, meta, payload = self.ch.recv_message()
      trace(f"recv.{kind}", meta.get("request_id"))
      if kind == KIND_STOP:
        self.stopped.set()
        return
      self.loop.call_soon_threadsafe(self._dispatch, kind, meta, payload)

  def _dispatch(self, kind: int, meta: dict, payload):
    node = self.node
    trace(f"dispatch.{kind}", meta.get("request_id"))
    if kind == KIND_RESULT:
      node.on_token.trigger_all(meta["request_id"], meta["result"], meta["is_finished"])
      return
    shard = Shard.from_dict(meta["shard"])
    if kind == KIND_PROMPT:
      asyncio.create_task(node.process_prompt(shard, meta["prompt"], meta["request_id"], meta.get("state")))
    elif kind == KIND_TENSOR:
      if isinstance(payload, np.ndarray):
        tensor = payload
      elif meta.get("landing") is not None:
        tensor = DeviceTensorRef(payload, landing=(int(meta["landing"][0]), int(meta["landing"][1])))
      else:
        tensor = DeviceTensorRef(payload, recv=(self.ch.prev_rank,))
      asyncio.create_task(node.process_tensor(shard, tensor, meta["request_id"], meta.get("state")))
