"""B200ShardInferenceEngine -- drop-in for the reference's TorchDynamicShardInferenceEngine
(``xotorch/inference/torch/sharded_inference_engine.py:37-424``) behind the same
``InferenceEngine`` methods, backed by libxb200 instead of eager torchtune modules.

Differences that are deliberate ("known defects"):
* state is keyed by ``request_id`` (one KV slot per sequence of a request) instead of a
  single mutable ``self.state`` -- concurrent requests do not corrupt each other;
* the ``inference_state`` dict that travels between nodes holds only ``curr_pos`` and
  the token history (ints), not the ``[1, L+1024, C]`` boolean mask: causality follows
  from positions;
* a non-last shard returns the output of its last layer (the reference's off-by-one
  hand-off is not reproduced), so any split reproduces the full-shard result;
* out-of-memory and every other native failure raise (the reference returns ``None``);
* a prompt that holds the pad id (the reference's padded branch) is run compacted: same result for
  every real token, padded slots get no KV entry (``_forward_padded_prompt``).

Host round trip.  The plugin API makes every decode step cross the host twice (logits out,
sampled ids back in).  When this engine holds the whole model (first AND last layer) it
therefore enqueues the next decode forward from ``sample()`` itself, on the ids still in device
memory, before handing those ids to the caller; ``infer_tensor`` called with exactly those ids at
exactly that position returns the result already in flight, anything else discards it (a
forward at the same position overwrites the KV entry, so nothing has to be undone).  The
results are those of the unspeculated path bit for bit; ``speculate_next=False`` /
``XB_API_SPECULATE=0`` turns it off.  With ONE request in flight the sampler itself is run ahead as
well (``_speculate_sample``: the parameters of the request's previous ``sample()`` call), so that in
a steady infer_tensor / sample loop the host only collects results.  The LAST stage of a ring does the
same for every request in flight, in the order of their forwards (``_speculate_ring_sample``).

Inside a box the engine is also the data plane of the ring (``connect_next`` / ``bind_ring`` /
``connect_ring_landing``): hidden rows leave as device references, decode steps through the next
stage's landing buffers (same process: device pointers; other process: CUDA IPC).
"""
import asyncio
import collections
import functools
import threading
import warnings
import os
import uuid
from concurrent.futures import ThreadPoolExecutor
from typing import Dict, List, Optional

import numpy as np
import torch

from .inference_engine import InferenceEngine
from .model_config import load_model_config
from . import networking as _nw
from .networking import DeviceTensorRef, trace
from .shard import Shard
from .stage import StageRuntime
from .synth import SyntheticTokenizer

# results are handed out as read-only numpy views of pinned buffers; feeding one back in is fine
warnings.filterwarnings("ignore", message="The given NumPy array is not writable")

# from the reference (sharded_inference_engine.py:34-35)
TEMP = 0.6
TOP_K = 35


def as_shard(shard) -> Shard:
  """This package's ``Shard`` from anything with the four fields (a reference checkout passes its own
  dataclass, ``xotorch/inference/shard.py``; the two compare unequal although they mean the same)."""
  return shard if type(shard) is Shard else Shard(shard.model_id, int(shard.start_layer), int(shard.end_layer), int(shard.n_layers))


def state_position(inference_state: Optional[dict]) -> Optional[int]:
  """``curr_pos`` of the state dict that travels with a request (the reference's
  ``ShardInferenceState.to_dict``, llm_utils.py:497-503; its ``input_pos`` / ``mask`` entries, when a
  reference node made the dict, are ignored: causality follows from the position)."""
  if not inference_state:
    return None
  p = inference_state.get("curr_pos")
  return None if p is None else int(p)


def state_to_dict(tokens: Optional[List[List[int]]], curr_pos: int, copy: bool = True) -> dict:
  """The state this engine hands back: the reference's field names, types and nesting for the two
  entries it needs (JSON-able, survives the gRPC hop of grpc_peer_handle.py:209-230).  ``copy=False``
  (engines bound to an intra-box ring): the token history is the engine's live list instead of a snapshot --
  O(B x context) per hop is what the box-internal path exists to avoid; whoever serialises it copies."""
  out = {"curr_pos": int(curr_pos)}
  if tokens is not None:
    out["tokens"] = [row[:] for row in tokens] if copy else tokens   # snapshot, like ShardInferenceState.to_dict
  return out


class _PinnedPool:
  """Page-locked host buffers behind the numpy arrays the plugin API hands out.

  SURVEY 8b: results are host-owned arrays the caller may keep (the reference's own contract test
  holds two results and compares them, xotorch/inference/test_inference_engine.py:43-44; a node may
  hold a hidden state while it serialises it for the next peer).  So a buffer belongs to the array
  that was returned -- and to every view derived from it -- for as long as the caller keeps it:
  a ``weakref.finalize`` on the array puts the buffer back on the free list when the last reference
  goes away; nothing is ever overwritten under a live array, and in steady state (the caller drops
  each result before asking for the next few) no page-locked memory is allocated after warm-up."""

  def __init__(self):
    self.free: Dict[tuple, list] = {}
    self.allocated = 0

  def take(self, dtype: torch.dtype, shape) -> tuple:
    """(pinned tensor viewed as ``shape``, read-only numpy array over the same memory)."""
    import weakref
    n = 1
    for d in shape:
      n *= int(d)
    n = max(n, 1)
    cap = 1 << (n - 1).bit_length()           # power-of-two size classes
    lst = self.free.setdefault((dtype, cap), [])
    base = lst.pop() if lst else None
    if base is None:
      base = torch.empty(cap, dtype=dtype).pin_memory()
      self.allocated += 1
    buf = base[:n].view(tuple(int(d) for d in shape))
    arr = buf.numpy()
    arr.setflags(write=False)
    weakref.finalize(arr, lst.append, base)
    return buf, arr


class _Request:
  __slots__ = ("slots", "pos", "tokens", "skip", "last_sample")

  def __init__(self, slots: List[int]):
    self.slots = slots
    self.pos = 0          # the reference's curr_pos: counts every prompt slot, padded ones included
    self.skip = [0] * len(slots)   # padded prompt slots of each row (they occupy no KV position here)
    self.last_sample = None        # (temp, top_k, top_p) of this request's latest sample() call
    # token history per row as Python ints: a decode step appends one id per row, so the
    # ``tokens`` entry of the returned state costs O(B) + a pointer copy, not a fresh
    # ndarray.tolist() of the whole history (1.2 ms per step at 32 x 2100 ids)
    self.tokens: Optional[List[List[int]]] = None


class B200ShardInferenceEngine(InferenceEngine):
  def __init__(self, shard_downloader, device: Optional[int] = None, max_seqs: Optional[int] = None,
               max_tokens_per_seq: Optional[int] = None, max_batch_tokens: Optional[int] = None,
               page_tokens: Optional[int] = None, sample_mode: str = "reference",
               speculate_next: Optional[bool] = None):
    self.shard = None
    self.shard_downloader = shard_downloader
    self.stage: Optional[StageRuntime] = None
    self.tokenizer = None
    self.model_path = None
    self.model_config = None
    self.uuid = str(uuid.uuid4())
    self.executor = ThreadPoolExecutor(max_workers=1)  # strictly serialised like the reference (:46)
    self._waiter = ThreadPoolExecutor(max_workers=32)   # waits for GPU events, so that the worker above never blocks on one
    env = os.environ
    self.device_index = device if device is not None else int(env.get("XB_DEVICE", env.get("LOCAL_RANK", "0")))
    self.max_seqs = max_seqs or int(env.get("XB_MAX_SEQS", "32"))
    self.max_tokens_per_seq = max_tokens_per_seq or int(env.get("XB_MAX_TOKENS_PER_SEQ", "4096"))
    self.max_batch_tokens = max_batch_tokens or int(env.get("XB_MAX_BATCH_TOKENS", "4096"))
    self.page_tokens = page_tokens or int(env.get("XB_PAGE_TOKENS", "64"))
    self.sample_mode = sample_mode  # "reference": q from torch.Generator(seed 1234) like :68-69 ; "fast": counter noise
    self.requests: Dict[str, _Request] = {}
    self.free_slots: List[int] = []
    self.rng = None
    self.sample_calls = 0
    # host staging: the [B, vocab] fp32 logits the plugin API returns are tens of MB per step, so
    # they land in PINNED buffers that the returned numpy arrays own (recycled once the caller has
    # dropped the array), and the device copy is kept so that sample() on that very array skips the upload
    self._pinned = _PinnedPool()
    # device copies of the logits arrays handed out that are still alive, keyed by id(array): sample() on such an
    # array skips the 16 MB upload -- also when other requests' results came in between (several requests in flight)
    self._logits_dev: Dict[int, tuple] = {}   # id(host array) -> (weakref to it, device tensor, request id)
    self._last_logits = None   # newest entry (kept for generate(), which resets it)
    self.speculate_next = (env.get("XB_API_SPECULATE", "1") != "0") if speculate_next is None else bool(speculate_next)
    self.direct_host_logits = env.get("XB_API_DIRECT_LOGITS", "1") != "0"   # LM head stores the logits into pinned host memory itself
    self._spec = None          # the decode forward enqueued ahead of the caller's infer_tensor, see sample()
    self.spec_hits = self.spec_misses = 0
    self._spec_sample = None   # the sampler run enqueued ahead of the caller's sample(), see _speculate_sample()
    self._ring_specs = collections.deque()   # the same for a ring's last stage: one per request in flight, in forward order
    self._ring_spec_lock = threading.Lock()  # sample()'s fast path looks at the queue from the event-loop thread
    self.sample_spec_hits = 0
    # intra-box hand-off (networking.py): with device_handoff a non-last shard returns a DeviceTensorRef -- the hidden
    # rows stay in HBM -- instead of fp32 numpy; connect_next() / bind_ring() say where the next shard lives
    self.device_handoff = False
    self.landing_slots = 0          # > 0: this engine keeps landing buffers the previous stage's kernels may store into
    self.landing_rows = 0
    self._landing_epoch = 0         # bumped whenever the stage (and with it the landing buffers) is rebuilt
    self._next_engine = None        # next shard's engine when it lives in this process
    self._next_epoch = None         # epoch of the next engine's landing buffers this stage is connected to (None: not connected)
    self._ho_next = 0               # round-robin landing slot + the flag value last written to each slot
    self._ho_value: List[int] = []
    self._ring = None               # networking.RingChannel when the next shard is another process (torchrun)
    self._ring_landing = None       # (slots, rows) of the next rank's landing buffers once connect_ring_landing mapped them

  # ------------------------------------------------------------------ shard / weights
  async def ensure_shard(self, shard: Shard):
    shard = as_shard(shard)
    if self.shard == shard:
      return
    self.model_path = await self.shard_downloader.ensure_shard(shard, self.__class__.__name__)
    self.model_config = load_model_config(self.model_path / "config.json")
    self.tokenizer = await self._resolve_tokenizer(self.model_path)

    def start_model():
      if self.stage is not None:
        self.stage.close()
      st = StageRuntime(shard, self.model_config, device=self.device_index, max_batch_tokens=self.max_batch_tokens,
                        max_seqs=self.max_seqs, max_tokens_per_seq=self.max_tokens_per_seq, page_tokens=self.page_tokens)
      st.load_dir(self.model_path)
      self.stage = st
      self._stage_changed()
      self.requests.clear()
      self.free_slots = list(range(self.max_seqs - 1, -1, -1))
      self.rng = torch.Generator(device=st.device)
      self.rng.manual_seed(1234)
      self._stream = torch.cuda.Stream(device=st.device)  # non-default stream: decode steps replay CUDA graphs
      # sampled ids -> host beside the next step; on ring stages also the sampler itself.  High priority: with other
      # requests' forwards queued back to back, its blocks get the next free SMs instead of waiting for a gap
      self._copy_stream = torch.cuda.Stream(device=st.device, priority=-1)
      self._logits_stream = torch.cuda.Stream(device=st.device, priority=-1)

    await asyncio.get_running_loop().run_in_executor(self.executor, start_model)
    self.shard = shard

  def attach_stage(self, shard: Shard, stage: StageRuntime, config: dict, tokenizer=None):
    """Adopt an already-loaded StageRuntime (weights resident in HBM) instead of loading the
    shard from a checkpoint directory -- in-memory / synthetic weights, benchmarks."""
    self.stage, self.shard, self.model_config = stage, as_shard(shard), config
    self.max_seqs, self.max_batch_tokens = stage.max_seqs, stage.max_batch_tokens
    self._stage_changed()
    self.tokenizer = tokenizer or SyntheticTokenizer(config["vocab_size"])
    self.requests.clear()
    self.free_slots = list(range(self.max_seqs - 1, -1, -1))
    self.rng = torch.Generator(device=stage.device)
    self.rng.manual_seed(1234)
    self._stream = torch.cuda.Stream(device=stage.device)
    self._copy_stream = torch.cuda.Stream(device=stage.device, priority=-1)   # see ensure_shard
    self._logits_stream = torch.cuda.Stream(device=stage.device, priority=-1)

  # ------------------------------------------------------------------ intra-box hand-off wiring
  def enable_landing(self, n_slots: Optional[int] = None, rows: Optional[int] = None):
    """Keep landing buffers on this engine's GPU so that the PREVIOUS stage can deliver decode steps
    through its own kernels (xb_forward_handoff).  Call before the shard is loaded (or re-attach)."""
    self.landing_slots = int(n_slots or self.max_seqs)
    self.landing_rows = int(rows or min(128, self.max_seqs))
    if self.stage is not None:
      self._stage_changed()

  def connect_next(self, next_engine: "B200ShardInferenceEngine"):
    """The next shard of the ring is served by ``next_engine`` in THIS process (another GPU of the box):
    hidden rows are handed over as device references, decode steps through its landing buffers."""
    self._next_engine = next_engine
    self._next_epoch = None
    self.device_handoff = True

  def bind_ring(self, channel, sends: bool = True):
    """The next shard lives in another process of a torchrun ring: hidden rows leave over the channel's data
    plane (NCCL), sent by this engine on its own stream right behind the forward that made them -- or, once
    ``connect_ring_landing`` has mapped the next rank's landing buffers, decode steps are stored there by this
    stage's last kernel.  sends=False: the last stage, which only receives."""
    self._ring = channel
    self.device_handoff = bool(sends)

  def ring_landing_handles(self, n_slots: int, rows: int) -> Optional[tuple]:
    """CUDA IPC handles of this stage's landing buffers for the PREVIOUS rank of the ring (created on first use; a
    stage that already has buffers of another geometry, or none that can be exported, answers None)."""
    st = self.stage
    if st is None:
      return None
    if st.ho_geometry is None:
      st.handoff_create(n_slots, rows, ipc=True)
    if st.ho_geometry != (int(n_slots), int(rows)) or st.ho_handles is None:
      return None
    return st.ho_handles

  def connect_ring_landing(self, next_handles: Optional[tuple], n_slots: int, rows: int):
    """Map the NEXT rank's landing buffers (its ``ring_landing_handles``); decode steps of at most ``rows`` sequences
    then leave through them, one slot per request in flight (at most ``n_slots``), everything else over the data plane."""
    st = self.stage
    if st is None or st.is_last or next_handles is None:
      return
    if st.ho_geometry is None:      # the first stage receives nothing, but a producer needs its side of the hand-off state too
      st.handoff_create(n_slots, rows, ipc=False)
    if st.ho_geometry != (int(n_slots), int(rows)):
      return
    if not st.ho_connected:
      st.handoff_connect(*next_handles)
    self._ring_landing = (int(n_slots), int(rows))
    self._ring_slot_of: dict = {}      # request id -> its slot in the next rank's buffers
    self._ring_slots_free = list(range(int(n_slots)))[::-1]

  def _stage_changed(self):
    self._landing_epoch += 1
    self._ring_landing = None     # mapped buffers belonged to the old stage: back to the data plane until reconnected
    self._next_epoch = None       # a rebuilt producer stage is not connected to anything yet
    if self.landing_slots > 0 and self.stage is not None and not self.stage.is_first:
      self._ensure_handoff_buffers(self.stage, self.landing_slots, self.landing_rows)
      self._landing_ready = self._landing_epoch
    else:
      self._landing_ready = None

  @staticmethod
  def _ensure_handoff_buffers(st: StageRuntime, n_slots: int, rows: int):
    """xb_handoff_create once per stage: landing buffers + flags (consumer side) and completion counters (producer side)."""
    geo = getattr(st, "_ho_geometry", None)
    if geo is None:
      st.handoff_create(n_slots, rows, ipc=False)
      st._ho_geometry = (n_slots, rows)
    elif geo != (n_slots, rows):
      raise RuntimeError(f"hand-off geometry {geo} of this stage differs from the {(n_slots, rows)} the next stage uses")

  def _landing_out(self, bsz: int, request_id: Optional[str] = None):
    """(slot, value) in the next engine's landing buffers for a decode step of ``bsz`` rows, or None when that
    path is not available (then the rows travel as a device tensor).  Connects on first use."""
    nxt = self._next_engine
    st = self.stage
    if self._ring_landing is not None and st is not None and not st.is_last:
      # process-per-GPU ring: one slot per request (a request has one step in flight, so its slot is free again
      # by the time its next step arrives here); the flag value of a slot only grows
      if bsz > self._ring_landing[1]:
        return None
      slot = self._ring_slot_of.get(request_id)
      if slot is None:
        if not self._ring_slots_free:
          return None
        slot = self._ring_slot_of[request_id] = self._ring_slots_free.pop()
      return slot, st.ho_sent.get(slot, 0) + 1
    if nxt is None or st is None or st.is_last or nxt.stage is None or getattr(nxt, "_landing_ready", None) is None:
      return None
    if bsz > nxt.landing_rows or nxt.stage.dim != st.dim:
      return None
    if self._next_epoch is None:
      if getattr(self, "_connected_once", None) == self._landing_epoch:
        return None     # this stage was connected to buffers that no longer exist: stay on the tensor path
      self._ensure_handoff_buffers(st, nxt.landing_slots, nxt.landing_rows)   # completion counters, consumer's geometry
      st.handoff_connect_local(nxt.stage)
      self._next_epoch = nxt._landing_ready
      self._connected_once = self._landing_epoch
      self._ho_value = [0] * nxt.landing_slots
      self._ho_next = 0
    if self._next_epoch != nxt._landing_ready:
      return None       # the consumer rebuilt its stage
    slot = self._ho_next
    self._ho_next = (slot + 1) % len(self._ho_value)
    self._ho_value[slot] += 1
    return slot, self._ho_value[slot]

  async def _resolve_tokenizer(self, path):
    try:
      if any((path / f).exists() for f in ("tokenizer.json", "tokenizer.model", "tokenizer_config.json")):
        from transformers import AutoTokenizer
        return AutoTokenizer.from_pretrained(path)
    except Exception:
      pass
    return SyntheticTokenizer(self.model_config["vocab_size"])

  async def load_checkpoint(self, shard: Shard, path: str):
    await self.ensure_shard(shard)

  # ------------------------------------------------------------------ tokens <-> text
  async def encode(self, shard: Shard, prompt: str) -> np.ndarray:
    await self.ensure_shard(shard)

    def enc():
      ids = self.tokenizer.encode(prompt)
      return np.asarray(ids, dtype=np.int64).reshape(-1)

    return await asyncio.get_running_loop().run_in_executor(self.executor, enc)

  async def decode(self, shard: Shard, tokens: np.ndarray) -> str:
    await self.ensure_shard(shard)
    return await asyncio.get_running_loop().run_in_executor(
      self.executor, functools.partial(self.tokenizer.decode, np.asarray(tokens).reshape(-1).tolist()))

  # ------------------------------------------------------------------ host staging
  def _to_host(self, dev: torch.Tensor) -> np.ndarray:
    arr, ev = self._to_host_async(dev)
    ev.synchronize()
    return arr

  def _to_host_async(self, dev: torch.Tensor):
    """Device tensor -> numpy through a pinned buffer (one DMA, no pageable bounce).  The returned
    read-only array OWNS its buffer (see _PinnedPool); its contents are defined once the returned
    event has completed."""
    buf, arr = self._pinned.take(dev.dtype, dev.shape)
    buf.copy_(dev, non_blocking=True)
    ev = torch.cuda.Event()
    ev.record(torch.cuda.current_stream(dev.device))
    return arr, ev

  def _logits_to_host_beside(self, out: torch.Tensor):
    """fp32 [B, V] to the host on the side stream: whatever is queued behind the forward on the main stream -- the
    sampler and the next step run ahead, another request's forward -- does not wait behind a 16 MB copy."""
    done = torch.cuda.Event()
    done.record(torch.cuda.current_stream(out.device))
    # not the sampler's stream: this copy waits for a forward that may be queued behind other requests' forwards, and a
    # sampler call for an EARLIER request enqueued after it must not wait with it
    side = self._logits_stream
    side.wait_event(done)
    # (no record_stream: _remember_logits keeps ``out`` for as long as the caller keeps the host array, and the
    # caller only gets that array once this copy has completed)
    with torch.cuda.stream(side):
      return self._to_host_async(out)

  def _remember_logits(self, host: np.ndarray, dev: torch.Tensor, request_id: str):
    import weakref
    key = id(host)
    self._logits_dev[key] = (weakref.ref(host), dev, request_id)
    weakref.finalize(host, self._logits_dev.pop, key, None)   # the device copy lives exactly as long as the caller's array
    self._last_logits = (host, dev, request_id)

  def _forward_logits(self, st: StageRuntime, slots, q_lens, positions, **inp):
    """One native forward on a last shard whose logits go to the host: (device logits, numpy
    view, event).  With ``direct_host_logits`` the LM-head kernel stores them into the pinned
    buffer itself (xb_forward_host_logits) -- the shortest path to the host when this request is the only one,
    since the stores overlap the LM head's own weight stream.  With several requests in flight the next request's
    forward is already queued behind this one, and 16 MB of stores over PCIe inside the LM-head kernel would hold it
    back (measured 0.45 ms per hop on the last stage of a 2-GPU ring): then a copy on the side stream follows the
    forward instead and overlaps the next one."""
    if not self.direct_host_logits or len(self.requests) > 1:
      out = st.forward(slots, q_lens, positions, **inp)
      return (out,) + self._logits_to_host_beside(out)
    buf, arr = self._pinned.take(torch.float32, (len(slots), st.vocab))
    out = st.forward(slots, q_lens, positions, host_logits=buf, **inp)
    ev = torch.cuda.Event()
    ev.record(torch.cuda.current_stream(st.device))
    return out, arr, ev

  # ------------------------------------------------------------------ sampling
  async def sample(self, x: np.ndarray, temp=TEMP, top_k=TOP_K, top_p: float = 0.0) -> np.ndarray:
    def sample_wrapper():
      # Whole-model shard: the sampler stays on the forward stream, the next step is enqueued right behind it.
      # Otherwise (a ring stage) it gets a stream of its own: with several requests in flight the engine's queue
      # may already hold ANOTHER request's forward, and ids that only need this request's finished logits would
      # wait a whole stage time behind it (measured 3 ms per hop at 2 x 32 sequences on a 2-GPU ring).
      own = self.shard is not None and self.shard.is_first_layer() and self.shard.is_last_layer() and self.speculate_next
      trace("sample.worker.begin")
      with torch.cuda.stream(self._stream if own else self._copy_stream):
        r = sample_body()
      trace(f"sample.worker.end.hits{self.sample_spec_hits}")
      return r

    def sample_body():
      st = self.stage
      params = (float(temp), int(top_k or 0), float(top_p))
      ss, self._spec_sample = self._spec_sample, None
      if ss is not None:
        if ss["ref"]() is x and ss["params"] == params:
          self.sample_spec_hits += 1       # this very call was enqueued right behind the forward that made x
          return ss["toks"], ss["ev"]
        self._rewind_sample(ss)
      hit = self._take_ring_spec(x, params, drop_on_miss=True)
      if hit is not None:
        return hit
      owner = None
      ent = self._logits_dev.get(id(x))
      if ent is not None and ent[0]() is x:
        logits, owner = ent[1], ent[2]     # an array this engine returned (read-only): its values are still on the device
      else:
        logits = torch.as_tensor(np.ascontiguousarray(x, dtype=np.float32)).to(st.device)
      if logits.ndim == 1:
        logits = logits.unsqueeze(0)
      logits.record_stream(torch.cuda.current_stream(st.device))
      req = self.requests.get(owner) if owner is not None else None
      if req is not None:
        req.last_sample = params
      return self._enqueue_sample(logits, owner, *params)

    loop = asyncio.get_running_loop()
    # the common case on a ring's last stage: this call's sampler already ran (or runs) behind the forward -- nothing
    # to enqueue, so no trip through the worker either.  (A call for anything but the oldest run-ahead sampler takes
    # the worker's queue like every call that draws noise, which keeps the draws in call order.)
    fast = self._take_ring_spec(x, (float(temp), int(top_k or 0), float(top_p)), drop_on_miss=False) if self._ring_specs else None
    host_toks, ev = fast if fast is not None else await loop.run_in_executor(self.executor, sample_wrapper)
    if not ev.query():
      trace("sample.wait")
      await loop.run_in_executor(self._waiter, ev.synchronize)
    return np.array(host_toks, dtype=np.int32).reshape(-1, 1)

  def _enqueue_sample(self, logits: torch.Tensor, owner: Optional[str], temp: float, top_k: int, top_p: float):
    """Sampler + (whole-model shard) the next decode step on the current stream; (pinned ids, event)."""
    st = self.stage
    q = None
    if self.sample_mode == "reference":
      # the reference draws q ~ Exp(1) of shape [B, vocab] on EVERY call, also at temp 0
      q = torch.empty((logits.size(0), st.vocab), device=st.device).exponential_(1, generator=self.rng)
    self.sample_calls += 1
    # reference semantics at any temperature: logits / max(temp, 1e-5) -> top-k -> softmax -> argmax(p/q)
    toks = st.sample(logits, temp=max(float(temp), 1e-5), top_k=int(top_k or 0), top_p=top_p, q=q, seed=self.sample_calls)
    # The next decode step goes onto the stream right behind the sampler (the GPU never waits for
    # the host); the ids reach the host through a second stream that only waits for the sampler.
    sampled = torch.cuda.Event()
    sampled.record(torch.cuda.current_stream(st.device))
    self._speculate(owner, toks)
    side = self._copy_stream
    side.wait_event(sampled)
    toks.record_stream(side)
    with torch.cuda.stream(side):
      host_toks, ev = self._to_host_async(toks)
    if self._spec is not None and self._spec.get("toks") is None:
      self._spec["toks"] = host_toks
    return host_toks, ev

  def _speculate_sample(self, request_id: str, req: _Request, host: np.ndarray, logits: torch.Tensor):
    """Whole-model shard, one request in flight: run the sampler the caller is about to ask for -- with the
    parameters of its previous sample() call -- right behind the forward that produced these logits, and the next
    decode step behind that.  The host then only COLLECTS: sample(x) on this array with those parameters returns the
    ids already on their way, infer_tensor(ids) the forward already running; between a step's logits and the next
    step's kernels the GPU no longer waits for a host round trip.  Anything else the caller does first un-draws the
    noise (the reference draws q once per sample() call, in call order) and discards the speculated work, so results
    are those of the unspeculated path bit for bit.  With more than one request in flight the order of their
    sample() calls is not known in advance: no sampler speculation then."""
    import weakref
    if not self.speculate_next or req.last_sample is None or len(self.requests) != 1 or self.shard is None:
      return
    if not (self.shard.is_first_layer() and self.shard.is_last_layer()) or logits.size(0) != len(req.slots):
      return
    before = (self._noise_mark(), self.sample_calls)
    toks, ev = self._enqueue_sample(logits, request_id, *req.last_sample)
    self._spec_sample = {"ref": weakref.ref(host), "params": req.last_sample, "toks": toks, "ev": ev, "before": before}

  def _speculate_ring_sample(self, request_id: str, req: _Request, host: np.ndarray, logits: torch.Tensor):
    """The last stage of a ring, any number of requests in flight: their logits leave this engine in the order of
    their forwards, and a node samples each as soon as it has them -- so the sampler of every request is run ahead in
    that order (right behind its forward, beside the 16 MB logits copy instead of a host round trip after it) and
    sample() collects.  The first call that does not match the oldest run-ahead sampler gives ALL their noise back
    (draws happen in call order) and is computed as asked."""
    import weakref
    if not self.speculate_next or req.last_sample is None or self.shard is None or logits.size(0) != len(req.slots):
      return
    if self.shard.is_first_layer() or not self.shard.is_last_layer():
      return
    before = (self._noise_mark(), self.sample_calls)
    # on the forward's own stream, whose tail at this moment IS that forward: ~50 us there; beside another request's
    # forward (the sampler's stream) the same kernels were measured at ~600 us, starved of bandwidth
    if _nw.TRACE is not None:
      e0 = torch.cuda.Event(enable_timing=True); e0.record()
    toks, ev = self._enqueue_sample(logits, request_id, *req.last_sample)
    if _nw.TRACE is not None:
      e1 = torch.cuda.Event(enable_timing=True); e1.record()
      self._gpu_trace_s = getattr(self, "_gpu_trace_s", []) + [(e0, e1)]
    with self._ring_spec_lock:
      self._ring_specs.append({"ref": weakref.ref(host), "params": req.last_sample, "toks": toks, "ev": ev, "before": before,
                               "request_id": request_id})

  def _take_ring_spec(self, x, params, drop_on_miss: bool):
    """The oldest run-ahead sampler if it is the one for ``x`` with these parameters: (pinned ids, event).  Otherwise
    None -- and, when asked (the engine's worker, about to draw noise for this call), all run-ahead draws are given
    back first: the calls come in another order than the forwards did."""
    with self._ring_spec_lock:
      if not self._ring_specs:
        return None
      head = self._ring_specs[0]
      if head["ref"]() is x and head["params"] == params:
        self._ring_specs.popleft()
        self.sample_spec_hits += 1
        return head["toks"], head["ev"]
      if drop_on_miss:
        self._rewind_sample(self._ring_specs[0])
        self._ring_specs.clear()
      return None

  def _drop_ring_specs(self):
    with self._ring_spec_lock:
      if self._ring_specs:
        self._rewind_sample(self._ring_specs[0])
        self._ring_specs.clear()

  def _noise_mark(self):
    """Where the noise stream stands (reference mode): the Philox offset -- an int, no tensor like get_state()."""
    return self.rng.get_offset() if self.sample_mode == "reference" else None

  def _rewind_sample(self, ss: dict):
    """A speculated sampler run that the caller did not ask for: give its noise back, forget the step enqueued on its ids."""
    rng_state, calls = ss["before"]
    if rng_state is not None:
      self.rng.set_offset(rng_state)
    self.sample_calls = calls
    self._spec = None

  def _speculate(self, request_id: Optional[str], toks: torch.Tensor):
    """Enqueue the decode step the caller is about to ask for (see the module docstring).
    Only when these logits came from ``request_id``'s last forward on a whole-model shard."""
    if not self.speculate_next or request_id is None or self.shard is None:
      return   # logits that are not the engine's latest result: nothing known about their request
    self._spec = None   # one speculation at a time; a stale one for another request is simply dropped
    if not (self.shard.is_first_layer() and self.shard.is_last_layer()):
      return
    req = self.requests.get(request_id)
    st = self.stage
    if req is None or len(req.slots) != toks.numel() or req.pos + 1 > st.max_tokens_per_seq or len(req.slots) > st.max_batch_tokens:
      return
    try:
      out, host, ev = self._forward_logits(st, req.slots, [1] * len(req.slots), self._positions(req, req.pos), tokens=toks)
    except RuntimeError:
      return   # e.g. KV pool exhausted: the real call will report it
    self._spec = {"request_id": request_id, "pos": req.pos, "toks": None, "out": out, "host": host, "event": ev}   # "toks": filled by sample()

  def _take_spec(self, request_id: str, pos: int, x: np.ndarray):
    """The speculated result if (request, position, ids) are exactly the ones it was computed
    for, else None (and the speculation, if it was this request's, is forgotten)."""
    sp = self._spec
    if sp is None or sp["request_id"] != request_id:
      return None
    self._spec = None
    if x.ndim == 2 and x.shape[1] == 1 and sp["pos"] == pos and x.shape[0] == sp["toks"].shape[0] \
       and np.array_equal(x[:, 0].astype(np.int64), sp["toks"].astype(np.int64)):
      self.spec_hits += 1
      return sp
    self.spec_misses += 1
    return None

  # ------------------------------------------------------------------ forward
  def _start_request(self, request_id: str, bsz: int) -> _Request:
    old = self.requests.pop(request_id, None)
    if old is not None:
      for s in old.slots:
        self.stage.release_slot(s)
        self.free_slots.append(s)
    if len(self.free_slots) < bsz:
      # evict the oldest finished/abandoned request (the reference keeps a single state and
      # simply overwrites it on the next prompt)
      for rid in list(self.requests.keys()):
        if len(self.free_slots) >= bsz:
          break
        for s in self.requests.pop(rid).slots:
          self.stage.release_slot(s)
          self.free_slots.append(s)
    if len(self.free_slots) < bsz:
      raise RuntimeError(f"request needs {bsz} KV slots, engine has {self.max_seqs}")
    req = _Request([self.free_slots.pop() for _ in range(bsz)])
    self.requests[request_id] = req
    return req

  def end_request(self, request_id: str):
    req = self.requests.pop(request_id, None)
    if self._spec_sample is not None:
      self._rewind_sample(self._spec_sample)
      self._spec_sample = None
    if self._spec is not None and self._spec["request_id"] == request_id:
      self._spec = None
    if any(sp["request_id"] == request_id for sp in self._ring_specs):
      self._drop_ring_specs()
    if self._ring_landing is not None and request_id in self._ring_slot_of:
      self._ring_slots_free.append(self._ring_slot_of.pop(request_id))
    if req is not None:
      for s in req.slots:
        self.stage.release_slot(s)
        self.free_slots.append(s)

  def _pad_id(self) -> int:
    """The id the reference treats as padding (sharded_inference_engine.py:153-163): the tokenizer's
    ``pad_id``, else its ``pad_token_id``, else -- also when that is None -- 0."""
    tk = self.tokenizer
    if hasattr(tk, "pad_id"):
      return tk.pad_id
    pid = getattr(tk, "pad_token_id", None)
    return 0 if pid is None else pid

  def _positions(self, req: _Request, pos: int) -> List[int]:
    return [pos - k for k in req.skip]

  def _forward_chunked(self, st: StageRuntime, req: _Request, dev_in: torch.Tensor, bsz: int, s: int, pos: int,
                       is_tokens: bool, to_host: bool = True):
    """One logical forward over [bsz, s] rows, split so that no native call carries more
    than max_batch_tokens rows: whole sequences per call, or position chunks of one
    sequence (chunked prefill) when a single sequence is longer than that."""
    cap = st.max_batch_tokens
    last = self.shard.is_last_layer()
    positions = self._positions(req, pos)
    if bsz * s <= cap:
      flat = dev_in.reshape(-1) if is_tokens else dev_in.reshape(bsz * s, -1)
      inp = {"tokens": flat} if is_tokens else {"hidden": flat}
      if last and to_host:
        return self._forward_logits(st, req.slots, [s] * bsz, positions, **inp)   # (device logits, host view, event)
      return st.forward(req.slots, [s] * bsz, positions, **inp)
    outs = []
    if s <= cap:
      per = max(1, cap // s)
      for b0 in range(0, bsz, per):
        b1 = min(bsz, b0 + per)
        part = dev_in[b0:b1]
        flat = part.reshape(-1) if is_tokens else part.reshape((b1 - b0) * s, -1)
        outs.append(st.forward(req.slots[b0:b1], [s] * (b1 - b0), positions[b0:b1],
                               **({"tokens": flat} if is_tokens else {"hidden": flat})))
      return torch.cat(outs, dim=0)
    for b in range(bsz):
      outs.append(self._forward_one_sequence(st, req.slots[b], dev_in[b], positions[b], is_tokens))
    return torch.cat(outs, dim=0)

  def _forward_one_sequence(self, st: StageRuntime, slot: int, rows: torch.Tensor, pos: int, is_tokens: bool) -> torch.Tensor:
    """rows: ids [n] or hidden [n, D] of ONE sequence, run in position chunks of at most max_batch_tokens.
    Returns hidden [n, D] (non-last shard) or the logits [1, V] of the last row."""
    cap, n = st.max_batch_tokens, rows.shape[0]
    seq_out = []
    for c0 in range(0, n, cap):
      c1 = min(n, c0 + cap)
      part = rows[c0:c1].contiguous()
      flat = part.reshape(-1) if is_tokens else part.reshape(c1 - c0, -1)
      seq_out.append(st.forward([slot], [c1 - c0], [pos + c0], **({"tokens": flat} if is_tokens else {"hidden": flat})))
    return seq_out[-1] if self.shard.is_last_layer() else torch.cat(seq_out, dim=0)

  def _forward_padded_prompt(self, st: StageRuntime, req: _Request, dev_in: torch.Tensor, keep: np.ndarray, is_tokens: bool):
    """Prompt holding the pad id (the reference's padded branch, sharded_inference_engine.py:165-177:
    positions from the padding mask, padded keys never attended).  Real tokens never see a padded
    key and their positions are the running count of real tokens, so running ONLY the real tokens,
    compacted, gives them exactly the reference's result; the padded slots get no KV entry here.
    Differences, by construction invisible to real tokens: rows of padded slots in a returned hidden
    state are zero (the reference computes a row that attends itself), and the logits come from the
    last REAL prompt token (the reference takes row -1 even when that slot is padding)."""
    bsz, s = keep.shape
    last = self.shard.is_last_layer()
    outs = []
    for b in range(bsz):
      kb = keep[b].copy()
      if not kb.any():
        kb[-1] = True   # a row of nothing but padding: its last slot stands for the row (position 0, attends itself)
      idx = torch.as_tensor(np.nonzero(kb)[0], device=dev_in.device)
      req.skip[b] = int(s - idx.numel())
      o = self._forward_one_sequence(st, req.slots[b], dev_in[b].index_select(0, idx), 0, is_tokens)
      if last:
        outs.append(o)
      else:
        full = torch.zeros((s, o.shape[-1]), dtype=o.dtype, device=o.device)
        full.index_copy_(0, idx, o)
        outs.append(full)
    return torch.cat(outs, dim=0)

  # ------------------------------------------------------------------ intra-box hand-off: data path
  def _rows_from_ref(self, st: StageRuntime, ref: DeviceTensorRef, bsz: int, s: int):
    """What a DeviceTensorRef stands for, as input of this stage: (bf16 rows [bsz, s, D] on this GPU | None, landing)."""
    if ref.landing is not None:
      return None, ref.landing                        # already in this GPU's landing buffer (stored by the producer's kernel)
    if ref.recv is not None and ref.tensor is None:   # process-per-GPU ring: post the receive on this engine's stream
      buf = torch.empty((bsz * s, st.dim), dtype=torch.bfloat16, device=st.device)
      self._ring.recv_rows(buf)
      return buf.reshape(bsz, s, -1), None
    t = ref.tensor
    if ref.event is not None:
      torch.cuda.current_stream(st.device).wait_event(ref.event)
    t.record_stream(torch.cuda.current_stream(st.device))   # the producer may drop its reference before this stream is done
    if t.device != st.device:
      t = t.to(st.device, non_blocking=True)          # GPU -> GPU over NVLink, no host bounce
    return t.reshape(bsz, s, -1), None

  def _device_result(self, out: Optional[torch.Tensor], bsz: int, s: int, landing_out) -> DeviceTensorRef:
    st = self.stage
    shape = (bsz, s, st.dim)
    if landing_out is not None:
      return DeviceTensorRef(shape, landing=landing_out, recv=("landed",) if self._ring_landing is not None else None)
    if self._ring is not None:
      self._ring.send_rows(out.reshape(bsz * s, -1))    # enqueued on this stream, right behind the forward
      return DeviceTensorRef(shape, tensor=out, recv=("sent",))
    ev = torch.cuda.Event()
    ev.record(torch.cuda.current_stream(st.device))
    return DeviceTensorRef(shape, tensor=out, event=ev)

  def _forward_decode_hop(self, st: StageRuntime, req: _Request, bsz: int, pos: int, dev_in, is_tokens: bool, landing_in,
                          request_id: Optional[str] = None):
    """One decode step of a ring stage: input = ids / hidden rows / this GPU's landing slot, output = logits (last
    shard), hidden rows, or -- when the next engine's landing buffers are connected -- nothing but a (slot, value)."""
    positions = self._positions(req, pos)
    landing_out = None if self.shard.is_last_layer() else self._landing_out(bsz, request_id)
    inp = {}
    if landing_in is None:
      flat = dev_in.reshape(-1) if is_tokens else dev_in.reshape(bsz, -1)
      inp = {"tokens": flat} if is_tokens else {"hidden": flat.contiguous()}
    ho = None
    if landing_in is not None or landing_out is not None:
      si, vi = landing_in if landing_in is not None else (-1, 0)
      so, vo = landing_out if landing_out is not None else (-1, 0)
      ho = (si, vi, so, vo)
    if _nw.TRACE is not None:     # developer aid: device time of the hop's forward and the gap to the next one
      e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
      e0.record()
      out = st.forward(req.slots, [1] * bsz, positions, handoff=ho, **inp)
      e1.record()
      self._gpu_trace = getattr(self, "_gpu_trace", []) + [(e0, e1)]
      return out, landing_out
    return st.forward(req.slots, [1] * bsz, positions, handoff=ho, **inp), landing_out

  def _prompt_keep_mask(self, x: np.ndarray, inference_state: Optional[dict], bsz: int, s: int, pos: int) -> Optional[np.ndarray]:
    """bool [bsz, s] of the prompt slots that are NOT padding, or None when the prompt holds no pad
    id (the usual case: the reference's ``tril`` branch).  Only a prompt (multi-token input at
    position 0) is examined -- a generated id equal to the pad id is an ordinary token, as in the
    reference, whose padding mask is extended with True over the generated slots.  A later shard
    sees hidden rows, not ids: it finds the ids in the ``tokens`` entry of the travelling state."""
    if pos != 0 or s <= 1:
      return None
    if x.ndim == 2:
      ids = x
    else:
      hist = (inference_state or {}).get("tokens")
      if hist is None:
        return None
      ids = np.asarray(hist)
      if ids.ndim != 2 or ids.shape != (bsz, s):
        return None
    keep = ids != self._pad_id()
    return None if keep.all() else keep

  async def infer_tensor(self, request_id: str, shard: Shard, input_data: np.ndarray,
                         inference_state: Optional[dict] = None) -> tuple[np.ndarray, Optional[dict]]:
    await self.ensure_shard(shard)

    def infer_wrapper():
      trace("eng.worker.begin", request_id)
      with torch.cuda.stream(self._stream):
        r = infer_body()
      trace("eng.worker.end", request_id)
      return r

    def infer_body():
      st = self.stage
      ref = input_data if isinstance(input_data, DeviceTensorRef) else None
      x = ref if ref is not None else np.asarray(input_data)
      if x.ndim == 2:
        bsz, s = x.shape
      elif x.ndim == 3:
        bsz, s = x.shape[0], x.shape[1]
      else:
        raise ValueError(f"input_data must be 2-D token ids or 3-D hidden states, got ndim={x.ndim}")
      if self._spec_sample is not None:     # the caller went on without the sample() call that was run ahead
        ss, self._spec_sample = self._spec_sample, None
        spec_fwd = self._spec
        self._rewind_sample(ss)
        self._spec = spec_fwd                # ... the forward enqueued on those ids is judged by _take_spec as usual
      state_pos = state_position(inference_state)
      req = self.requests.get(request_id)
      # a new request starts with a multi-token input at position 0 (encode() resets the
      # reference's state the same way, sharded_inference_engine.py:131-147)
      if req is None or (state_pos == 0) or (state_pos is None and s > 1) or len(req.slots) != bsz:
        req = self._start_request(request_id, bsz)
      pos = req.pos if state_pos is None else int(state_pos)
      sp = self._take_spec(request_id, pos, x) if self._spec is not None else None
      if x.ndim == 2:
        if req.tokens is None or s > 1:
          req.tokens = x.astype(np.int64).tolist()
        else:
          for row, t in zip(req.tokens, x[:, 0].tolist()):
            row.append(int(t))
      elif inference_state and inference_state.get("tokens") is not None:
        # a later shard: the history arrives with the state (ShardInferenceState.from_dict, llm_utils.py:489-495)
        # and is handed on unchanged
        req.tokens = inference_state["tokens"]
      if x.ndim == 2:   # the caller's ids always go to the device, also when the step below is already in flight
        dev_in = torch.as_tensor(np.ascontiguousarray(x, dtype=np.int32)).to(st.device, non_blocking=True)
      if sp is not None:   # this very step was enqueued by sample() on the same ids (checked on the host)
        return finish(req, sp["out"], sp["host"], sp["event"], bsz, s, pos)
      landing_in = None
      if ref is not None:
        dev_in, landing_in = self._rows_from_ref(st, ref, bsz, s)
      elif x.ndim == 3:
        dev_in = torch.as_tensor(np.ascontiguousarray(x, dtype=np.float32)).to(st.device, non_blocking=True).to(torch.bfloat16)
      keep = self._prompt_keep_mask(x, inference_state, bsz, s, pos)
      if keep is not None:
        out = self._forward_padded_prompt(st, req, dev_in, keep, x.ndim == 2)
      elif s == 1 and (landing_in is not None or (self.device_handoff and not shard.is_last_layer())) and bsz <= st.max_batch_tokens:
        return finish_hop(req, *self._forward_decode_hop(st, req, bsz, pos, dev_in, x.ndim == 2, landing_in, request_id), bsz, pos)
      else:
        out = self._forward_chunked(st, req, dev_in, bsz, s, pos, x.ndim == 2)
      if isinstance(out, tuple):
        return finish(req, out[0], out[1], out[2], bsz, s, pos)
      return finish(req, out, None, None, bsz, s, pos)

    def finish_hop(req, out, landing_out, bsz, pos):
      """A decode step that arrived through / leaves through the box-internal hand-off."""
      if shard.is_last_layer():
        return finish(req, out, None, None, bsz, 1, pos)
      req.pos = pos + 1
      return self._device_result(out, bsz, 1, landing_out), state_to_dict(req.tokens, pos, copy=False), None

    def finish(req, out, host, ev, bsz, s, pos):
      req.pos = pos + s
      new_state = state_to_dict(req.tokens, pos if not shard.is_last_layer() else pos + s,
                                copy=not (self.device_handoff or self._ring is not None))
      if shard.is_last_layer():
        if host is None:
          host, ev = self._logits_to_host_beside(out)
        self._remember_logits(host, out, request_id)
        self._speculate_sample(request_id, req, host, out)
        self._speculate_ring_sample(request_id, req, host, out)
        return host, new_state, ev
      if self.device_handoff:
        return self._device_result(out, bsz, s, None), new_state, None   # hidden rows stay in HBM (networking.DeviceTensorRef)
      host, ev = self._to_host_async(out.float().reshape(bsz, s, -1))    # fp32 [B, s, D]
      return host, new_state, ev

    # The engine's single worker only ENQUEUES (strictly serialised, like the reference's executor); waiting for the
    # GPU happens on another thread, so that with several requests in flight the worker can already enqueue the next
    # request's forward or sampler while this result is on its way to the host.
    loop = asyncio.get_running_loop()
    fut = loop.create_future()

    def settle(value, exc=None):
      if not fut.cancelled():
        fut.set_exception(exc) if exc is not None else fut.set_result(value)

    def wait_gpu(result, new_state, ev):
      try:
        ev.synchronize()
        trace("eng.gpu.done", request_id)
        loop.call_soon_threadsafe(settle, (result, new_state))
      except BaseException as e:
        loop.call_soon_threadsafe(settle, None, e)

    def job():     # worker -> (waiter) -> event loop: one wake-up of the loop per call
      try:
        result, new_state, ev = infer_wrapper()
        if ev is None or ev.query():
          loop.call_soon_threadsafe(settle, (result, new_state))
        else:
          self._waiter.submit(wait_gpu, result, new_state, ev)
      except BaseException as e:
        loop.call_soon_threadsafe(settle, None, e)

    self.executor.submit(job)
    return await fut

  # ------------------------------------------------------------------ device-resident token loop
  async def generate(self, request_id: str, shard: Shard, input_data: np.ndarray, inference_state: Optional[dict] = None, *,
                     max_generate_tokens: int = 1024, temp: float = TEMP, top_k: int = TOP_K, top_p: float = 0.0,
                     eos_token_id: Optional[int] = None, on_token=None, lookahead: int = 8) -> tuple[np.ndarray, dict]:
    """The ring walk of a one-node ring without leaving the device.

    In the reference every generated token costs ``process_inference_result`` -> ``sample`` ->
    ``create_task(forward_tensor)`` -> ``process_tensor`` -> ``infer_tensor`` (node.py:109-147,
    424-443,361-380): numpy logits out, numpy ids in, one asyncio task and one state dict per
    token.  Here sample -> forward -> sample is chained on the stream with the ids staying in
    device memory; each step's ids are also DMA-ed into a pinned table that the host drains
    ``lookahead`` steps behind the device, applying the reference's stop rule (eos, or
    ``max_generate_tokens`` buffered -- node.py:119,124) and firing ``on_token(request_id,
    [ids of the step], is_finished)`` on the event-loop thread like
    ``trigger_on_token_callbacks`` (node.py:136).  Up to ``lookahead`` steps computed past the
    stop are discarded: positions and the noise stream are rewound to the accepted prefix, so the
    ids and the state equal those of the step-by-step ``infer_tensor``/``sample`` calls.

    input_data: int ids [B, s] -- a prompt (new request) or the ids to continue a request with.
    Returns (ids [B, n] int32, state to continue from with ``infer_tensor(ids[:, -1:], state)``).
    With B > 1 generation stops when every row has produced eos (or at the cap)."""
    if not (shard.is_first_layer() and shard.is_last_layer()):
      raise ValueError("generate() needs the whole model on this engine (first and last layer); "
                       "a multi-stage ring keeps its token loop on the devices through PipelineRunner")
    await self.ensure_shard(shard)
    loop = asyncio.get_running_loop()

    def gen_wrapper():
      with torch.cuda.stream(self._stream):
        return gen_body()

    def gen_body():
      st = self.stage
      x = np.asarray(input_data)
      if x.ndim != 2:
        raise ValueError(f"generate() takes 2-D token ids, got ndim={x.ndim}")
      bsz, s = x.shape
      state_pos = state_position(inference_state)
      req = self.requests.get(request_id)
      if req is None or (state_pos == 0) or (state_pos is None and s > 1) or len(req.slots) != bsz:
        req = self._start_request(request_id, bsz)
      pos = req.pos if state_pos is None else int(state_pos)
      if self._spec_sample is not None:
        self._rewind_sample(self._spec_sample)
      self._spec = self._last_logits = self._spec_sample = None
      self._logits_dev.clear()
      if req.tokens is None or s > 1:
        req.tokens = x.astype(np.int64).tolist()
      else:
        for row, t in zip(req.tokens, x[:, 0].tolist()):
          row.append(int(t))
      dev_in = torch.as_tensor(np.ascontiguousarray(x, dtype=np.int32)).to(st.device, non_blocking=True)
      keep = self._prompt_keep_mask(x, inference_state, bsz, s, pos)
      if keep is not None:
        logits = self._forward_padded_prompt(st, req, dev_in, keep, True)
      else:
        logits = self._forward_chunked(st, req, dev_in, bsz, s, pos, True, to_host=False)
      pos += s
      # id n is drawn from the forward at position pos+n-2, which must exist in the KV slot
      cap = max(1, min(int(max_generate_tokens), st.max_tokens_per_seq - pos + 1))
      depth = max(1, int(lookahead))
      table, table_np = self._pinned_ids(cap, bsz)
      eos = self.tokenizer.eos_token_id if (eos_token_id is None and self.tokenizer is not None) else eos_token_id
      t_eff = max(float(temp), 1e-5)
      events, rewind = [], []
      done_rows = np.zeros(bsz, dtype=bool)
      n_enq = n_acc = 0
      stop = False
      while not stop:
        # keep the device up to `depth` steps ahead of what the host has looked at
        while n_enq < cap and n_enq - n_acc < depth:
          q = None
          if self.sample_mode == "reference":
            q = torch.empty((bsz, st.vocab), device=st.device).exponential_(1, generator=self.rng)
          self.sample_calls += 1
          toks = st.sample(logits, temp=t_eff, top_k=int(top_k or 0), top_p=top_p, q=q, seed=self.sample_calls)
          table[n_enq].copy_(toks, non_blocking=True)
          ev = torch.cuda.Event()
          ev.record(torch.cuda.current_stream(st.device))
          events.append(ev)
          rewind.append((self.rng.get_state() if q is not None else None, self.sample_calls))
          n_enq += 1
          if n_enq < cap:   # id number n_enq enters the KV slot at position pos + n_enq - 1
            logits = st.forward(req.slots, [1] * bsz, self._positions(req, pos + n_enq - 1), tokens=toks)
        events[n_acc].synchronize()
        ids = table_np[n_acc]
        n_acc += 1
        if eos is not None:
          done_rows |= (ids == eos)
        stop = bool(done_rows.all()) or n_acc >= cap
        if on_token is not None:
          loop.call_soon_threadsafe(on_token, request_id, ids.tolist(), stop)
      # forget what was computed past the stop: noise stream, call counter, positions
      rng_state, calls = rewind[n_acc - 1]
      if rng_state is not None:
        self.rng.set_state(rng_state)
      self.sample_calls = calls
      out_ids = np.ascontiguousarray(table_np[:n_acc].T, dtype=np.int32)      # [B, n]
      for row, fed in zip(req.tokens, out_ids[:, :n_acc - 1].tolist()):        # all but the last id went through the model
        row.extend(fed)
      req.pos = pos + n_acc - 1
      return out_ids, state_to_dict(req.tokens, req.pos)

    return await loop.run_in_executor(self.executor, gen_wrapper)

  def _pinned_ids(self, rows: int, bsz: int):
    """[rows, bsz] int32 pinned table the device loop of generate() writes its ids into."""
    t = getattr(self, "_ids_table", None)
    if t is None or t.shape[0] < rows or t.shape[1] != bsz:
      t = torch.empty((max(rows, 256), bsz), dtype=torch.int32).pin_memory()
      self._ids_table = t
    return t, t.numpy()
