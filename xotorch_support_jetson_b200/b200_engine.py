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
* out-of-memory and every other native failure raise (the reference returns ``None``).
"""
import asyncio
import functools
import warnings
import os
import uuid
from concurrent.futures import ThreadPoolExecutor
from typing import Dict, List, Optional

import numpy as np
import torch

from .inference_engine import InferenceEngine
from .model_config import load_model_config
from .shard import Shard
from .stage import StageRuntime
from .synth import SyntheticTokenizer

# results are handed out as read-only numpy views of pinned buffers; feeding one back in is fine
warnings.filterwarnings("ignore", message="The given NumPy array is not writable")

# from the reference (sharded_inference_engine.py:34-35)
TEMP = 0.6
TOP_K = 35


class _Request:
  __slots__ = ("slots", "pos", "tokens")

  def __init__(self, slots: List[int]):
    self.slots = slots
    self.pos = 0
    self.tokens: Optional[np.ndarray] = None


class B200ShardInferenceEngine(InferenceEngine):
  def __init__(self, shard_downloader, device: Optional[int] = None, max_seqs: Optional[int] = None,
               max_tokens_per_seq: Optional[int] = None, max_batch_tokens: Optional[int] = None,
               page_tokens: Optional[int] = None, sample_mode: str = "reference"):
    self.shard = None
    self.shard_downloader = shard_downloader
    self.stage: Optional[StageRuntime] = None
    self.tokenizer = None
    self.model_path = None
    self.model_config = None
    self.uuid = str(uuid.uuid4())
    self.executor = ThreadPoolExecutor(max_workers=1)  # strictly serialised like the reference (:46)
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
    # they land in a small ring of PINNED buffers (a numpy view is returned, valid until the
    # ring wraps) and the device copy is kept so that sample() on that very array skips the upload
    self._pin_rings: Dict[torch.dtype, list] = {}   # dtype -> [buffers, next index]
    self._last_logits = None   # (numpy view handed to the caller, device tensor it mirrors)

  # ------------------------------------------------------------------ shard / weights
  async def ensure_shard(self, shard: Shard):
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
      self.requests.clear()
      self.free_slots = list(range(self.max_seqs - 1, -1, -1))
      self.rng = torch.Generator(device=st.device)
      self.rng.manual_seed(1234)
      self._stream = torch.cuda.Stream(device=st.device)  # non-default stream: decode steps replay CUDA graphs

    await asyncio.get_running_loop().run_in_executor(self.executor, start_model)
    self.shard = shard

  def attach_stage(self, shard: Shard, stage: StageRuntime, config: dict, tokenizer=None):
    """Adopt an already-loaded StageRuntime (weights resident in HBM) instead of loading the
    shard from a checkpoint directory -- in-memory / synthetic weights, benchmarks."""
    self.stage, self.shard, self.model_config = stage, shard, config
    self.max_seqs, self.max_batch_tokens = stage.max_seqs, stage.max_batch_tokens
    self.tokenizer = tokenizer or SyntheticTokenizer(config["vocab_size"])
    self.requests.clear()
    self.free_slots = list(range(self.max_seqs - 1, -1, -1))
    self.rng = torch.Generator(device=stage.device)
    self.rng.manual_seed(1234)
    self._stream = torch.cuda.Stream(device=stage.device)

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
  PIN_RING = 4

  def _to_host(self, dev: torch.Tensor) -> np.ndarray:
    """Device tensor -> numpy through a pinned ring buffer (one DMA, no pageable bounce).  The
    returned array is a read-only view that stays valid for the next PIN_RING-1 results."""
    n = dev.numel()
    ring = self._pin_rings.get(dev.dtype)
    if ring is None or ring[0][0].numel() < n:
      ring = [[torch.empty(max(n, 1), dtype=dev.dtype).pin_memory() for _ in range(self.PIN_RING)], 0]
      self._pin_rings[dev.dtype] = ring
    buf = ring[0][ring[1]][:n].view(dev.shape)
    ring[1] = (ring[1] + 1) % self.PIN_RING
    buf.copy_(dev, non_blocking=True)
    torch.cuda.current_stream(dev.device).synchronize()
    arr = buf.numpy()
    arr.setflags(write=False)
    return arr

  # ------------------------------------------------------------------ sampling
  async def sample(self, x: np.ndarray, temp=TEMP, top_k=TOP_K, top_p: float = 0.0) -> np.ndarray:
    def sample_wrapper():
      with torch.cuda.stream(self._stream):
        return sample_body()

    def sample_body():
      st = self.stage
      if self._last_logits is not None and x is self._last_logits[0]:
        logits = self._last_logits[1]      # the array we just returned (read-only): still on the device
      else:
        logits = torch.as_tensor(np.ascontiguousarray(x, dtype=np.float32)).to(st.device)
      if logits.ndim == 1:
        logits = logits.unsqueeze(0)
      q = None
      if self.sample_mode == "reference":
        # the reference draws q ~ Exp(1) of shape [B, vocab] on EVERY call, also at temp 0
        q = torch.empty((logits.size(0), st.vocab), device=st.device).exponential_(1, generator=self.rng)
      self.sample_calls += 1
      # reference semantics at any temperature: logits / max(temp, 1e-5) -> top-k -> softmax -> argmax(p/q)
      toks = st.sample(logits, temp=max(float(temp), 1e-5), top_k=int(top_k or 0), top_p=top_p, q=q, seed=self.sample_calls)
      return np.array(self._to_host(toks), dtype=np.int32).reshape(-1, 1)

    return await asyncio.get_running_loop().run_in_executor(self.executor, sample_wrapper)

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
    if req is not None:
      for s in req.slots:
        self.stage.release_slot(s)
        self.free_slots.append(s)

  def _forward_chunked(self, st: StageRuntime, req: _Request, dev_in: torch.Tensor, bsz: int, s: int, pos: int,
                       is_tokens: bool) -> torch.Tensor:
    """One logical forward over [bsz, s] rows, split so that no native call carries more
    than max_batch_tokens rows: whole sequences per call, or position chunks of one
    sequence (chunked prefill) when a single sequence is longer than that."""
    cap = st.max_batch_tokens
    last = self.shard.is_last_layer()
    if bsz * s <= cap:
      flat = dev_in.reshape(-1) if is_tokens else dev_in.reshape(bsz * s, -1)
      return st.forward(req.slots, [s] * bsz, [pos] * bsz, **({"tokens": flat} if is_tokens else {"hidden": flat}))
    outs = []
    if s <= cap:
      per = max(1, cap // s)
      for b0 in range(0, bsz, per):
        b1 = min(bsz, b0 + per)
        part = dev_in[b0:b1]
        flat = part.reshape(-1) if is_tokens else part.reshape((b1 - b0) * s, -1)
        outs.append(st.forward(req.slots[b0:b1], [s] * (b1 - b0), [pos] * (b1 - b0),
                               **({"tokens": flat} if is_tokens else {"hidden": flat})))
      return torch.cat(outs, dim=0)
    for b in range(bsz):
      seq_out = []
      for c0 in range(0, s, cap):
        c1 = min(s, c0 + cap)
        part = dev_in[b, c0:c1]
        flat = part.reshape(-1) if is_tokens else part.reshape(c1 - c0, -1)
        o = st.forward([req.slots[b]], [c1 - c0], [pos + c0], **({"tokens": flat} if is_tokens else {"hidden": flat}))
        seq_out.append(o)
      outs.append(seq_out[-1] if last else torch.cat(seq_out, dim=0))
    return torch.cat(outs, dim=0)

  async def infer_tensor(self, request_id: str, shard: Shard, input_data: np.ndarray,
                         inference_state: Optional[dict] = None) -> tuple[np.ndarray, Optional[dict]]:
    await self.ensure_shard(shard)

    def infer_wrapper():
      with torch.cuda.stream(self._stream):
        return infer_body()

    def infer_body():
      st = self.stage
      x = np.asarray(input_data)
      if x.ndim == 2:
        bsz, s = x.shape
      elif x.ndim == 3:
        bsz, s = x.shape[0], x.shape[1]
      else:
        raise ValueError(f"input_data must be 2-D token ids or 3-D hidden states, got ndim={x.ndim}")
      state_pos = None if not inference_state else inference_state.get("curr_pos")
      req = self.requests.get(request_id)
      # a new request starts with a multi-token input at position 0 (encode() resets the
      # reference's state the same way, sharded_inference_engine.py:131-147)
      if req is None or (state_pos == 0) or (state_pos is None and s > 1) or len(req.slots) != bsz:
        req = self._start_request(request_id, bsz)
      pos = req.pos if state_pos is None else int(state_pos)
      if x.ndim == 2:
        dev_in = torch.as_tensor(np.ascontiguousarray(x, dtype=np.int32)).to(st.device, non_blocking=True)
        req.tokens = x.astype(np.int64) if (req.tokens is None or s > 1) else np.concatenate([req.tokens, x.astype(np.int64)], axis=-1)
      else:
        dev_in = torch.as_tensor(np.ascontiguousarray(x, dtype=np.float32)).to(st.device, non_blocking=True).to(torch.bfloat16)
      out = self._forward_chunked(st, req, dev_in, bsz, s, pos, x.ndim == 2)
      req.pos = pos + s
      new_state = {"curr_pos": pos if not shard.is_last_layer() else pos + s}
      if req.tokens is not None:
        new_state["tokens"] = req.tokens.tolist()
      if shard.is_last_layer():
        host = self._to_host(out)                      # fp32 [B, V]
        self._last_logits = (host, out)
        return host, new_state
      return self._to_host(out.float().reshape(bsz, s, -1)), new_state  # fp32 [B, s, D]

    return await asyncio.get_running_loop().run_in_executor(self.executor, infer_wrapper)
