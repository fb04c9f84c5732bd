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

    await asyncio.get_running_loop().run_in_executor(self.executor, start_model)
    self.shard = shard

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

  # ------------------------------------------------------------------ sampling
  async def sample(self, x: np.ndarray, temp=TEMP, top_k=TOP_K, top_p: float = 0.0) -> np.ndarray:
    def sample_wrapper():
      st = self.stage
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
      return toks.cpu().numpy().astype(np.int32).reshape(-1, 1)

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

  async def infer_tensor(self, request_id: str, shard: Shard, input_data: np.ndarray,
                         inference_state: Optional[dict] = None) -> tuple[np.ndarray, Optional[dict]]:
    await self.ensure_shard(shard)

    def infer_wrapper():
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
        toks = torch.as_tensor(np.ascontiguousarray(x, dtype=np.int32)).to(st.device).reshape(-1)
        out = st.forward(req.slots, [s] * bsz, [pos] * bsz, tokens=toks)
        req.tokens = x.astype(np.int64) if (req.tokens is None or s > 1) else np.concatenate([req.tokens, x.astype(np.int64)], axis=-1)
      else:
        hid = torch.as_tensor(np.ascontiguousarray(x, dtype=np.float32)).to(st.device).to(torch.bfloat16).reshape(bsz * s, -1)
        out = st.forward(req.slots, [s] * bsz, [pos] * bsz, hidden=hid)
      req.pos = pos + s
      new_state = {"curr_pos": pos if not shard.is_last_layer() else pos + s}
      if req.tokens is not None:
        new_state["tokens"] = req.tokens.tolist()
      if shard.is_last_layer():
        return out.cpu().numpy(), new_state            # fp32 [B, V]
      return out.float().reshape(bsz, s, -1).cpu().numpy(), new_state  # fp32 [B, s, D]

    return await asyncio.get_running_loop().run_in_executor(self.executor, infer_wrapper)
