"""One pipeline stage = one layer range of one model on one B200, over the C ABI.

``StageRuntime`` is the synchronous, device-tensor level under the async
``B200ShardInferenceEngine``: it owns the native engine handle, loads the shard's
weights (the work of ``load_model_weights_torchtune``,
``xotorch/inference/torch/models/llm_utils.py:136-284`` -- here only the tensors of
the layer range are read from the safetensors files), and runs forward / sample calls
on torch CUDA tensors without any numpy round trip.
"""
import ctypes as C
import json
from pathlib import Path
from typing import Dict, List, Optional, Sequence

import torch

from . import _cabi
from .model_config import has_qkv_bias, load_model_config, model_family, rope_table, uses_tied_head
from .shard import Shard


def _permute(t: torch.Tensor, n_heads: int, head_dim: int, dim: int) -> torch.Tensor:
  """HF half-split q/k rows -> torchtune interleaved rows (reference llm_utils.py:126-134)."""
  return t.view(n_heads, 2, head_dim // 2, dim).transpose(1, 2).reshape(head_dim * n_heads, dim)


class _SafetensorsIndex:
  """name -> file lookup over every ``*.safetensors`` in a directory, reading lazily."""

  def __init__(self, path: Path):
    from safetensors import safe_open
    self.files = sorted(Path(path).glob("*.safetensors"))
    if not self.files:
      raise FileNotFoundError(f"No safetensors files found in {path}")
    self._open = safe_open
    self.where: Dict[str, Path] = {}
    for f in self.files:
      with safe_open(str(f), framework="pt") as h:
        for k in h.keys():
          self.where[k] = f
    self._handles = {}

  def has(self, name: str) -> bool:
    return name in self.where

  def get(self, name: str) -> torch.Tensor:
    f = self.where[name]
    h = self._handles.get(f)
    if h is None:
      h = self._open(str(f), framework="pt")
      self._handles[f] = h
    return h.get_tensor(name)


class StageRuntime:
  def __init__(self, shard: Shard, config: dict, device: int = 0, max_batch_tokens: int = 4096,
               max_seqs: int = 32, max_tokens_per_seq: int = 4096, page_tokens: int = 64,
               rope_table_bf16: bool = True, rmsnorm_two_round: bool = True, max_positions: Optional[int] = None,
               first_from_mlp: bool = False, last_to_attn: bool = False):
    if not torch.cuda.is_available():
      raise RuntimeError("B200 engine needs a CUDA device: there is no CPU fallback")
    self.lib = _cabi.load()
    self.shard = shard
    self.config = config
    self.device = torch.device("cuda", device)
    self.family = model_family(shard.model_id)
    self.tied = uses_tied_head(shard.model_id)
    self.qkv_bias = has_qkv_bias(shard.model_id, config)
    self.max_seqs = max_seqs
    self.max_batch_tokens = max_batch_tokens
    self.max_tokens_per_seq = max_tokens_per_seq
    if max_positions is None:
      max_positions = min(config["max_seq_len"], max(max_tokens_per_seq, 8))
    self.max_positions = max_positions
    d = _cabi.model_desc(
      n_layers=shard.n_layers, start_layer=shard.start_layer, end_layer=shard.end_layer,
      dim=config["embed_dim"], n_heads=config["num_heads"], n_kv_heads=config["num_kv_heads"],
      head_dim=config["head_dim"], inter_dim=config["intermediate_dim"], vocab=config["vocab_size"],
      norm_eps=config["norm_eps"],
      rope_kind=_cabi.ROPE_HALF_SPLIT if self.family == "qwen" else _cabi.ROPE_INTERLEAVED,
      rmsnorm_two_round=int(rmsnorm_two_round), qkv_bias=int(self.qkv_bias), tied_head=int(self.tied),
      max_positions=max_positions, max_batch_tokens=max_batch_tokens, max_seqs=max_seqs,
      first_from_mlp=int(first_from_mlp), last_to_attn=int(last_to_attn))
    # half-layer stage boundaries (xb200.h): the attention block of start_layer / the MLP block of
    # end_layer may belong to the neighbouring pipeline stage
    self.first_from_mlp, self.last_to_attn = bool(first_from_mlp), bool(last_to_attn)
    self.is_first = shard.is_first_layer() and not self.first_from_mlp    # holds the embedding
    self.is_last = shard.is_last_layer() and not self.last_to_attn        # holds the final norm + LM head
    self.desc = d
    self._h = C.c_void_p()
    _cabi.check(self.lib.xb_engine_create(C.byref(d), device, C.byref(self._h)))
    tab = rope_table(self.family, config, max_positions, bf16_rounded=rope_table_bf16)
    self._load(-1, _cabi.W_ROPE_TABLE, tab)
    _cabi.check(self.lib.xb_kv_reserve(self._h, max_seqs, max_tokens_per_seq, page_tokens))
    self.page_tokens = page_tokens
    self.dim = config["embed_dim"]
    self.vocab = config["vocab_size"]
    # hand-off state: (slots, rows) of this stage's landing buffers, their IPC handles, whether the NEXT stage's buffers
    # are mapped, and the last flag value this stage stored into each of the next stage's slots
    self.ho_geometry: Optional[tuple] = None
    self.ho_handles: Optional[tuple] = None
    self.ho_connected = False
    self.ho_sent: Dict[int, int] = {}

  # -- lifetime
  def close(self):
    if getattr(self, "_h", None) is not None and self._h.value:
      self.lib.xb_engine_destroy(self._h)
      self._h = C.c_void_p()

  def __del__(self):
    try:
      self.close()
    except Exception:
      pass

  # -- weights
  def _load(self, layer: int, kind: int, t: torch.Tensor):
    t = t.contiguous()
    if kind != _cabi.W_ROPE_TABLE:
      t = t.to(torch.bfloat16)
    _cabi.check(self.lib.xb_load_weight(self._h, layer, kind, C.c_void_p(t.data_ptr()), t.numel() * t.element_size()))

  def load_hf_tensors(self, get, has):
    """Feed HF-named tensors (callable ``get(name)``) for this shard's layer range."""
    cfg, sh = self.config, self.shard
    H, KV, hd, D = cfg["num_heads"], cfg["num_kv_heads"], cfg["head_dim"], cfg["embed_dim"]
    llama = self.family == "llama"
    for l in range(sh.start_layer, sh.end_layer + 1):
      p = f"model.layers.{l}."
      has_attn = not (l == sh.start_layer and self.first_from_mlp)
      has_mlp = not (l == sh.end_layer and self.last_to_attn)
      if has_attn:
        q, k = get(p + "self_attn.q_proj.weight"), get(p + "self_attn.k_proj.weight")
        if llama:  # reference permutes q/k for Llama only (llm_utils.py:213-231)
          q, k = _permute(q, H, hd, D), _permute(k, KV, hd, D)
        self._load(l, _cabi.W_Q, q)
        self._load(l, _cabi.W_K, k)
        self._load(l, _cabi.W_V, get(p + "self_attn.v_proj.weight"))
        self._load(l, _cabi.W_O, get(p + "self_attn.o_proj.weight"))
        self._load(l, _cabi.W_SA_NORM, get(p + "input_layernorm.weight"))
      if has_mlp:
        self._load(l, _cabi.W_GATE, get(p + "mlp.gate_proj.weight"))
        self._load(l, _cabi.W_UP, get(p + "mlp.up_proj.weight"))
        self._load(l, _cabi.W_DOWN, get(p + "mlp.down_proj.weight"))
        self._load(l, _cabi.W_MLP_NORM, get(p + "post_attention_layernorm.weight"))
      if self.qkv_bias and has_attn:
        for nm, kind in (("q_proj", _cabi.W_Q_BIAS), ("k_proj", _cabi.W_K_BIAS), ("v_proj", _cabi.W_V_BIAS)):
          if has(p + f"self_attn.{nm}.bias"):
            self._load(l, kind, get(p + f"self_attn.{nm}.bias"))
    if self.is_first or (self.is_last and self.tied):
      self._load(-1, _cabi.W_EMBED, get("model.embed_tokens.weight"))
    if self.is_last:
      self._load(-1, _cabi.W_FINAL_NORM, get("model.norm.weight"))
      if not self.tied:
        self._load(-1, _cabi.W_LM_HEAD, get("lm_head.weight"))

  def load_dir(self, path: Path, native: bool = True) -> dict:
    """Load this shard's layer range from the ``*.safetensors`` files of a checkpoint directory.
    native=True (default): libxb200 maps each file and streams only this shard's tensors to the
    device (``xb_load_safetensors``); native=False: through the safetensors Python package and
    ``load_hf_tensors`` (same result, used to cross-check).  Returns {files, tensors, bytes, seconds}."""
    import time
    t0 = time.perf_counter()
    if not native:
      idx = _SafetensorsIndex(path)
      self.load_hf_tensors(idx.get, idx.has)
      return {"files": len(idx.files), "tensors": None, "bytes": None, "seconds": time.perf_counter() - t0}
    files = sorted(Path(path).glob("*.safetensors"))
    if not files:
      raise FileNotFoundError(f"No safetensors files found in {path}")
    tensors, nbytes = 0, 0
    for f in files:
      n, b = C.c_int(), C.c_uint64()
      _cabi.check(self.lib.xb_load_safetensors(self._h, str(f).encode(), int(self.family == "llama"), C.byref(n), C.byref(b)))
      tensors += n.value
      nbytes += b.value
    missing = self.weights_missing()
    if missing:
      raise RuntimeError(f"checkpoint {path} lacks weights of {self.shard}: {missing}")
    return {"files": len(files), "tensors": tensors, "bytes": nbytes, "seconds": time.perf_counter() - t0}

  def weights_missing(self) -> str:
    """Readable list of this shard's weights that have not been supplied yet ('' = complete)."""
    buf, n = C.create_string_buffer(1024), C.c_int()
    _cabi.check(self.lib.xb_weights_missing(self._h, buf, 1024, C.byref(n)))
    return f"{n.value} missing: {buf.value.decode()}" if n.value else ""

  def load_state(self, state: Dict[str, torch.Tensor]):
    self.load_hf_tensors(state.__getitem__, state.__contains__)

  # -- compute
  def forward(self, slot_ids: Sequence[int], q_lens: Sequence[int], positions: Sequence[int],
              tokens: Optional[torch.Tensor] = None, hidden: Optional[torch.Tensor] = None,
              out: Optional[torch.Tensor] = None, taps: Optional[List[torch.Tensor]] = None,
              handoff: Optional[tuple] = None, host_logits: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
    """tokens: int32 [T] (first shard) / hidden: bf16 [T, D].  Returns bf16 [T, D] hidden
    or fp32 [n_seqs, V] logits (last shard).  Enqueued on the current torch stream.
    host_logits (last shard): pinned fp32 [n_seqs, V] CPU tensor the LM-head kernel also stores the
    logits into (xb_forward_host_logits); valid once the stream has reached this point."""
    n = len(slot_ids)
    T = int(sum(q_lens))
    arr = (C.c_int32 * n)
    b = _cabi.Batch(n_seqs=n, slot_ids=arr(*slot_ids), q_lens=arr(*q_lens), positions=arr(*positions),
                    tokens=None, hidden_in=None)
    if tokens is not None:
      assert tokens.dtype == torch.int32 and tokens.is_cuda and tokens.numel() == T
      b.tokens = tokens.data_ptr()
    if hidden is not None:
      assert hidden.dtype == torch.bfloat16 and hidden.is_cuda and hidden.numel() == T * self.dim and hidden.is_contiguous()
      b.hidden_in = hidden.data_ptr()
    sends = handoff is not None and handoff[2] >= 0
    if out is None and not sends:
      if self.is_last:
        out = torch.empty((n, self.vocab), dtype=torch.float32, device=self.device)
      else:
        out = torch.empty((T, self.dim), dtype=torch.bfloat16, device=self.device)
    tap_ptrs = None
    if taps is not None:
      tap_ptrs = (C.c_void_p * len(taps))(*[t.data_ptr() if t is not None else None for t in taps])   # None: layer not tapped
    stream = torch.cuda.current_stream(self.device).cuda_stream
    optr = C.c_void_p(out.data_ptr()) if out is not None else None
    if handoff is not None:
      # (slot_in, wait_value, slot_out, signal_value): input from / output to the NVLink landing buffers
      _cabi.check(self.lib.xb_forward_handoff(self._h, C.byref(b), int(handoff[0]), int(handoff[1]), int(handoff[2]),
                                              int(handoff[3]), optr, C.c_void_p(stream)))
      if sends:
        self.ho_sent[int(handoff[2])] = int(handoff[3])   # flag value of the NEXT stage's slot once this step has run
      return None if sends else out
    if host_logits is not None:
      # (page-locked, device-addressable memory is checked by the library: Tensor.is_pinned() costs ~30 us here)
      assert taps is None and host_logits.dtype == torch.float32 and host_logits.is_contiguous() \
        and host_logits.numel() >= n * self.vocab
      _cabi.check(self.lib.xb_forward_host_logits(self._h, C.byref(b), optr, C.c_void_p(host_logits.data_ptr()), C.c_void_p(stream)))
      return out
    _cabi.check(self.lib.xb_forward(self._h, C.byref(b), optr, tap_ptrs, C.c_void_p(stream)))
    return out

  def handoff_create(self, n_slots: int, rows: int, ipc: bool = True) -> Optional[tuple]:
    """Allocate this stage's landing buffers; returns the two CUDA IPC handles (bytes) for stages in
    other processes, or None with ipc=False (stages of one process connect with handoff_connect_local)."""
    self.ho_geometry = (int(n_slots), int(rows))
    if not ipc:
      _cabi.check(self.lib.xb_handoff_create(self._h, int(n_slots), int(rows), None, None))
      return None
    hl, hf = C.create_string_buffer(64), C.create_string_buffer(64)
    _cabi.check(self.lib.xb_handoff_create(self._h, int(n_slots), int(rows), hl, hf))
    self.ho_handles = (bytes(hl.raw), bytes(hf.raw))
    return self.ho_handles

  def handoff_connect_local(self, next_stage: "StageRuntime"):
    """The NEXT stage lives in this process (another GPU, or the same one): store into its landing
    buffers through their device addresses (xb_handoff_connect_local)."""
    _cabi.check(self.lib.xb_handoff_connect_local(self._h, next_stage._h))
    self._next_stage = next_stage   # keeps the consumer's buffers alive as long as this producer

  def reserve_sms(self, n: int):
    """Keep n SMs free of the decode Linears (xb_reserve_sms): for stages with a resident communication kernel."""
    _cabi.check(self.lib.xb_reserve_sms(self._h, int(n)))

  def handoff_clear_error(self):
    _cabi.check(self.lib.xb_handoff_clear_error(self._h))

  def handoff_connect(self, landing_handle: bytes, flag_handle: bytes):
    """Map the NEXT stage's landing buffers (handles from its handoff_create)."""
    hl, hf = C.create_string_buffer(landing_handle, 64), C.create_string_buffer(flag_handle, 64)
    _cabi.check(self.lib.xb_handoff_connect(self._h, hl, hf))
    self.ho_connected = True

  def sample(self, logits: torch.Tensor, temp: float = 0.0, top_k: int = 0, top_p: float = 0.0,
             q: Optional[torch.Tensor] = None, seed: int = 0, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    assert logits.dtype == torch.float32 and logits.is_cuda and logits.is_contiguous()
    n = logits.shape[0]
    if out is None:
      out = torch.empty((n,), dtype=torch.int32, device=self.device)
    stream = torch.cuda.current_stream(self.device).cuda_stream
    _cabi.check(self.lib.xb_sample(self._h, C.c_void_p(logits.data_ptr()), n, float(temp), int(top_k), float(top_p),
                                   C.c_void_p(q.data_ptr()) if q is not None else None, int(seed),
                                   C.c_void_p(out.data_ptr()), C.c_void_p(stream)))
    return out

  def bench_linears(self, T: int, iters: int) -> int:
    """Enqueue the Linear kernels of one forward `iters` times (xb_bench_linears)."""
    n = C.c_int()
    stream = torch.cuda.current_stream(self.device).cuda_stream
    _cabi.check(self.lib.xb_bench_linears(self._h, int(T), int(iters), C.byref(n), C.c_void_p(stream)))
    return n.value

  def bench_attention(self, n_seqs: int, max_ctx: int, n_splits: int, iters: int) -> int:
    """Enqueue the decode attention of the first local layer `iters` times (xb_bench_attention)."""
    used = C.c_int()
    stream = torch.cuda.current_stream(self.device).cuda_stream
    _cabi.check(self.lib.xb_bench_attention(self._h, int(n_seqs), int(max_ctx), int(n_splits), int(iters), C.byref(used),
                                            C.c_void_p(stream)))
    return used.value

  def timeline(self):
    """Stamps of the Linear launches since the last call (engine created with XB_GEMM_TS=1): uint64 numpy
    array [n_launches, blocks, 16] of globaltimer nanoseconds (0 = stamp not reached)."""
    import numpy as np
    n, blocks = C.c_int(), C.c_int()
    slots = 16
    cap = 1024 * 160 * slots
    buf = np.zeros(cap, dtype=np.uint64)
    _cabi.check(self.lib.xb_debug_timeline(self._h, buf.ctypes.data_as(C.c_void_p), cap, C.byref(n), C.byref(blocks)))
    return buf[: n.value * blocks.value * slots].reshape(n.value, blocks.value, slots)

  def release_slot(self, slot: int):
    _cabi.check(self.lib.xb_kv_release_slot(self._h, slot))

  def slot_len(self, slot: int) -> int:
    v = C.c_int()
    _cabi.check(self.lib.xb_kv_slot_len(self._h, slot, C.byref(v)))
    return v.value

  def counters(self) -> dict:
    c = _cabi.Counters()
    _cabi.check(self.lib.xb_get_counters(self._h, C.byref(c)))
    return {k: int(getattr(c, k)) for k, _ in c._fields_}


def stage_from_dir(path: Path, shard: Shard, **kw) -> StageRuntime:
  cfg = load_model_config(Path(path) / "config.json")
  st = StageRuntime(shard, cfg, **kw)
  st.load_dir(Path(path))
  return st


def debug_linear(x: torch.Tensor, w: torch.Tensor, use_tensor_cores: bool = True, force_splits: int = 0,
                 round_bf16: bool = False) -> torch.Tensor:
  """Kernel-level hook (xb_debug_linear): fp32 [T, N] = x[T, K] @ w[N, K]^T."""
  lib = _cabi.load()
  T, K = x.shape
  N = w.shape[0]
  y = torch.empty((T, N), dtype=torch.float32, device=x.device)
  scratch = torch.empty((16 * T * N,), dtype=torch.float32, device=x.device)
  stream = torch.cuda.current_stream(x.device).cuda_stream
  _cabi.check(lib.xb_debug_linear(C.c_void_p(x.data_ptr()), C.c_void_p(w.data_ptr()), T, N, K, C.c_void_p(y.data_ptr()),
                                  C.c_void_p(scratch.data_ptr()), int(use_tensor_cores), int(force_splits), int(round_bf16),
                                  C.c_void_p(stream)))
  return y


def linear_noalloc(x: torch.Tensor, w: torch.Tensor, y: torch.Tensor, round_bf16: bool = False) -> torch.Tensor:
  """xb_linear_noalloc: fp32 ``y[T, N] = x[T, K] @ w[N, K]^T`` into the CALLER's ``y``.

  For callers that must stay away from GPU allocators while they run: a decode loop where a
  cudaMalloc/cudaFree (or a framework allocator growing its pool) would stall every stream of
  the device, or a process whose HBM is already fully committed to weights and KV pages.  This
  wrapper therefore creates no tensor either -- inputs and the output come from the caller and
  the split-K workspace is an 8 MiB array inside libxb200.  T <= 128; shapes whose workspace
  does not fit are rejected with a RuntimeError naming the required size.  Calls share that one
  array: do not let two of them overlap on the GPU."""
  lib = _cabi.load()
  T, K = x.shape
  N = w.shape[0]
  assert x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and y.dtype == torch.float32
  assert y.shape == (T, N) and x.is_contiguous() and w.is_contiguous() and y.is_contiguous()
  assert w.device == x.device and y.device == x.device
  with torch.cuda.device(x.device):   # the library uses the CURRENT device's copy of its built-in workspace
    stream = torch.cuda.current_stream(x.device).cuda_stream
    _cabi.check(lib.xb_linear_noalloc(C.c_void_p(x.data_ptr()), C.c_void_p(w.data_ptr()), T, N, K, C.c_void_p(y.data_ptr()),
                                      int(round_bf16), C.c_void_p(stream)))
  return y
