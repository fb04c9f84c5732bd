"""Layer-range pipeline over the GPUs of one box: one process (rank) per GPU.

The reference walks its ring of peers once per generated token, shipping the hidden
state as fp32 numpy + a JSON state blob over gRPC (``Node.forward_tensor``,
``xotorch/orchestration/node.py:424-443``; ``GRPCPeerHandle.send_tensor``,
``xotorch/networking/grpc/grpc_peer_handle.py:117-136``) and the sampled token back from
the last shard to the first (``node.py:126,145``), strictly one request in flight.

Here the stages are the layer ranges the reference's own ring strategy assigns to N
equal peers (``topology.pipeline_shards``); the hand-off is a bf16 ``[B, dim]`` /
``[T, dim]`` NCCL send/recv on the stage's stream, the return edge an int32 ``[B]`` send,
and N micro-batches are kept in flight so every stage is busy in steady state.  Each
stage reads its weight slice once per micro-batch.
"""
import os
import zlib
from typing import Callable, Dict, List, Optional

import torch

from .model_config import load_model_config  # noqa: F401  (re-exported for callers)
from .shard import Shard
from .stage import StageRuntime
from .synth import hf_param_shapes
from .topology import half_layer_stages, pipeline_shards


def gpu_random_tensor(name: str, shape, seed: int, std: float, device, dtype=torch.bfloat16) -> torch.Tensor:
  """Seeded random-init tensor generated ON the GPU (norm scales = 1): the same (seed,
  name) gives the same values on every rank, so a layer-range subset equals the same rows
  of the full model."""
  if name.endswith("layernorm.weight") or name == "model.norm.weight":
    return torch.ones(shape, dtype=dtype, device=device)
  g = torch.Generator(device=device).manual_seed((seed * 1000003 + zlib.crc32(name.encode())) & 0x7FFFFFFF)
  t = torch.empty(shape, dtype=torch.float32, device=device).normal_(0.0, std, generator=g)
  return t.to(dtype)


def gpu_random_getter(hf_cfg: dict, seed: int, std: float, device) -> tuple:
  shapes = hf_param_shapes(hf_cfg)

  def get(name: str) -> torch.Tensor:
    return gpu_random_tensor(name, shapes[name], seed, std, device)

  def has(name: str) -> bool:
    return name in shapes

  return get, has


class PipelineRunner:
  """One rank of an N-stage pipeline.  ``dist`` is ``torch.distributed`` (already
  initialised with the NCCL backend for world > 1) or None for a single stage."""

  def __init__(self, model_id: str, hf_cfg: dict, config: dict, rank: int, world: int, device: int, batch: int,
               micro_batches: int, max_ctx: int, max_batch_tokens: int = 8192, page_tokens: int = 64,
               seed: int = 0, std: float = 0.02, dist=None, weights: Optional[tuple] = None,
               stage_factory: Optional[Callable] = None, ret_group=None,
               balance_head: bool = True, half_layers: Optional[bool] = None):
    """``stage_factory(shard) -> stage`` lets the host-side protocol (ordering of the
    send/recv pairs, micro-batching, token return edge) be exercised without a GPU (the
    world_size-2 gloo tests plug the CPU oracle in; the product never does)."""
    self.rank, self.world, self.dist = rank, world, dist
    self.ret_group = ret_group   # separate communicator for the last->first token edge
    self.B, self.M = batch, micro_batches
    self.tok_work = [None] * micro_batches
    n_layers = config["num_layers"]
    # the last stage also streams the LM head every step: weigh it in layer units
    D, I = config["embed_dim"], config["intermediate_dim"]
    layer_params = 2 * D * config["num_heads"] * config["head_dim"] + 2 * D * config["num_kv_heads"] * config["head_dim"] + 3 * D * I
    head_equiv = config["vocab_size"] * D / layer_params if balance_head else 0.0
    # stage boundaries: half-layer granularity by bytes streamed per decode step (weights + the KV cache
    # of one micro-batch at the expected context), or the reference's whole-layer ring strategy
    self.first_from_mlp = self.last_to_attn = False
    if half_layers is None:
      half_layers = os.environ.get("XB_HALF_LAYERS", "1") != "0"
    if half_layers and world > 1 and stage_factory is None:
      H, KV, hd = config["num_heads"], config["num_kv_heads"], config["head_dim"]
      attn_cost = 2.0 * (D * H * hd * 2 + 2 * D * KV * hd) + batch * max_ctx * (2 * KV * hd * 2)
      mlp_cost = 2.0 * 3 * D * I
      head_cost = 2.0 * config["vocab_size"] * D if balance_head else 0.0
      self.shard, self.first_from_mlp, self.last_to_attn = half_layer_stages(model_id, n_layers, world, attn_cost, mlp_cost,
                                                                              head_cost)[rank]
    else:
      self.shard: Shard = pipeline_shards(model_id, n_layers, world, head_equiv)[rank]
    self.first, self.last = self.shard.is_first_layer(), self.shard.is_last_layer()
    if self.first_from_mlp:
      self.first = False
    if self.last_to_attn:
      self.last = False
    if stage_factory is not None:
      self.stage = stage_factory(self.shard)
      self.device = torch.device("cpu")
      self.stream = None
    else:
      self.device = torch.device("cuda", device)
      self.stage = StageRuntime(self.shard, config, device=device, max_batch_tokens=max_batch_tokens,
                                max_seqs=batch * micro_batches, max_tokens_per_seq=max_ctx, page_tokens=page_tokens,
                                first_from_mlp=self.first_from_mlp, last_to_attn=self.last_to_attn)
      get, has = weights if weights is not None else gpu_random_getter(hf_cfg, seed, std, self.device)
      self.stage.load_hf_tensors(get, has)
      self.stream = torch.cuda.Stream(device=self.device)
    self.dim, self.vocab = config["embed_dim"], config["vocab_size"]
    self.max_batch_tokens = max_batch_tokens
    # decode hand-off over NVLink peer memory (fused into the stages' last / first kernels);
    # prefill chunks and the token return edge stay on NCCL.  XB_PEER_HANDOFF=0 forces NCCL.
    self.peer = False
    self.ho_step = [0] * micro_batches      # per-slot flag value of the last hand-off
    if world > 1 and stage_factory is None and batch <= 128 and os.environ.get("XB_PEER_HANDOFF", "1") != "0":
      handles = self.stage.handoff_create(micro_batches, batch)
      gathered = [None] * world
      dist.all_gather_object(gathered, handles)
      if rank + 1 < world:
        self.stage.handoff_connect(*gathered[rank + 1])
      self.peer = True
    B, M = batch, micro_batches
    self.tok = [torch.zeros(B, dtype=torch.int32, device=self.device) for _ in range(M)]
    self.h_in = torch.empty((max_batch_tokens, self.dim), dtype=torch.bfloat16, device=self.device)
    self.h_out = torch.empty((max_batch_tokens, self.dim), dtype=torch.bfloat16, device=self.device)
    self.logits = torch.empty((B, self.vocab), dtype=torch.float32, device=self.device) if self.last else None
    self.pos = [0] * M  # context length of each micro-batch (all its sequences share it)

  def _on_stream(self):
    import contextlib
    return torch.cuda.stream(self.stream) if self.stream is not None else contextlib.nullcontext()

  def slots(self, m: int) -> List[int]:
    return list(range(m * self.B, (m + 1) * self.B))

  def reset(self):
    for s in range(self.B * self.M):
      self.stage.release_slot(s)
    self.pos = [0] * self.M

  # ------------------------------------------------------------------ comms
  # Forward edge (hidden states, stage i -> i+1): blocking send/recv in micro-batch order on the
  # default group -- an acyclic chain.  Return edge (sampled ids, last -> first stage): async
  # isend/irecv on its OWN group, posted ahead of use.  With both edges on one communicator the
  # last stage's "send ids(m)" and the first stage's "send hidden(m+1)" wait on each other
  # (p2p ops rendezvous), which deadlocks once more than one micro-batch is in flight.
  def _recv(self, t: torch.Tensor, src: int):
    self.dist.recv(t, src=src)

  def _send(self, t: torch.Tensor, dst: int):
    self.dist.send(t, dst=dst)

  def _post_tok_recv(self, m: int):
    self.tok_work[m] = self.dist.irecv(self.tok[m], src=self.world - 1, group=self.ret_group)

  def _post_tok_send(self, m: int):
    self.tok_work[m] = self.dist.isend(self.tok[m], dst=0, group=self.ret_group)

  def _wait_tok(self, m: int):
    w = self.tok_work[m]
    if w is not None:
      w.wait()
      self.tok_work[m] = None

  # ------------------------------------------------------------------ prefill
  def prefill(self, prompts: List[torch.Tensor]):
    """prompts[m]: int32 [B, L] on this rank's device (only rank 0 reads the values).
    Runs every micro-batch through the pipeline in chunks of whole sequences and leaves the
    first sampled token of each sequence in ``self.tok[m]`` on rank 0 (sent back by the last
    stage like the reference's last->first ``send_tensor``)."""
    B = self.B
    with self._on_stream():
      for m in range(self.M):
        L = prompts[m].shape[1]
        per = max(1, self.max_batch_tokens // L)
        if self.last and self.world > 1:
          self._wait_tok(m)
        for c0 in range(0, B, per):
          c1 = min(B, c0 + per)
          n, T = c1 - c0, (c1 - c0) * L
          slots = self.slots(m)[c0:c1]
          hid_in = None
          if not self.first:
            hid_in = self.h_in[:T]
            self._recv(hid_in, self.rank - 1)
          if self.last:
            lg = self.stage.forward(slots, [L] * n, [0] * n,
                                    tokens=prompts[m][c0:c1].reshape(-1) if self.first else None, hidden=hid_in,
                                    out=self.logits[:n])
            self.stage.sample(lg, temp=0.0, out=self.tok[m][c0:c1])
          else:
            out = self.stage.forward(slots, [L] * n, [0] * n,
                                     tokens=prompts[m][c0:c1].reshape(-1) if self.first else None, hidden=hid_in,
                                     out=self.h_out[:T])
            self._send(out, self.rank + 1)
        self.pos[m] = L
        if self.world > 1:
          if self.last:
            self._post_tok_send(m)
          elif self.first:
            self._post_tok_recv(m)

  # ------------------------------------------------------------------ decode
  def decode(self, steps: int, host_tokens: Optional[List[torch.Tensor]] = None, host_out: Optional[List[torch.Tensor]] = None):
    """`steps` decode steps for every micro-batch, greedy sampling on the last stage.
    With ``host_tokens`` / ``host_out`` (pinned CPU tensors) rank 0 uploads each step's input
    ids from host memory and the last rank downloads each step's sampled ids (end-to-end
    mode); otherwise ids stay on the devices."""
    B, M = self.B, self.M
    ones = [1] * B
    with self._on_stream():
      for k in range(steps):
        for m in range(M):
          pos = [self.pos[m]] * B
          slots = self.slots(m)
          if self.first:
            if self.world > 1:
              self._wait_tok(m)            # ids sampled by the last stage for this micro-batch
            if host_tokens is not None:
              # the caller's ids live in (pinned) host memory: read back, then upload as this
              # step's input -- what a host-driven caller pays every step
              host_tokens[m].copy_(self.tok[m], non_blocking=True)
              if self.stream is not None:
                self.stream.synchronize()
              self.tok[m].copy_(host_tokens[m], non_blocking=True)
          if self.first and self.last:
            lg = self.stage.forward(slots, ones, pos, tokens=self.tok[m], out=self.logits)
            self.stage.sample(lg, temp=0.0, out=self.tok[m])
            if host_out is not None:
              host_out[m].copy_(self.tok[m], non_blocking=True)
          elif self.peer:
            # hidden rows travel GPU -> GPU inside the stages' own kernels (no send/recv launch)
            self.ho_step[m] += 1
            v = self.ho_step[m]
            ho = (-1 if self.first else m, v, -1 if self.last else m, v)
            if self.last:
              lg = self.stage.forward(slots, ones, pos, out=self.logits, handoff=ho)
              self._wait_tok(m)            # previous send of this buffer has completed
              self.stage.sample(lg, temp=0.0, out=self.tok[m])
              if host_out is not None:
                host_out[m].copy_(self.tok[m], non_blocking=True)
              self._post_tok_send(m)
            else:
              self.stage.forward(slots, ones, pos, tokens=self.tok[m] if self.first else None, handoff=ho)
              if self.first:
                self._post_tok_recv(m)     # next step's ids, posted ahead
          elif self.first:
            out = self.stage.forward(slots, ones, pos, tokens=self.tok[m], out=self.h_out[:B])
            self._send(out, 1)
            self._post_tok_recv(m)         # next step's ids, posted ahead
          else:
            hid = self.h_in[:B]
            self._recv(hid, self.rank - 1)
            if self.last:
              lg = self.stage.forward(slots, ones, pos, hidden=hid, out=self.logits)
              self._wait_tok(m)            # previous send of this buffer has completed
              self.stage.sample(lg, temp=0.0, out=self.tok[m])
              if host_out is not None:
                host_out[m].copy_(self.tok[m], non_blocking=True)
              self._post_tok_send(m)
            else:
              out = self.stage.forward(slots, ones, pos, hidden=hid, out=self.h_out[:B])
              self._send(out, self.rank + 1)
          self.pos[m] += 1

  def drain(self):
    """Complete the outstanding return-edge transfers (rank 0 ends up holding the last ids)."""
    if self.world > 1 and (self.first or self.last):
      with self._on_stream():
        for m in range(self.M):
          self._wait_tok(m)

  def close(self):
    self.stage.close()
