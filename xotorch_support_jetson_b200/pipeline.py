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
import zlib
from typing import Callable, Dict, List, Optional

import torch

from .model_config import load_model_config  # noqa: F401  (re-exported for callers)
from .shard import Shard
from .stage import StageRuntime
from .synth import hf_param_shapes
from .topology import pipeline_shards


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
               seed: int = 0, std: float = 0.02, dist=None, weights: Optional[tuple] = None):
    self.rank, self.world, self.dist = rank, world, dist
    self.B, self.M = batch, micro_batches
    n_layers = config["num_layers"]
    self.shard: Shard = pipeline_shards(model_id, n_layers, world)[rank]
    self.first, self.last = self.shard.is_first_layer(), self.shard.is_last_layer()
    self.device = torch.device("cuda", device)
    self.stage = StageRuntime(self.shard, config, device=device, max_batch_tokens=max_batch_tokens,
                              max_seqs=batch * micro_batches, max_tokens_per_seq=max_ctx, page_tokens=page_tokens)
    get, has = weights if weights is not None else gpu_random_getter(hf_cfg, seed, std, self.device)
    self.stage.load_hf_tensors(get, has)
    self.dim, self.vocab = config["embed_dim"], config["vocab_size"]
    self.max_batch_tokens = max_batch_tokens
    self.stream = torch.cuda.Stream(device=self.device)
    B, M = batch, micro_batches
    self.tok = [torch.zeros(B, dtype=torch.int32, device=self.device) for _ in range(M)]
    self.h_in = torch.empty((max_batch_tokens, self.dim), dtype=torch.bfloat16, device=self.device)
    self.h_out = torch.empty((max_batch_tokens, self.dim), dtype=torch.bfloat16, device=self.device)
    self.logits = torch.empty((B, self.vocab), dtype=torch.float32, device=self.device) if self.last else None
    self.pos = [0] * M  # context length of each micro-batch (all its sequences share it)

  def slots(self, m: int) -> List[int]:
    return list(range(m * self.B, (m + 1) * self.B))

  def reset(self):
    for s in range(self.B * self.M):
      self.stage.release_slot(s)
    self.pos = [0] * self.M

  # ------------------------------------------------------------------ comms
  def _recv(self, t: torch.Tensor, src: int):
    self.dist.recv(t, src=src)

  def _send(self, t: torch.Tensor, dst: int):
    self.dist.send(t, dst=dst)

  # ------------------------------------------------------------------ prefill
  def prefill(self, prompts: List[torch.Tensor]):
    """prompts[m]: int32 [B, L] on this rank's device (only rank 0 reads the values).
    Runs every micro-batch through the pipeline in chunks of whole sequences and leaves the
    first sampled token of each sequence in ``self.tok[m]`` on rank 0 (sent back by the last
    stage like the reference's last->first ``send_tensor``)."""
    B = self.B
    with torch.cuda.stream(self.stream):
      for m in range(self.M):
        L = prompts[m].shape[1]
        per = max(1, self.max_batch_tokens // L)
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
            self._send(self.tok[m], 0)
          elif self.first:
            pass  # received at the top of the first decode step

  # ------------------------------------------------------------------ decode
  def decode(self, steps: int, host_tokens: Optional[List[torch.Tensor]] = None, host_out: Optional[List[torch.Tensor]] = None):
    """`steps` decode steps for every micro-batch, greedy sampling on the last stage.
    With ``host_tokens`` / ``host_out`` (pinned CPU tensors) rank 0 uploads each step's input
    ids from host memory and the last rank downloads each step's sampled ids (end-to-end
    mode); otherwise ids stay on the devices."""
    B, M = self.B, self.M
    ones = [1] * B
    with torch.cuda.stream(self.stream):
      for k in range(steps):
        for m in range(M):
          pos = [self.pos[m]] * B
          slots = self.slots(m)
          if self.first:
            if self.world > 1:
              self._recv(self.tok[m], self.world - 1)
            if host_tokens is not None:
              # the caller's ids live in (pinned) host memory: read back, then upload as this
              # step's input -- what a host-driven caller pays every step
              host_tokens[m].copy_(self.tok[m], non_blocking=True)
              self.stream.synchronize()
              self.tok[m].copy_(host_tokens[m], non_blocking=True)
          if self.first and self.last:
            lg = self.stage.forward(slots, ones, pos, tokens=self.tok[m], out=self.logits)
            self.stage.sample(lg, temp=0.0, out=self.tok[m])
            if host_out is not None:
              host_out[m].copy_(self.tok[m], non_blocking=True)
          elif self.first:
            out = self.stage.forward(slots, ones, pos, tokens=self.tok[m], out=self.h_out[:B])
            self._send(out, 1)
          else:
            hid = self.h_in[:B]
            self._recv(hid, self.rank - 1)
            if self.last:
              lg = self.stage.forward(slots, ones, pos, hidden=hid, out=self.logits)
              self.stage.sample(lg, temp=0.0, out=self.tok[m])
              if host_out is not None:
                host_out[m].copy_(self.tok[m], non_blocking=True)
              self._send(self.tok[m], 0)
            else:
              out = self.stage.forward(slots, ones, pos, hidden=hid, out=self.h_out[:B])
              self._send(out, self.rank + 1)
          self.pos[m] += 1

  def drain(self):
    """Match the last stage's final token sends (rank 0 receives them as the next inputs)."""
    if self.world > 1 and self.first:
      with torch.cuda.stream(self.stream):
        for m in range(self.M):
          self._recv(self.tok[m], self.world - 1)

  def close(self):
    self.stage.close()
