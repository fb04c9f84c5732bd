"""Device capabilities and layer partitioning for one box of B200s.

Mirrors the parts of the reference's ``xotorch/topology`` the hot path depends on:
``DeviceCapabilities`` / ``DeviceFlops`` (``device_capabilities.py:22-52``), ``Topology``
(``topology.py:22-45``), ``Partition`` + ``map_partitions_to_shards``
(``partitioning_strategy.py:11-42``) and ``RingMemoryWeightedPartitioningStrategy``
(``ring_memory_weighted_partitioning_strategy.py:7-18``).  The one change north_star asks
for: a multi-GPU host is exposed as ONE NODE PER GPU (the reference collapses all GPUs of
a host into one node, ``device_capabilities.py:214-252``) and ``CHIP_FLOPS`` knows B200.
"""
from dataclasses import dataclass, field
from typing import Dict, List, Optional

from .shard import Shard

TFLOPS = 1.00


@dataclass
class DeviceFlops:
  fp32: float
  fp16: float
  int8: float

  def to_dict(self):
    return {"fp32": self.fp32, "fp16": self.fp16, "int8": self.int8}


@dataclass
class DeviceCapabilities:
  model: str
  chip: str
  memory: int  # MB
  flops: DeviceFlops

  def __post_init__(self):
    if isinstance(self.flops, dict):
      self.flops = DeviceFlops(**self.flops)

  def to_dict(self):
    return {"model": self.model, "chip": self.chip, "memory": self.memory, "flops": self.flops.to_dict()}


# dense peaks (TFLOPS); fp16 = bf16/fp16 tensor, int8 slot carries the fp8 tensor figure
CHIP_FLOPS = {
  "NVIDIA B200": DeviceFlops(fp32=80.0 * TFLOPS, fp16=2250.0 * TFLOPS, int8=4500.0 * TFLOPS),
}
UNKNOWN_DEVICE_CAPABILITIES = DeviceCapabilities(model="Unknown Model", chip="Unknown Chip", memory=0,
                                                 flops=DeviceFlops(fp32=0, fp16=0, int8=0))


def b200_device_capabilities(memory_mb: int = 183359) -> DeviceCapabilities:
  return DeviceCapabilities(model="Linux Box (NVIDIA B200)", chip="NVIDIA B200", memory=memory_mb,
                            flops=CHIP_FLOPS["NVIDIA B200"])


def local_gpu_capabilities() -> List[DeviceCapabilities]:
  """One DeviceCapabilities per visible GPU (one node per GPU)."""
  import torch
  out = []
  for i in range(torch.cuda.device_count()):
    p = torch.cuda.get_device_properties(i)
    name = p.name
    out.append(DeviceCapabilities(model=f"Linux Box ({name})", chip=name, memory=p.total_memory // 2**20,
                                  flops=CHIP_FLOPS.get(name, DeviceFlops(fp32=0, fp16=0, int8=0))))
  return out


class Topology:
  """Nodes + directed peer edges, as ``xotorch/topology/topology.py:22-77`` keeps them (an edge is keyed
  by its two ends; ``merge`` takes over only what the peer says about ITSELF)."""

  def __init__(self):
    self.nodes: Dict[str, DeviceCapabilities] = {}
    self.peer_graph: Dict[str, Dict[str, Optional[str]]] = {}   # from_id -> {to_id: description}
    self.active_node_id: Optional[str] = None

  def update_node(self, node_id: str, device_capabilities: DeviceCapabilities):
    self.nodes[node_id] = device_capabilities

  def get_node(self, node_id: str) -> DeviceCapabilities:
    return self.nodes.get(node_id)

  def all_nodes(self):
    return self.nodes.items()

  def add_edge(self, from_id: str, to_id: str, description: Optional[str] = None):
    self.peer_graph.setdefault(from_id, {}).setdefault(to_id, description)

  def merge(self, peer_node_id: str, other: "Topology"):
    if peer_node_id in other.nodes:
      self.update_node(peer_node_id, other.nodes[peer_node_id])
    for to_id, desc in other.peer_graph.get(peer_node_id, {}).items():
      self.add_edge(peer_node_id, to_id, desc)

  def to_json(self) -> dict:
    return {"nodes": {i: c.to_dict() for i, c in self.nodes.items()},
            "peer_graph": {f: [{"from_id": f, "to_id": t, "description": d} for t, d in tos.items()] for f, tos in self.peer_graph.items()},
            "active_node_id": self.active_node_id}


@dataclass
class Partition:
  node_id: str
  start: float
  end: float


class RingMemoryWeightedPartitioningStrategy:
  """Same rule as the reference's strategy
  (``xotorch/topology/ring_memory_weighted_partitioning_strategy.py:7-18``): order the nodes by
  (memory, node id) descending and give each a slice of [0, 1) proportional to its memory, the
  running boundary rounded to 5 decimals."""

  def partition(self, topology: Topology) -> List[Partition]:
    ring = sorted(topology.all_nodes(), key=lambda kv: (kv[1].memory, kv[0]), reverse=True)
    total = float(sum(caps.memory for _, caps in ring))
    out: List[Partition] = []
    lo = 0
    for node_id, caps in ring:
      hi = round(lo + caps.memory / total, 5)
      out.append(Partition(node_id, lo, hi))
      lo = hi
    return out


def map_partitions_to_shards(partitions: List[Partition], num_layers: int, model_id: str) -> List[Shard]:
  """Fractions of [0, 1) -> inclusive layer ranges, with the reference's conventions
  (``xotorch/topology/partitioning_strategy.py:23-42``): boundaries are truncated
  (``int(x * L)``), the last partition always runs to the final layer, partitions that end up
  empty are dropped, and whatever comes last is stretched to cover the model."""
  n_parts = len(partitions)
  shards: List[Shard] = []
  for idx, part in enumerate(partitions):
    first = int(part.start * num_layers)
    last = num_layers - 1 if idx == n_parts - 1 else int(part.end * num_layers) - 1
    if last >= first:
      shards.append(Shard(model_id, first, last, num_layers))
  if shards and shards[-1].end_layer != num_layers - 1:
    tail = shards[-1]
    if tail.end_layer < num_layers - 1:
      shards[-1] = Shard(model_id, tail.start_layer, num_layers - 1, num_layers)
  return shards


def box_topology(n_gpus: int, memory_mb: int = 183359, last_stage_reserved: float = 0.0) -> Topology:
  """n B200 nodes named so that the ring order (memory desc, node id desc) is gpu0, gpu1, ...

  ``last_stage_reserved`` (fraction of one node's share, 0..1) is the part of the LAST GPU's
  capacity that is spoken for by the LM head, which the last shard holds and streams every
  decode step on top of its layers.  The memory-weighted strategy then hands that GPU
  proportionally fewer layers, so the stages of the pipeline stream equal bytes per step."""
  t = Topology()
  for i in range(n_gpus):
    mem = memory_mb
    if i == n_gpus - 1 and n_gpus > 1 and last_stage_reserved > 0:
      mem = max(1, int(round(memory_mb * (1.0 - last_stage_reserved))))
    t.update_node(f"gpu{n_gpus - 1 - i:02d}-rank{i}", b200_device_capabilities(mem))
  return t


def pipeline_shards(model_id: str, n_layers: int, n_gpus: int, head_layer_equiv: float = 0.0) -> List[Shard]:
  """Layer ranges of an n_gpus pipeline: what the reference's ring strategy +
  map_partitions_to_shards assign to the n per-GPU nodes; entry i belongs to rank i.

  ``head_layer_equiv`` = bytes of the LM head / bytes of one decoder layer.  With 0 the nodes
  are equal (the reference's view: 32 layers -> 8 x 4).  With the real value the last node's
  capacity is reduced so that (its layers + head) matches the other stages' layers."""
  def shards_for(reserved: float) -> List[Shard]:
    parts = RingMemoryWeightedPartitioningStrategy().partition(box_topology(n_gpus, last_stage_reserved=reserved))
    return map_partitions_to_shards(parts, n_layers, model_id)

  if head_layer_equiv <= 0 or n_gpus <= 1:
    return shards_for(0.0)
  # the strategy rounds fractions to whole layers, so scan the last GPU's reserved share and keep
  # the assignment whose slowest stage (layers + head on the last one) is cheapest
  best, best_cost = None, None
  for step in range(0, 91):
    sh = shards_for(step / 100.0)
    if len(sh) != n_gpus:
      continue
    cost = max(s.get_layer_count() + (head_layer_equiv if i == n_gpus - 1 else 0.0) for i, s in enumerate(sh))
    if best_cost is None or cost < best_cost - 1e-9:
      best, best_cost = sh, cost
  return best if best is not None else shards_for(0.0)


def half_layer_stages(model_id: str, n_layers: int, n_gpus: int, attn_cost: float, mlp_cost: float,
                      head_cost: float) -> List[tuple]:
  """Stage boundaries at HALF-layer granularity: the per-step cost of a decode stage is bytes
  streamed, and a decoder layer is two blocks of similar weight -- attention (q/k/v/o matrices +
  the KV cache of the batch) and MLP (gate/up/down) -- plus the LM head on the last stage, so
  whole-layer ranges (``pipeline_shards``, the reference's ring strategy) leave up to a layer of
  imbalance (32 layers + head over 8 GPUs: slowest stage 5 layers against a mean of 4.3).

  Minimises the slowest stage over contiguous splits of [attn_0, mlp_0, ..., attn_{n-1}, mlp_{n-1} + head].
  Returns, per rank, ``(Shard, first_from_mlp, last_to_attn)``: a layer cut in the middle appears
  in both neighbouring Shards (end_layer of one, start_layer of the next)."""
  units = []
  for _ in range(n_layers):
    units += [float(attn_cost), float(mlp_cost)]
  units[-1] += float(head_cost)
  U = len(units)
  n = max(1, min(n_gpus, U))
  pre = [0.0]
  for u in units:
    pre.append(pre[-1] + u)
  INF = float("inf")
  # best[k][i] = smallest possible slowest-stage cost when the first i units form k stages
  best = [[INF] * (U + 1) for _ in range(n + 1)]
  cut = [[0] * (U + 1) for _ in range(n + 1)]
  best[0][0] = 0.0
  for k in range(1, n + 1):
    for i in range(k, U + 1):
      for j in range(k - 1, i):
        c = max(best[k - 1][j], pre[i] - pre[j])
        if c < best[k][i] - 1e-12:
          best[k][i], cut[k][i] = c, j
  bounds, i = [], U
  for k in range(n, 0, -1):
    j = cut[k][i]
    bounds.append((j, i))
    i = j
  bounds.reverse()
  out = []
  for (a, b) in bounds:            # units [a, b): unit 2l = attention of layer l, 2l+1 = its MLP
    start_layer, end_layer = a // 2, (b - 1) // 2
    out.append((Shard(model_id, start_layer, end_layer, n_layers), a % 2 == 1, (b - 1) % 2 == 0))
  return out

