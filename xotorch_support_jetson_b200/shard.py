"""Shard key of the layer-range forward.

Same fields, predicates and dict form as the reference's
``xotorch/inference/shard.py:4-39`` so that ``Node``/partitioning code can hand either
class to the engine (the engine only reads the four fields).
"""
from dataclasses import dataclass


@dataclass(frozen=True)
class Shard:
  model_id: str
  start_layer: int
  end_layer: int   # inclusive
  n_layers: int

  def __hash__(self):
    return hash((self.model_id, self.start_layer, self.end_layer, self.n_layers))

  def is_first_layer(self) -> bool:
    return self.start_layer == 0

  def is_last_layer(self) -> bool:
    return self.end_layer == self.n_layers - 1

  def get_layer_count(self) -> int:
    return self.end_layer - self.start_layer + 1

  def to_dict(self) -> dict:
    return {"model_id": self.model_id, "start_layer": self.start_layer,
            "end_layer": self.end_layer, "n_layers": self.n_layers}

  @staticmethod
  def from_dict(data: dict) -> "Shard":
    return Shard(**data)

  def overlaps(self, other: "Shard") -> bool:
    return shards_overlap(self, other)


def shards_overlap(a: Shard, b: Shard) -> bool:
  return a.model_id == b.model_id and max(a.start_layer, b.start_layer) <= min(a.end_layer, b.end_layer)
