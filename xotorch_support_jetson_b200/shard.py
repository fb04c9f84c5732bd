"""The key of the layer-range forward: which contiguous, inclusive range of decoder layers of
which model a stage holds.

Field names, meaning and the dict form are those of the reference's ``Shard``
(``xotorch/inference/shard.py:4-39``) so that orchestration code can pass either class; the
engine itself only reads the four fields.
"""
from typing import NamedTuple


class Shard(NamedTuple):
  model_id: str
  start_layer: int   # first layer held (absolute index)
  end_layer: int     # last layer held, INCLUSIVE
  n_layers: int      # layers of the whole model

  # -- position in the pipeline
  def is_first_layer(self) -> bool:
    """Holds layer 0, i.e. consumes token ids (embedding lives here)."""
    return not self.start_layer

  def is_last_layer(self) -> bool:
    """Holds the final layer, i.e. produces logits (final norm + LM head live here)."""
    return self.end_layer + 1 == self.n_layers

  def get_layer_count(self) -> int:
    return 1 + self.end_layer - self.start_layer

  # -- wire form (JSON-able)
  def to_dict(self) -> dict:
    return dict(self._asdict())

  @classmethod
  def from_dict(cls, data: dict) -> "Shard":
    return cls(data["model_id"], int(data["start_layer"]), int(data["end_layer"]), int(data["n_layers"]))

  def overlaps(self, other: "Shard") -> bool:
    return shards_overlap(self, other)


def shards_overlap(a: Shard, b: Shard) -> bool:
  """Two shards of the same model overlap when their inclusive ranges share a layer."""
  if a.model_id != b.model_id:
    return False
  return not (a.end_layer < b.start_layer or b.end_layer < a.start_layer)
