"""The plugin surface of the reference, kept signature-for-signature.

``InferenceEngine`` mirrors ``xotorch/inference/inference_engine.py:11-51`` (abstract
encode / sample / decode / infer_tensor / load_checkpoint, concrete infer_prompt), and
``get_inference_engine`` mirrors the factory at ``inference_engine.py:54-70`` with one
new key, ``"b200"``.  A reference checkout registers the engine by adding that key to
its own ``inference_engine_classes`` / ``get_inference_engine``.
"""
from abc import ABC, abstractmethod
from typing import Optional

import numpy as np

from .shard import Shard


class InferenceEngine(ABC):
  session = {}

  @abstractmethod
  async def encode(self, shard: Shard, prompt: str) -> np.ndarray:
    pass

  @abstractmethod
  async def sample(self, x: np.ndarray) -> np.ndarray:
    pass

  @abstractmethod
  async def decode(self, shard: Shard, tokens: np.ndarray) -> str:
    pass

  @abstractmethod
  async def infer_tensor(self, request_id: str, shard: Shard, input_data: np.ndarray,
                         inference_state: Optional[dict] = None) -> tuple[np.ndarray, Optional[dict]]:
    pass

  @abstractmethod
  async def load_checkpoint(self, shard: Shard, path: str):
    pass

  async def save_checkpoint(self, shard: Shard, path: str):
    pass

  async def save_session(self, key, value):
    self.session[key] = value

  async def clear_session(self):
    self.session.clear()

  async def infer_prompt(self, request_id: str, shard: Shard, prompt: str,
                         inference_state: Optional[dict] = None) -> tuple[np.ndarray, Optional[dict]]:
    tokens = await self.encode(shard, prompt)
    x = tokens.reshape(1, -1)
    output_data, inference_state = await self.infer_tensor(request_id, shard, x, inference_state)
    return output_data, inference_state


inference_engine_classes = {
  "b200": "B200ShardInferenceEngine",
}


def get_inference_engine(inference_engine_name: str, shard_downloader):
  if inference_engine_name == "b200":
    from .b200_engine import B200ShardInferenceEngine
    return B200ShardInferenceEngine(shard_downloader)
  raise ValueError(f"Unsupported inference engine: {inference_engine_name}")
