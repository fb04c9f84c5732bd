"""A scripted InferenceEngine -- TEST INFRASTRUCTURE (used by oracle/gen_ref_fixtures.py to record what the
reference's ``Node`` does, and by tests/test_orchestration.py to replay it against this package's node).

It has the reference plugin surface (encode / infer_prompt / infer_tensor / sample / decode / ensure_shard,
``.tokenizer.eos_token_id``, ``.shard``) and a toy 'model': hidden rows carry (token id, position), the last
shard's logits are one-hot at ``(3 * id + position + 1) % vocab``.  Every call is appended to ``log``."""
import types
from typing import Optional

import numpy as np

VOCAB, HID = 97, 4


def next_token_rule(tok: int, pos: int) -> int:
  return (3 * int(tok) + int(pos) + 1) % VOCAB


class ScriptedEngine:
  def __init__(self, name: str, log: list, eos_token_id: int = VOCAB - 1):
    self.name, self.log = name, log
    self.shard = None
    self.tokenizer = types.SimpleNamespace(eos_token_id=eos_token_id)
    self.requests = {}

  async def ensure_shard(self, shard):
    self.shard = shard

  async def encode(self, shard, prompt: str) -> np.ndarray:
    await self.ensure_shard(shard)
    return np.array([int(t) for t in prompt.split()], dtype=np.int64)

  async def decode(self, shard, tokens) -> str:
    return " ".join(str(int(t)) for t in np.asarray(tokens).reshape(-1))

  async def infer_prompt(self, request_id, shard, prompt, inference_state=None):
    ids = await self.encode(shard, prompt)
    return await self.infer_tensor(request_id, shard, ids.reshape(1, -1), inference_state)

  async def infer_tensor(self, request_id, shard, input_data, inference_state: Optional[dict] = None):
    await self.ensure_shard(shard)
    x = np.asarray(input_data)
    pos = int((inference_state or {}).get("curr_pos", 0) or 0)
    if x.ndim == 2:
      B, s = x.shape
      hidden = np.zeros((B, s, HID), dtype=np.float32)
      hidden[..., 0] = x
      hidden[..., 1] = pos + np.arange(s)[None, :]
      toks = self.requests.get(request_id)
      toks = x.astype(np.int64).tolist() if (toks is None or s > 1) else [r + [int(t)] for r, t in zip(toks, x[:, 0])]
      self.requests[request_id] = toks
    else:
      B, s = x.shape[:2]
      hidden = x.astype(np.float32)
      toks = (inference_state or {}).get("tokens")
    last = shard.is_last_layer()
    state = {"curr_pos": pos + s if last else pos}
    if toks is not None:
      state["tokens"] = [list(map(int, r)) for r in toks]
    if last:
      out = np.zeros((B, VOCAB), dtype=np.float32)
      for b in range(B):
        out[b, next_token_rule(hidden[b, -1, 0], hidden[b, -1, 1])] = 1.0
    else:
      out = hidden
    self.log.append({"node": self.name, "call": "infer_tensor", "request_id": request_id,
                     "shard": [int(shard.start_layer), int(shard.end_layer), int(shard.n_layers)],
                     "in_ndim": int(x.ndim), "in_shape": [int(d) for d in x.shape], "curr_pos_in": pos,
                     "first_in": int(x.reshape(-1)[0]), "out_shape": [int(d) for d in out.shape],
                     "curr_pos_out": state["curr_pos"]})
    return out, state

  async def sample(self, x, temp=0.0, top_k=35):
    tok = np.asarray(x).argmax(axis=-1).astype(np.int64).reshape(-1, 1)
    self.log.append({"node": self.name, "call": "sample", "token": int(tok[0, 0]), "temp": float(temp)})
    return tok

  async def load_checkpoint(self, shard, path):
    await self.ensure_shard(shard)
