"""CPU stand-in for StageRuntime built on the oracle -- TEST INFRASTRUCTURE ONLY.
It lets the host-side pipeline protocol run under gloo without a GPU.  Same forward /
sample / release_slot surface as xotorch_support_jetson_b200.stage.StageRuntime."""
import torch

from oracle import torch_oracle as O


class OracleStage:
  def __init__(self, shard, cfg, state, max_ctx=64):
    self.shard, self.cfg, self.state, self.max_ctx = shard, cfg, state, max_ctx
    self.models = {}   # slot -> OracleShardModel with batch-1 caches
    self.dim, self.vocab = cfg["embed_dim"], cfg["vocab_size"]
    self.mask = torch.tril(torch.ones(max_ctx, max_ctx, dtype=torch.bool)).unsqueeze(0)

  def _model(self, slot):
    m = self.models.get(slot)
    if m is None:
      m = O.OracleShardModel(self.cfg, self.shard.model_id, self.shard.start_layer, self.shard.end_layer, self.state,
                             O.OracleOptions(max_generated_tokens=0))
      m.setup_caches(1, self.max_ctx)
      self.models[slot] = m
    return m

  def forward(self, slot_ids, q_lens, positions, tokens=None, hidden=None, out=None, taps=None):
    rows, r = [], 0
    for slot, ql, pos in zip(slot_ids, q_lens, positions):
      m = self._model(slot)
      assert m.cache_pos == pos, (m.cache_pos, pos)
      ip = torch.arange(pos, pos + ql)
      mask = self.mask[:, pos:pos + ql]
      if tokens is not None:
        y = m.forward(tokens[r:r + ql].reshape(1, ql).long(), mask=mask, input_pos=ip)
      else:
        y = m.forward(None, mask=mask, input_pos=ip, hidden_state=hidden[r:r + ql].reshape(1, ql, -1))
      rows.append(y[0, -1:].float() if self.shard.is_last_layer() else y[0])
      r += ql
    res = torch.cat(rows, dim=0)
    if out is not None:
      out.copy_(res)
      return out
    return res

  def sample(self, logits, temp=0.0, out=None, **kw):
    t = logits.argmax(-1).to(torch.int32)
    if out is not None:
      out.copy_(t)
      return out
    return t

  def release_slot(self, slot):
    self.models.pop(slot, None)

  def close(self):
    pass
