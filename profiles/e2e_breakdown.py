"""Where the time of one API decode step goes (B200ShardInferenceEngine.infer_tensor + sample with
host buffers, Llama-3.1-8B, batch 32, 2048-token contexts).  Host wall-clock around the pieces of
the step; prints the mean per step in microseconds.  Run: python profiles/e2e_breakdown.py"""
import asyncio
import json
import os
import sys
import tempfile
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import __graft_entry__ as graft  # noqa: E402


def main():
  graft.build()
  from xotorch_support_jetson_b200 import model_config as mc
  from xotorch_support_jetson_b200.b200_engine import B200ShardInferenceEngine
  from xotorch_support_jetson_b200.models import model_cards
  from xotorch_support_jetson_b200.pipeline import PipelineRunner
  model, B, L, K, W = "llama-3.1-8b", 32, 2048, 48, 6
  hf = model_cards[model]["hf_config"]
  with tempfile.TemporaryDirectory() as td:
    (Path(td) / "config.json").write_text(json.dumps(hf))
    cfg = mc.load_model_config(Path(td) / "config.json")
  runner = PipelineRunner(model, hf, cfg, 0, 1, 0, B, 1, L + 2 * (K + W) + 80)
  eng = B200ShardInferenceEngine(shard_downloader=None)
  eng.attach_stage(runner.shard, runner.stage, cfg)
  g = torch.Generator(device=runner.device).manual_seed(1234)
  toks = torch.randint(1, cfg["vocab_size"], (B, L), generator=g, device=runner.device, dtype=torch.int32).cpu().numpy().astype(np.int64)
  acc = {}

  def wrap(obj, name, key):
    fn = getattr(obj, name)

    def timed(*a, **k):
      t = time.perf_counter()
      r = fn(*a, **k)
      acc[key] = acc.get(key, 0.0) + time.perf_counter() - t
      return r
    setattr(obj, name, timed)

  async def go():
    logits, state = await eng.infer_tensor("r", runner.shard, toks, None)
    tok = await eng.sample(logits, temp=0.0)
    for _ in range(W):
      logits, state = await eng.infer_tensor("r", runner.shard, tok.reshape(B, 1), state)
      tok = await eng.sample(logits, temp=0.0)
    # device timeline: events around the forward and the sampler on the engine's stream
    marks = []
    st = runner.stage
    fwd, smp = st.forward, st.sample

    def ev():
      e = torch.cuda.Event(enable_timing=True)
      e.record(torch.cuda.current_stream())
      return e

    def fwd_t(*a, **k):
      e0 = ev()
      r = fwd(*a, **k)
      marks.append(("forward", e0, ev()))
      return r

    def smp_t(*a, **k):
      e0 = ev()
      r = smp(*a, **k)
      marks.append(("sample", e0, ev()))
      return r
    st.forward, st.sample = fwd_t, smp_t
    wrap(eng, "_to_host_async", "to_host_async (host)")
    wrap(eng, "_speculate", "speculative forward launch (host)")
    torch.cuda.synchronize()
    t_inf = t_smp = 0.0
    t0 = time.perf_counter()
    for _ in range(K):
      t = time.perf_counter()
      logits, state = await eng.infer_tensor("r", runner.shard, tok.reshape(B, 1), state)
      t_inf += time.perf_counter() - t
      t = time.perf_counter()
      tok = await eng.sample(logits, temp=0.0)
      t_smp += time.perf_counter() - t
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"step {dt / K * 1e6:.0f} us = infer_tensor {t_inf / K * 1e6:.0f} + sample {t_smp / K * 1e6:.0f}  ->  {B * K / dt:.0f} tok/s")
    for k, v in acc.items():
      print(f"  {k}: {v / K * 1e6:.0f} us/step")
    dur = {"forward": 0.0, "sample": 0.0}
    gap_before = {"forward": 0.0, "sample": 0.0}
    for i, (kind, e0, e1) in enumerate(marks):
      dur[kind] += e0.elapsed_time(e1)
      if i:
        gap_before[kind] += marks[i - 1][2].elapsed_time(e0)
    print(f"  device: forward {dur['forward'] / K * 1e3:.0f} us, sampler {dur['sample'] / K * 1e3:.0f} us; stream time between the end "
          f"of the forward and the start of the sampler (logits D2H + host round trip) {gap_before['sample'] / K * 1e3:.0f} us, "
          f"between sampler and next forward {gap_before['forward'] / K * 1e3:.0f} us; speculation hits {eng.spec_hits}")

  asyncio.run(go())
  # device-only step for comparison
  x = torch.from_numpy(np.zeros((B,), dtype=np.int32)).to(runner.device)
  # D2H bandwidth of the logits block alone
  dev = torch.empty((B, cfg["vocab_size"]), dtype=torch.float32, device=runner.device)
  pin = torch.empty((B, cfg["vocab_size"]), dtype=torch.float32).pin_memory()
  torch.cuda.synchronize()
  t = time.perf_counter()
  for _ in range(20):
    pin.copy_(dev, non_blocking=True)
    torch.cuda.synchronize()
  dt = (time.perf_counter() - t) / 20
  print(f"logits D2H alone: {dt * 1e6:.0f} us for {dev.numel() * 4 / 1e6:.1f} MB = {dev.numel() * 4 / dt / 1e9:.1f} GB/s")


if __name__ == "__main__":
  main()
