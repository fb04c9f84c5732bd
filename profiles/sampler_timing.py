"""Device time of the sampling pieces at the bench shape (32 rows x 128256 logits), CUDA events,
200 iterations each: the Exp(1) noise draw (torch), xb_sample (top-k 35, supplied noise),
xb_sample (counter noise), xb_sample with a nucleus cut, xb_argmax path (temp 0, top_k 0)."""
import json
import sys
import tempfile
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import __graft_entry__ as graft  # noqa: E402


def main():
  graft.build()
  from xotorch_support_jetson_b200 import model_config as mc
  from xotorch_support_jetson_b200.models import model_cards
  from xotorch_support_jetson_b200.shard import Shard
  from xotorch_support_jetson_b200.stage import StageRuntime
  hf = dict(model_cards["llama-3.2-tiny"]["hf_config"])
  hf["vocab_size"] = 128256
  with tempfile.TemporaryDirectory() as td:
    (Path(td) / "config.json").write_text(json.dumps(hf))
    cfg = mc.load_model_config(Path(td) / "config.json")
  nl = hf["num_hidden_layers"]
  st = StageRuntime(Shard("llama-3.2-tiny", 0, nl - 1, nl), cfg, device=0, max_batch_tokens=64, max_seqs=32, max_tokens_per_seq=64)
  B, V = 32, 128256
  g = torch.Generator(device="cuda").manual_seed(1)
  logits = torch.randn((B, V), device="cuda", generator=g) * 3.0
  q = torch.empty((B, V), device="cuda").exponential_(1, generator=g)
  out = torch.empty((B,), dtype=torch.int32, device="cuda")

  def timed(fn, n=200):
    for _ in range(10):
      fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
      fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3

  s = torch.cuda.Stream()
  with torch.cuda.stream(s):
    print(f"exponential_ [32,128256]        : {timed(lambda: q.exponential_(1, generator=g)):7.1f} us")
    print(f"xb_sample top-k 35, given noise : {timed(lambda: st.sample(logits, temp=0.6, top_k=35, q=q, out=out)):7.1f} us")
    print(f"xb_sample top-k 35, temp 1e-5   : {timed(lambda: st.sample(logits, temp=1e-5, top_k=35, q=q, out=out)):7.1f} us")
    print(f"xb_sample top-k 35, counter     : {timed(lambda: st.sample(logits, temp=0.6, top_k=35, seed=3, out=out)):7.1f} us")
    print(f"xb_sample top-k 35 + top-p 0.9  : {timed(lambda: st.sample(logits, temp=0.6, top_k=35, top_p=0.9, q=q, out=out)):7.1f} us")
    print(f"xb_sample no top-k              : {timed(lambda: st.sample(logits, temp=0.6, top_k=0, q=q, out=out)):7.1f} us")
    print(f"argmax (temp 0, top_k 0)        : {timed(lambda: st.sample(logits, temp=0.0, top_k=0, out=out)):7.1f} us")


if __name__ == "__main__":
  main()
