"""Sweep the kv-split count of the decode attention kernel (Llama-3.1-8B geometry, 2 layers)."""
import argparse, json, sys, tempfile
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch
import __graft_entry__ as graft
ap = argparse.ArgumentParser()
ap.add_argument("--ctx", type=int, default=2048)
a = ap.parse_args()
graft.build()
from xotorch_support_jetson_b200 import model_config as mc
from xotorch_support_jetson_b200.models import model_cards
from xotorch_support_jetson_b200.pipeline import gpu_random_getter
from xotorch_support_jetson_b200.shard import Shard
from xotorch_support_jetson_b200.stage import StageRuntime
hf = dict(model_cards["llama-3.1-8b"]["hf_config"]); hf["num_hidden_layers"] = 2
with tempfile.TemporaryDirectory() as td:
  (Path(td) / "config.json").write_text(json.dumps(hf)); cfg = mc.load_model_config(Path(td) / "config.json")
dev = torch.device("cuda", 0)
for B in (32, 1, 8):
  st = StageRuntime(Shard("llama-3.1-8b", 0, 1, 2), cfg, device=0, max_batch_tokens=8192, max_seqs=B, max_tokens_per_seq=a.ctx + 64)
  st.load_hf_tensors(*gpu_random_getter(hf, 0, 0.02, dev))
  s = torch.cuda.Stream()
  with torch.cuda.stream(s):
    g = torch.Generator(device=dev).manual_seed(1)
    per = max(1, 8192 // a.ctx)
    for c0 in range(0, B, per):
      n = min(per, B - c0)
      st.forward(list(range(c0, c0 + n)), [a.ctx] * n, [0] * n, tokens=torch.randint(1, 1000, (n * a.ctx,), generator=g, device=dev, dtype=torch.int32))
    tok = torch.ones(B, dtype=torch.int32, device=dev)
    st.forward(list(range(B)), [1] * B, [a.ctx] * B, tokens=tok)   # leaves the decode metadata on the device
    kv_bytes = B * (a.ctx + 1) * cfg["num_kv_heads"] * cfg["head_dim"] * 2 * 2
    res = {}
    for ns in (0, 1, 2, 3, 4, 5, 6, 8, 10, 12, 16, 24, 32):
      st.bench_attention(B, a.ctx + 1, ns, 3)
      e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
      e0.record(s); used = st.bench_attention(B, a.ctx + 1, ns, 50); e1.record(s); s.synchronize()
      us = e0.elapsed_time(e1) * 1e3 / 50
      res[f"{ns}->{used}"] = (round(us, 2), round(kv_bytes / us / 1e3, 1))
  print(f"B={B} ctx={a.ctx} kv_bytes={kv_bytes/1e6:.1f}MB  splits: (us, GB/s)  {res}", flush=True)
  st.close()
