"""A/B timing of the graph-replayed decode step (Llama-3.1-8B, ctx 2048) under engine switches.

  python profiles/decode_ab.py --variants "XB_FUSE_QKV=0,XB_ATTN_CLUSTER=0;XB_FUSE_QKV=1,XB_ATTN_CLUSTER=1" --batches 1,32

One engine per (variant, batch): real prefill, W warm steps, then K timed steps with CUDA events on the
runner's stream, repeated R times (median reported).  Prints ms/step, tok/s and the whole-step HBM fraction."""
import argparse, json, os, statistics, sys, tempfile
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch
import __graft_entry__ as graft

ap = argparse.ArgumentParser()
ap.add_argument("--variants", default="XB_FUSE_QKV=0,XB_ATTN_CLUSTER=0;XB_FUSE_QKV=1,XB_ATTN_CLUSTER=1")
ap.add_argument("--batches", default="1,32")
ap.add_argument("--ctx", type=int, default=2048)
ap.add_argument("--steps", type=int, default=32)
ap.add_argument("--warmup", type=int, default=4)
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--model", default="llama-3.1-8b")
a = ap.parse_args()
graft.build()
from xotorch_support_jetson_b200 import model_config as mc
from xotorch_support_jetson_b200.models import model_cards
from xotorch_support_jetson_b200.pipeline import PipelineRunner
hf = model_cards[a.model]["hf_config"]
with tempfile.TemporaryDirectory() as td:
  (Path(td) / "config.json").write_text(json.dumps(hf))
  cfg = mc.load_model_config(Path(td) / "config.json")
peak = 6453.4
p = ROOT / "MEASURED_PEAKS.json"
if p.exists():
  peak = json.loads(p.read_text())["hbm_gbs"]
rows = []
for var in a.variants.split(";"):
  for kv in [x for x in var.split(",") if x]:
    k, v = kv.split("=")
    os.environ[k] = v
  for B in [int(x) for x in a.batches.split(",")]:
    r = PipelineRunner(a.model, hf, cfg, 0, 1, 0, B, 1, a.ctx + a.warmup + a.steps * a.reps + 16, max_batch_tokens=8192)
    g = torch.Generator(device=r.device).manual_seed(1)
    prompts = [torch.randint(1, cfg["vocab_size"], (B, a.ctx), generator=g, device=r.device, dtype=torch.int32)]
    r.prefill(prompts)
    r.decode(a.warmup)
    torch.cuda.synchronize()
    l0 = r.stage.counters()["launches"]
    ts = []
    for _ in range(a.reps):
      e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
      e0.record(r.stream)
      r.decode(a.steps)
      e1.record(r.stream)
      torch.cuda.synchronize()
      ts.append(e0.elapsed_time(e1) / a.steps)
    ms = statistics.median(ts)
    c = r.stage.counters()
    launches = (c["launches"] - l0) / (a.steps * a.reps)
    ctx_avg = a.ctx + a.warmup + (a.steps * a.reps - 1) / 2
    frac = (c["weight_bytes"] + B * ctx_avg * c["kv_bytes_per_pos"]) / (ms * 1e-3) / 1e9 / peak
    rows.append((var, B, ms, B / ms * 1e3, frac, launches))
    print(f"{var:40s} B={B:3d}  {ms:7.4f} ms/step  {B / ms * 1e3:9.1f} tok/s  step HBM frac {frac:.4f}  launches/step {launches:.0f}  reps {['%.4f' % t for t in ts]}", flush=True)
    r.close()
    del r
    torch.cuda.empty_cache()
