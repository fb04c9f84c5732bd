"""Profiling driver: Llama-3.1-8B, real prefill to --ctx, then eager decode steps; the last
`--steps` decode steps run between cudaProfilerStart/Stop (use ncu --profile-from-start off)."""
import argparse, json, os, sys, tempfile
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
os.environ.setdefault("XB_GRAPHS", "0")
import torch
import __graft_entry__ as graft

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=32)
ap.add_argument("--ctx", type=int, default=2048)
ap.add_argument("--steps", type=int, default=1)
ap.add_argument("--model", default="llama-3.1-8b")
ap.add_argument("--prefill-only", action="store_true")
a = ap.parse_args()
graft.build()
from xotorch_support_jetson_b200 import model_config as mc
from xotorch_support_jetson_b200.models import model_cards
from xotorch_support_jetson_b200.pipeline import PipelineRunner
hf = model_cards[a.model]["hf_config"]
with tempfile.TemporaryDirectory() as td:
  (Path(td) / "config.json").write_text(json.dumps(hf))
  cfg = mc.load_model_config(Path(td) / "config.json")
r = PipelineRunner(a.model, hf, cfg, 0, 1, 0, a.batch, 1, a.ctx + 64, max_batch_tokens=8192)
g = torch.Generator(device=r.device).manual_seed(1)
prompts = [torch.randint(1, cfg["vocab_size"], (a.batch, a.ctx), generator=g, device=r.device, dtype=torch.int32)]
if a.prefill_only:
  r.prefill(prompts); torch.cuda.synchronize(); r.reset()
  torch.cuda.profiler.start()
  r.prefill([prompts[0][:4]] if a.batch >= 4 else prompts) if False else r.prefill(prompts)
  torch.cuda.synchronize(); torch.cuda.profiler.stop()
else:
  r.prefill(prompts)
  r.decode(2)
  torch.cuda.synchronize()
  torch.cuda.profiler.start()
  r.decode(a.steps)
  torch.cuda.synchronize()
  torch.cuda.profiler.stop()
print("launches", r.stage.counters()["launches"])
