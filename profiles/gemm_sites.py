"""Per-site timing of the decode Linear kernel, back to back over all 32 layers (Llama-3.1-8B)."""
import json, os, sys, tempfile
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch
import __graft_entry__ as graft
graft.build()
from xotorch_support_jetson_b200 import model_config as mc
from xotorch_support_jetson_b200.models import model_cards
from xotorch_support_jetson_b200.pipeline import gpu_random_getter
from xotorch_support_jetson_b200.shard import Shard
from xotorch_support_jetson_b200.stage import StageRuntime
hf = model_cards["llama-3.1-8b"]["hf_config"]
with tempfile.TemporaryDirectory() as td:
  (Path(td) / "config.json").write_text(json.dumps(hf)); cfg = mc.load_model_config(Path(td) / "config.json")
dev = torch.device("cuda", 0)
st = StageRuntime(Shard("llama-3.1-8b", 0, 31, 32), cfg, device=0, max_batch_tokens=256, max_seqs=32, max_tokens_per_seq=128)
st.load_hf_tensors(*gpu_random_getter(hf, 0, 0.02, dev))
s = torch.cuda.Stream()
MB = {"0": 50.33 * 32, "1": 33.55 * 32, "2": 234.88 * 32, "3": 117.44 * 32, "4": 1050.67, "-1": 15009.0}
names = {"0": "qkv", "1": "o", "2": "gate_up", "3": "down", "4": "head", "-1": "all"}
for T in (1, 32):
  for site in ("-1", "0", "1", "2", "3", "4"):
    os.environ["XB_BENCH_SITE"] = site
    with torch.cuda.stream(s):
      st.bench_linears(T, 2)
      e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
      e0.record(s); n = st.bench_linears(T, 10); e1.record(s); s.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(f"T={T:2d} {names[site]:8s} {ms*1e3/ (n/10):7.2f} us/launch  {MB[site]/ms/1e3:6.2f} TB/s  ({MB[site]/ms/1e3/6.4534:.2f} of measured HBM peak)", flush=True)
