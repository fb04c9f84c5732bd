import asyncio, json, sys, tempfile, time, cProfile, pstats
from pathlib import Path
import numpy as np, torch
sys.path.insert(0, "/root/repo")
import __graft_entry__ as graft
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
class Inline:
  def submit(self, fn, *a, **k):
    from concurrent.futures import Future
    f = Future()
    try: f.set_result(fn(*a, **k))
    except BaseException as e: f.set_exception(e)
    return f
inline = "--inline" in sys.argv
if inline: eng.executor = Inline()
toks = np.random.randint(1, 1000, (B, L)).astype(np.int64)
async def go():
  logits, state = await eng.infer_tensor("r", runner.shard, toks, None)
  tok = await eng.sample(logits, temp=0.0)
  for _ in range(W):
    logits, state = await eng.infer_tensor("r", runner.shard, tok.reshape(B, 1), state)
    tok = await eng.sample(logits, temp=0.0)
  torch.cuda.synchronize()
  pr = cProfile.Profile() if inline else None
  t0 = time.perf_counter()
  if pr: pr.enable()
  for _ in range(K):
    logits, state = await eng.infer_tensor("r", runner.shard, tok.reshape(B, 1), state)
    tok = await eng.sample(logits, temp=0.0)
  if pr: pr.disable()
  torch.cuda.synchronize()
  dt = time.perf_counter() - t0
  print(f"inline={inline} step {dt/K*1e6:.0f} us")
  if pr: pstats.Stats(pr).sort_stats("tottime").print_stats(18)
asyncio.run(go())
