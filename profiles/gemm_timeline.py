"""Where the time between two decode Linears goes: per-block %globaltimer stamps (XB_GEMM_TS=1, eager launches
with programmatic dependent launch, Llama-3.1-8B, ctx 2048).

  python profiles/gemm_timeline.py --batch 32 [--mode step|linears]

mode step: one whole decode step (the small kernels and the attention run between the Linears);
mode linears: the step's 129 Linears back to back (what bench.py's roofline leg times).
For every site (qkv, o, gate/up, down) averaged over layers 1..31, times in microseconds relative to the moment
the previous Linear's last block exited."""
import argparse, json, os, sys, tempfile
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
os.environ["XB_GEMM_TS"] = "1"
import numpy as np
import torch
import __graft_entry__ as graft

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=32)
ap.add_argument("--ctx", type=int, default=2048)
ap.add_argument("--mode", default="step")
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
r = PipelineRunner(a.model, hf, cfg, 0, 1, 0, a.batch, 1, a.ctx + 64, max_batch_tokens=8192)
g = torch.Generator(device=r.device).manual_seed(1)
prompts = [torch.randint(1, cfg["vocab_size"], (a.batch, a.ctx), generator=g, device=r.device, dtype=torch.int32)]
r.prefill(prompts)
r.decode(3)
torch.cuda.synchronize()
r.stage.timeline()
if a.mode == "linears":
  with torch.cuda.stream(r.stream):
    r.stage.bench_linears(a.batch, 1)
else:
  r.decode(1)
torch.cuda.synchronize()
ts = r.stage.timeline().astype(np.int64)          # [launches, blocks, 8]
n = ts.shape[0]
print(f"mode {a.mode} batch {a.batch}: {n} Linear launches recorded")
names = ["qkv", "o", "gate/up", "down"]
rows = {k: [] for k in names}
for i in range(5, n - 1):        # skip layer 0 and the LM head (last launch)
  site = names[i % 4]
  cur, prev = ts[i], ts[i - 1]
  act = cur[:, 0] > 0
  pact = prev[:, 0] > 0
  prev_exit = prev[pact, 7].max()
  prev_mma_end = prev[pact, 4].max()
  prev_epi_end = prev[pact, 6].max()
  d = lambda v: (v - prev_exit) / 1e3
  rows[site].append([
    d(prev_mma_end), d(prev_epi_end),                    # previous Linear: last MMA committed, last epilogue store (<= 0)
    d(cur[act, 0].max()),                                # this Linear: last block resident
    d(cur[act, 1].max()),                                # weight stages issued everywhere
    d(cur[act, 2].min()), d(cur[act, 2].max()),          # dependency resolved (first / last block)
    d(np.median(cur[act, 3])),                           # first MMA (median block)
    d(cur[act, 4].min()), d(cur[act, 4].max()),          # last MMA committed (first / last block)
    d(cur[act, 7].max()),                                # last block exit
    (cur[act, 4].max() - np.median(cur[act, 3])) / 1e3,  # MMA window
  ])
hdr = ["prev lastMMA", "prev epi end", "resident", "w issued", "dep first", "dep last", "1st MMA med", "lastMMA first", "lastMMA last", "exit", "MMA window"]
print(f"{'site':8s} " + " ".join(f"{h:>13s}" for h in hdr))
for k in names:
  m = np.mean(np.array(rows[k]), axis=0)
  print(f"{k:8s} " + " ".join(f"{v:13.2f}" for v in m))

# the tail of a launch: per block, relative to its own "last stage landed" stamp
print("tail per block (us, median over blocks and layers): last load issued -> last stage landed -> last MMA issued -> accumulator complete -> epilogue stored -> exit | staged, barrier, bulk issued, bulk read")
for k_i, k in enumerate(names):
  acc = []
  for i in range(5 + (k_i - 5) % 4, n - 1, 4):
    cur = ts[i]
    act = (cur[:, 0] > 0) & (cur[:, 10] > 0)
    ref = cur[act, 10]
    acc.append([np.median(cur[act, 9] - ref), 0.0, np.median(cur[act, 4] - ref), np.median(cur[act, 8] - ref),
                np.median(cur[act, 6] - ref), np.median(cur[act, 7] - ref)] + [np.median(cur[act, k2] - ref) for k2 in (11, 12, 13, 14)])
  m = np.mean(np.array(acc), axis=0) / 1e3
  print(f"{k:8s} " + " ".join(f"{v:9.2f}" for v in m))

# per-block detail of one gate/up launch (layer 8) and one down launch: which SMs finish late?
for i in ():
  cur = ts[i]
  act = cur[:, 0] > 0
  t0 = cur[act, 2].min()
  rowsb = sorted([(int(cur[b, 4] - t0), int(cur[b, 5]), b, int(cur[b, 3] - t0)) for b in np.nonzero(act)[0]])
  print(f"launch {i} ({names[i % 4]}): blocks sorted by last-MMA time after the dependency resolved (ns): (t_lastMMA, smid, block, t_firstMMA)")
  print(" ".join(f"({t},{sm},{b},{f})" for t, sm, b, f in rowsb))


# how uneven is the finish across blocks, and is it the same SMs every time?  (last MMA committed, relative to the
# dependency release of the launch; normalised by the launch's median)
print("finish spread per site: percentiles over blocks of (last MMA - dep release) in us, mean over layers")
per_sm = {}
for k_i, k in enumerate(names):
  pct = []
  for i in range(5 + (k_i - 5) % 4, n - 1, 4):
    cur = ts[i]
    act = cur[:, 0] > 0
    t0 = cur[act, 2].min()
    fin = (cur[act, 4] - t0) / 1e3
    pct.append(np.percentile(fin, [0, 5, 25, 50, 75, 95, 100]))
    med = np.median(fin)
    for sm, f in zip(cur[act, 5], fin):
      per_sm.setdefault((k, int(sm)), []).append(f / med)
  m = np.mean(np.array(pct), axis=0)
  print(f"{k:8s} min {m[0]:6.2f}  p5 {m[1]:6.2f}  p25 {m[2]:6.2f}  p50 {m[3]:6.2f}  p75 {m[4]:6.2f}  p95 {m[5]:6.2f}  max {m[6]:6.2f}")
for k in names:
  stats = sorted((float(np.mean(v)), float(np.std(v)), sm) for (kk, sm), v in per_sm.items() if kk == k)
  print(f"{k}: relative finish time by SM (mean over layers, std, smid): fastest " +
        " ".join(f"({m:.2f},{sd:.2f},{sm})" for m, sd, sm in stats[:8]) + " ... slowest " +
        " ".join(f"({m:.2f},{sd:.2f},{sm})" for m, sd, sm in stats[-8:]))
  print(f"   all SMs, sorted by smid: " + " ".join(f"{np.mean(per_sm[(k, sm)]):.2f}" for sm in sorted(sm for (kk, sm) in per_sm if kk == k)))
