"""Size-independent properties at BASELINE.json's full size (Llama-3.1-8B and Qwen2.5-7B bf16, 2048-token prompt),
where the CPU oracle is too slow to be the checker:
  * a 2-shard split (layers 0-15 | 16-31, hidden state handed over in bf16) gives EXACTLY the
    full-shard logits and greedy tokens (the reference's contract,
    xotorch/inference/test_inference_engine.py:11,43-44);
  * chunked prefill (2 x 1024) equals the one-shot prefill within the stated tolerance;
  * a cached decode step and the last position of a prefill over the extended prompt are both as
    close to an fp32 evaluation of the same weights as the bf16 oracle is (the oracle, in fp32 and
    in bf16, is run on the GPU through torch here purely as the checker: on the CPU a 2048-token
    8B prefill takes tens of minutes)."""
import json
import tempfile
from pathlib import Path

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
REL_TOL = 0.04
# Engine vs the bf16 oracle at FULL size, stated at about twice the worst case measured on the B200 (2048-token
# prefill + a cached decode step; 1B / 3B / 8B / Qwen-7B / 70B width):
#   logits: max-abs error / max-abs of the logits -- measured 0.021 .. 0.054 (8B: 0.046, absolute 0.26 on +-5.6)
#   residual stream after the first / a middle / the last layer -- measured max-abs/max-abs 0.006 .. 0.044,
#   relative Frobenius error 0.004 .. 0.046 (8B layer 31: 0.044 / 0.046; the bf16 oracle itself sits 0.059 / 0.062
#   from the fp32 evaluation there, the engine 0.060 / 0.062)
# This is larger than the 4 % of the tiny models because the error of each bf16 evaluation against the exact
# answer grows with depth and width (0.03 at 16 layers of 2048, 0.06 at 32 layers of 4096) and two evaluations that
# round in different orders can be up to the sum of their errors apart.
FULL_LOGIT_TOL = 0.10
FULL_HIDDEN_TOL = 0.09
FULL_HIDDEN_FRO_TOL = 0.09


def _rel(a, b):
  return float((a.float() - b.float()).abs().max() / b.float().abs().max())


@pytest.mark.parametrize("mid,n_layers", [
  ("llama-3.1-8b", None), ("qwen-2.5-7b", None),
  ("llama-3.2-1b", None),     # tied head, head_dim 64
  ("llama-3.2-3b", None),     # tied head, 3 query heads per kv head
  ("llama-3.1-70b", 8),       # 70B geometry (D 8192, 8 query heads per kv head, I 28672), first 8 layers as the model
])
def test_fullsize_properties(built_lib, mid, n_layers):
  from xotorch_support_jetson_b200 import model_config as mc
  from xotorch_support_jetson_b200.models import model_cards
  from xotorch_support_jetson_b200.pipeline import gpu_random_getter
  from xotorch_support_jetson_b200.shard import Shard
  from xotorch_support_jetson_b200.stage import StageRuntime
  if torch.cuda.get_device_properties(0).total_memory < 60e9:
    pytest.skip("needs ~40 GB of HBM")
  hf = dict(model_cards[mid]["hf_config"])
  if n_layers is not None:
    hf["num_hidden_layers"] = n_layers
  with tempfile.TemporaryDirectory() as td:
    (Path(td) / "config.json").write_text(json.dumps(hf))
    cfg = mc.load_model_config(Path(td) / "config.json")
  dev = torch.device("cuda", 0)
  L, nl = 2048, hf["num_hidden_layers"]
  cut = nl // 2
  kw = dict(device=0, max_batch_tokens=4096, max_seqs=3, max_tokens_per_seq=L + 64)
  full = StageRuntime(Shard(mid, 0, nl - 1, nl), cfg, **kw)
  s1 = StageRuntime(Shard(mid, 0, cut - 1, nl), cfg, **kw)
  s2 = StageRuntime(Shard(mid, cut, nl - 1, nl), cfg, **kw)
  for st in (full, s1, s2):
    st.load_hf_tensors(*gpu_random_getter(hf, 0, 0.02, dev))
  g = torch.Generator(device=dev).manual_seed(4321)
  prompt = torch.randint(1, cfg["vocab_size"], (L + 1,), generator=g, device=dev, dtype=torch.int32)
  with torch.cuda.stream(torch.cuda.Stream()):
    # one-shot prefill, full shard vs 2-shard ring; residual stream tapped after the first, a middle and the last layer
    tap_layers = (0, nl // 2 - 1, nl - 1)
    taps = [torch.empty((L, cfg["embed_dim"]), dtype=torch.bfloat16, device=dev) if l in tap_layers else None for l in range(nl)]
    lg_full = full.forward([0], [L], [0], tokens=prompt[:L], taps=taps)
    eng_taps = {l: taps[l].float() for l in tap_layers}
    h = s1.forward([0], [L], [0], tokens=prompt[:L])
    lg_ring = s2.forward([0], [L], [0], hidden=h)
    assert torch.equal(lg_full, lg_ring), "2-shard split differs from the full shard"
    # chunked prefill in slot 1
    full.forward([1], [1024], [0], tokens=prompt[:1024])
    lg_chunk = full.forward([1], [1024], [1024], tokens=prompt[1024:L])
    e_chunk = _rel(lg_chunk, lg_full)
    # decode step at position L vs prefill of L+1 tokens (slot 2)
    lg_dec = full.forward([0], [1], [L], tokens=prompt[L:L + 1])
    lg_ext = full.forward([2], [L + 1], [0], tokens=prompt)
    e_dec = _rel(lg_dec, lg_ext)
    # greedy: full shard vs ring, 8 tokens (decode path: CUDA-graph replays on both)
    full.release_slot(1)
    full.release_slot(2)
    full.release_slot(0)
    s1.release_slot(0)
    s2.release_slot(0)
    lg_a = full.forward([0], [L], [0], tokens=prompt[:L])
    lg_b = s2.forward([0], [L], [0], hidden=s1.forward([0], [L], [0], tokens=prompt[:L]))
    ta, tb = [], []
    for k in range(8):
      xa = full.sample(lg_a, temp=0.0)
      xb = s2.sample(lg_b, temp=0.0)
      ta.append(int(xa.item()))
      tb.append(int(xb.item()))
      lg_a = full.forward([0], [1], [L + k], tokens=xa)
      lg_b = s2.forward([0], [1], [L + k], hidden=s1.forward([0], [1], [L + k], tokens=xb))
  torch.cuda.synchronize()
  print(f"\n[{mid} properties] chunked-vs-one-shot prefill rel err {e_chunk:.4f}; decode-vs-prefill rel err {e_dec:.4f}; "
        f"tokens full {ta} ring {tb}")
  assert ta == tb
  assert torch.isfinite(lg_full).all() and e_chunk < REL_TOL / 2 and e_dec < 2 * REL_TOL
  for st in (full, s1, s2):
    st.close()
  del full, s1, s2
  torch.cuda.empty_cache()

  # ---- the oracle at full size (fp32 = truth, bf16 = the reference's own precision)
  import sys
  sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
  from oracle import torch_oracle as O
  from xotorch_support_jetson_b200.synth import hf_param_shapes
  torch.backends.cuda.matmul.allow_bf16_reduced_precision_reduction = False
  torch.backends.cuda.matmul.allow_tf32 = False
  get, _ = gpu_random_getter(hf, 0, 0.02, dev)
  state = {n: get(n) for n in hf_param_shapes(hf)}
  ocfg = O.model_config_from_hf(hf)
  cpu_prompt = prompt.cpu().long().reshape(1, -1)

  def oracle_run(dtype, opts, n_tok, want_taps=False):
    """Last-position logits (and the residual stream after the tapped layers) of an n_tok-token prefill."""
    m = O.OracleShardModel(ocfg, mid, 0, nl - 1, state, opts, dtype=dtype, device="cuda")
    e = O.OracleEngine(m)
    e.start_request(cpu_prompt[:, :n_tok].to("cuda"))
    taps = {} if want_taps else None
    lg = e.infer_tensor(cpu_prompt[:, :n_tok].numpy(), taps=taps)
    hs = {l: taps[l][0].float() for l in tap_layers} if want_taps else None
    del m, e, taps
    torch.cuda.empty_cache()
    return torch.from_numpy(np.asarray(lg[0], dtype=np.float32)), hs

  o32 = O.OracleOptions(rope_table_dtype="fp32")
  truth_L, truth_h = oracle_run(torch.float32, o32, L, True)
  truth_L1, _ = oracle_run(torch.float32, o32, L + 1)
  ref_L, ref_h = oracle_run(torch.bfloat16, O.OracleOptions(), L, True)
  ref_L1, _ = oracle_run(torch.bfloat16, O.OracleOptions(), L + 1)
  pre, dec, ext = lg_full.float().cpu()[0], lg_dec.float().cpu()[0], lg_ext.float().cpu()[0]
  e = {"prefill": _rel(pre, truth_L), "decode": _rel(dec, truth_L1), "prefill+1": _rel(ext, truth_L1),
       "oracle bf16 L": _rel(ref_L, truth_L), "oracle bf16 L+1": _rel(ref_L1, truth_L1),
       "prefill vs oracle bf16": _rel(pre, ref_L), "decode vs oracle bf16": _rel(dec, ref_L1)}
  print(f"[{mid} vs oracle] logits, max-abs error / max-abs: " + "; ".join(f"{k} {v:.4f}" for k, v in e.items()))
  print(f"[{mid} vs oracle] logits, max-abs error: prefill vs oracle bf16 {float((pre - ref_L).abs().max()):.4f} "
        f"(logits span +-{float(ref_L.abs().max()):.3f}); decode vs oracle bf16 {float((dec - ref_L1).abs().max()):.4f}")
  # per-layer hidden states (north_star: max-abs and relative error reported): engine vs the bf16 oracle, and each
  # of them vs the fp32 evaluation, over all 2048 rows of the prefill
  fro = lambda a, b: float((a - b).norm() / b.norm())
  worst_h_rel, worst_h_fro = 0.0, 0.0
  for l in tap_layers:
    a, r, t = eng_taps[l], ref_h[l], truth_h[l]
    h_rel, h_fro = _rel(a, r), fro(a, r)
    worst_h_rel, worst_h_fro = max(worst_h_rel, h_rel), max(worst_h_fro, h_fro)
    print(f"[{mid} vs oracle] hidden after layer {l:2d}: engine vs oracle bf16 max-abs {float((a - r).abs().max()):.4f} "
          f"(/max-abs {h_rel:.4f}, rel-Frobenius {h_fro:.4f}); vs fp32: engine {_rel(a, t):.4f} / {fro(a, t):.4f}, "
          f"oracle bf16 {_rel(r, t):.4f} / {fro(r, t):.4f}")
  worst_ref = max(e["oracle bf16 L"], e["oracle bf16 L+1"])
  assert e["prefill"] < 1.5 * worst_ref + 0.005
  assert e["decode"] < 1.5 * worst_ref + 0.005
  assert e["prefill+1"] < 1.5 * worst_ref + 0.005
  # Stated tolerance at full size: two bf16 evaluations that each sit ~0.06 of max-abs from
  # the exact answer after 28-32 layers and 2048 positions can be ~sqrt(2) x that apart from each other
  assert e["prefill vs oracle bf16"] < FULL_LOGIT_TOL and e["decode vs oracle bf16"] < FULL_LOGIT_TOL
  assert worst_h_rel < FULL_HIDDEN_TOL and worst_h_fro < FULL_HIDDEN_FRO_TOL
