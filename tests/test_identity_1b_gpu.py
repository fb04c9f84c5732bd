"""BASELINE.json configs[0]: Llama-3.2-1B random-init, full shard 0-15, 128-token prompt + 64
greedy tokens, CUDA engine against the oracle (the reference-equivalent CPU torch engine).

Random-init logits over V = 128256 are bf16 values with frequent EXACT top-2 ties (measured
here: the oracle's min top-2 margin over 64 steps is 0.0 for every init scale tried, SURVEY.md
section 7 predicted 98.5 %), so "identical 64 tokens" is only well defined on a checkpoint with
peaked logits.  Two checks:
  1. peaked checkpoint (embedding table scaled so the tied head keeps the last token dominant;
     oracle min margin ~50 bf16 ulps): 64 free-running greedy tokens must be IDENTICAL;
  2. plain N(0, 0.02^2) checkpoint: teacher-forced on the oracle's tokens, every step's
     hidden states / logits within the stated bf16 tolerance, the CUDA token identical wherever
     the oracle's top-2 margin exceeds 4 bf16 ulps of the top logit, and otherwise one of the
     oracle's near-tied candidates.  Max-abs / relative errors and margins are printed."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

REL_TOL = 0.04   # same stated tolerance as tests/test_engine_gpu.py
L, N = 128, 64


def _setup(embed_std):
  from oracle import torch_oracle as O
  from xotorch_support_jetson_b200.models import model_cards
  from xotorch_support_jetson_b200.shard import Shard
  from xotorch_support_jetson_b200.stage import StageRuntime
  from xotorch_support_jetson_b200.synth import random_hf_state
  mid = "llama-3.2-1b"
  hf_cfg = model_cards[mid]["hf_config"]
  cfg = O.model_config_from_hf(hf_cfg)
  state = random_hf_state(hf_cfg, seed=0, std=0.02, embed_std=embed_std)
  nl = cfg["num_layers"]
  g = torch.Generator().manual_seed(1234)
  prompt = torch.randint(1, cfg["vocab_size"], (1, L), generator=g)
  ref = O.OracleShardModel(cfg, mid, 0, nl - 1, state, O.OracleOptions())
  st = StageRuntime(Shard(mid, 0, nl - 1, nl), cfg, device=0, max_batch_tokens=256, max_seqs=2, max_tokens_per_seq=256)
  st.load_state(state)
  return O, cfg, ref, st, prompt


def test_llama32_1b_free_running_64_tokens_identical_on_peaked_checkpoint(built_lib):
  O, cfg, ref, st, prompt = _setup(embed_std=0.6)
  ref_toks, ref_logits, _ = O.greedy_generate([ref], prompt, N, collect_logits=True)
  with torch.cuda.stream(torch.cuda.Stream()):
    logits = st.forward([0], [L], [0], tokens=prompt.to(torch.int32).cuda().reshape(-1))
    toks = []
    for i in range(N):
      t = st.sample(logits, temp=0.0)
      toks.append(int(t.item()))
      if i < N - 1:
        logits = st.forward([0], [1], [L + i], tokens=t)
  torch.cuda.synchronize()
  srt = np.sort(np.stack([x[0] for x in ref_logits]), axis=-1)
  ulps = ((srt[:, -1] - srt[:, -2]) / (np.abs(srt[:, -1]) * 2.0 ** -8)).min()
  print(f"\n[1B peaked] oracle min top-2 margin {ulps:.1f} bf16 ulps; tokens {toks[:6]}...")
  assert toks == ref_toks[0].tolist()
  st.close()


def test_llama32_1b_teacher_forced_64_steps_within_tolerance(built_lib):
  O, cfg, ref, st, prompt = _setup(embed_std=None)
  nl, D = cfg["num_layers"], cfg["embed_dim"]
  ref_toks, ref_logits, ref_taps = O.greedy_generate([ref], prompt, N, collect_logits=True, taps_at=(0, 1))
  ref_l = np.stack([x[0] for x in ref_logits])
  got = []
  with torch.cuda.stream(torch.cuda.Stream()):
    taps = [torch.empty((L, D), dtype=torch.bfloat16, device="cuda") for _ in range(nl)]
    logits = st.forward([0], [L], [0], tokens=prompt.to(torch.int32).cuda().reshape(-1), taps=taps)
    hid_abs = max(float(np.abs(taps[l].float().cpu().numpy() - ref_taps[0][l][0]).max()) for l in range(nl))
    hid_rel = max(float(np.abs(taps[l].float().cpu().numpy() - ref_taps[0][l][0]).max() / np.abs(ref_taps[0][l]).max()) for l in range(nl))
    got.append(logits.cpu().numpy()[0])
    free_prefix, free_ok = 0, True
    for i in range(N - 1):
      if free_ok and int(st.sample(logits, temp=0.0).item()) == int(ref_toks[0, i]):
        free_prefix += 1
      else:
        free_ok = False
      x = torch.tensor([int(ref_toks[0, i])], dtype=torch.int32, device="cuda")
      if i == 0:
        taps1 = [torch.empty((1, D), dtype=torch.bfloat16, device="cuda") for _ in range(nl)]
        logits = st.forward([0], [1], [L + i], tokens=x, taps=taps1)
        hid_rel = max(hid_rel, max(float(np.abs(taps1[l].float().cpu().numpy() - ref_taps[1][l][0]).max() /
                                         np.abs(ref_taps[1][l]).max()) for l in range(nl)))
      else:
        logits = st.forward([0], [1], [L + i], tokens=x)
      got.append(logits.cpu().numpy()[0])
  torch.cuda.synchronize()
  got = np.stack(got)
  max_abs = float(np.abs(got - ref_l).max())
  rel = float((np.abs(got - ref_l).max(axis=-1) / np.abs(ref_l).max(axis=-1)).max())
  srt = np.sort(ref_l, axis=-1)
  top, margin = srt[:, -1], srt[:, -1] - srt[:, -2]
  ulp = np.abs(top) * 2.0 ** -8
  mine = got.argmax(-1)
  strict = margin > 4 * ulp
  n_strict_ok = int((mine[strict] == ref_toks[0][strict]).sum())
  print(f"\n[1B plain] logits max-abs err {max_abs:.4f} rel {rel:.4f}; hidden max-abs {hid_abs:.4f} rel {hid_rel:.4f}; "
        f"oracle margins: min {margin.min():.4f} ({(margin / ulp).min():.1f} ulps), {int(strict.sum())}/{N} steps above 4 ulps, "
        f"{n_strict_ok} of those identical; all-steps identical {int((mine == ref_toks[0]).sum())}/{N}; "
        f"free-running identical prefix {free_prefix}")
  # accuracy against an fp32 evaluation of the same weights: the CUDA engine must be no further
  # from the fp32 answer than the bf16 reference path itself is (prefill logits)
  from xotorch_support_jetson_b200.models import model_cards
  from xotorch_support_jetson_b200.synth import random_hf_state
  state32 = random_hf_state(model_cards["llama-3.2-1b"]["hf_config"], seed=0, std=0.02)
  ref32 = O.OracleShardModel(cfg, "llama-3.2-1b", 0, nl - 1, state32, O.OracleOptions(rope_table_dtype="fp32"), dtype=torch.float32)
  _, l32, _ = O.greedy_generate([ref32], prompt, 1, collect_logits=True)
  truth = l32[0][0]
  e_gpu = float(np.abs(got[0] - truth).max() / np.abs(truth).max())
  e_ref = float(np.abs(ref_l[0] - truth).max() / np.abs(truth).max())
  print(f"[1B plain] prefill logits vs fp32 evaluation: CUDA engine rel err {e_gpu:.4f}, bf16 oracle rel err {e_ref:.4f}")
  assert e_gpu < 1.5 * e_ref + 0.005
  assert hid_rel < REL_TOL and rel < REL_TOL
  assert n_strict_ok == int(strict.sum())
  for i in range(N):   # near-ties: the CUDA pick must be one of the oracle's near-top candidates
    assert ref_l[i, mine[i]] >= top[i] - 4 * ulp[i] - 1e-6, (i, mine[i], ref_toks[0, i])
  st.close()
