"""The decode hand-off kernels (xb_resid_send_kernel / xb_recv_norm_kernel and their flag protocol) and
the other boundary rules of the C ABI, on ONE GPU: two pipeline stages of one process connected with
xb_handoff_connect_local -- the same kernels and flags a multi-GPU ring runs over NVLink, with the
next stage's landing buffer on the same device -- against the plain ``hidden=`` path bit for bit.
On a box with two GPUs the same test also runs with the stages on different devices (peer stores)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup(mid):
  import json, tempfile
  from pathlib import Path
  from xotorch_support_jetson_b200 import model_config as mc
  from xotorch_support_jetson_b200.models import model_cards
  from xotorch_support_jetson_b200.synth import random_hf_state
  hf_cfg = model_cards[mid]["hf_config"]
  with tempfile.TemporaryDirectory() as td:
    (Path(td) / "config.json").write_text(json.dumps(hf_cfg))
    cfg = mc.load_model_config(Path(td) / "config.json")
  return cfg, random_hf_state(hf_cfg, seed=0, std=0.05, norm_jitter=0.1)


def _stage(mid, cfg, state, a, b, device=0, **kw):
  from xotorch_support_jetson_b200.shard import Shard
  from xotorch_support_jetson_b200.stage import StageRuntime
  st = StageRuntime(Shard(mid, a, b, cfg["num_layers"]), cfg, device=device, max_batch_tokens=64, max_seqs=8,
                    max_tokens_per_seq=128, page_tokens=16, **kw)
  st.load_state(state)
  return st


def _move(t, dev):
  """Copy across devices and wait on both, so the consumer stream may use the result at once."""
  if t.device == dev:
    return t
  out = t.to(dev)
  for d in (t.device, dev):
    if d.type == "cuda":
      torch.cuda.synchronize(d)
  return out


def _devices():
  return [(0, 0)] + ([(0, 1)] if torch.cuda.device_count() >= 2 else [])


@pytest.mark.parametrize("mid,cut", [
  ("llama-3.1-tiny", dict(a=(0, 2), b=(3, 5))),
  ("qwen-2.5-tiny", dict(a=(0, 1), b=(2, 3))),
  # half-layer cut: stage A stops after the attention block of layer 2, stage B starts at its MLP block
  ("llama-3.1-tiny", dict(a=(0, 2), b=(2, 5), half=True)),
])
@pytest.mark.parametrize("devs", _devices(), ids=lambda d: f"gpu{d[0]}-gpu{d[1]}")
def test_local_handoff_equals_hidden_path_bit_for_bit(built_lib, mid, cut, devs):
  cfg, state = _setup(mid)
  half = cut.get("half", False)
  B, L, K, M = 3, 19, 5, 2
  g = torch.Generator().manual_seed(5)
  prompts = [torch.randint(1, cfg["vocab_size"], (B, L), generator=g, dtype=torch.int32) for _ in range(M)]

  def build():
    sa = _stage(mid, cfg, state, *cut["a"], device=devs[0], last_to_attn=half)
    sb = _stage(mid, cfg, state, *cut["b"], device=devs[1], first_from_mlp=half)
    return sa, sb

  def run(handoff: bool):
    sa, sb = build()
    if handoff:
      sa.handoff_create(M, B, ipc=False)
      sb.handoff_create(M, B, ipc=False)
      sa.handoff_connect_local(sb)
    stream_a, stream_b = torch.cuda.Stream(device=devs[0]), torch.cuda.Stream(device=devs[1])
    logits_out, toks_out = [], []
    toks = []
    for m in range(M):     # prefill through the plain path (the hand-off carries decode steps only)
      slots = list(range(m * B, (m + 1) * B))
      with torch.cuda.stream(stream_a):
        h = sa.forward(slots, [L] * B, [0] * B, tokens=_move(prompts[m], sa.device).reshape(-1))
      stream_a.synchronize()
      with torch.cuda.stream(stream_b):
        lg = sb.forward(slots, [L] * B, [0] * B, hidden=_move(h, sb.device))
        toks.append(sb.sample(lg, temp=0.0))
      stream_b.synchronize()
    for k in range(K):
      for m in range(M):
        slots = list(range(m * B, (m + 1) * B))
        pos = [L + k] * B
        tin = _move(toks[m], sa.device)
        if handoff:
          v = k + 1                                   # per-slot flag value: only grows
          with torch.cuda.stream(stream_a):
            sa.forward(slots, [1] * B, pos, tokens=tin, handoff=(-1, v, m, v))
          with torch.cuda.stream(stream_b):           # no host synchronisation between the stages: the flag orders them
            lg = sb.forward(slots, [1] * B, pos, handoff=(m, v, -1, v))
            toks[m] = sb.sample(lg, temp=0.0)
        else:
          with torch.cuda.stream(stream_a):
            h = sa.forward(slots, [1] * B, pos, tokens=tin)
          stream_a.synchronize()
          with torch.cuda.stream(stream_b):
            lg = sb.forward(slots, [1] * B, pos, hidden=_move(h, sb.device))
            toks[m] = sb.sample(lg, temp=0.0)
        stream_b.synchronize()
        logits_out.append(lg.float().cpu().numpy().copy())
        toks_out.append(toks[m].cpu().numpy().copy())
    n_kernels = sa.counters()["launches"] + sb.counters()["launches"]
    sa.close()
    sb.close()
    return logits_out, toks_out, n_kernels

  ref_l, ref_t, _ = run(False)
  got_l, got_t, _ = run(True)
  assert len(ref_l) == K * M
  for a, b in zip(got_l, ref_l):
    assert np.array_equal(a, b)          # same arithmetic, other plumbing: identical bits
  for a, b in zip(got_t, ref_t):
    assert np.array_equal(a, b)


def test_handoff_timeout_is_reported_not_fatal(built_lib, monkeypatch):
  """A consumer whose producer never signals: the step finishes after the (here: 50 ms) time-out, the
  NEXT call reports it, xb_handoff_clear_error() makes the engine usable again -- the CUDA context
  survives (a trapped kernel would have poisoned every later call of the process)."""
  monkeypatch.setenv("XB_HANDOFF_TIMEOUT_S", "0.05")
  mid = "llama-3.1-tiny"
  cfg, state = _setup(mid)
  sb = _stage(mid, cfg, state, 3, 5)
  sb.handoff_create(1, 2, ipc=False)
  s = torch.cuda.Stream()
  h = torch.zeros((2, cfg["embed_dim"]), dtype=torch.bfloat16, device="cuda")
  with torch.cuda.stream(s):
    sb.forward([0, 1], [1, 1], [0, 0], hidden=h)                      # plain step: fine
    sb.forward([0, 1], [1, 1], [1, 1], handoff=(0, 7, -1, 7))           # waits for a flag nobody sets
  s.synchronize()
  with pytest.raises(RuntimeError, match="timed out waiting for landing slot 0 to reach 7"):
    with torch.cuda.stream(s):
      sb.forward([0, 1], [1, 1], [2, 2], hidden=h)
  sb.handoff_clear_error()
  with torch.cuda.stream(s):
    out = sb.forward([0, 1], [1, 1], [2, 2], hidden=h)
  s.synchronize()
  assert torch.isfinite(out).all()
  sb.close()


def test_failed_batch_leaves_slots_untouched(built_lib):
  """A batch that is refused changes nothing: the sequences examined before the failing one keep their
  cached length and pages, so a later forward cannot attend rows that were never written."""
  from xotorch_support_jetson_b200.shard import Shard
  from xotorch_support_jetson_b200.stage import StageRuntime
  mid = "llama-3.2-tiny"
  cfg, state = _setup(mid)
  st = StageRuntime(Shard(mid, 0, 3, 4), cfg, device=0, max_batch_tokens=64, max_seqs=2, max_tokens_per_seq=32, page_tokens=16)
  st.load_state(state)
  t = lambda n: torch.arange(1, n + 1, dtype=torch.int32, device="cuda")
  st.forward([0], [8], [0], tokens=t(8))
  free0 = st.counters()["kv_pages_free"]
  # both sequences in bounds individually, but the position gap of the second makes the call fail after
  # the first was examined
  with pytest.raises(RuntimeError, match="gap"):
    st.forward([0, 1], [20, 4], [8, 5], tokens=t(24))
  assert st.slot_len(0) == 8 and st.slot_len(1) == 0 and st.counters()["kv_pages_free"] == free0
  with pytest.raises(RuntimeError, match="gap"):
    st.forward([0], [1], [28], tokens=t(1))      # would have been accepted had slot 0 been advanced to 28
  st.forward([0], [1], [8], tokens=t(1))
  assert st.slot_len(0) == 9
  st.close()


def test_stale_descriptor_is_refused(built_lib):
  import ctypes as C
  from xotorch_support_jetson_b200 import _cabi
  d = _cabi.model_desc(n_layers=4, start_layer=0, end_layer=3, dim=256, n_heads=8, n_kv_heads=2, head_dim=64, inter_dim=512,
                       vocab=1024, norm_eps=1e-5, max_positions=64, max_batch_tokens=16, max_seqs=2)
  d.struct_size = 76     # the ABI-v2 layout
  h = C.c_void_p()
  assert built_lib.xb_engine_create(C.byref(d), 0, C.byref(h)) != 0
  assert b"struct_size is 76" in built_lib.xb_last_error()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs at least 2 GPUs")
def test_two_engines_on_two_devices_in_one_process(built_lib):
  """Kernel attributes (dynamic shared memory above 48 KB) are per device: a second engine on another
  GPU of the same process must run every kernel class -- prefill, graph-replayed decode, sampler --
  and agree with the first bit for bit."""
  mid = "llama-3.1-tiny"
  cfg, state = _setup(mid)
  outs = []
  for dev in (0, 1):
    st = _stage(mid, cfg, state, 0, 5, device=dev)
    with torch.cuda.device(dev):
      s = torch.cuda.Stream(device=dev)
      g = torch.Generator().manual_seed(3)
      prompt = torch.randint(1, cfg["vocab_size"], (2, 30), generator=g, dtype=torch.int32).to(st.device)
      with torch.cuda.stream(s):
        lg = st.forward([0, 1], [30, 30], [0, 0], tokens=prompt.reshape(-1))
        seq = [lg.float().cpu()]
        tok = st.sample(lg, temp=0.7, top_k=20, seed=1)
        for k in range(4):
          lg = st.forward([0, 1], [1, 1], [30 + k, 30 + k], tokens=tok)
          tok = st.sample(lg, temp=0.7, top_k=20, seed=2 + k)
          seq.append(lg.float().cpu())
          seq.append(tok.cpu().float())
      s.synchronize()
    outs.append(seq)
    st.close()
  for a, b in zip(*outs):
    assert torch.equal(a, b)
