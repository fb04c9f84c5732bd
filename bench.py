#!/usr/bin/env python
"""bench.py -- decode / prefill throughput of the layer-range forward on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

Workload (BASELINE.json configs[2], the one the metric is quoted on): Llama-3.1-8B bf16,
random-init weights, synthetic 2048-token prompts, decode at context 2048.., batch 32 per
pipeline micro-batch.  A "step" is one decode step of every in-flight sequence.  At N GPUs
the model is an N-stage layer pipeline (stage boundaries at half-layer granularity balanced by
bytes per step; XB_HALF_LAYERS=0: the whole-layer ranges the reference's ring strategy assigns)
with N micro-batches of 32 sequences in flight: per-GPU work is fixed -> "weak" scaling.

One JSON line on stdout (rank 0).  `value` = decode tokens/s with ids resident in HBM;
`e2e` = the same metric through the public API with host buffers; `roofline` = the
dominant kernel (tcgen05 weight-streaming Linear) against the measured HBM peak;
`cpu_baseline` = the oracle (reference-equivalent eager PyTorch) on the host cores;
`extra` carries decode bs=1 and prefill tok/s with their roofline fractions.

`--impl reference` times the reference's own CPU path (the oracle port -- torchtune is not
installable here) on the same config with a bounded per-step sample.
"""
import argparse
import asyncio
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

MODEL = "llama-3.1-8b"
PROMPT_LEN = 2048
BATCH = 32


def measured_peaks():
  p = ROOT / "MEASURED_PEAKS.json"
  if p.exists():
    d = json.loads(p.read_text())
    return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"], "bf16_tflops_sustained": d["bf16_tflops_sustained"],
            "source": "measured"}
  return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
  """nvidia-smi clocks / throttle reasons during the timed region."""
  Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
       "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

  def __init__(self, gpu_index: int):
    self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
    self.p = None
    self.idx = gpu_index

  def start(self):
    try:
      self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                 "-i", str(self.idx)], stdout=self.f, stderr=subprocess.DEVNULL)
    except Exception:
      self.p = None

  def stop(self) -> dict:
    if self.p is None:
      return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
    time.sleep(0.25)
    self.p.terminate()
    try:
      self.p.wait(timeout=5)
    except Exception:
      self.p.kill()
    self.f.flush()
    rows = [r.split(",") for r in Path(self.f.name).read_text().strip().splitlines() if r.strip()]
    os.unlink(self.f.name)
    sm, smax, reasons = [], [], set()
    names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    for r in rows:
      try:
        sm.append(float(r[1]))
        smax.append(float(r[2]))
        for n, v in zip(names, r[5:9]):
          if "Active" in v and "Not" not in v:
            reasons.add(n)
      except Exception:
        pass
    busy = [x for x in sm if x > 0]
    return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": max(smax) if smax else None,
            "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------- reference / CPU baseline
class CpuDecodeSample:
  """The oracle (reference-exact eager PyTorch: per-op bf16 modules, contiguous
  [B, KV, prompt+1024, hd] cache, GQA expand, full-cache masked SDPA, top-k-35 sampler) on the
  host cores: decode steps of `sample_batch` sequences at context `ctx`.  The cache is filled
  with synthetic K/V instead of running the 2048-token prefill on the CPU (the decode cost does
  not depend on the cached values)."""

  def __init__(self, hf_cfg, model_id: str, sample_batch: int, ctx: int, seed: int = 0):
    import numpy as np
    import torch
    from oracle import torch_oracle as O
    from xotorch_support_jetson_b200.synth import random_hf_state
    torch.set_num_threads(os.cpu_count() or 1)
    cfg = O.model_config_from_hf(hf_cfg)
    nl = cfg["num_layers"]
    weights_cpu = random_hf_state(hf_cfg, seed=seed)
    self.model = O.OracleShardModel(cfg, model_id, 0, nl - 1, weights_cpu, O.OracleOptions())
    del weights_cpu
    self.eng = O.OracleEngine(self.model)
    self.B = sample_batch
    g = torch.Generator().manual_seed(1234)
    prompt = torch.randint(1, cfg["vocab_size"], (self.B, ctx), generator=g)
    self.eng.start_request(prompt)   # allocates the [B, KV, ctx+1024, hd] caches, mask and input_pos
    # one random [1, KV, ctx, hd] block broadcast into every layer's cache: drawing 13 GB of normals at batch 32
    # took minutes of set-up, and the cost of a decode step does not depend on the cached values
    kv0 = next(iter(self.model.k_cache.values()))
    blk_k = torch.empty((1, kv0.shape[1], ctx, kv0.shape[3]), dtype=torch.float32).normal_(0, 0.5, generator=g).to(kv0.dtype)
    blk_v = torch.empty((1, kv0.shape[1], ctx, kv0.shape[3]), dtype=torch.float32).normal_(0, 0.5, generator=g).to(kv0.dtype)
    for l in self.model.k_cache:
      self.model.k_cache[l][:, :, :ctx].copy_(blk_k)
      self.model.v_cache[l][:, :, :ctx].copy_(blk_v)
    self.model.cache_pos = ctx
    self.eng.curr_pos = ctx
    self.tok = np.ones((self.B, 1), dtype=np.int64)

  def steps(self, n: int):
    """Runs n decode steps (infer_tensor + sample), returns the list of per-step seconds."""
    import numpy as np
    times = []
    for _ in range(n):
      t0 = time.perf_counter()
      logits = self.eng.infer_tensor(self.tok)
      self.tok = self.eng.sample(logits, temp=0.0, top_k=35).astype(np.int64).reshape(self.B, 1)
      times.append(time.perf_counter() - t0)
    return times


def cpu_decode_sample(hf_cfg, model_id, sample_batch: int, ctx: int, steps: int, warmup: int):
  s = CpuDecodeSample(hf_cfg, model_id, sample_batch, ctx)
  s.steps(warmup)
  t = sum(s.steps(steps))
  return sample_batch * steps / t, t / steps


def run_reference(args):
  """`--impl reference`: the reference's CPU implementation of the path (oracle port) on this
  box's host cores, same config/metric/unit; each step = a bounded sample of the batch."""
  rank = int(os.environ.get("RANK", "0"))
  if rank != 0:
    return
  from xotorch_support_jetson_b200.models import model_cards
  hf_cfg = model_cards[args.model]["hf_config"]
  cores = os.cpu_count() or 1
  sb = args.ref_batch
  steps, warm = args.steps, min(args.warmup, 2)
  # keep the whole run within a few minutes: one model, calibrate on the warm-up steps
  import torch  # noqa: F401
  t0 = time.perf_counter()
  smp = CpuDecodeSample(hf_cfg, args.model, sb, args.prompt_len)
  setup_s = time.perf_counter() - t0
  per_step = max(smp.steps(max(1, warm)))
  budget = 150.0
  if per_step * steps > budget:
    steps = max(2, int(budget / per_step))
  times = smp.steps(steps)
  val, per_step = sb * steps / sum(times), sum(times) / steps
  sample = (f"{steps} decode steps of {sb} sequences (one micro-batch of {args.batch}"
            + (f"; the {args.gpus}-GPU arm keeps {micro_batches(args)} such micro-batches in flight" if args.gpus > 1 else "")
            + f") at context {args.prompt_len} (synthetic KV cache, no CPU prefill), oracle port in reference-exact mode")
  line = {
    "impl": "reference", "metric": "decode_tok_s", "value": val, "unit": "tok/s", "n_gpus": args.gpus, "steps": steps,
    "warmup": warm, "ms_per_step": per_step * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
    "dtype": "bf16", "data": "synthetic",
    "config": workload_config(args),
    "cpu_baseline": {"value": val, "unit": "tok/s", "cores": cores, "kind": "port", "sample": sample},
    "e2e": {"value": val, "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    "setup_s": setup_s,
  }
  print(json.dumps(line), flush=True)


def micro_batches(args) -> int:
  """Micro-batches of `batch` sequences kept in flight: one per pipeline stage unless --micro-batches says otherwise."""
  return args.micro_batches if args.micro_batches > 0 else args.gpus


def workload_config(args):
  return {"workload": f"{args.model} bf16 random-init, {args.prompt_len}-token synthetic prompts, greedy decode from context "
                      f"{args.prompt_len}, batch {args.batch} per micro-batch x {micro_batches(args)} micro-batches in flight "
                      f"({args.gpus}-stage layer pipeline, "
                      + ("stage boundaries at half-layer granularity balanced by bytes per step)" if args.gpus > 1 and os.environ.get("XB_HALF_LAYERS", "1") != "0"
                         else "whole-layer ring partition)"),
          "global_batch": args.batch * micro_batches(args), "prompt_len": args.prompt_len, "parallelism": f"pp{args.gpus}",
          "l2_policy": "inputs larger than L2 (weights 15 GB per pass >> 126 MB L2)"}


# ----------------------------------------------------------------------------- B200 arm
def run_b200(args):
  import torch
  import torch.distributed as dist
  import __graft_entry__ as graft
  rank = int(os.environ.get("RANK", "0"))
  world = int(os.environ.get("WORLD_SIZE", "1"))
  local = int(os.environ.get("LOCAL_RANK", "0"))
  if world != args.gpus:
    if world == 1 and args.gpus > 1:
      raise SystemExit(f"--gpus {args.gpus} needs torchrun with {args.gpus} ranks")
    args.gpus = world
  torch.cuda.set_device(local)
  ret_group = None
  if world > 1:
    # lazy communicator init (no device_id): point-to-point ops then get per-peer communicators
    # instead of being serialised as collectives on the default group
    os.environ.setdefault("TORCH_NCCL_SHOW_EAGER_INIT_P2P_SERIALIZATION_WARNING", "false")
    dist.init_process_group("nccl")
    ret_group = dist.new_group(ranks=list(range(world)))   # own communicator for the token return edge
  graft.build()
  from xotorch_support_jetson_b200.model_config import load_model_config  # noqa: F401
  from xotorch_support_jetson_b200.models import model_cards
  from xotorch_support_jetson_b200.pipeline import PipelineRunner
  from xotorch_support_jetson_b200 import model_config as mc

  hf_cfg = model_cards[args.model]["hf_config"]
  # same mapping load_model_config applies to a config.json on disk
  with tempfile.TemporaryDirectory() as td:
    (Path(td) / "config.json").write_text(json.dumps(hf_cfg))
    cfg = mc.load_model_config(Path(td) / "config.json")
  B, M, L = args.batch, micro_batches(args), args.prompt_len
  K, W = args.steps, args.warmup
  peaks = measured_peaks()
  R = max(1, args.reps)
  max_ctx = L + W + 2 * R * K + 16 + 80
  runner = PipelineRunner(args.model, hf_cfg, cfg, rank, world, local, B, M, max_ctx, max_batch_tokens=args.max_batch_tokens,
                          dist=dist if world > 1 else None, seed=0, ret_group=ret_group)
  dev = runner.device
  g = torch.Generator(device=dev).manual_seed(1234)
  prompts = [torch.randint(1, cfg["vocab_size"], (B, L), generator=g, device=dev, dtype=torch.int32) for _ in range(M)]

  def barrier():
    torch.cuda.synchronize()
    if world > 1:
      dist.barrier()
    torch.cuda.synchronize()

  def timed(fn, reduce_max=True):
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(runner.stream)
    fn()
    e1.record(runner.stream)
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1 and reduce_max:
      dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())

  def timed_median(fn, reps):
    """The K-step region timed `reps` times (each bracketed by barrier + synchronize, max over ranks);
    a 0.1 s region on a freshly started box is not at steady state, the median of several is."""
    ts = [timed(fn) for _ in range(max(1, reps))]
    return statistics.median(ts), ts

  c0 = runner.stage.counters()
  # ---- prefill: one untimed pass (builds launch plans), then a timed pass on fresh slots.  At N > 1 the
  # untimed pass is also the correctness check: the pipeline's greedy ids against a one-GPU full-shard run
  tokens_check, ids_n1 = None, None
  if world > 1:
    tokens_check = pipeline_token_check(runner, args.model, hf_cfg, cfg, prompts, 8, local, dist)
    ids_n1 = tokens_check.pop("_ids_n1", None)
  else:
    runner.prefill(prompts)
    runner.drain()
  barrier()
  runner.reset()
  prefill_ms = timed(lambda: runner.prefill(prompts))
  prefill_tok_s = B * M * L / (prefill_ms * 1e-3)

  # ---- decode: W warm-up steps then exactly K timed steps
  runner.decode(W)
  clocks = ClockSampler(local)
  if rank == 0:
    clocks.start()
  l0 = runner.stage.counters()["launches"]
  decode_ms, decode_all = timed_median(lambda: runner.decode(K), R)
  launches = (runner.stage.counters()["launches"] - l0) // R
  clk = clocks.stop() if rank == 0 else None
  value = B * M * K / (decode_ms * 1e-3)
  ctx_avg = L + W + (R * K - 1) / 2.0

  # ---- end-to-end leg with host buffers
  e2e = None
  runner.drain()

  # ---- roofline of the dominant kernel: the tcgen05 Linear streaming this stage's weights
  st = runner.stage
  cnt = st.counters()
  with torch.cuda.stream(runner.stream):
    st.bench_linears(B, 2)
  iters = 20
  nl_box = [0]

  def lin():
    with torch.cuda.stream(runner.stream):
      nl_box[0] = st.bench_linears(B, iters)

  lin_ms = timed(lin, reduce_max=False)     # THIS rank's time: the fraction is per stage (stages hold different matrices)
  n_lin = nl_box[0]
  wbytes = cnt["weight_bytes"]
  per_rank = [(lin_ms, wbytes, n_lin, cnt["kv_bytes_per_pos"])]
  if world > 1:
    per_rank = [None] * world
    dist.all_gather_object(per_rank, (lin_ms, wbytes, n_lin, cnt["kv_bytes_per_pos"]))
  fracs = [(wb * iters / (ms * 1e-3) / 1e9) / peaks["hbm_gbs"] for ms, wb, _, _ in per_rank]
  # the stage whose Linears take longest per step governs the pipeline: report that one, list the others
  br = max(range(world), key=lambda r: per_rank[r][0])
  lin_ms_b, wbytes_b, n_lin_b, _ = per_rank[br]
  achieved = wbytes_b * iters / (lin_ms_b * 1e-3) / 1e9
  roofline = {"bound": "hbm", "kernel": "xb_gemm_tcgen05_kernel (skinny weight-streaming Linear)", "achieved": achieved,
              "peak": peaks["hbm_gbs"], "peak_source": peaks["source"] + " (MEASURED_PEAKS.json hbm_gbs)" if peaks["source"] == "measured" else "fallback",
              "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"], "traffic": None,
              "bytes_per_launch": wbytes_b / (n_lin_b / iters), "launch_us": lin_ms_b * 1e3 / n_lin_b,
              "launches_per_step": n_lin_b // iters}
  if world > 1:
    roofline["rank"] = br
    roofline["frac_per_rank"] = fracs
  else:
    # DRAM bytes per launch of this kernel from the committed `ncu --set full` capture of the N=1 bs=32 step
    # (evidence from a profile run, not re-measured here; stages of an N>1 pipeline hold other matrices: null)
    for name in ("r02_ncu_traffic.json", "r01_ncu_traffic.json"):
      prof = ROOT / "profiles" / name
      if prof.exists():
        try:
          roofline["traffic"] = json.loads(prof.read_text()).get("gemm_dram_bytes_per_launch")
          roofline["traffic_source"] = f"ncu --set full capture at N=1, committed as profiles/{name} (not measured in this run)"
          break
        except Exception:
          pass
  # whole decode step against the HBM roofline (weights once per micro-batch + KV of every sequence)
  # (per GPU: the stage that streams most bytes per step)
  step_bytes = max(M * wb + B * M * ctx_avg * kvb for _, wb, _, kvb in per_rank)
  step_frac = (step_bytes / (decode_ms * 1e-3 / K) / 1e9) / peaks["hbm_gbs"]

  extra = {"decode_ms_per_step_all_reps": [t / K for t in decode_all], "prefill_tok_s": prefill_tok_s, "prefill_ms": prefill_ms,
           "decode_step_hbm_frac": step_frac, "decode_step_bytes_per_gpu": step_bytes,
           "decode_ctx_avg": ctx_avg}
  # prefill against the tensor roofline (algorithmic FLOPs)
  D, H, KV, hd, I, V = cfg["embed_dim"], cfg["num_heads"], cfg["num_kv_heads"], cfg["head_dim"], cfg["intermediate_dim"], cfg["vocab_size"]
  nl_total = cfg["num_layers"]
  T_all = B * M * L
  flops = 2.0 * T_all * nl_total * (2 * D * H * hd + 2 * D * KV * hd + 3 * D * I) + 2.0 * B * M * L * L * H * hd * nl_total + 2.0 * B * M * V * D
  extra["prefill_tflops_per_gpu"] = flops / world / (prefill_ms * 1e-3) / 1e12
  extra["prefill_tensor_frac"] = extra["prefill_tflops_per_gpu"] / peaks["bf16_tflops_sustained"]
  if tokens_check is not None:
    extra.update(tokens_check)

  # ---- N=1: end-to-end through the InferenceEngine API with host numpy buffers
  cpu_base = None
  if world == 1:
    runner.reset()
    e2e = e2e_engine_api(runner, cfg, prompts[0], K, W)
    if rank == 0 and not args.no_cpu_baseline:
      t0 = time.perf_counter()
      val, per_step = cpu_decode_sample(hf_cfg, args.model, args.ref_batch, L, args.cpu_steps, 1)
      cpu_base = {"value": val, "unit": "tok/s", "cores": os.cpu_count() or 1, "kind": "port",
                  "sample": f"{args.cpu_steps} decode steps of {args.ref_batch} sequences (the named batch) at context {L} (synthetic KV cache), "
                            f"oracle port in reference-exact mode, {time.perf_counter() - t0:.0f}s incl. weight init"}

  # ---- decode bs=1 (latency-bound single stream), same pipeline; median of R x K steps like the main leg.  (At N=1 it
  # comes after the CPU-baseline leg, i.e. on a GPU that has been idle for a while, not one still at its power cap
  # from the batch-32 and prefill legs: a single stream draws far less power than they do.)
  runner.reset()
  barrier()
  r1 = PipelineRunnerView(runner)
  bs1 = r1.run_bs1(prompts[0][:1], W, K, lambda f: timed_median(f, R)[0], dist if world > 1 else None)
  extra.update(bs1)
  if world == 1:
    b1_bytes = wbytes + (L + W + (R * K - 1) / 2.0) * cnt["kv_bytes_per_pos"]
    extra["decode_bs1_hbm_frac"] = (b1_bytes / (1.0 / bs1["decode_bs1_tok_s"])) / 1e9 / peaks["hbm_gbs"]

  # ---- N > 1: end-to-end through the plugin API and the ring: one BoxNode per rank, RingPeerHandle / RingServer
  # between them, infer_tensor + sample per hop with host buffers at both ends of the ring
  if world > 1:
    runner.reset()
    barrier()
    e2e = e2e_ring(runner, args.model, cfg, prompts, K, W, dist, rank, world, ids_n1)

  if rank == 0:
    line = {
      "metric": "decode_tok_s", "value": value, "unit": "tok/s", "n_gpus": world, "steps": K, "warmup": W,
      "ms_per_step": decode_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
      "data": "synthetic", "config": workload_config(args), "clocks": clk, "e2e": e2e, "gpu_launches": launches,
      "roofline": roofline, "cpu_baseline": cpu_base, "extra": extra,
    }
    print(json.dumps(line), flush=True)
  runner.close()
  if world > 1:
    dist.destroy_process_group()


def pipeline_token_check(runner, model_id, hf_cfg, cfg, prompts, n_steps, local, dist):
  """N > 1: `n_steps` greedy decode steps of every micro-batch through the pipeline (whatever hand-off
  and stage partition the run uses: peer stores / NCCL, half-layer / whole-layer stages) compared, id for
  id, with a ONE-GPU full-shard engine on the same seeded weights and prompts (built on the last rank,
  same micro-batch shape, so both run the same kernels).  Leaves the runner reset.  Returns the `extra`
  entries {tokens_equal_n1, tokens_crc, tokens_crc_n1, tokens_checked} on every rank."""
  import zlib
  import torch
  from xotorch_support_jetson_b200.pipeline import gpu_random_getter
  from xotorch_support_jetson_b200.shard import Shard
  from xotorch_support_jetson_b200.stage import StageRuntime
  B, M, world, rank = runner.B, runner.M, runner.world, runner.rank
  L = prompts[0].shape[1]
  runner.reset()
  runner.prefill(prompts)
  hist = []
  for k in range(n_steps + 1):
    runner.drain()
    torch.cuda.synchronize()
    if rank == world - 1:
      hist.append(torch.stack([t.clone() for t in runner.tok]).cpu())   # ids sampled after k decode steps, [M, B]
    if k < n_steps:
      runner.decode(1)
  res = [None]
  if rank == world - 1:
    got = torch.stack(hist)                                              # [n_steps + 1, M, B]
    nl = cfg["num_layers"]
    st = StageRuntime(Shard(model_id, 0, nl - 1, nl), cfg, device=local, max_batch_tokens=runner.max_batch_tokens, max_seqs=B,
                      max_tokens_per_seq=L + n_steps + 8)
    st.load_hf_tensors(*gpu_random_getter(hf_cfg, 0, 0.02, runner.device))
    per = max(1, runner.max_batch_tokens // L)
    want = torch.zeros_like(got)
    ties = {}
    with torch.cuda.stream(torch.cuda.Stream(device=runner.device)):
      for m in range(M):
        tok = torch.zeros(B, dtype=torch.int32, device=runner.device)
        for c0 in range(0, B, per):                                      # same prefill chunking as PipelineRunner.prefill
          c1 = min(B, c0 + per)
          lg = st.forward(list(range(c0, c1)), [L] * (c1 - c0), [0] * (c1 - c0), tokens=prompts[m][c0:c1].reshape(-1))
          st.sample(lg, temp=0.0, out=tok[c0:c1])
          tied0 = (lg == lg.max(dim=-1, keepdim=True).values)
          for b in torch.nonzero(tied0.sum(-1) > 1).reshape(-1).tolist():
            ties[f"0,{m},{c0 + b}"] = torch.nonzero(tied0[b]).reshape(-1).tolist()
        want[0, m] = tok.cpu()
        for k in range(n_steps):
          lg = st.forward(list(range(B)), [1] * B, [L + k] * B, tokens=tok)
          tok = st.sample(lg, temp=0.0)
          want[k + 1, m] = tok.cpu()
          # exact ties of the (bf16-rounded) maximum: argmax takes the first index, the reference sampler lets its noise decide
          tied = (lg == lg.max(dim=-1, keepdim=True).values)
          for b in torch.nonzero(tied.sum(-1) > 1).reshape(-1).tolist():
            ties[f"{k + 1},{m},{b}"] = torch.nonzero(tied[b]).reshape(-1).tolist()
        for b in range(B):
          st.release_slot(b)
    torch.cuda.synchronize()
    st.close()
    del st
    torch.cuda.empty_cache()
    crc = lambda t: zlib.crc32(t.to(torch.int32).numpy().tobytes())
    res = [{"tokens_equal_n1": bool(torch.equal(got, want)), "tokens_crc": crc(got), "tokens_crc_n1": crc(want),
            "tokens_checked": f"{n_steps + 1} greedy ids x {M} micro-batches x {B} sequences (first id after the prefill + {n_steps} decode steps), "
                              f"pipeline vs one-GPU full shard; mismatching ids: {int((got != want).sum())}",
            "_ids_n1": (want.tolist(), ties)}]   # kept for the ring's end-to-end check, not printed
  dist.broadcast_object_list(res, src=world - 1)
  runner.reset()
  return res[0]


class PipelineRunnerView:
  """bs=1 single-stream decode on the slots of an existing runner (micro-batch 0, slot 0)."""

  def __init__(self, runner):
    self.r = runner

  def run_bs1(self, prompt, W, K, timed, dist):
    import torch
    r = self.r
    L = prompt.shape[1]
    tok = torch.zeros(1, dtype=torch.int32, device=r.device)
    logits = torch.empty((1, r.vocab), dtype=torch.float32, device=r.device) if r.last else None
    pos = [0]

    def fwd(q_len, tokens):
      with torch.cuda.stream(r.stream):
        hid = None
        if not r.first:
          hid = r.h_in[:q_len]
          dist.recv(hid, src=r.rank - 1)
        if r.last:
          lg = r.stage.forward([0], [q_len], [pos[0]], tokens=tokens if r.first else None, hidden=hid, out=logits)
          r.stage.sample(lg, temp=0.0, out=tok)
          if r.world > 1:
            dist.send(tok, dst=0)
        else:
          out = r.stage.forward([0], [q_len], [pos[0]], tokens=tokens if r.first else None, hidden=hid, out=r.h_out[:q_len])
          dist.send(out, dst=r.rank + 1)
          if r.first:
            dist.recv(tok, src=r.world - 1)
        pos[0] += q_len

    fwd(L, prompt.reshape(-1))

    def steps(n):
      for _ in range(n):
        fwd(1, tok)

    steps(W)
    ms = timed(lambda: steps(K))
    return {"decode_bs1_tok_s": K / (ms * 1e-3), "decode_bs1_ms_per_tok": ms / K}


def e2e_ring(runner, model_id, cfg, prompts, K, W, dist, rank, world, ids_n1=None):
  """N > 1 end to end: every rank runs a BoxNode around a B200ShardInferenceEngine that adopts the rank's resident
  stage; the entry node (rank 0) gets one request per micro-batch as HOST token ids ``[B, L]`` (process_tokens) and
  receives every step's sampled ids back as host lists through on_token; between ranks the request travels the way
  Node.forward_tensor sends it -- RingPeerHandle.send_tensor: header over the gloo control channel, hidden rows over
  NCCL posted by the engines -- and on the last rank through infer_tensor -> fp32 logits on the host -> sample().
  Rate = ids delivered to rank 0's callback between step W and step W+K of every request."""
  import numpy as np
  import torch
  from xotorch_support_jetson_b200.b200_engine import B200ShardInferenceEngine
  from xotorch_support_jetson_b200.orchestration import build_rank_node
  from xotorch_support_jetson_b200.shard import Shard
  B, M = runner.B, runner.M
  L = prompts[0].shape[1]
  ctrl = dist.new_group(ranks=list(range(world)), backend="gloo")
  sync = dist.new_group(ranks=list(range(world)), backend="gloo")
  shards = [None] * world
  dist.all_gather_object(shards, (tuple(runner.shard), bool(runner.first), bool(runner.last)))
  for i, (sh, first, last) in enumerate(shards):
    s_ = Shard(*sh)
    # the node decides by the Shard where ids enter and where logits leave: that must agree with what the stage holds
    assert s_.is_first_layer() == first and s_.is_last_layer() == last, f"stage {i}: half-layer cut inside the first / last layer"
  shard_map = [Shard(*sh) for sh, _, _ in shards]
  eng = B200ShardInferenceEngine(shard_downloader=None, device=runner.device.index)
  eng.attach_stage(runner.shard, runner.stage, cfg)
  n_gen = W + K + 1
  host_prompts = [p.cpu().numpy().astype(np.int64) for p in prompts] if rank == 0 else None
  stamps, ids_seen = {}, {}
  # where a hop's host time goes: wall time inside the engine calls (they include waiting for the GPU)
  prof = {"infer_tensor": [], "sample": []}

  def timed_call(name, fn):
    async def wrapped(*a, **k):
      t = time.perf_counter()
      try:
        return await fn(*a, **k)
      finally:
        prof[name].append(time.perf_counter() - t)
    return wrapped

  eng.infer_tensor = timed_call("infer_tensor", eng.infer_tensor)
  eng.sample = timed_call("sample", eng.sample)

  async def run():
    loop = asyncio.get_running_loop()
    node = build_rank_node(eng, dist, rank, world, ctrl_group=ctrl, data_group=None, max_generate_tokens=n_gen, shard_map=shard_map,
                           landing=(M, B) if os.environ.get("XB_RING_LANDING", "1") != "0" else None)
    await node.start()
    if rank == 0:
      left = set(range(M))
      done = asyncio.Event()

      def on_token(rid, toks, fin):
        stamps.setdefault(rid, []).append(time.perf_counter())
        if len(ids_seen.setdefault(rid, [])) < 16:
          ids_seen[rid].append([int(t) for t in toks])
        if fin:
          left.discard(int(rid[2:]))
          if not left:
            done.set()

      node.on_token.register("bench").on_next(on_token)
      base = Shard(model_id, 0, 0, cfg["num_layers"])
      for m in range(M):
        asyncio.create_task(node.process_tokens(base, host_prompts[m], f"mb{m}"))
      await asyncio.wait_for(done.wait(), 600)
    await loop.run_in_executor(None, lambda: dist.barrier(group=sync))   # the other ranks keep serving until rank 0 is done
    await node.stop()

  asyncio.run(run())
  torch.cuda.synchronize()
  from xotorch_support_jetson_b200 import networking as _nw
  if _nw.TRACE is not None:
    (ROOT / "gpurun_out").mkdir(exist_ok=True)
    (ROOT / "gpurun_out" / f"ring_trace_rank{rank}.json").write_text(json.dumps(_nw.TRACE))
    gt = getattr(eng, "_gpu_trace", [])
    rows = [(round(a.elapsed_time(b) * 1e3, 1), round(b.elapsed_time(gt[i + 1][0]) * 1e3, 1) if i + 1 < len(gt) else None) for i, (a, b) in enumerate(gt)]
    (ROOT / "gpurun_out" / f"ring_gpu_rank{rank}.json").write_text(json.dumps(rows))
    gs = getattr(eng, "_gpu_trace_s", [])
    if gs:
      (ROOT / "gpurun_out" / f"ring_gpu_sample_rank{rank}.json").write_text(json.dumps([round(a.elapsed_time(b) * 1e3, 1) for a, b in gs]))
  per_rank = [None] * world
  dist.all_gather_object(per_rank, {k: (round(statistics.median(v) * 1e3, 3) if v else None) for k, v in prof.items()})
  if rank != 0:
    return None
  t0 = max(stamps[f"mb{m}"][W] for m in range(M))
  t1 = max(stamps[f"mb{m}"][W + K] for m in range(M))
  V = cfg["vocab_size"]
  # the ids the ring delivered against the one-GPU full-shard run of pipeline_token_check (same prompts, greedy)
  same = None
  if ids_n1 is not None:
    want, ties = ids_n1
    n_chk = min(len(want), min(len(v) for v in ids_seen.values()))
    at_tie, bad = 0, 0
    for m in range(M):
      for b in range(B):
        for k in range(n_chk):
          g_, w_ = ids_seen[f"mb{m}"][k][b], want[k][m][b]
          if g_ != w_:
            # the plugin API samples with the reference's rule, argmax(softmax(logits / 1e-5) / q): among EXACTLY tied
            # bf16 logits its noise decides, where the device loop's argmax takes the first index.  From such a step
            # on the two runs are different (equally valid) continuations; anything else is a mismatch.
            if g_ in ties.get(f"{k},{m},{b}", ()):
              at_tie += 1
            else:
              bad += 1
            break
    same = {"tokens_equal_n1": bad == 0,
            "tokens_checked": f"{n_chk} ids x {M} requests x {B} sequences delivered to rank 0's on_token vs one-GPU full shard: "
                              f"{M * B - at_tie - bad} sequences identical, {at_tie} leave it at an exact tie of the maximum logit "
                              f"(reference sampler: noise decides; device loop: first index), {bad} unexplained"}
  return {"value": B * M * K / (t1 - t0), "unit": "tok/s", "h2d_bytes_per_step": 8 * B * M, "d2h_bytes_per_step": B * M * (V * 4 + 4),
          "round_trip_ms": (t1 - t0) / K * 1e3, "engine_call_median_ms_per_rank": per_rank, **(same or {}),
          "api": "one BoxNode per rank (the reference Node's walk): process_tokens(host ids [32, 2048]) per micro-batch on rank 0 -> "
                 "B200ShardInferenceEngine.infer_tensor per stage -> Node.forward_tensor -> RingPeerHandle.send_tensor (header over gloo, "
                 "bf16 hidden rows over NCCL posted by the engines) -> last rank: fp32 logits [32, V] to the host, sample(), ids back "
                 "to rank 0 as host arrays (next step's input) and through send_result / on_token (counted here)"}


def e2e_engine_api(runner, cfg, prompt, K, W):
  """decode tok/s through B200ShardInferenceEngine.infer_tensor + sample with HOST numpy
  buffers: every step uploads the ids, downloads the [B, vocab] fp32 logits the reference's
  API returns, uploads them again for sample() and downloads the sampled ids."""
  import numpy as np
  import torch
  from xotorch_support_jetson_b200.b200_engine import B200ShardInferenceEngine
  eng = B200ShardInferenceEngine(shard_downloader=None)
  eng.attach_stage(runner.shard, runner.stage, cfg)
  B = prompt.shape[0]
  toks = prompt.cpu().numpy().astype(np.int64)

  async def go():
    logits, state = await eng.infer_tensor("bench", runner.shard, toks, None)
    tok = await eng.sample(logits, temp=0.0)
    for _ in range(W):
      logits, state = await eng.infer_tensor("bench", runner.shard, tok.reshape(B, 1), state)
      tok = await eng.sample(logits, temp=0.0)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(K):
      logits, state = await eng.infer_tensor("bench", runner.shard, tok.reshape(B, 1), state)
      tok = await eng.sample(logits, temp=0.0)
    torch.cuda.synchronize()
    return time.perf_counter() - t0

  dt = asyncio.run(go())
  V = cfg["vocab_size"]
  eng.end_request("bench")

  # the same loop kept on the device (generate(): SURVEY 8f.1); rate from the arrival times of the
  # ids at the host callback, first W steps (and the prefill) excluded
  stamps = []

  async def go_loop():
    await eng.generate("bench", runner.shard, toks, max_generate_tokens=W + K + 1, temp=0.0, eos_token_id=-1,
                       on_token=lambda rid, ids, fin: stamps.append(time.perf_counter()))
    await asyncio.sleep(0)

  asyncio.run(go_loop())
  eng.end_request("bench")
  loop_rate = B * K / (stamps[W + K] - stamps[W]) if len(stamps) > W + K else None
  return {"value": B * K / dt, "unit": "tok/s", "h2d_bytes_per_step": B * 4, "d2h_bytes_per_step": B * V * 4 + B * 4,
          "api": "B200ShardInferenceEngine.infer_tensor(np ids) -> np fp32 logits [B,V] (written into pinned host memory by the "
                 "LM-head kernel itself); .sample(that array) -> np ids (reference plugin API, host buffers; the logits array "
                 "handed back to sample() is recognised and not re-uploaded; with one request in flight the engine runs the sampler "
                 "(parameters of the previous sample() call) and the next decode step ahead, right behind the forward: "
                 "sample() and infer_tensor() with the ids it returned collect work already in flight, anything else discards it)",
          "device_loop": {"value": loop_rate, "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": B * 4,
                          "api": "B200ShardInferenceEngine.generate(np prompt ids, on_token=cb): sample->forward chained on the "
                                 "device, ids streamed to a pinned table, reference-semantics sampler (top-k 35, q noise) every step"}}


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument("--gpus", type=int, default=1)
  ap.add_argument("--steps", type=int, default=64)
  ap.add_argument("--warmup", type=int, default=4)
  ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
  ap.add_argument("--model", default=MODEL)
  ap.add_argument("--batch", type=int, default=BATCH)
  ap.add_argument("--prompt-len", type=int, default=PROMPT_LEN)
  ap.add_argument("--max-batch-tokens", type=int, default=8192)
  ap.add_argument("--ref-batch", type=int, default=BATCH, help="sequences per step of the CPU arm (the named batch)")
  ap.add_argument("--cpu-steps", type=int, default=3)
  ap.add_argument("--micro-batches", type=int, default=int(os.environ.get("XB_BENCH_MICRO_BATCHES", "0")),
                  help="micro-batches in flight (default: one per GPU)")
  ap.add_argument("--reps", type=int, default=5, help="the K-step timed region is repeated this often; the median is reported")
  ap.add_argument("--no-cpu-baseline", action="store_true")
  args = ap.parse_args()
  if args.warmup < 3 and args.impl == "b200":
    args.warmup = 3
  if args.impl == "reference":
    run_reference(args)
  else:
    run_b200(args)


if __name__ == "__main__":
  main()
