"""world_size-2 gloo run of the pipeline protocol on CPU: PipelineRunner's ordering of
send/recv pairs, micro-batching and the last->first token edge, with the oracle standing in
for the GPU stage.  The 2-stage, 2-micro-batch result must equal single-stage greedy runs."""
import os
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent


def _worker(rank, world, port, ret):
  sys.path.insert(0, str(ROOT))
  sys.path.insert(0, str(ROOT / "tests"))
  os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
  torch.set_num_threads(2)
  dist.init_process_group("gloo", rank=rank, world_size=world)
  from oracle import torch_oracle as O
  from oracle_stage import OracleStage
  from xotorch_support_jetson_b200.models import model_cards
  from xotorch_support_jetson_b200.pipeline import PipelineRunner
  from xotorch_support_jetson_b200.synth import random_hf_state
  mid = "llama-3.1-tiny"
  hf_cfg = model_cards[mid]["hf_config"]
  cfg = O.model_config_from_hf(hf_cfg)
  state = random_hf_state(hf_cfg, seed=0, std=0.05, norm_jitter=0.1)
  B, M, L, K = 2, 2, 9, 4
  ret_pg = dist.new_group(ranks=list(range(world)))
  r = PipelineRunner(mid, hf_cfg, cfg, rank, world, 0, B, M, 64, max_batch_tokens=32, dist=dist, ret_group=ret_pg,
                     stage_factory=lambda sh: OracleStage(sh, cfg, state, max_ctx=64))
  g = torch.Generator().manual_seed(5)
  prompts = [torch.randint(1, cfg["vocab_size"], (B, L), generator=g, dtype=torch.int32) for _ in range(M)]
  r.prefill(prompts)
  hist = []
  for k in range(K):
    r.decode(1)
    if rank == world - 1:
      r.drain()
      hist.append(torch.stack([t.clone() for t in r.tok]))
  r.drain()
  if rank == world - 1:
    ret["tokens"] = torch.stack(hist).numpy()   # [K, M, B] tokens sampled AFTER each decode step
  if rank == 0:
    ret["first"] = torch.stack([t.clone() for t in r.tok]).numpy()
  dist.barrier()
  dist.destroy_process_group()


def test_two_stage_pipeline_equals_single_stage():
  sys.path.insert(0, str(ROOT / "tests"))
  port = 29500 + os.getpid() % 2000
  mgr = mp.Manager()
  ret = mgr.dict()
  mp.spawn(_worker, args=(2, port, ret), nprocs=2, join=True)
  got = ret["tokens"]       # [K, M, B]
  from oracle import torch_oracle as O
  from xotorch_support_jetson_b200.models import model_cards
  from xotorch_support_jetson_b200.synth import random_hf_state
  mid = "llama-3.1-tiny"
  hf_cfg = model_cards[mid]["hf_config"]
  cfg = O.model_config_from_hf(hf_cfg)
  state = random_hf_state(hf_cfg, seed=0, std=0.05, norm_jitter=0.1)
  B, M, L, K = 2, 2, 9, 4
  g = torch.Generator().manual_seed(5)
  prompts = [torch.randint(1, cfg["vocab_size"], (B, L), generator=g, dtype=torch.int32) for _ in range(M)]
  nl = cfg["num_layers"]
  for m in range(M):
    for b in range(B):
      model = O.OracleShardModel(cfg, mid, 0, nl - 1, state, O.OracleOptions(max_generated_tokens=55))
      toks, _, _ = O.greedy_generate([model], prompts[m][b:b + 1].long(), K + 1)
      # pipeline: prefill samples token 0, decode step k samples token k+1
      assert got[:, m, b].tolist() == toks[0, 1:].tolist(), (m, b, got[:, m, b], toks)
  # rank 0 ends up holding the last sampled ids (the return edge delivered them)
  assert np.array_equal(ret["first"], got[-1])
