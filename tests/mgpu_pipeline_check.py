"""Run under torchrun on >= 2 GPUs (tests/test_mgpu_gpu.py launches it): the NCCL layer
pipeline must produce exactly the tokens of a single-GPU full-shard run on the same weights
(the reference's contract "an inference engine should work the same for any number of
shards", xotorch/inference/test_inference_engine.py:11,43-44)."""
import json
import os
import sys
import tempfile
from pathlib import Path

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
  rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
  torch.cuda.set_device(local)
  dist.init_process_group("nccl", device_id=torch.device("cuda", local))
  ret = dist.new_group(ranks=list(range(world)))
  from xotorch_support_jetson_b200 import model_config as mc
  from xotorch_support_jetson_b200.models import model_cards
  from xotorch_support_jetson_b200.pipeline import PipelineRunner
  from xotorch_support_jetson_b200.shard import Shard
  from xotorch_support_jetson_b200.stage import StageRuntime
  from xotorch_support_jetson_b200.synth import random_hf_state
  mid = sys.argv[1] if len(sys.argv) > 1 else "llama-3.1-tiny"
  hf_cfg = model_cards[mid]["hf_config"]
  with tempfile.TemporaryDirectory() as td:
    (Path(td) / "config.json").write_text(json.dumps(hf_cfg))
    cfg = mc.load_model_config(Path(td) / "config.json")
  state = random_hf_state(hf_cfg, seed=0, std=0.05, norm_jitter=0.1)
  B, M, L, K = 3, world, 21, 6
  r = PipelineRunner(mid, hf_cfg, cfg, rank, world, local, B, M, 128, max_batch_tokens=64, page_tokens=16, dist=dist,
                     ret_group=ret, weights=(state.__getitem__, state.__contains__))
  g = torch.Generator().manual_seed(11)
  prompts = [torch.randint(1, cfg["vocab_size"], (B, L), generator=g, dtype=torch.int32).to(r.device) for _ in range(M)]
  r.prefill(prompts)
  hist = []
  for k in range(K):
    r.decode(1)
    r.drain()
    torch.cuda.synchronize()   # the runner works on its own stream; read the ids only when it is done
    if rank == world - 1:
      hist.append(torch.stack([t.clone() for t in r.tok]).cpu())
  torch.cuda.synchronize()
  ok = True
  if rank == world - 1:
    got = torch.stack(hist)            # [K, M, B] ids sampled after decode step k
    nl = cfg["num_layers"]
    st = StageRuntime(Shard(mid, 0, nl - 1, nl), cfg, device=local, max_batch_tokens=64, max_seqs=B, max_tokens_per_seq=128,
                      page_tokens=16)
    st.load_state(state)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
      for m in range(M):
        lg = st.forward(list(range(B)), [L] * B, [0] * B, tokens=prompts[m].reshape(-1))
        tok = st.sample(lg, temp=0.0)
        for k in range(K):
          lg = st.forward(list(range(B)), [1] * B, [L + k] * B, tokens=tok)
          tok = st.sample(lg, temp=0.0)
          if not torch.equal(tok.cpu(), got[k, m]):
            ok = False
            print(f"MISMATCH micro-batch {m} step {k}: pipeline {got[k, m].tolist()} single {tok.cpu().tolist()}")
        for b in range(B):
          st.release_slot(b)
    torch.cuda.synchronize()
    print("MGPU_PIPELINE_OK" if ok else "MGPU_PIPELINE_FAIL", f"world={world} model={mid}", flush=True)
  dist.barrier()
  dist.destroy_process_group()
  sys.exit(0 if ok else 1)


if __name__ == "__main__":
  main()
