"""Run under torchrun on >= 2 GPUs (tests/test_mgpu_gpu.py launches it): one BoxNode per rank around a
B200ShardInferenceEngine, requests entering at rank 0 as host token ids and walking the ring the way
Node.forward_tensor sends them (RingPeerHandle: header over gloo; hidden rows over NCCL for the prompt, decode
steps through the next rank's landing buffers over CUDA IPC when argv[2] == "landing", over NCCL otherwise).
The ids delivered to rank 0's on_token must equal a single full-shard engine driven call by call
(xotorch/inference/test_inference_engine.py:11,43-44: same result for any number of shards)."""
import asyncio
import os
import sys
import tempfile
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
  import faulthandler
  faulthandler.dump_traceback_later(75, exit=True)     # a hung ring must not hold the GPU box: stacks, then exit
  rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
  torch.cuda.set_device(local)
  dist.init_process_group("nccl", device_id=torch.device("cuda", local))
  ctrl = dist.new_group(ranks=list(range(world)), backend="gloo")
  sync = dist.new_group(ranks=list(range(world)), backend="gloo")
  from xotorch_support_jetson_b200.b200_engine import B200ShardInferenceEngine
  from xotorch_support_jetson_b200.orchestration import build_rank_node
  from xotorch_support_jetson_b200.shard import Shard
  from xotorch_support_jetson_b200.synth import LocalShardDownloader
  mid = sys.argv[1] if len(sys.argv) > 1 else "llama-3.1-tiny"
  landing = len(sys.argv) > 2 and sys.argv[2] == "landing"
  nl = {"llama-3.1-tiny": 6, "qwen-2.5-tiny": 4, "llama-3.2-tiny": 4, "mistral-tiny": 4}[mid]
  td = tempfile.mkdtemp(prefix=f"xb_ring_{rank}_")

  def engine():
    dl = LocalShardDownloader(root=Path(td), seed=0, std=0.05, norm_jitter=0.1)
    return B200ShardInferenceEngine(dl, device=local, max_seqs=8, max_tokens_per_seq=128, max_batch_tokens=64, page_tokens=16)

  g = np.random.default_rng(5)
  prompts = {"a": g.integers(1, 1000, size=(1, 19)).astype(np.int64), "b": g.integers(1, 1000, size=(3, 11)).astype(np.int64),
             "c": g.integers(1, 1000, size=(2, 7)).astype(np.int64)}
  n_new = 7
  got = {rid: [] for rid in prompts}
  eng = engine()
  seen = []

  async def run():
    loop = asyncio.get_running_loop()
    # layer ranges as the memory-weighted strategy cuts them for equal GPUs; every rank loads its own before the ring starts
    node = build_rank_node(eng, dist, rank, world, ctrl_group=ctrl, data_group=None, max_generate_tokens=n_new)
    base = Shard(mid, 0, 0, nl)
    await node.start()
    await node.collect_topology(set())
    await eng.ensure_shard(node.get_current_shard(base))
    if landing:     # a collective step, after every rank holds its stage: export / map the landing buffers
      from xotorch_support_jetson_b200.orchestration import connect_rank_landing
      connect_rank_landing(eng, dist, rank, world, sync, (4, 8))    # (not the group the RingServer thread is listening on)
    inner = eng.infer_tensor

    async def spy(request_id, shard, x, state=None):
      from xotorch_support_jetson_b200.networking import DeviceTensorRef
      seen.append("landing" if isinstance(x, DeviceTensorRef) and x.landing is not None else type(x).__name__)
      return await inner(request_id, shard, x, state)

    eng.infer_tensor = spy
    await loop.run_in_executor(None, lambda: dist.barrier(group=sync))
    if rank == 0:
      left, done = set(prompts), asyncio.Event()

      def on_token(rid, toks, fin):
        got[rid].append([int(t) for t in toks])
        if fin:
          left.discard(rid)
          if not left:
            done.set()

      node.on_token.register("t").on_next(on_token)
      await asyncio.gather(*[node.process_tokens(base, p, rid) for rid, p in prompts.items()])
      await asyncio.wait_for(done.wait(), 120)
    await loop.run_in_executor(None, lambda: dist.barrier(group=sync))
    await node.stop()

  asyncio.run(run())
  torch.cuda.synchronize()
  ok = True
  n_landing = [None] * world
  dist.all_gather_object(n_landing, seen.count("landing"), group=sync)
  if rank == 0:
    if eng.stage is not None:
      eng.stage.close()
      eng.stage = eng.shard = None
    ref = engine()
    full = Shard(mid, 0, nl - 1, nl)

    async def direct():
      want = {}
      for rid, p in prompts.items():
        out, st = await ref.infer_tensor(rid, full, p, None)
        steps = []
        for _ in range(n_new):
          tok = await ref.sample(out, temp=0.0)
          steps.append([int(t) for t in tok.reshape(-1)])
          out, st = await ref.infer_tensor(rid, full, tok.reshape(-1, 1), st)
        want[rid] = steps
      return want

    want = asyncio.run(direct())
    for rid in prompts:
      if got[rid] != want[rid]:
        ok = False
        print(f"MISMATCH request {rid}: ring {got[rid]} direct {want[rid]}")
    # decode steps of ranks 1.. must have arrived through the landing buffers when asked to (and never otherwise)
    exp = sum(n_landing[1:])
    if landing and exp < (world - 1) * len(prompts) * (n_new - 2):
      ok = False
      print(f"too few landing hand-offs: {n_landing}")
    if not landing and exp != 0:
      ok = False
    print("MGPU_RING_OK" if ok else "MGPU_RING_FAIL", f"world={world} model={mid} landing={landing} hops through landing buffers {n_landing}", flush=True)
  dist.barrier()
  dist.destroy_process_group()
  sys.exit(0 if ok else 1)


if __name__ == "__main__":
  main()
