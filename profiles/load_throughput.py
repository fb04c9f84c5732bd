"""Checkpoint load time of one stage: native loader (xb_load_safetensors: mmap -> pinned ring ->
device, layout kernel per chunk) against the safetensors-package path (get_tensor -> CPU permute ->
pageable H2D -> repack).  Synthetic Llama-3.2-1B checkpoint (2.5 GB bf16, 2 files) written to a temp
directory first, so both paths read from the page cache.  Run: python profiles/load_throughput.py"""
import json
import sys
import tempfile
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import __graft_entry__ as graft  # noqa: E402


def main():
  graft.build()
  from xotorch_support_jetson_b200 import model_config as mc
  from xotorch_support_jetson_b200.models import model_cards
  from xotorch_support_jetson_b200.shard import Shard
  from xotorch_support_jetson_b200.stage import StageRuntime
  from xotorch_support_jetson_b200.synth import write_random_checkpoint
  mid = sys.argv[1] if len(sys.argv) > 1 else "llama-3.2-1b"
  hf = model_cards[mid]["hf_config"]
  nl = hf["num_hidden_layers"]
  with tempfile.TemporaryDirectory() as td:
    t0 = time.perf_counter()
    ck = write_random_checkpoint(Path(td) / mid, hf, seed=0, std=0.02, shard_bytes=2 << 30)
    size = sum(f.stat().st_size for f in ck.glob("*.safetensors"))
    print(f"checkpoint {mid}: {size / 1e9:.2f} GB in {len(list(ck.glob('*.safetensors')))} files, written in {time.perf_counter() - t0:.1f}s")
    cfg = mc.load_model_config(ck / "config.json")
    for label, shard, native in (("native, whole model", Shard(mid, 0, nl - 1, nl), True),
                                 ("native, second half of the layers", Shard(mid, nl // 2, nl - 1, nl), True),
                                 ("safetensors package, whole model", Shard(mid, 0, nl - 1, nl), False),
                                 ("native, whole model (again)", Shard(mid, 0, nl - 1, nl), True)):
      st = StageRuntime(shard, cfg, device=0, max_batch_tokens=256, max_seqs=2, max_tokens_per_seq=256)
      torch.cuda.synchronize()
      t0 = time.perf_counter()
      info = st.load_dir(ck, native=native)
      torch.cuda.synchronize()
      dt = time.perf_counter() - t0
      nbytes = info["bytes"] if info["bytes"] else size
      print(f"  {label}: {dt:.2f} s, {nbytes / 1e9:.2f} GB -> {nbytes / dt / 1e9:.2f} GB/s")
      st.close()


if __name__ == "__main__":
  main()
