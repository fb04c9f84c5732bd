"""Generates the committed golden vectors under tests/golden/ (TEST INFRASTRUCTURE).

Run in the build container:  python -m oracle.gen_golden

1. ``oracle_<model>.npz`` -- for each small same-architecture model: seeded prompt ids, the
   residual stream after every layer (last prompt row of the prefill and of the first decode
   step), the last-position logits of the prefill and of N_DEC-1 decode steps, and the greedy
   token sequence, all from ``oracle.torch_oracle`` in reference-exact mode (CPU, bf16, bf16
   RoPE table, two-rounding RMSNorm, full shard).  Weights are NOT stored: they are
   regenerated from (seed, std, norm_jitter) by ``synth.random_hf_state``.
2. ``partition_golden.json`` -- layer ranges produced by the REFERENCE's own
   ``RingMemoryWeightedPartitioningStrategy`` + ``map_partitions_to_shards`` imported from
   /root/reference (pure Python; a 2-line ``scapy`` stub satisfies ``xotorch/__init__``),
   for equal-memory rings of 1..8 nodes over the layer counts of the benchmark models, plus
   the expectations of the reference's test ``xotorch/topology/test_map_partitions.py:9-50``.
"""
import json
import sys
import types
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
GOLD = ROOT / "tests" / "golden"

MODELS = ("llama-3.2-tiny", "llama-3.1-tiny", "qwen-2.5-tiny", "mistral-tiny")
SEED, STD, JITTER = 0, 0.05, 0.1
PROMPT_LEN, N_DEC = 37, 8


def gen_oracle_vectors():
  from oracle import torch_oracle as O
  from xotorch_support_jetson_b200.models import model_cards
  from xotorch_support_jetson_b200.synth import random_hf_state
  for mid in MODELS:
    hf_cfg = model_cards[mid]["hf_config"]
    cfg = O.model_config_from_hf(hf_cfg)
    state = random_hf_state(hf_cfg, seed=SEED, std=STD, norm_jitter=JITTER)
    nl = cfg["num_layers"]
    g = torch.Generator().manual_seed(1234)
    prompt = torch.randint(1, cfg["vocab_size"], (1, PROMPT_LEN), generator=g)
    model = O.OracleShardModel(cfg, mid, 0, nl - 1, state, O.OracleOptions(max_generated_tokens=32))
    toks, logits, taps = O.greedy_generate([model], prompt, N_DEC, collect_logits=True, taps_at=(0, 1))
    out = {
      "prompt": prompt.numpy().astype(np.int32),
      "tokens": toks.astype(np.int32),
      "logits": np.stack([l[0] for l in logits]).astype(np.float32),            # [N_DEC, V]
      "prefill_hidden_last": np.stack([taps[0][l][0, -1] for l in range(nl)]),   # [L, D]
      "decode_hidden": np.stack([taps[1][l][0, 0] for l in range(nl)]),          # [L, D]
      "meta": np.array(json.dumps({"model": mid, "seed": SEED, "std": STD, "norm_jitter": JITTER,
                                   "torch": torch.__version__, "threads": torch.get_num_threads()})),
    }
    np.savez_compressed(GOLD / f"oracle_{mid}.npz", **out)
    top2 = np.sort(np.stack([l[0] for l in logits]), axis=-1)[:, -2:]
    print(f"{mid}: tokens {toks[0].tolist()}  min top-2 margin {float((top2[:, 1] - top2[:, 0]).min()):.4f}")


def gen_partition_golden():
  ref = Path("/root/reference")
  if not ref.exists():
    print("reference not mounted: partition_golden.json left untouched")
    return
  sys.modules.setdefault("scapy", types.ModuleType("scapy"))
  sa = types.ModuleType("scapy.all")
  sa.get_if_addr = sa.get_if_list = lambda *a, **k: []
  sys.modules.setdefault("scapy.all", sa)
  sys.path.insert(0, str(ref))
  import os
  cwd = os.getcwd()
  os.chdir("/tmp")  # device_capabilities.py drops a run_<date>.log into the cwd on import
  try:
    from xotorch.topology.device_capabilities import DeviceCapabilities, DeviceFlops
    from xotorch.topology.partitioning_strategy import Partition, map_partitions_to_shards
    from xotorch.topology.ring_memory_weighted_partitioning_strategy import RingMemoryWeightedPartitioningStrategy
    from xotorch.topology.topology import Topology
  finally:
    os.chdir(cwd)
  out = {"equal_rings": {}, "ref_test_map_partitions": [], "weighted": []}
  for n_layers in (16, 28, 32, 80, 4, 6):
    for n in range(1, 9):
      t = Topology()
      for i in range(n):
        t.update_node(f"gpu{n - 1 - i:02d}-rank{i}", DeviceCapabilities(model="b", chip="NVIDIA B200", memory=183359,
                                                                        flops=DeviceFlops(fp32=0, fp16=0, int8=0)))
      parts = RingMemoryWeightedPartitioningStrategy().partition(t)
      shards = map_partitions_to_shards(parts, n_layers, "m")
      out["equal_rings"][f"{n_layers}/{n}"] = {"nodes": [p.node_id for p in parts],
                                               "ranges": [[s.start_layer, s.end_layer] for s in shards]}
  # the reference's own test vectors (test_map_partitions.py:9-50), recomputed through its code
  for parts, n_layers in (
      ([("node1", 0.0, 0.42857), ("node2", 0.42857, 0.71428), ("node3", 0.71428, 0.99999)], 32),
      ([("node1", 0.0, 0.1), ("node2", 0.1, 0.2), ("node3", 0.2, 1.0)], 32),
      ([("node1", 0.0, 0.9), ("node2", 0.9, 0.99), ("node3", 0.99, 1.0)], 32),
      ([("node1", 0.0, 0.5), ("node2", 0.5, 1.0)], 1)):
    ps = [Partition(*p) for p in parts]
    shards = map_partitions_to_shards(ps, n_layers, "model")
    out["ref_test_map_partitions"].append({"partitions": parts, "n_layers": n_layers,
                                           "ranges": [[s.start_layer, s.end_layer] for s in shards]})
  for mems in ([3000, 1000, 6000], [128 * 1024, 192 * 1024, 64 * 1024], [1, 1, 1, 5]):
    t = Topology()
    for i, m in enumerate(mems):
      t.update_node(f"node{i + 1}", DeviceCapabilities(model="t", chip="t", memory=m, flops=DeviceFlops(fp32=0, fp16=0, int8=0)))
    parts = RingMemoryWeightedPartitioningStrategy().partition(t)
    out["weighted"].append({"memories": mems, "partitions": [[p.node_id, p.start, p.end] for p in parts]})
  (GOLD / "partition_golden.json").write_text(json.dumps(out, indent=1))
  print("partition_golden.json written from the reference's own code")


if __name__ == "__main__":
  GOLD.mkdir(parents=True, exist_ok=True)
  gen_oracle_vectors()
  gen_partition_golden()
