"""Multi-GPU parity (needs >= 2 GPUs; skipped on a 1-GPU box, where tests/test_handoff_gpu.py and
tests/test_box_ring_gpu.py run the same kernels and protocols with both stages on one GPU): the layer pipeline
under torchrun must give exactly the tokens of a single-GPU full-shard run -- with the peer-store hand-off and
with NCCL, with half-layer and with whole-layer stage boundaries."""
import os
import subprocess
import sys
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs at least 2 GPUs")
@pytest.mark.parametrize("mid,peer,half", [
  ("llama-3.1-tiny", "1", "1"), ("qwen-2.5-tiny", "1", "1"),
  ("llama-3.1-tiny", "0", "1"), ("llama-3.1-tiny", "1", "0"), ("qwen-2.5-tiny", "0", "0"),
])
def test_pipeline_equals_single_gpu(built_lib, mid, peer, half):
  n = 2
  env = dict(os.environ, XB_PEER_HANDOFF=peer, XB_HALF_LAYERS=half)
  cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
         "--master-port", "29533", str(ROOT / "tests" / "mgpu_pipeline_check.py"), mid]
  out = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env)
  assert out.returncode == 0 and "MGPU_PIPELINE_OK" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs at least 2 GPUs")
@pytest.mark.parametrize("mid,mode", [("llama-3.1-tiny", "landing"), ("llama-3.1-tiny", "nccl"), ("qwen-2.5-tiny", "landing")])
def test_process_ring_equals_direct_engine_loop(built_lib, mid, mode):
  """One BoxNode + B200ShardInferenceEngine per rank (torchrun), three requests in flight, ids delivered to rank 0:
  equal to a single full-shard engine -- decode steps through the next rank's landing buffers over CUDA IPC, and over
  NCCL send / recv."""
  n = 2
  cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
         "--master-port", "29537", str(ROOT / "tests" / "mgpu_ring_check.py"), mid, mode]
  out = subprocess.run(cmd, capture_output=True, text=True, timeout=120)
  assert out.returncode == 0 and "MGPU_RING_OK" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]
