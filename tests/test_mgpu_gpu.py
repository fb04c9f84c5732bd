"""Multi-GPU parity (needs >= 2 GPUs; skipped on a 1-GPU box): NCCL pipeline == single GPU."""
import subprocess
import sys
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs at least 2 GPUs")
@pytest.mark.parametrize("mid", ["llama-3.1-tiny", "qwen-2.5-tiny"])
def test_nccl_pipeline_equals_single_gpu(built_lib, mid):
  n = 2
  cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
         "--master-port", "29533", str(ROOT / "tests" / "mgpu_pipeline_check.py"), mid]
  out = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
  assert out.returncode == 0 and "MGPU_PIPELINE_OK" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]
