"""Kernel-level parity of the tcgen05 Linear (through the C ABI hook xb_debug_linear)
against a plain PyTorch fp32 reference of the same op."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(x, w):
  return x.float() @ w.float().t()


@pytest.mark.parametrize("T,N,K,splits", [
  (1, 256, 256, 1), (1, 768, 512, 2), (5, 384, 1024, 3), (16, 1024, 2048, 0), (17, 640, 512, 0),
  (32, 6144, 4096, 0), (33, 512, 8192, 4), (64, 1536, 384, 0), (100, 2048, 1024, 0), (128, 4096, 4096, 0),
  (1, 128256, 2048, 1), (3, 1000, 320, 1),
])
def test_skinny_linear_matches_fp32(built_lib, T, N, K, splits):
  from xotorch_support_jetson_b200.stage import debug_linear
  g = torch.Generator(device="cuda").manual_seed(T * 1000 + N + K)
  x = (torch.randn(T, K, device="cuda", generator=g) * 0.5).to(torch.bfloat16)
  w = (torch.randn(N, K, device="cuda", generator=g) * 0.05).to(torch.bfloat16)
  y = debug_linear(x, w, use_tensor_cores=True, force_splits=splits)
  ref = _ref(x, w)
  chk = debug_linear(x, w, use_tensor_cores=False)
  torch.cuda.synchronize()
  scale = ref.abs().max().item()
  # fp32 accumulation in a different order: tolerance 1e-4 of the output scale
  assert (chk - ref).abs().max().item() <= 1e-4 * scale
  assert (y - ref).abs().max().item() <= 1e-4 * scale


@pytest.mark.parametrize("T,N,K", [(129, 512, 256), (256, 768, 512), (300, 1000, 320), (2048, 6144, 4096), (1024, 4096, 14336)])
def test_wide_linear_matches_fp32(built_lib, T, N, K):
  from xotorch_support_jetson_b200.stage import debug_linear
  g = torch.Generator(device="cuda").manual_seed(T + N + K)
  x = (torch.randn(T, K, device="cuda", generator=g) * 0.5).to(torch.bfloat16)
  w = (torch.randn(N, K, device="cuda", generator=g) * 0.05).to(torch.bfloat16)
  y = debug_linear(x, w, use_tensor_cores=True)
  ref = _ref(x, w)
  torch.cuda.synchronize()
  # output is rounded to bf16 once (nn.Linear semantics): compare against the rounded reference
  ref_b = ref.to(torch.bfloat16).float()
  scale = ref.abs().max().item()
  assert (y - ref_b).abs().max().item() <= 2 ** -7 * scale
  # and almost every element must be the identical bf16 value
  assert (y != ref_b).float().mean().item() < 0.02
