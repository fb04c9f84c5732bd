"""GPU test of the allocation-free entry point xb_linear_noalloc (collected last on purpose:
a run that stops at the first failure has then already run every other test).

Who needs the entry point: code that has to keep clear of GPU allocators while it runs -- a
decode loop that cannot take the device-wide stall of a cudaMalloc/cudaFree between steps, or a
process whose HBM is already fully handed to weights and KV pages, where asking for a few MiB
of split-K workspace can fail in the middle of a request.  The workspace therefore lives in an
8 MiB array inside libxb200 and the caller supplies only inputs and the output.

What is checked, through the C ABI: the result equals the existing entry point's
(xb_debug_linear with caller-supplied scratch and the same tcgen05 kernel) within the
tolerance of tests/test_gemm_gpu.py, on consecutive calls of different shapes that reuse the
built-in array (so stale contents of an earlier, larger call must not leak into a later one),
and a shape that does not fit is refused with a clear message."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _pair(T, N, K, seed):
  g = torch.Generator(device="cuda").manual_seed(seed)
  x = (torch.randn(T, K, device="cuda", generator=g) * 0.5).to(torch.bfloat16)
  w = (torch.randn(N, K, device="cuda", generator=g) * 0.05).to(torch.bfloat16)
  return x, w


def test_noalloc_linear_matches_existing_entry_point_on_consecutive_calls(built_lib):
  from xotorch_support_jetson_b200.stage import debug_linear, linear_noalloc
  shapes = [(32, 4096, 4096), (5, 384, 1024), (32, 4096, 4096), (1, 1536, 8192), (128, 512, 256)]
  for i, (T, N, K) in enumerate(shapes):
    x, w = _pair(T, N, K, seed=100 + i)
    y = torch.full((T, N), float("nan"), dtype=torch.float32, device="cuda")   # the caller's output buffer
    linear_noalloc(x, w, y)
    ref = debug_linear(x, w, use_tensor_cores=True)
    fp32 = x.float() @ w.float().t()
    torch.cuda.synchronize()
    scale = fp32.abs().max().item()
    assert torch.isfinite(y).all()
    assert (y - ref).abs().max().item() <= 1e-4 * scale, (i, T, N, K)
    assert (y - fp32).abs().max().item() <= 1e-4 * scale, (i, T, N, K)


def test_noalloc_linear_allocates_nothing(built_lib):
  """The point of the entry point: once it has run on a device, a call changes neither the
  driver's free-memory figure nor the framework allocator's counters."""
  from xotorch_support_jetson_b200.stage import linear_noalloc
  x, w = _pair(32, 4096, 4096, seed=1)
  y = torch.empty((32, 4096), dtype=torch.float32, device="cuda")
  linear_noalloc(x, w, y)            # first call on the device: resolves the built-in array's address
  torch.cuda.synchronize()
  free0, _ = torch.cuda.mem_get_info()
  stats0 = torch.cuda.memory_stats()
  for _ in range(3):
    linear_noalloc(x, w, y)
  torch.cuda.synchronize()
  free1, _ = torch.cuda.mem_get_info()
  stats1 = torch.cuda.memory_stats()
  assert free1 == free0
  for k in ("allocation.all.allocated", "segment.all.allocated", "allocated_bytes.all.current", "reserved_bytes.all.current"):
    assert stats1.get(k) == stats0.get(k), k


def test_noalloc_linear_rounding_mode_and_rejections(built_lib):
  from xotorch_support_jetson_b200.stage import debug_linear, linear_noalloc
  x, w = _pair(16, 768, 512, seed=7)
  y = torch.empty((16, 768), dtype=torch.float32, device="cuda")
  linear_noalloc(x, w, y, round_bf16=True)
  ref = debug_linear(x, w, use_tensor_cores=True, round_bf16=True)
  torch.cuda.synchronize()
  assert torch.equal(y, y.to(torch.bfloat16).float())          # values are bf16-representable
  assert (y != ref).float().mean().item() < 0.01               # same kernel, at most a rare 1-ulp flip
  # more than 128 token rows is outside this entry point
  xb, wb = _pair(129, 256, 256, seed=8)
  with pytest.raises(RuntimeError, match="128 token rows"):
    linear_noalloc(xb, wb, torch.empty((129, 256), dtype=torch.float32, device="cuda"))
  # workspace beyond the built-in 2Mi floats: 128 x 32768 = 4Mi floats even with one split
  xb, wb = _pair(128, 32768, 64, seed=9)
  with pytest.raises(RuntimeError, match="built-in buffer holds"):
    linear_noalloc(xb, wb, torch.empty((128, 32768), dtype=torch.float32, device="cuda"))
  # and the entry point still works after a refusal
  linear_noalloc(x, w, y)
  torch.cuda.synchronize()
  assert (y - (x.float() @ w.float().t())).abs().max().item() <= 1e-4 * y.abs().max().item()
