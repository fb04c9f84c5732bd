"""ctypes binding of libxb200 (include/xb200.h).  Thin: one Python function per C entry
point, non-zero status -> RuntimeError(xb_last_error()), exactly the error convention the
reference's Python engine has (exceptions)."""
import ctypes as C
import os
from pathlib import Path

_LIB_NAME = "libxb200.so"

# symbols include/xb200.h declares (tests check the built library exports every one)
SYMBOLS = (
  "xb_last_error", "xb_abi_version", "xb_engine_create", "xb_engine_destroy", "xb_load_weight",
  "xb_load_safetensors", "xb_weights_missing", "xb_safetensors_list",
  "xb_kv_reserve", "xb_kv_release_slot", "xb_kv_slot_len", "xb_forward", "xb_forward_host_logits", "xb_sample",
  "xb_handoff_create", "xb_handoff_connect", "xb_forward_handoff", "xb_handoff_connect_local", "xb_handoff_clear_error", "xb_reserve_sms",
  "xb_get_counters", "xb_debug_linear", "xb_bench_linears", "xb_bench_attention", "xb_linear_noalloc", "xb_debug_timeline",
)

# weight kinds (enum in xb200.h)
(W_Q, W_K, W_V, W_O, W_GATE, W_UP, W_DOWN, W_SA_NORM, W_MLP_NORM, W_Q_BIAS, W_K_BIAS, W_V_BIAS,
 W_EMBED, W_FINAL_NORM, W_LM_HEAD, W_ROPE_TABLE) = range(16)

ROPE_INTERLEAVED, ROPE_HALF_SPLIT = 0, 1


class ModelDesc(C.Structure):
  _fields_ = [
    ("struct_size", C.c_uint32),   # sizeof(xb_model_desc) as this binding knows it (filled in by model_desc())
    ("n_layers", C.c_int32), ("start_layer", C.c_int32), ("end_layer", C.c_int32),
    ("dim", C.c_int32), ("n_heads", C.c_int32), ("n_kv_heads", C.c_int32), ("head_dim", C.c_int32),
    ("inter_dim", C.c_int32), ("vocab", C.c_int32),
    ("norm_eps", C.c_float),
    ("rope_kind", C.c_int32), ("rmsnorm_two_round", C.c_int32), ("qkv_bias", C.c_int32), ("tied_head", C.c_int32),
    ("max_positions", C.c_int32), ("max_batch_tokens", C.c_int32), ("max_seqs", C.c_int32),
    ("first_from_mlp", C.c_int32), ("last_to_attn", C.c_int32),
  ]


ABI_VERSION = 3


def model_desc(**fields) -> ModelDesc:
  """ModelDesc with ``struct_size`` set: xb_engine_create rejects a descriptor laid out by another header."""
  return ModelDesc(struct_size=C.sizeof(ModelDesc), **fields)


class Batch(C.Structure):
  _fields_ = [
    ("n_seqs", C.c_int32),
    ("slot_ids", C.POINTER(C.c_int32)), ("q_lens", C.POINTER(C.c_int32)), ("positions", C.POINTER(C.c_int32)),
    ("tokens", C.c_void_p), ("hidden_in", C.c_void_p),
  ]


class Counters(C.Structure):
  _fields_ = [
    ("launches", C.c_uint64), ("forwards", C.c_uint64), ("weight_bytes", C.c_uint64),
    ("kv_bytes_per_pos", C.c_uint64), ("kv_pages_total", C.c_uint64), ("kv_pages_free", C.c_uint64),
  ]


def lib_path() -> Path:
  return Path(os.environ.get("XB200_LIB", Path(__file__).resolve().parent / _LIB_NAME))


_lib = None


def load():
  """Load the shared library.  Fails loudly when it has not been built: there is no
  Python/torch fallback for the compute path."""
  global _lib
  if _lib is not None:
    return _lib
  p = lib_path()
  if not p.exists():
    raise RuntimeError(f"{p} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                       "(libxb200 is the only compute path; there is no fallback)")
  lib = C.CDLL(str(p))
  lib.xb_last_error.restype = C.c_char_p
  lib.xb_abi_version.restype = C.c_int
  lib.xb_engine_create.argtypes = [C.POINTER(ModelDesc), C.c_int, C.POINTER(C.c_void_p)]
  lib.xb_engine_destroy.argtypes = [C.c_void_p]
  lib.xb_load_weight.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_size_t]
  lib.xb_load_safetensors.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_uint64)]
  lib.xb_weights_missing.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t, C.POINTER(C.c_int)]
  lib.xb_safetensors_list.argtypes = [C.c_char_p, C.c_char_p, C.c_size_t, C.POINTER(C.c_int)]
  lib.xb_kv_reserve.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
  lib.xb_kv_release_slot.argtypes = [C.c_void_p, C.c_int]
  lib.xb_kv_slot_len.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int)]
  lib.xb_forward.argtypes = [C.c_void_p, C.POINTER(Batch), C.c_void_p, C.POINTER(C.c_void_p), C.c_void_p]
  lib.xb_forward_host_logits.argtypes = [C.c_void_p, C.POINTER(Batch), C.c_void_p, C.c_void_p, C.c_void_p]
  lib.xb_handoff_create.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
  lib.xb_handoff_connect.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
  lib.xb_handoff_connect_local.argtypes = [C.c_void_p, C.c_void_p]
  lib.xb_handoff_clear_error.argtypes = [C.c_void_p]
  lib.xb_reserve_sms.argtypes = [C.c_void_p, C.c_int]
  lib.xb_forward_handoff.argtypes = [C.c_void_p, C.POINTER(Batch), C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
  lib.xb_sample.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_int, C.c_float, C.c_void_p, C.c_uint64,
                            C.c_void_p, C.c_void_p]
  lib.xb_get_counters.argtypes = [C.c_void_p, C.POINTER(Counters)]
  lib.xb_debug_linear.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                  C.c_int, C.c_int, C.c_void_p]
  lib.xb_linear_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p]
  lib.xb_bench_attention.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_void_p]
  lib.xb_debug_timeline.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_int), C.POINTER(C.c_int)]
  lib.xb_bench_linears.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_void_p]
  for name in SYMBOLS:
    fn = getattr(lib, name)
    if name not in ("xb_last_error",):
      fn.restype = C.c_int
  if lib.xb_abi_version() != ABI_VERSION:
    raise RuntimeError(f"{p} exports ABI v{lib.xb_abi_version()}, this binding is for v{ABI_VERSION}: rebuild the library")
  _lib = lib
  return lib


def check(status: int):
  if status != 0:
    msg = load().xb_last_error()
    raise RuntimeError(f"libxb200: {msg.decode() if msg else 'unknown error'} (status {status})")
