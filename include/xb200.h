/* xb200.h -- C ABI of libxb200, the B200-native layer-range forward + sampling engine.
 *
 * The reference (satoutahhaithem/xotorch_support_jetson) is 100% Python and has no FFI;
 * its plugin boundary is the Python class `InferenceEngine`
 * (xotorch/inference/inference_engine.py:11-51).  This header is the native layer that a
 * drop-in `InferenceEngine` subclass binds with ctypes; each entry
 * point cites the reference code whose work it replaces.  Paths are relative to the
 * reference repository root.
 *
 * Conventions: every function returns 0 on success, non-zero on failure;
 * xb_last_error() returns a thread-local message for the last failure on this thread.
 * No C++ exception crosses this boundary.  All device work is enqueued on the
 * caller-supplied CUDA stream (`stream` is a cudaStream_t passed as void*) and the calls
 * do not synchronise unless the name ends in _sync.  One engine handle serves one
 * (GPU, layer range); calls on a handle must come from one thread at a time (the
 * reference serialises engine work on a 1-worker executor,
 * xotorch/inference/torch/sharded_inference_engine.py:46).
 * There is NO CPU fallback: without a CUDA device every compute entry point fails.
 */
#ifndef XB200_H
#define XB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define XB_ABI_VERSION 3

typedef struct xb_engine xb_engine;

/* RoPE flavour (xotorch/inference/torch/models/general_mha.py:33-63) */
enum {
  XB_ROPE_INTERLEAVED = 0, /* Llama3ScaledRoPE / RotaryPositionalEmbeddings: pairs (2i,2i+1), fp32 math */
  XB_ROPE_HALF_SPLIT = 1   /* Qwen2RotaryPositionalEmbeddings: pairs (i,i+hd/2), bf16 math            */
};

/* Static description of one shard of one model: the fields the reference derives in
 * load_model_config (xotorch/inference/torch/models/llm_utils.py:30-77) plus the Shard key
 * (xotorch/inference/shard.py:4-39: start_layer/end_layer inclusive). */
typedef struct xb_model_desc {
  uint32_t struct_size;      /* = sizeof(xb_model_desc) of the header the CALLER was built against; xb_engine_create
                                refuses a descriptor whose size differs from the library's (stale binding) */
  int32_t n_layers, start_layer, end_layer;
  int32_t dim, n_heads, n_kv_heads, head_dim, inter_dim, vocab;
  float norm_eps;
  int32_t rope_kind;         /* XB_ROPE_* */
  int32_t rmsnorm_two_round; /* 1: torchtune 0.6 RMSNorm (normalise, round, scale, round); 0: single rounding */
  int32_t qkv_bias;          /* Qwen: q/k/v projections carry a bias (general_mha.py:52) */
  int32_t tied_head;         /* LM head shares the embedding matrix (general_mha.py:43-44,56-57) */
  int32_t max_positions;     /* rows of the RoPE table uploaded with XB_W_ROPE_TABLE */
  int32_t max_batch_tokens;  /* most token rows one xb_forward call may carry */
  int32_t max_seqs;          /* most sequences one xb_forward call may carry */
  /* Half-layer stage boundaries (no counterpart in the reference, whose Shard is whole layers): a
   * pipeline stage may begin at the MLP block of start_layer (its attention block ran on the previous
   * stage) and/or end after the attention block of end_layer.  The residual stream is what crosses
   * the boundary, as between whole layers.  Lets a ring of GPUs balance layers + LM head to half a
   * layer.  A stage with first_from_mlp is never a first shard (no embedding), one with last_to_attn
   * never a last shard (no final norm / head). */
  int32_t first_from_mlp, last_to_attn;
} xb_model_desc;

/* Weight slots.  Row order is the reference's torchtune layout AFTER its HF->torchtune
 * remap (llm_utils.py:181-258; Llama q/k rows permuted by _permute, llm_utils.py:126-134).
 * All matrices are bf16, row-major [out_features, in_features]. */
enum {
  XB_W_Q = 0, XB_W_K, XB_W_V, XB_W_O, XB_W_GATE, XB_W_UP, XB_W_DOWN,
  XB_W_SA_NORM, XB_W_MLP_NORM, XB_W_Q_BIAS, XB_W_K_BIAS, XB_W_V_BIAS, /* per layer */
  XB_W_EMBED, XB_W_FINAL_NORM, XB_W_LM_HEAD,                          /* layer = -1 */
  XB_W_ROPE_TABLE /* layer = -1: fp32 [max_positions, head_dim/2, 2] (cos, sin) */
};

/* One forward call over a ragged batch: sequence i contributes q_lens[i] consecutive
 * token rows starting at absolute position positions[i] of KV slot slot_ids[i].
 * Replaces the (tokens|hidden, mask, input_pos, curr_pos) plumbing of
 * ShardedGeneralModel.generate (general_mha.py:178-254) and the per-request
 * ShardInferenceState (llm_utils.py:473-511): causality and cache extent follow from the
 * positions, no [L+1024, C] mask is built or shipped. */
typedef struct xb_batch {
  int32_t n_seqs;
  const int32_t* slot_ids;  /* host [n_seqs] */
  const int32_t* q_lens;    /* host [n_seqs] */
  const int32_t* positions; /* host [n_seqs] */
  const int32_t* tokens;    /* device int32 [sum q_lens]; first shard only, else NULL */
  const void* hidden_in;    /* device bf16 [sum q_lens, dim]; non-first shard only, else NULL */
} xb_batch;

typedef struct xb_counters {
  uint64_t launches;          /* kernels launched by this engine since creation */
  uint64_t forwards;          /* xb_forward calls */
  uint64_t weight_bytes;      /* bf16 bytes of this shard's matrices read by one decode step */
  uint64_t kv_bytes_per_pos;  /* K+V bytes per cached position over this shard's layers */
  uint64_t kv_pages_total, kv_pages_free;
} xb_counters;

const char* xb_last_error(void);
int xb_abi_version(void);

/* Replaces ShardedGeneralModel.__init__ / GeneralMHA (general_mha.py:23-176): allocates
 * only what the layer range needs (embedding on a first shard; final norm + head on a
 * last shard -- the reference allocates all three on every shard, general_mha.py:124-130). */
int xb_engine_create(const xb_model_desc* desc, int device, xb_engine** out);
int xb_engine_destroy(xb_engine* e);

/* Replaces load_model_weights_torchtune + load_state_dict (llm_utils.py:136-284).
 * `src` may be host or device memory; `layer` is the absolute layer index or -1. */
int xb_load_weight(xb_engine* e, int layer, int kind, const void* src, size_t nbytes);

/* Replaces the safetensors walk of load_model_weights_torchtune (llm_utils.py:136-160: every file
 * opened, every tensor read) for ONE checkpoint file: the file is mapped, the header parsed, and
 * only the tensors of this shard (its layer range; embedding / final norm / head where the shard
 * holds them) are streamed page cache -> pinned ring -> device, where one kernel per chunk casts
 * (BF16/F16/F32 -> bf16), applies the reference's q/k row permutation when permute_qk != 0
 * (llm_utils.py:126-134,213-231: Llama yes, Qwen no) and writes the engine's fused, tile-major
 * layout.  HF tensor names (model.layers.N.self_attn.q_proj.weight ...).  *n_tensors / *n_bytes
 * (either may be NULL) receive what was taken from this file.  Call once per file of a sharded
 * checkpoint, then xb_weights_missing. */
int xb_load_safetensors(xb_engine* e, const char* path, int permute_qk, int* n_tensors, uint64_t* n_bytes);

/* *n_missing = weights of this shard that neither xb_load_weight nor xb_load_safetensors has
 * supplied yet (0 = ready to run); buf (may be NULL) receives a readable list.  The reference
 * surfaces the same condition as load_state_dict(strict=False) missing keys (llm_utils.py:262-284). */
int xb_weights_missing(xb_engine* e, char* buf, size_t buflen, int* n_missing);

/* The header of a .safetensors file as text, one tensor per line: "name dtype begin end dim0 dim1 ..."
 * (offsets relative to the data section) -- what xb_load_safetensors parses; needs no GPU.  The
 * reference reads the same information through safetensors' safe_open (llm_utils.py:151-160).
 * buf may be NULL to get the count only. */
int xb_safetensors_list(const char* path, char* buf, size_t buflen, int* n_tensors);

/* Paged KV cache.  Replaces setup_caches / reset_caches / KVCache.update
 * (llm_utils.py:320-365; sharded_inference_engine.py:71-82,139-147): pages of
 * `page_tokens` positions are handed to a slot on demand as it grows and returned by
 * xb_kv_release_slot; nothing is zero-filled. */
int xb_kv_reserve(xb_engine* e, int n_slots, int max_tokens_per_slot, int page_tokens);
int xb_kv_release_slot(xb_engine* e, int slot);
int xb_kv_slot_len(xb_engine* e, int slot, int* len);

/* Replaces ShardTransformerDecoder.forward (llm_utils.py:367-440) with the intended
 * hand-off semantics (a non-last shard returns the OUTPUT of its last layer).
 * Decode-only batches (every q_len == 1) on a non-default stream are replayed from a CUDA
 * graph captured on the second call of each (n_seqs, kv-split) shape.
 * out: device bf16 [T, dim] on a non-last shard; device fp32 [n_seqs, vocab] logits of the
 * last position of each sequence on a last shard (bf16-rounded values, as the reference's
 * `model_logits[:, -1].float()`, sharded_inference_engine.py:366).
 * taps: NULL, or (end_layer-start_layer+1) device pointers, each bf16 [T, dim], that
 * receive the residual stream after each layer (parity instrumentation). */
int xb_forward(xb_engine* e, const xb_batch* batch, void* out, void* const* taps, void* stream);

/* xb_forward on a last shard whose logits must reach the host every step, as the plugin API
 * requires (infer_tensor returns them as a numpy array, sharded_inference_engine.py:361-367:
 * `.float().numpy(force=True)` = a blocking device-to-host copy after the forward).  The LM-head
 * kernel stores the fp32 [n_seqs, vocab] logits into `host_logits` (page-locked, device-addressable
 * host memory) from its own epilogue while it streams the weights, in addition to the device copy
 * in `out` (may be NULL); they are complete when the work enqueued on `stream` has finished. */
int xb_forward_host_logits(xb_engine* e, const xb_batch* batch, void* out, float* host_logits, void* stream);

/* Intra-box hand-off between pipeline stages over NVLink peer memory, replacing the per-token gRPC
 * SendTensor hop of the reference (Node.forward_tensor, xotorch/orchestration/node.py:424-443;
 * GRPCPeerHandle.send_tensor, xotorch/networking/grpc/grpc_peer_handle.py:117-136) for decode steps.
 * Every stage calls xb_handoff_create (n_slots landing buffers of `rows` x dim bf16 + one flag
 * each; the two 64-byte CUDA IPC handles are exchanged by the host, e.g. all_gather) and every
 * non-last stage xb_handoff_connect with the handles of the NEXT stage.  xb_forward_handoff is
 * xb_forward for a decode-only batch where
 *   slot_in  >= 0: the hidden rows are taken from this stage's landing[slot_in] once its flag has
 *                  reached wait_value (the stage's first kernel spins on the flag; batch->hidden_in
 *                  is ignored);
 *   slot_out >= 0: the stage's last kernel stores the hidden rows into the next stage's
 *                  landing[slot_out] and then sets that flag to signal_value (`out` is ignored).
 * Flags only grow; use a per-slot step counter.  A slot may be rewritten once its consumer has run
 * (in a ring of stages this follows from the token that travels back to the first stage). */
int xb_handoff_create(xb_engine* e, int n_slots, int rows, void* landing_handle, void* flag_handle);
int xb_handoff_connect(xb_engine* e, const void* next_landing_handle, const void* next_flag_handle);
int xb_forward_handoff(xb_engine* e, const xb_batch* batch, int slot_in, int wait_value, int slot_out, int signal_value,
                       void* out, void* stream);
/* xb_handoff_connect for two stages that live in ONE process (an in-process ring of per-GPU nodes, or two
 * stages on one GPU in a test): `e` stores straight into `next`'s landing buffers through their device
 * addresses, no CUDA IPC handle involved (xb_handoff_create may then be given NULL for both handle
 * buffers); peer access from e's device to next's is enabled when they differ.  Both stages must have
 * called xb_handoff_create with the same geometry.  `next` must outlive the connection. */
int xb_handoff_connect_local(xb_engine* e, xb_engine* next);
/* A consumer stage whose producer never signalled does not kill the CUDA context: after the time-out
 * (XB_HANDOFF_TIMEOUT_S, default 30 s) the waiting kernel records the slot in a host-visible error word
 * and lets the step finish on whatever the landing buffer holds; every later xb_forward* call on that
 * engine fails with the recorded message until xb_handoff_clear_error is called. */
int xb_handoff_clear_error(xb_engine* e);
/* Keep `n` SMs out of the decode Linears' grids (one persistent block per SM otherwise; default 1, XB_RESERVE_SMS at
 * engine creation).  A stage of a ring has a communication kernel resident for whole steps -- an NCCL receive posted
 * ahead for the ids the last stage will send: a Linear block that finds no SM free of it runs as a second wave after
 * all the others and doubles that launch (measured at 2 GPUs: 4.9 -> 5.45 ms per step once the blocks had grown to
 * 124 KB of shared memory; 4.70 ms with one SM set aside).  The stream-K split follows the block count, so all stages
 * of a ring -- and a single-GPU engine whose ids they are compared with -- must use the same value to agree bit for
 * bit.  Drops the captured decode steps and launch plans; call it before serving. */
int xb_reserve_sms(xb_engine* e, int n);

/* Replaces TorchDynamicShardInferenceEngine.sample + torchtune.generation.sample
 * (sharded_inference_engine.py:208-228).  logits: device fp32 [n, vocab].
 * temp <= 0 selects first-index argmax.  top_k <= 0 and top_p <= 0 (or >= 1) disable
 * the respective filter.  q: NULL (counter-based noise from `seed`) or device fp32
 * [n, vocab] Exp(1) noise (reference-exact replay of its torch.Generator stream).
 * tokens_out: device int32 [n]. */
int xb_sample(xb_engine* e, const float* logits, int n, float temp, int top_k, float top_p, const float* q,
              uint64_t seed, int32_t* tokens_out, void* stream);

int xb_get_counters(xb_engine* e, xb_counters* out);

/* Kernel-level test hook: y[T,N] (fp32) = x[T,K] * w[N,K]^T on device buffers.
 * use_tensor_cores = 1 runs the tcgen05 kernel the engine uses (force_splits > 0 pins the
 * K split count; round_bf16 rounds like nn.Linear), 0 runs the CUDA-core checker kernel.
 * scratch: device fp32, at least 16*T*N elements. */
int xb_debug_linear(const void* x, const void* w, int T, int N, int K, float* y, float* scratch,
                    int use_tensor_cores, int force_splits, int round_bf16, void* stream);

/* Measurement hook (bench.py roofline leg): enqueue the Linear kernels of one forward over
 * T token rows -- q/k/v, o, gate/up, down per layer (general_mha.py:83-102,
 * llm_utils.py:513-522) and the LM head on a last shard (llm_utils.py:425-435) -- `iters`
 * times back to back on `stream`, on the engine's own weights.  The caller brackets the
 * call with CUDA events; *n_launches receives the number of kernels enqueued.
 */
int xb_bench_linears(xb_engine* e, int T, int iters, int* n_launches, void* stream);

/* Measurement hook: the decode attention kernel(s) of this shard's first layer
 * (torchtune MultiHeadAttention's SDPA over the KV cache, general_mha.py:78-106) for the batch of
 * the last xb_forward call, `iters` times back to back, with `n_splits` kv splits (0 = the
 * engine's choice, returned in *splits_used). */
int xb_bench_attention(xb_engine* e, int n_seqs, int max_ctx, int n_splits, int iters, int* splits_used, void* stream);

/* Measurement hook: per-block timeline of the Linear launches (llm_utils.py:513-522 / general_mha.py:83-102 sites) of
 * an engine created with XB_GEMM_TS=1 (eager launches; every Linear then records %globaltimer stamps per block:
 * entry, weight prefetch issued, dependency resolved, first / last MMA, epilogue start / end, exit).  Copies
 * [n_launches][blocks][16] uint64 nanosecond stamps recorded since the previous call into host_out (may be NULL to
 * just reset) and starts a new recording. */
int xb_debug_timeline(xb_engine* e, unsigned long long* host_out, size_t cap_u64, int* n_launches, int* blocks);

/* Allocation-free Linear (no counterpart in the reference, which never manages device memory
 * itself): y[T,N] (fp32) = x[T,K] * w[N,K]^T on device buffers through the engine's split-K
 * tcgen05 kernel, with the split-K workspace in an 8 MiB array built into the library instead
 * of caller- or allocator-provided memory.  For callers that must not reach a GPU allocator:
 * a decode loop that cannot afford the device-wide stall of cudaMalloc/cudaFree between steps,
 * or a process whose HBM is already fully committed to weights and KV pages.  Performs no
 * cudaMalloc/cudaFree and creates no framework tensor.  Limits: T <= 128 and
 * splits*T*N <= 2Mi floats (rejected with the required size otherwise).  The first call on a
 * device resolves the array's device address: make it before any stream capture.  All calls on
 * a device share the array (its old contents are never read before being rewritten);
 * calls that can overlap on the GPU race on it. */
int xb_linear_noalloc(const void* x, const void* w, int T, int N, int K, float* y, int round_bf16, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* XB200_H */
