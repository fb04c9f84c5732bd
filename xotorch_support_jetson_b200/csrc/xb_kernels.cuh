// Memory-bound kernels of the layer-range forward: embedding gather, fused
// (split-K reduce +) residual + RMSNorm, QKV finalize (bias + RoPE + paged-KV append),
// SiLU*up and on-device sampling (attention lives in xb_attn_mma.cuh).
//
// Every "rounding point" follows the reference's eager bf16 graph:
// each nn.Linear output, RoPE output, attention output, residual sum, RMSNorm output,
// SiLU output and the gate*up product are rounded to bf16 exactly once.
#pragma once
#include "xb_common.cuh"
#include <math.h>

// Per-forward metadata living in device memory at fixed addresses (so captured graphs can
// be replayed after the host refreshes the contents).
struct BatchMeta {
  const int* tok_seq;     // [T]   sequence index of each token row
  const int* tok_pos;     // [T]   absolute position of each token row
  const int* seq_slot;    // [n_seqs] KV slot of each sequence
  const int* seq_last;    // [n_seqs] row index of the last token of each sequence
  const int* page_table;  // [n_slots, max_pages] page ids
  const int* tile_start;  // [n_tiles] first token row of each prefill query tile
  const int* tile_len;    // [n_tiles] tokens in the tile (<= 128 / group size, one sequence)
  int max_pages;
  int page_tokens;
};

// ----------------------------------------------------------------------------- reductions
template <int NT>
__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  float t = (threadIdx.x < NT / 32) ? red[threadIdx.x] : 0.f;
  if (w == 0) {
    t = warp_sum(t);
    if (l == 0) red[0] = t;
  }
  __syncthreads();
  return red[0];
}

// torchtune RMSNorm (reference builds it at general_mha.py:116-117,130):
//   x_normed = (x_fp32 * rsqrt(mean(x_fp32^2) + eps)).type_as(x);  out = x_normed * scale
// two_round = 1 reproduces both bf16 roundings; 0 = F.rms_norm-style single rounding.
__device__ __forceinline__ float rms_apply(float x, float inv, float w, int two_round) {
  return two_round ? rbf(rbf(x * inv) * w) : rbf(x * inv * w);
}

// ----------------------------------------------------------------------------- embed + norm
// h[t] = tok_embeddings[token[t]] (llm_utils.py:399-400); hn[t] = sa_norm(h[t]).
template <int NT>
__global__ void __launch_bounds__(NT) xb_embed_norm_kernel(const int* __restrict__ tokens, const bf16* __restrict__ emb,
                                                           const bf16* __restrict__ nw, bf16* __restrict__ h,
                                                           bf16* __restrict__ hn, int D, int V, float eps, int two_round) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float red[32];
  const int t = blockIdx.x;
  int tok = tokens[t];
  tok = tok < 0 ? 0 : (tok >= V ? V - 1 : tok);
  const bf16* src = emb + (long long)tok * D;
  float ss = 0.f;
  for (int i = threadIdx.x * 2; i < D; i += NT * 2) {
    const uint32_t u = *reinterpret_cast<const uint32_t*>(src + i);
    *reinterpret_cast<uint32_t*>(h + (long long)t * D + i) = u;
    const float2 f = bf2_to_f2(u);
    ss += f.x * f.x + f.y * f.y;
  }
  ss = block_sum<NT>(ss, red);
  const float inv = 1.0f / sqrtf(ss / (float)D + eps);
  for (int i = threadIdx.x * 2; i < D; i += NT * 2) {
    const float2 f = bf2_to_f2(*reinterpret_cast<const uint32_t*>(src + i));
    const float2 w = bf2_to_f2(*reinterpret_cast<const uint32_t*>(nw + i));
    *reinterpret_cast<uint32_t*>(hn + (long long)t * D + i) =
        f2_to_bf2(rms_apply(f.x, inv, w.x, two_round), rms_apply(f.y, inv, w.y, two_round));
  }
}

// ----------------------------------------------------------------------------- residual + norm
// y   = Linear output: bf16(sum over splits of fp32 partials)   (SRC = 0)
//       or an already-rounded bf16 tensor                         (SRC = 1)
//       or absent (plain norm of h, stage input)                  (SRC = 2)
// h   = bf16(h + y)            (TransformerSelfAttentionLayer residual)
// hn  = RMSNorm(h) * nw        (the next consumer's norm; skipped when nw == nullptr)
template <int NT, int SRC>
__global__ void __launch_bounds__(NT) xb_resid_norm_kernel(const float* __restrict__ part, SplitInfo si, long long split_stride,
                                                           const bf16* __restrict__ y, bf16* __restrict__ h,
                                                           const bf16* __restrict__ nw, bf16* __restrict__ hn, int D,
                                                           float eps, int two_round) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float red[32];
  extern __shared__ float row[];  // D floats: the new residual row
  const int t = blockIdx.x;
  float ss = 0.f;
  for (int i = threadIdx.x * 2; i < D; i += NT * 2) {
    const long long o = (long long)t * D + i;
    float2 hv = bf2_to_f2(*reinterpret_cast<const uint32_t*>(h + o));
    if (SRC == 0) {
      float a = 0.f, b = 0.f;
      const int splits = xb_n_segments(si, i);
      for (int s = 0; s < splits; ++s) {
        const float2 p = *reinterpret_cast<const float2*>(part + s * split_stride + o);
        a += p.x;
        b += p.y;
      }
      hv.x = rbf(hv.x + rbf(a));
      hv.y = rbf(hv.y + rbf(b));
    } else if (SRC == 1) {
      const float2 yv = bf2_to_f2(*reinterpret_cast<const uint32_t*>(y + o));
      hv.x = rbf(hv.x + yv.x);
      hv.y = rbf(hv.y + yv.y);
    }
    if (SRC != 2) *reinterpret_cast<uint32_t*>(h + o) = f2_to_bf2(hv.x, hv.y);
    row[i] = hv.x;
    row[i + 1] = hv.y;
    ss += hv.x * hv.x + hv.y * hv.y;
  }
  if (nw == nullptr) return;
  ss = block_sum<NT>(ss, red);
  const float inv = 1.0f / sqrtf(ss / (float)D + eps);
  for (int i = threadIdx.x * 2; i < D; i += NT * 2) {
    const float2 w = bf2_to_f2(*reinterpret_cast<const uint32_t*>(nw + i));
    *reinterpret_cast<uint32_t*>(hn + (long long)t * D + i) =
        f2_to_bf2(rms_apply(row[i], inv, w.x, two_round), rms_apply(row[i + 1], inv, w.y, two_round));
  }
}

// Decode variant: one 1024-thread block per token row, 4 consecutive columns per thread (float4
// reads of the split-K planes, all planes' loads issued before the adds), row sum of squares by
// shuffle + one shared-memory round.  Same arithmetic and rounding points as
// xb_resid_norm_kernel<., 0>; planes are summed in plane order.  (An 8-block cluster per row with a
// DSMEM exchange measured 5.7-6.0 us per launch at T <= 32: cluster scheduling and the cluster
// barrier cost more than the extra SMs bring for 100 KB of input.)
// Sum of squares of 4 row values and the RMSNorm store of a row held as v[VPT][4] per thread (columns
// (it * NT + tid) * 4 ..): the ONE accumulation order every decode-path norm kernel uses -- after a Linear
// (xb_resid_norm_row_kernel), at a stage's input (xb_norm_row_kernel), behind the peer hand-off (xb_recv_norm_row_kernel) --
// so that a pipeline cut anywhere in the model reproduces the unsplit result bit for bit, not just almost always.
__device__ __forceinline__ float row_ss4(const float* v) { return v[0] * v[0] + v[1] * v[1] + v[2] * v[2] + v[3] * v[3]; }

template <int NT, int VPT>
__device__ __forceinline__ void row_norm_store(const float (&v)[VPT][4], float ss, const bf16* __restrict__ nw, bf16* __restrict__ hn_row,
                                               int D, float eps, int two_round, float* red) {
  ss = block_sum<NT>(ss, red);
  const float inv = 1.0f / sqrtf(ss / (float)D + eps);
#pragma unroll
  for (int it = 0; it < VPT; ++it) {
    const int c = (it * NT + threadIdx.x) * 4;
    if (c < D) {
      const uint2 wu = *reinterpret_cast<const uint2*>(nw + c);
      const float2 w0 = bf2_to_f2(wu.x), w1 = bf2_to_f2(wu.y);
      uint2 r;
      r.x = f2_to_bf2(rms_apply(v[it][0], inv, w0.x, two_round), rms_apply(v[it][1], inv, w0.y, two_round));
      r.y = f2_to_bf2(rms_apply(v[it][2], inv, w1.x, two_round), rms_apply(v[it][3], inv, w1.y, two_round));
      *reinterpret_cast<uint2*>(hn_row + c) = r;
    }
  }
}

// stage input of a decode step: hn = RMSNorm(h) (h handed in by the caller; NCCL / plugin-API path)
template <int NT, int VPT>
__global__ void __launch_bounds__(NT) xb_norm_row_kernel(const bf16* __restrict__ h, const bf16* __restrict__ nw, bf16* __restrict__ hn,
                                                         int D, float eps, int two_round) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float red[32];
  const int t = blockIdx.x;
  float v[VPT][4];
  float ss = 0.f;
#pragma unroll
  for (int it = 0; it < VPT; ++it) {
    const int c = (it * NT + threadIdx.x) * 4;
    v[it][0] = v[it][1] = v[it][2] = v[it][3] = 0.f;
    if (c < D) {
      const uint2 hu = *reinterpret_cast<const uint2*>(h + (long long)t * D + c);
      const float2 h0 = bf2_to_f2(hu.x), h1 = bf2_to_f2(hu.y);
      v[it][0] = h0.x; v[it][1] = h0.y; v[it][2] = h1.x; v[it][3] = h1.y;
      ss += row_ss4(v[it]);
    }
  }
  row_norm_store<NT, VPT>(v, ss, nw, hn + (long long)t * D, D, eps, two_round, red);
}

template <int NT, int VPT>
__global__ void __launch_bounds__(NT) xb_resid_norm_row_kernel(const float* __restrict__ part, SplitInfo si,
                                                               long long split_stride, bf16* __restrict__ h,
                                                               const bf16* __restrict__ nw, bf16* __restrict__ hn, int D,
                                                               float eps, int two_round) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float red[32];
  const int t = blockIdx.x;
  float v[VPT][4];
  float ss = 0.f;
#pragma unroll
  for (int it = 0; it < VPT; ++it) {
    const int c = (it * NT + threadIdx.x) * 4;
    v[it][0] = v[it][1] = v[it][2] = v[it][3] = 0.f;
    if (c < D) {
      const long long o = (long long)t * D + c;
      const int planes = xb_n_segments(si, c);
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int s0 = 0; s0 < planes; s0 += 4) {
        float4 p[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
          p[u] = (s0 + u < planes) ? *reinterpret_cast<const float4*>(part + (s0 + u) * split_stride + o)
                                   : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int u = 0; u < 4; ++u) { a.x += p[u].x; a.y += p[u].y; a.z += p[u].z; a.w += p[u].w; }
      }
      const uint2 hu = *reinterpret_cast<const uint2*>(h + o);
      const float2 h0 = bf2_to_f2(hu.x), h1 = bf2_to_f2(hu.y);
      v[it][0] = rbf(h0.x + rbf(a.x));
      v[it][1] = rbf(h0.y + rbf(a.y));
      v[it][2] = rbf(h1.x + rbf(a.z));
      v[it][3] = rbf(h1.y + rbf(a.w));
      uint2 ho;
      ho.x = f2_to_bf2(v[it][0], v[it][1]);
      ho.y = f2_to_bf2(v[it][2], v[it][3]);
      *reinterpret_cast<uint2*>(h + o) = ho;
      ss += row_ss4(v[it]);
    }
  }
  if (nw == nullptr) return;
  row_norm_store<NT, VPT>(v, ss, nw, hn + (long long)t * D, D, eps, two_round, red);
}

// ----------------------------------------------------------------------------- peer hand-off (pipeline, decode)
// Replaces the gRPC SendTensor hop of the reference (Node.forward_tensor, orchestration/node.py:
// 424-443; GRPCPeerHandle.send_tensor, networking/grpc/grpc_peer_handle.py:117-136) inside one box:
// the producing stage's LAST kernel stores its hidden rows straight into the next GPU's landing
// buffer through an NVLink peer mapping and then releases a flag; the consuming stage's FIRST
// kernel acquires the flag and fuses the read with its RMSNorm.  No copy kernel, no NCCL launch.
struct HandoffCtl {      // lives in device memory next to the batch metadata (graph replays re-read it)
  int slot_in, wait_value, slot_out, signal_value;
};

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// consumer head: wait for landing[slot_in] (flag >= wait_value), h = landing row, hn = RMSNorm(h)
template <int NT>
__global__ void __launch_bounds__(NT) xb_recv_norm_kernel(const bf16* __restrict__ landing, const uint32_t* __restrict__ flags,
                                                          const HandoffCtl* __restrict__ ctl, long long slot_elems,
                                                          bf16* __restrict__ h, const bf16* __restrict__ nw,
                                                          bf16* __restrict__ hn, int D, float eps, int two_round,
                                                          long long timeout_cycles, int* __restrict__ err_word) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float red[32];
  extern __shared__ float row[];
  const int t = blockIdx.x;
  const int slot = ctl->slot_in;
  const uint32_t want = (uint32_t)ctl->wait_value;
  if (threadIdx.x == 0) {
    // A producer that is merely slow (its first call of a shape captures a graph, a long prefill chunk is
    // queued ahead of it) must not cost the consumer its CUDA context: on time-out the slot is recorded in
    // a host-visible word (checked by the next xb_forward* call) and the step runs on to completion.
    long long t0 = clock64();
    while ((int32_t)(ld_acquire_sys(flags + slot) - want) < 0) {
      if (clock64() - t0 > timeout_cycles) {
        if (t == 0) {
          err_word[1] = (int)want;
          __threadfence_system();
          *reinterpret_cast<volatile int*>(err_word) = slot + 1;
        }
        break;
      }
    }
  }
  __syncthreads();
  const bf16* src = landing + (long long)slot * slot_elems + (long long)t * D;
  float ss = 0.f;
  for (int i = threadIdx.x * 2; i < D; i += NT * 2) {
    uint32_t u;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(u) : "l"(src + i) : "memory");   // written by a peer GPU
    *reinterpret_cast<uint32_t*>(h + (long long)t * D + i) = u;
    const float2 f = bf2_to_f2(u);
    row[i] = f.x;
    row[i + 1] = f.y;
    ss += f.x * f.x + f.y * f.y;
  }
  ss = block_sum<NT>(ss, red);
  const float inv = 1.0f / sqrtf(ss / (float)D + eps);
  for (int i = threadIdx.x * 2; i < D; i += NT * 2) {
    const float2 w = bf2_to_f2(*reinterpret_cast<const uint32_t*>(nw + i));
    *reinterpret_cast<uint32_t*>(hn + (long long)t * D + i) =
        f2_to_bf2(rms_apply(row[i], inv, w.x, two_round), rms_apply(row[i + 1], inv, w.y, two_round));
  }
}

// the same consumer head with the decode path's row-kernel layout and accumulation order (see row_ss4)
template <int NT, int VPT>
__global__ void __launch_bounds__(NT) xb_recv_norm_row_kernel(const bf16* __restrict__ landing, const uint32_t* __restrict__ flags,
                                                              const HandoffCtl* __restrict__ ctl, long long slot_elems,
                                                              bf16* __restrict__ h, const bf16* __restrict__ nw,
                                                              bf16* __restrict__ hn, int D, float eps, int two_round,
                                                              long long timeout_cycles, int* __restrict__ err_word) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float red[32];
  const int t = blockIdx.x;
  const int slot = ctl->slot_in;
  const uint32_t want = (uint32_t)ctl->wait_value;
  if (threadIdx.x == 0) {
    long long t0 = clock64();
    while ((int32_t)(ld_acquire_sys(flags + slot) - want) < 0) {
      if (clock64() - t0 > timeout_cycles) {   // reported through the host-visible word, see xb_recv_norm_kernel
        if (t == 0) {
          err_word[1] = (int)want;
          __threadfence_system();
          *reinterpret_cast<volatile int*>(err_word) = slot + 1;
        }
        break;
      }
    }
  }
  __syncthreads();
  const bf16* src = landing + (long long)slot * slot_elems + (long long)t * D;
  float v[VPT][4];
  float ss = 0.f;
#pragma unroll
  for (int it = 0; it < VPT; ++it) {
    const int c = (it * NT + threadIdx.x) * 4;
    v[it][0] = v[it][1] = v[it][2] = v[it][3] = 0.f;
    if (c < D) {
      uint2 hu;
      asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(hu.x), "=r"(hu.y) : "l"(src + c) : "memory");   // written by a peer GPU
      *reinterpret_cast<uint2*>(h + (long long)t * D + c) = hu;
      const float2 h0 = bf2_to_f2(hu.x), h1 = bf2_to_f2(hu.y);
      v[it][0] = h0.x; v[it][1] = h0.y; v[it][2] = h1.x; v[it][3] = h1.y;
      ss += row_ss4(v[it]);
    }
  }
  row_norm_store<NT, VPT>(v, ss, nw, hn + (long long)t * D, D, eps, two_round, red);
}

// producer tail: split-K reduce + Linear rounding + residual add (the stage's final hidden rows),
// stored locally AND into the next GPU's landing[slot_out]; the last block to finish releases the flag.
template <int NT>
__global__ void __launch_bounds__(NT) xb_resid_send_kernel(const float* __restrict__ part, SplitInfo si, long long split_stride,
                                                           bf16* __restrict__ h, bf16* __restrict__ peer_landing,
                                                           uint32_t* __restrict__ peer_flags, unsigned int* __restrict__ done_cnt,
                                                           const HandoffCtl* __restrict__ ctl, long long slot_elems, int D) {
  pdl_launch_dependents();
  pdl_wait();
  const int t = blockIdx.x;
  const int slot = ctl->slot_out;
  bf16* dst = peer_landing + (long long)slot * slot_elems + (long long)t * D;
  for (int i = threadIdx.x * 4; i < D; i += NT * 4) {
    const long long o = (long long)t * D + i;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    const int splits = xb_n_segments(si, i);
    for (int s = 0; s < splits; ++s) {
      const float4 p = *reinterpret_cast<const float4*>(part + s * split_stride + o);
      a.x += p.x; a.y += p.y; a.z += p.z; a.w += p.w;
    }
    const uint2 hu = *reinterpret_cast<const uint2*>(h + o);
    const float2 h0 = bf2_to_f2(hu.x), h1 = bf2_to_f2(hu.y);
    uint2 r;
    r.x = f2_to_bf2(h0.x + rbf(a.x), h0.y + rbf(a.y));
    r.y = f2_to_bf2(h1.x + rbf(a.z), h1.y + rbf(a.w));
    *reinterpret_cast<uint2*>(h + o) = r;
    *reinterpret_cast<uint2*>(dst + i) = r;          // NVLink peer store
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(done_cnt + slot, 1u);
    if (prev == gridDim.x - 1) {
      done_cnt[slot] = 0;
      __threadfence_system();
      st_release_sys(peer_flags + slot, (uint32_t)ctl->signal_value);
    }
  }
}

// ----------------------------------------------------------------------------- QKV finalize
// Decode path (the prefill GEMM does the same work in its epilogue, xb_gemm.cuh epi_wide_qkv).
// One block per token over the fused [q | k | v] projection output, 4 consecutive elements per
// work item (float4 reads of the split-K planes):
//   value = bf16(sum_planes(partial) + bias)        (nn.Linear rounding)
//   q,k heads: RoPE at the token's position, table = [pos][hd/2][cos,sin] fp32
//     rope_kind 0: interleaved pairs (2i,2i+1), fp32 math, one rounding (Llama3ScaledRoPE /
//                  RotaryPositionalEmbeddings as built at general_mha.py:35-40,58-63)
//     rope_kind 1: half-split pairs (i,i+hd/2), every product and the sum rounded to bf16
//                  (Qwen2RotaryPositionalEmbeddings, general_mha.py:47-51)
//   q -> q_out[t][h*hd..]; k,v -> the paged cache (KVCache.update, llm_utils.py:320-345)
__device__ __forceinline__ void qkv_load4(const float* __restrict__ part, const SplitInfo& si, long long split_stride,
                                          long long base, int col, const bf16* __restrict__ bias, float* x) {
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
  const int planes = xb_n_segments(si, col);
  for (int s = 0; s < planes; ++s) {
    const float4 p = *reinterpret_cast<const float4*>(part + s * split_stride + base + col);
    a.x += p.x; a.y += p.y; a.z += p.z; a.w += p.w;
  }
  if (bias != nullptr) {
    const uint2 bu = *reinterpret_cast<const uint2*>(bias + col);
    const float2 b0 = bf2_to_f2(bu.x), b1 = bf2_to_f2(bu.y);
    a.x += b0.x; a.y += b0.y; a.z += b1.x; a.w += b1.y;
  }
  x[0] = rbf(a.x); x[1] = rbf(a.y); x[2] = rbf(a.z); x[3] = rbf(a.w);
}

// One work item of the finalize: 4 consecutive elements [c, c+4) of fused-projection head `head` (interleaved
// RoPE: two pairs) or, for half-split RoPE, those 4 plus their partners [half+c, half+c+4).  x (and y) receive the
// finished bf16-representable values.  Shared by the stand-alone kernel below and by the decode attention kernel,
// which does this work in its prologue (xb_attn_mma.cuh) -- one arithmetic, bit-identical results.
__device__ __forceinline__ void qkv_item(const float* __restrict__ part, const SplitInfo& si, long long split_stride,
                                         long long base, int head, int c, int hd, bool rot, const bf16* __restrict__ bias,
                                         const float* __restrict__ rope_row, int rope_kind, float* x, float* y) {
  const int half = hd >> 1;
  qkv_load4(part, si, split_stride, base, head * hd + c, bias, x);
  if (rope_kind == 0) {
    if (rot) {
      const float4 cs = *reinterpret_cast<const float4*>(rope_row + c);   // (cos,sin) of pairs c/2, c/2+1
      const float a0 = x[0], b0 = x[1], a1 = x[2], b1 = x[3];
      x[0] = __fsub_rn(__fmul_rn(a0, cs.x), __fmul_rn(b0, cs.y));
      x[1] = __fadd_rn(__fmul_rn(b0, cs.x), __fmul_rn(a0, cs.y));
      x[2] = __fsub_rn(__fmul_rn(a1, cs.z), __fmul_rn(b1, cs.w));
      x[3] = __fadd_rn(__fmul_rn(b1, cs.z), __fmul_rn(a1, cs.w));
    }
  } else {
    qkv_load4(part, si, split_stride, base, head * hd + half + c, bias, y);
    if (rot) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 cs = *reinterpret_cast<const float2*>(rope_row + (c + j) * 2);
        const float cc = rbf(cs.x), ss = rbf(cs.y);
        const float a = x[j], b = y[j];
        x[j] = __fadd_rn(rbf(__fmul_rn(a, cc)), rbf(__fmul_rn(-b, ss)));
        y[j] = __fadd_rn(rbf(__fmul_rn(b, cc)), rbf(__fmul_rn(a, ss)));
      }
    }
  }
}
__device__ __forceinline__ void qkv_store4(bf16* dst, const float* x) {
  uint2 o;
  o.x = f2_to_bf2(x[0], x[1]);
  o.y = f2_to_bf2(x[2], x[3]);
  *reinterpret_cast<uint2*>(dst) = o;
}

__global__ void __launch_bounds__(256) xb_qkv_finalize_kernel(const float* __restrict__ part, SplitInfo si, long long split_stride,
                                                              const bf16* __restrict__ bias, const float* __restrict__ rope,
                                                              bf16* __restrict__ q_out, bf16* __restrict__ k_pool,
                                                              bf16* __restrict__ v_pool, BatchMeta m, int H, int KV, int hd,
                                                              int rope_kind) {
  pdl_launch_dependents();
  pdl_wait();
  const int t = blockIdx.x;
  const int heads = H + 2 * KV;
  const int half = hd >> 1;
  const long long base = (long long)t * heads * hd;
  const int pos = m.tok_pos[t];
  const int slot = m.seq_slot[m.tok_seq[t]];
  const int page = m.page_table[(long long)slot * m.max_pages + pos / m.page_tokens];
  const int off = pos % m.page_tokens;
  const float* rope_row = rope + (long long)pos * half * 2;
  // work items per head: hd/4 (interleaved: 4 consecutive elements = 2 pairs) or hd/8 (half-split:
  // 4 elements of the low half + their 4 partners of the high half)
  const int iph = rope_kind == 0 ? hd / 4 : hd / 8;
  // grid = (tokens, ceil(items / 256)): one work item per thread, so a 1-token decode step still spreads out
  const int it = blockIdx.y * blockDim.x + threadIdx.x;
  if (it < heads * iph) {
    const int head = it / iph, c = (it % iph) * 4;
    bf16* dst;
    if (head < H) {
      dst = q_out + (long long)t * H * hd + (long long)head * hd;
    } else {
      const int kvh = (head - H) % KV;
      bf16* pool = (head < H + KV) ? k_pool : v_pool;
      dst = pool + (((long long)page * KV + kvh) * m.page_tokens + off) * hd;
    }
    float x[4], y[4];
    qkv_item(part, si, split_stride, base, head, c, hd, head < H + KV, bias, rope_row, rope_kind, x, y);
    qkv_store4(dst + c, x);
    if (rope_kind != 0) qkv_store4(dst + half + c, y);
  }
}

// ----------------------------------------------------------------------------- SiLU(gate) * up
// FeedForward (llm_utils.py:513-522): w2(silu(w1 x) * w3 x); silu output and the product
// are separate bf16 ops in the reference.  Input = split-K partial planes of the fused gate/up
// projection (decode path; the prefill GEMM does this in its epilogue, xb_gemm.cuh epi_wide_silu).
// Fused gate/up columns are interleaved in 16-feature blocks; 4 consecutive features per thread.
__global__ void __launch_bounds__(256) xb_silu_mul_kernel(const float* __restrict__ part, SplitInfo si, long long split_stride,
                                                          bf16* __restrict__ act, int T, int I) {
  pdl_launch_dependents();
  pdl_wait();
  const int i4 = I >> 2;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)T * i4) return;
  const int t = (int)(idx / i4), i = (int)(idx % i4) * 4;
  const int cg = (i >> 4) * 32 + (i & 15), cu = cg + 16;     // same 128-column tile: (i>>4)*32 is a multiple of 32
  const long long row = (long long)t * 2 * I;
  const int planes = xb_n_segments(si, cg);                  // cg and cu lie in one 128-column tile
  float4 g = make_float4(0.f, 0.f, 0.f, 0.f), u = g;
  for (int s = 0; s < planes; ++s) {
    const float4 pg = *reinterpret_cast<const float4*>(part + s * split_stride + row + cg);
    const float4 pu = *reinterpret_cast<const float4*>(part + s * split_stride + row + cu);
    g.x += pg.x; g.y += pg.y; g.z += pg.z; g.w += pg.w;
    u.x += pu.x; u.y += pu.y; u.z += pu.z; u.w += pu.w;
  }
  const float gv[4] = {rbf(g.x), rbf(g.y), rbf(g.z), rbf(g.w)};
  const float uv[4] = {rbf(u.x), rbf(u.y), rbf(u.z), rbf(u.w)};
  float o[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) o[j] = rbf(gv[j] / (1.0f + expf(-gv[j]))) * uv[j];
  uint2 w;
  w.x = f2_to_bf2(o[0], o[1]);
  w.y = f2_to_bf2(o[2], o[3]);
  *reinterpret_cast<uint2*>(act + (long long)t * I + i) = w;
}

// ----------------------------------------------------------------------------- misc
// rows[i] = src[idx[i]]  (last-position gather before the LM head: the reference runs the
// head on all s positions and slices [:, -1], sharded_inference_engine.py:366)
__global__ void xb_gather_rows_kernel(const bf16* __restrict__ src, const int* __restrict__ idx, bf16* __restrict__ dst, int D) {
  pdl_launch_dependents();
  pdl_wait();
  const int r = blockIdx.x;
  const bf16* s = src + (long long)idx[r] * D;
  for (int i = threadIdx.x * 2; i < D; i += blockDim.x * 2)
    *reinterpret_cast<uint32_t*>(dst + (long long)r * D + i) = *reinterpret_cast<const uint32_t*>(s + i);
}

// ----------------------------------------------------------------------------- sampling
// Restates torchtune.generation.sample as the reference calls it
// (sharded_inference_engine.py:208-228): logits / max(temp, 1e-5); values below the
// k-th largest -> -inf; softmax; argmax(probs / q), q ~ Exp(1).
__device__ __forceinline__ uint32_t float_order_key(float f) {
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);   // monotone: larger float -> larger key
}
__device__ __forceinline__ float key_to_float(uint32_t k) {
  const uint32_t u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
  return __uint_as_float(u);
}
__device__ __forceinline__ uint32_t mix32(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
  return (uint32_t)(x >> 11);
}

// f(i, v) for every element of a row, one block per row.  A 128 K-entry fp32 row is read from L2 by a
// single SM, so the loop is latency bound unless many loads are in flight: four float4 loads per
// thread per trip (64 B x 1024 threads) instead of one scalar.  (Even so one SM pulls only ~40 GB/s of
// a row out of L2 -- ~14 us per pass over 128256 logits -- which is why the sampler proper keeps its
// slice of the row in shared memory across passes, xb_sample_cluster.cuh.)
template <int NT, typename F>
__device__ __forceinline__ void row_for_each(const float* __restrict__ row, int V, F&& f) {
  if ((V & 3) == 0 && (reinterpret_cast<uintptr_t>(row) & 15) == 0) {
    const float4* r4 = reinterpret_cast<const float4*>(row);
    const int n4 = V >> 2;
    int i = threadIdx.x;
    for (; i + 3 * NT < n4; i += 4 * NT) {
      const float4 a = r4[i], b = r4[i + NT], c = r4[i + 2 * NT], d = r4[i + 3 * NT];
      f(4 * i, a.x); f(4 * i + 1, a.y); f(4 * i + 2, a.z); f(4 * i + 3, a.w);
      f(4 * (i + NT), b.x); f(4 * (i + NT) + 1, b.y); f(4 * (i + NT) + 2, b.z); f(4 * (i + NT) + 3, b.w);
      f(4 * (i + 2 * NT), c.x); f(4 * (i + 2 * NT) + 1, c.y); f(4 * (i + 2 * NT) + 2, c.z); f(4 * (i + 2 * NT) + 3, c.w);
      f(4 * (i + 3 * NT), d.x); f(4 * (i + 3 * NT) + 1, d.y); f(4 * (i + 3 * NT) + 2, d.z); f(4 * (i + 3 * NT) + 3, d.w);
    }
    for (; i < n4; i += NT) {
      const float4 a = r4[i];
      f(4 * i, a.x); f(4 * i + 1, a.y); f(4 * i + 2, a.z); f(4 * i + 3, a.w);
    }
  } else {
    for (int i = threadIdx.x; i < V; i += NT) f(i, row[i]);
  }
}
// first-index argmax of each row (greedy): one block per row
template <int NT>
__global__ void __launch_bounds__(NT) xb_argmax_kernel(const float* __restrict__ logits, int V, int* __restrict__ out) {
  __shared__ float sv[NT / 32];
  __shared__ int si[NT / 32];
  const float* row = logits + (long long)blockIdx.x * V;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  row_for_each<NT>(row, V, [&](int i, float v) {
    if (v > best || (v == best && i < bi)) { best = v; bi = i; }
  });
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
  }
  if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = best; si[threadIdx.x >> 5] = bi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < NT / 32; ++w)
      if (sv[w] > best || (sv[w] == best && si[w] < bi)) { best = sv[w]; bi = si[w]; }
    out[blockIdx.x] = bi == 0x7fffffff ? 0 : bi;
  }
}

// temperature / top-k / top-p sampling: one block per row.
//  1. radix-select (4 passes of 8 bits over order-preserving keys) the k-th largest scaled logit
//  2. softmax over the kept set; optional nucleus cut: keep the smallest set of largest
//     probabilities whose mass reaches top_p (second radix descent, on mass instead of count)
//  3. argmax(p / q), q = caller-supplied Exp(1) noise (reference-exact) or counter-hash noise
template <int NT>
__global__ void __launch_bounds__(NT) xb_sample_kernel(const float* __restrict__ logits, int V, float temp, int top_k,
                                                       float top_p, const float* __restrict__ q_noise,
                                                       unsigned long long seed, int* __restrict__ out) {
  __shared__ unsigned int hist[256];
  __shared__ float fhist[256];
  __shared__ uint32_t s_prefix;
  __shared__ int s_need;
  __shared__ float s_needf;
  __shared__ float red[32];
  __shared__ float sv[NT / 32];
  __shared__ int si[NT / 32];
  const int rowi = blockIdx.x;
  const float* row = logits + (long long)rowi * V;
  const float inv_t = 1.0f / fmaxf(temp, 1e-5f);

  // ---- 1. pivot = k-th largest scaled logit
  uint32_t pivot_key = 0;
  if (top_k > 0 && top_k < V) {
    uint32_t prefix = 0;
    int need = top_k;
    for (int pass = 0; pass < 4; ++pass) {
      const int shift = 24 - 8 * pass;
      for (int i = threadIdx.x; i < 256; i += NT) hist[i] = 0;
      __syncthreads();
      const uint32_t mask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
      row_for_each<NT>(row, V, [&](int, float x) {
        const uint32_t key = float_order_key(x * inv_t);
        if ((key & mask) == (prefix & mask)) atomicAdd(&hist[(key >> shift) & 255], 1u);
      });
      __syncthreads();
      if (threadIdx.x == 0) {
        int n = need;
        int b = 255;
        for (; b > 0; --b) {
          if ((int)hist[b] >= n) break;
          n -= (int)hist[b];
        }
        s_prefix = prefix | ((uint32_t)b << shift);
        s_need = n;
      }
      __syncthreads();
      prefix = s_prefix;
      need = s_need;
      __syncthreads();
    }
    pivot_key = prefix;
  }
  const float pivot = (top_k > 0 && top_k < V) ? key_to_float(pivot_key) : -INFINITY;

  // ---- 2. softmax statistics over kept values (logits >= pivot)
  float mxv = -INFINITY;
  row_for_each<NT>(row, V, [&](int, float x) {
    const float v = x * inv_t;
    if (v >= pivot) mxv = fmaxf(mxv, v);
  });
  mxv = warp_max(mxv);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mxv;
  __syncthreads();
  mxv = red[0];
  for (int w = 1; w < NT / 32; ++w) mxv = fmaxf(mxv, red[w]);
  __syncthreads();
  float sum = 0.f;
  row_for_each<NT>(row, V, [&](int, float x) {
    const float v = x * inv_t;
    if (v >= pivot) sum += expf(v - mxv);
  });
  sum = block_sum<NT>(sum, red);

  // ---- 2b. nucleus: largest probability threshold whose upper set has mass >= top_p
  float p_cut = 0.f;
  if (top_p > 0.f && top_p < 1.f) {
    uint32_t prefix = 0;
    float needf = top_p * sum;   // in un-normalised mass
    for (int pass = 0; pass < 4; ++pass) {
      const int shift = 24 - 8 * pass;
      for (int i = threadIdx.x; i < 256; i += NT) fhist[i] = 0.f;
      __syncthreads();
      const uint32_t mask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
      row_for_each<NT>(row, V, [&](int, float x) {
        const float v = x * inv_t;
        if (v >= pivot) {
          const float e = expf(v - mxv);
          const uint32_t key = float_order_key(e);
          if ((key & mask) == (prefix & mask)) atomicAdd(&fhist[(key >> shift) & 255], e);
        }
      });
      __syncthreads();
      if (threadIdx.x == 0) {
        float n = needf;
        int b = 255;
        for (; b > 0; --b) {
          if (fhist[b] >= n) break;
          n -= fhist[b];
        }
        s_prefix = prefix | ((uint32_t)b << shift);
        s_needf = n;
      }
      __syncthreads();
      prefix = s_prefix;
      needf = s_needf;
      __syncthreads();
    }
    p_cut = key_to_float(prefix);
  }

  // ---- 3. argmax(p / q)
  float best = -INFINITY;
  int bi = 0x7fffffff;
  row_for_each<NT>(row, V, [&](int i, float x) {
    const float v = x * inv_t;
    if (v < pivot) return;
    const float e = expf(v - mxv);
    if (e < p_cut) return;
    float qv;
    if (q_noise != nullptr) {
      qv = q_noise[(long long)rowi * V + i];
    } else {
      const uint32_t r = mix32(seed * 0x9E3779B97F4A7C15ULL + ((unsigned long long)rowi << 32) + (unsigned)i);
      const float u = ((float)(r & 0xffffff) + 0.5f) * (1.0f / 16777216.0f);
      qv = -logf(u);
    }
    const float sc = (e / sum) / qv;
    if (sc > best || (sc == best && i < bi)) { best = sc; bi = i; }
  });
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
  }
  if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = best; si[threadIdx.x >> 5] = bi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < NT / 32; ++w)
      if (sv[w] > best || (sv[w] == best && si[w] < bi)) { best = sv[w]; bi = si[w]; }
    out[rowi] = bi == 0x7fffffff ? 0 : bi;
  }
}
