// Tensor-core attention over the paged KV cache (replaces torchtune MultiHeadAttention's
// GQA-expand + F.scaled_dot_product_attention, built at the reference's
// xotorch/inference/torch/models/general_mha.py:78-106).
//
// K/V tiles of 64 keys x head_dim arrive by TMA (128B-swizzled boxes cut from the paged pool,
// one producer warp, mbarrier full/empty ring); the math is mma.sync.m16n8k16 bf16 -> fp32
// with an online softmax in registers (flash-attention style).  Query rows of a tile are
// (token, head-in-group) pairs of ONE kv head, so K/V are read once per GQA group.
//   * decode kernel: the G heads of one token form one 16-row tile (zero padded); the 4
//     compute warps split the KEYS of each tile; blocks split the context (split-KV) and a
//     combine kernel merges the partial (max, sum, acc).
//   * prefill kernel: 128 rows = (128/G) consecutive tokens x G heads; the 8 compute warps
//     split the ROWS; causal masking by each row's absolute position.
#pragma once
#include "xb_common.cuh"
#include "xb_kernels.cuh"

#define XB_ATT_TILE 64   // keys per smem tile

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16_16816(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// byte offset of (key row r, 16-byte chunk c16 of the head dim) inside a [64 x HD] tile made of
// HD/64 swizzled boxes of [64 rows x 128 B]
__device__ __forceinline__ uint32_t att_tile_off(int r, int c16) {
  return (uint32_t)((c16 >> 3) * (XB_ATT_TILE * 128) + r * 128 + (((c16 & 7) ^ (r & 7)) << 4));
}

// One warp: online-softmax update of its 16 query rows against NKB*16 keys of a staged tile.
//   qf      : A fragments of the 16 x HD query block (HD/16 k-steps x 4 regs), exact bf16 q;
//             scores are scaled by scale_log2 = softmax_scale * log2(e) in fp32
//   sK, sV  : shared-memory byte addresses of the K / V tile
//   krow0   : first key row (inside the tile) this warp handles; it handles NKB*16 rows
//   key_abs0: absolute position of tile row 0;  lim0/lim1: last visible key of the thread's two rows
template <int HD, int NKB>
__device__ __forceinline__ void att_warp_block(const uint32_t (*qf)[4], uint32_t sK, uint32_t sV, int krow0, int key_abs0,
                                               int lim0, int lim1, float scale_log2, float (&mrow)[2],
                                               float (&lrow)[2], float (*o)[4]) {
  constexpr int KS = HD / 16;   // k-steps over the head dim
  constexpr int DT = HD / 8;    // output n-tiles over the head dim
  constexpr int NT = NKB * 2;   // score n-tiles (8 keys each)
  const int lane = threadIdx.x & 31;
  const int mi = lane >> 3, lr = lane & 7;
  float s[NT][4];
#pragma unroll
  for (int j = 0; j < NT; ++j) s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
  // ---- S = Q K^T
#pragma unroll
  for (int kb = 0; kb < NKB; ++kb) {
    const int krow = krow0 + kb * 16 + ((mi >> 1) << 3) + lr;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      uint32_t b0, b1, b2, b3;
      ldsm_x4(sK + att_tile_off(krow, ks * 2 + (mi & 1)), b0, b1, b2, b3);
      mma_bf16_16816(s[kb * 2], qf[ks], b0, b1);
      mma_bf16_16816(s[kb * 2 + 1], qf[ks], b2, b3);
    }
  }
  // ---- causal / length mask + row max
  const int kq = key_abs0 + krow0 + ((lane & 3) << 1);
  float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
  for (int j = 0; j < NT; ++j) {
    const int k0 = kq + j * 8;
    s[j][0] *= scale_log2; s[j][1] *= scale_log2; s[j][2] *= scale_log2; s[j][3] *= scale_log2;
    if (k0 > lim0) s[j][0] = -INFINITY;
    if (k0 + 1 > lim0) s[j][1] = -INFINITY;
    if (k0 > lim1) s[j][2] = -INFINITY;
    if (k0 + 1 > lim1) s[j][3] = -INFINITY;
    mx0 = fmaxf(mx0, fmaxf(s[j][0], s[j][1]));
    mx1 = fmaxf(mx1, fmaxf(s[j][2], s[j][3]));
  }
  mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
  mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
  mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
  mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
  const float mn0 = fmaxf(mrow[0], mx0), mn1 = fmaxf(mrow[1], mx1);
  const float ref0 = mn0 == -INFINITY ? 0.f : mn0, ref1 = mn1 == -INFINITY ? 0.f : mn1;
  const float c0 = exp2f(mrow[0] - ref0), c1 = exp2f(mrow[1] - ref1);   // exp2(-inf) = 0 on the first block
  mrow[0] = mn0;
  mrow[1] = mn1;
  float sum0 = 0.f, sum1 = 0.f;
  uint32_t pa[NKB][4];
#pragma unroll
  for (int j = 0; j < NT; ++j) {
    const float p0 = exp2f(s[j][0] - ref0), p1 = exp2f(s[j][1] - ref0);
    const float p2 = exp2f(s[j][2] - ref1), p3 = exp2f(s[j][3] - ref1);
    sum0 += p0 + p1;
    sum1 += p2 + p3;
    pa[j >> 1][(j & 1) * 2] = f2_to_bf2(p0, p1);
    pa[j >> 1][(j & 1) * 2 + 1] = f2_to_bf2(p2, p3);
  }
  lrow[0] = lrow[0] * c0 + sum0;
  lrow[1] = lrow[1] * c1 + sum1;
#pragma unroll
  for (int d = 0; d < DT; ++d) {
    o[d][0] *= c0; o[d][1] *= c0;
    o[d][2] *= c1; o[d][3] *= c1;
  }
  // ---- O += P V
#pragma unroll
  for (int kb = 0; kb < NKB; ++kb) {
    const int krow = krow0 + kb * 16 + ((mi & 1) << 3) + lr;
#pragma unroll
    for (int d = 0; d < DT; d += 2) {
      uint32_t b0, b1, b2, b3;
      ldsm_x4_t(sV + att_tile_off(krow, d + (mi >> 1)), b0, b1, b2, b3);
      mma_bf16_16816(o[d], pa[kb], b0, b1);
      mma_bf16_16816(o[d + 1], pa[kb], b2, b3);
    }
  }
}

// producer: stage one 64-key tile of K and V of kv head `kvh` (absolute tile index kt of the slot)
template <int HD>
__device__ __forceinline__ void att_issue_tile(const CUtensorMap* tmK, const CUtensorMap* tmV, uint8_t* sK, uint8_t* sV,
                                               uint64_t* bar, const int* pt, int max_pages, int kt, int kvh, int KV, int P) {
  constexpr int TILE_BYTES = XB_ATT_TILE * HD * 2;
  mbar_expect_tx(bar, 2 * TILE_BYTES);
  const int BR = P < XB_ATT_TILE ? P : XB_ATT_TILE;   // rows per TMA box
  for (int rg = 0; rg < XB_ATT_TILE / BR; ++rg) {
    const int key = kt * XB_ATT_TILE + rg * BR;
    const int pi = key / P;
    const int page = pt[pi < max_pages ? pi : max_pages - 1];   // rows past the context are masked
    const int row = (page * KV + kvh) * P + (key % P);
#pragma unroll
    for (int db = 0; db < HD / 64; ++db) {
      tma_load_2d(sK + db * (XB_ATT_TILE * 128) + rg * BR * 128, tmK, bar, db * 64, row);
      tma_load_2d(sV + db * (XB_ATT_TILE * 128) + rg * BR * 128, tmV, bar, db * 64, row);
    }
  }
}

// ----------------------------------------------------------------------------- decode
template <int HD, int G, int NST>
struct AttDecCfg {
  static constexpr int TILE_BYTES = XB_ATT_TILE * HD * 2;
  static constexpr int STAGE_BYTES = 2 * TILE_BYTES;
  static constexpr int MERGE_BYTES = 4 * G * HD * 4 + 4 * 16 * 2 * 4;
  static constexpr int SMEM_BYTES = NST * STAGE_BYTES + MERGE_BYTES + 1024 + 128;
};

template <int HD, int G, int NST>
__global__ void __launch_bounds__(160) xb_attn_decode_mma_kernel(
    const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV, const bf16* __restrict__ q,
    BatchMeta m, float* __restrict__ o_part, float* __restrict__ ml_part, bf16* __restrict__ out, int H, int KV,
    int n_splits, float scale_log2) {
  using Cfg = AttDecCfg<HD, G, NST>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* stages = smem;
  float* mg_o = reinterpret_cast<float*>(smem + NST * Cfg::STAGE_BYTES);   // [4 warps][G][HD]
  float* mg_ml = mg_o + 4 * G * HD;                                       // [4 warps][16][2]
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(mg_ml + 4 * 16 * 2);
  uint64_t* empty_bar = full_bar + NST;

  const int t = blockIdx.x, kvh = blockIdx.y, sp = blockIdx.z;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_launch_dependents();
  pdl_wait();
  const int pos = m.tok_pos[t];
  const int slot = m.seq_slot[m.tok_seq[t]];
  const int* pt = m.page_table + (long long)slot * m.max_pages;
  const int n_tiles = pos / XB_ATT_TILE + 1;
  const int kt0 = (int)(((long long)sp * n_tiles) / n_splits);
  const int kt1 = (int)(((long long)(sp + 1) * n_tiles) / n_splits);

  if (threadIdx.x == 0) {
    for (int i = 0; i < NST; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 4);
    }
    fence_mbar_init();
  }
  __syncthreads();

  if (warp == 4) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      tma_prefetch_desc(&tmK);
      tma_prefetch_desc(&tmV);
      int stage = 0;
      uint32_t phase = 0;
      for (int kt = kt0; kt < kt1; ++kt) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sK = stages + stage * Cfg::STAGE_BYTES;
        att_issue_tile<HD>(&tmK, &tmV, sK, sK + Cfg::TILE_BYTES, &full_bar[stage], pt, m.max_pages, kt, kvh, KV, m.page_tokens);
        if (++stage == NST) { stage = 0; phase ^= 1; }
      }
    }
    return;
  }

  // ===================== 4 compute warps: each takes 16 keys of every tile =====================
  constexpr int KS = HD / 16, DT = HD / 8;
  uint32_t qf[KS][4];
  {
    // A fragment rows: r0 = lane/4 (a head of the group when < G), r1 = r0 + 8 (always padding)
    const int r0 = lane >> 2;
    const bf16* qrow = q + ((long long)t * H + (kvh * G + (r0 < G ? r0 : 0))) * HD + ((lane & 3) << 1);
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      uint32_t lo = 0, hi = 0;
      if (r0 < G) {
        lo = *reinterpret_cast<const uint32_t*>(qrow + ks * 16);
        hi = *reinterpret_cast<const uint32_t*>(qrow + ks * 16 + 8);
      }
      qf[ks][0] = lo; qf[ks][1] = 0u; qf[ks][2] = hi; qf[ks][3] = 0u;
    }
  }
  float mrow[2] = {-INFINITY, -INFINITY}, lrow[2] = {0.f, 0.f};
  float o[DT][4];
#pragma unroll
  for (int d = 0; d < DT; ++d) o[d][0] = o[d][1] = o[d][2] = o[d][3] = 0.f;

  int stage = 0;
  uint32_t phase = 0;
  for (int kt = kt0; kt < kt1; ++kt) {
    mbar_wait(&full_bar[stage], phase);
    const int key_abs0 = kt * XB_ATT_TILE;
    if (key_abs0 + warp * 16 <= pos) {
      const uint32_t sK = smem_u32(stages + stage * Cfg::STAGE_BYTES);
      att_warp_block<HD, 1>(qf, sK, sK + Cfg::TILE_BYTES, warp * 16, key_abs0, pos, pos, scale_log2, mrow, lrow, o);
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty_bar[stage]);
    if (++stage == NST) { stage = 0; phase ^= 1; }
  }

  // ---- merge the 4 warps (each saw a disjoint key subset); only rows < G are real
  lrow[0] += __shfl_xor_sync(0xffffffffu, lrow[0], 1);
  lrow[0] += __shfl_xor_sync(0xffffffffu, lrow[0], 2);
  {
    const int r0 = lane >> 2;
    if (r0 < G) {
      if ((lane & 3) == 0) {
        mg_ml[(warp * 16 + r0) * 2] = mrow[0];
        mg_ml[(warp * 16 + r0) * 2 + 1] = lrow[0];
      }
#pragma unroll
      for (int d = 0; d < DT; ++d)
        *reinterpret_cast<float2*>(&mg_o[(warp * G + r0) * HD + d * 8 + ((lane & 3) << 1)]) = make_float2(o[d][0], o[d][1]);
    }
  }
  named_bar_sync(1, 128);
  for (int idx = threadIdx.x; idx < G * HD; idx += 128) {
    const int g = idx / HD, d = idx % HD;
    float M = -INFINITY;
#pragma unroll
    for (int w = 0; w < 4; ++w) M = fmaxf(M, mg_ml[(w * 16 + g) * 2]);
    float L = 0.f, A = 0.f;
    if (M > -INFINITY) {
#pragma unroll
      for (int w = 0; w < 4; ++w) {
        const float mw = mg_ml[(w * 16 + g) * 2];
        const float wt = mw == -INFINITY ? 0.f : exp2f(mw - M);
        L += mg_ml[(w * 16 + g) * 2 + 1] * wt;
        A += mg_o[(w * G + g) * HD + d] * wt;
      }
    }
    const int hh = kvh * G + g;
    if (n_splits == 1) {
      out[((long long)t * H + hh) * HD + d] = f2bf(L > 0.f ? A / L : 0.f);
    } else {
      const long long oo = ((long long)t * H + hh) * n_splits + sp;
      o_part[oo * HD + d] = A;
      if (d == 0) { ml_part[oo * 2] = M; ml_part[oo * 2 + 1] = L; }   // M in log2 units
    }
  }
}

// merge the kv splits of one (token, head); statistics are in log2 units (exp2 domain).
// One warp per (token, head); lane s holds split s (n_splits <= 32).
template <int HD>
__global__ void __launch_bounds__(128) xb_attn_combine2_kernel(const float* __restrict__ o_part,
                                                               const float* __restrict__ ml_part, bf16* __restrict__ out,
                                                               int n_splits, int n_rows) {
  const long long th = (long long)blockIdx.x * 4 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  pdl_launch_dependents();
  pdl_wait();
  if (th >= n_rows) return;
  float ms = -INFINITY, ls = 0.f;
  if (lane < n_splits) {
    const float2 v = *reinterpret_cast<const float2*>(ml_part + (th * n_splits + lane) * 2);
    ms = v.x; ls = v.y;
  }
  const float M = warp_max(ms);
  const float w = (ms == -INFINITY) ? 0.f : exp2f(ms - M);
  const float L = warp_sum(ls * w);
  constexpr int VPL = HD / 32;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int s0 = 0; s0 < n_splits; s0 += 8) {
    float4 v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {          // 8 independent loads in flight per lane
      v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (s0 + u < n_splits) {
        const float* src = o_part + (th * n_splits + s0 + u) * HD + lane * VPL;
        if (VPL == 4) v[u] = *reinterpret_cast<const float4*>(src);
        else { const float2 t2 = *reinterpret_cast<const float2*>(src); v[u].x = t2.x; v[u].y = t2.y; }
      }
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const float ws = __shfl_sync(0xffffffffu, w, (s0 + u) & 31);
      if (s0 + u < n_splits) {
        acc[0] = fmaf(ws, v[u].x, acc[0]); acc[1] = fmaf(ws, v[u].y, acc[1]);
        acc[2] = fmaf(ws, v[u].z, acc[2]); acc[3] = fmaf(ws, v[u].w, acc[3]);
      }
    }
  }
  const float invL = L > 0.f ? 1.0f / L : 0.f;
  bf16* dst = out + th * HD + lane * VPL;
#pragma unroll
  for (int j = 0; j < VPL; j += 2)
    *reinterpret_cast<uint32_t*>(dst + j) = f2_to_bf2(acc[j] * invL, acc[j + 1] * invL);
}

// ----------------------------------------------------------------------------- prefill
// Work item = (query tile, kv head).  A query tile is up to 128/G consecutive tokens of ONE
// sequence (tile table built on the host): tile_start[i] = first token row, tile_len[i] = #tokens.
template <int HD, int NST>
struct AttPreCfg {
  static constexpr int TILE_BYTES = XB_ATT_TILE * HD * 2;
  static constexpr int STAGE_BYTES = 2 * TILE_BYTES;
  static constexpr int SMEM_BYTES = NST * STAGE_BYTES + 1024 + 128;
};

template <int HD, int G, int NST>
__global__ void __launch_bounds__(288) xb_attn_prefill_mma_kernel(
    const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV, const bf16* __restrict__ q,
    BatchMeta m, const int* __restrict__ tile_start, const int* __restrict__ tile_len, bf16* __restrict__ out, int H,
    int KV, float scale_log2) {
  using Cfg = AttPreCfg<HD, NST>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* stages = smem;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + NST * Cfg::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + NST;

  const int tile = blockIdx.x, kvh = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_launch_dependents();
  pdl_wait();
  const int t0 = tile_start[tile];
  const int nt = tile_len[tile];
  const int pos0 = m.tok_pos[t0];                       // tokens of a tile have consecutive positions
  const int slot = m.seq_slot[m.tok_seq[t0]];
  const int* pt = m.page_table + (long long)slot * m.max_pages;
  const int last_pos = pos0 + nt - 1;
  const int n_tiles = last_pos / XB_ATT_TILE + 1;

  if (threadIdx.x == 0) {
    for (int i = 0; i < NST; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 8);
    }
    fence_mbar_init();
  }
  __syncthreads();

  if (warp == 8) {
    if (lane == 0) {
      tma_prefetch_desc(&tmK);
      tma_prefetch_desc(&tmV);
      int stage = 0;
      uint32_t phase = 0;
      for (int kt = 0; kt < n_tiles; ++kt) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sK = stages + stage * Cfg::STAGE_BYTES;
        att_issue_tile<HD>(&tmK, &tmV, sK, sK + Cfg::TILE_BYTES, &full_bar[stage], pt, m.max_pages, kt, kvh, KV, m.page_tokens);
        if (++stage == NST) { stage = 0; phase ^= 1; }
      }
    }
    return;
  }

  // ===================== 8 compute warps: 16 query rows each; row = (token, head-in-group) ====
  constexpr int KS = HD / 16, DT = HD / 8;
  const int ra = warp * 16 + (lane >> 2), rb = ra + 8;
  const int ta = ra / G, ga = ra % G, tb = rb / G, gb = rb % G;
  const bool va = ta < nt, vb = tb < nt;
  uint32_t qf[KS][4];
  {
    const bf16* qa = q + ((long long)(t0 + (va ? ta : 0)) * H + (kvh * G + ga)) * HD + ((lane & 3) << 1);
    const bf16* qb = q + ((long long)(t0 + (vb ? tb : 0)) * H + (kvh * G + gb)) * HD + ((lane & 3) << 1);
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      uint32_t a0 = 0, a1 = 0, a2 = 0, a3 = 0;
      if (va) {
        a0 = *reinterpret_cast<const uint32_t*>(qa + ks * 16);
        a2 = *reinterpret_cast<const uint32_t*>(qa + ks * 16 + 8);
      }
      if (vb) {
        a1 = *reinterpret_cast<const uint32_t*>(qb + ks * 16);
        a3 = *reinterpret_cast<const uint32_t*>(qb + ks * 16 + 8);
      }
      qf[ks][0] = a0; qf[ks][1] = a1; qf[ks][2] = a2; qf[ks][3] = a3;
    }
  }
  const int lim0 = va ? pos0 + ta : -1, lim1 = vb ? pos0 + tb : -1;
  // last key any row of this WARP may see (warp-uniform): rows of the warp are consecutive
  const int wt_last = (warp * 16 + 15) / G;
  const int warp_lim = pos0 + (wt_last < nt ? wt_last : nt - 1);
  const bool warp_active = (warp * 16) / G < nt;
  float mrow[2] = {-INFINITY, -INFINITY}, lrow[2] = {0.f, 0.f};
  float o[DT][4];
#pragma unroll
  for (int d = 0; d < DT; ++d) o[d][0] = o[d][1] = o[d][2] = o[d][3] = 0.f;

  int stage = 0;
  uint32_t phase = 0;
  for (int kt = 0; kt < n_tiles; ++kt) {
    mbar_wait(&full_bar[stage], phase);
    const int key_abs0 = kt * XB_ATT_TILE;
    if (warp_active && key_abs0 <= warp_lim) {
      const uint32_t sK = smem_u32(stages + stage * Cfg::STAGE_BYTES);
#pragma unroll 1
      for (int kb = 0; kb < XB_ATT_TILE / 32; ++kb) {
        if (key_abs0 + kb * 32 <= warp_lim)
          att_warp_block<HD, 2>(qf, sK, sK + Cfg::TILE_BYTES, kb * 32, key_abs0, lim0, lim1, scale_log2, mrow, lrow, o);
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty_bar[stage]);
    if (++stage == NST) { stage = 0; phase ^= 1; }
  }

  lrow[0] += __shfl_xor_sync(0xffffffffu, lrow[0], 1);
  lrow[0] += __shfl_xor_sync(0xffffffffu, lrow[0], 2);
  lrow[1] += __shfl_xor_sync(0xffffffffu, lrow[1], 1);
  lrow[1] += __shfl_xor_sync(0xffffffffu, lrow[1], 2);
  const float i0 = lrow[0] > 0.f ? 1.0f / lrow[0] : 0.f, i1 = lrow[1] > 0.f ? 1.0f / lrow[1] : 0.f;
  if (va) {
    bf16* dst = out + ((long long)(t0 + ta) * H + (kvh * G + ga)) * HD + ((lane & 3) << 1);
#pragma unroll
    for (int d = 0; d < DT; ++d) *reinterpret_cast<uint32_t*>(dst + d * 8) = f2_to_bf2(o[d][0] * i0, o[d][1] * i0);
  }
  if (vb) {
    bf16* dst = out + ((long long)(t0 + tb) * H + (kvh * G + gb)) * HD + ((lane & 3) << 1);
#pragma unroll
    for (int d = 0; d < DT; ++d) *reinterpret_cast<uint32_t*>(dst + d * 8) = f2_to_bf2(o[d][2] * i1, o[d][3] * i1);
  }
}
