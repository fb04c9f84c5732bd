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
//   * prefill kernel (tcgen05, further down): 128 rows = (128/G) consecutive tokens x G heads of
//     ONE sequence (tile table built on the host: tile_start[i] = first token row, tile_len[i] =
//     #tokens); S and O live in TMEM; causal masking by each row's absolute position.
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
#include <cooperative_groups.h>

template <int HD, int G, int NST>
struct AttDecCfg {
  static constexpr int TILE_BYTES = XB_ATT_TILE * HD * 2;
  static constexpr int STAGE_BYTES = 2 * TILE_BYTES;
  static constexpr int MERGE_BYTES = 4 * G * HD * 4 + 4 * 16 * 2 * 4;   // aliases the K/V ring after the key loop
  static constexpr int XCHG_OFF = (MERGE_BYTES + 1023) & ~1023;         // cluster exchange area (also in the ring), past the merge buffers
  static constexpr int XCHG_BYTES = G * HD * 4 + G * 2 * 4;
  static_assert(MERGE_BYTES <= XCHG_OFF && XCHG_OFF + XCHG_BYTES <= NST * STAGE_BYTES, "merge / exchange buffers must fit in the K/V ring");
  static constexpr int ROWS_OFF = NST * STAGE_BYTES + 128;               // after the barriers: q of the group + the step's own k / v row
  static constexpr int ROWS_BYTES = (G + 2) * HD * 2;
  static constexpr int SMEM_BYTES = NST * STAGE_BYTES + 128 + ((ROWS_BYTES + 127) & ~127) + 1024;
};

// FUSED: the q/k/v projection of the step, still in the split-K partial planes its Linear left behind.  The
// attention block finishes the q rows of its own head group and -- in the block that holds the newest tile --
// the step's k / v row (bias, bf16 rounding, RoPE; arithmetic of xb_qkv_finalize_kernel), appends that row to
// the paged cache and uses it from shared memory: one launch and one dependency hop less per layer.
struct QkvFin {
  const float* part;
  SplitInfo si;
  long long split_stride;
  const bf16* bias;
  const float* rope;
  bf16 *k_pool, *v_pool;
  int rope_kind;
};

template <int HD, int G, int NST, bool FUSED>
__global__ void __launch_bounds__(160) xb_attn_decode_mma_kernel(
    const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV, const bf16* __restrict__ q,
    BatchMeta m, float* __restrict__ o_part, float* __restrict__ ml_part, bf16* __restrict__ out, int H, int KV,
    int n_splits, float scale_log2, QkvFin f, int cluster_merge) {
  namespace cg = cooperative_groups;
  using Cfg = AttDecCfg<HD, G, NST>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* stages = smem;
  float* mg_o = reinterpret_cast<float*>(smem);                            // [4 warps][G][HD], reuses the ring
  float* mg_ml = mg_o + 4 * G * HD;                                       // [4 warps][16][2]
  float* xc_o = reinterpret_cast<float*>(smem + Cfg::XCHG_OFF);            // [G][HD] this split's merged accumulator
  float* xc_ml = xc_o + G * HD;                                           // [G][2]  (max, sum)
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + NST * Cfg::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + NST;
  bf16* sQ = reinterpret_cast<bf16*>(smem + Cfg::ROWS_OFF);                // [G][HD]
  bf16* sKrow = sQ + G * HD;                                              // [HD]
  bf16* sVrow = sKrow + HD;                                               // [HD]

  const int t = blockIdx.x, kvh = blockIdx.y, sp = blockIdx.z;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_launch_dependents();
  // The batch metadata and page table were uploaded before the step's first kernel started, and
  // every K/V row below position `pos` was written by an EARLIER step: only q and the row at `pos`
  // come from the kernel just upstream.  So the producer streams the older tiles before
  // griddepcontrol.wait -- the K/V reads overlap the q/k/v Linear.
  // (Safe because every forward call begins with the metadata H2D copy, a full stream dependency:
  // no kernel of an earlier call can still be running, and within a call nothing else writes this
  // layer's pool.)  FUSED: the row at `pos` is made by THIS block and patched into the staged tile,
  // so the producer never waits at all -- what the TMA reads for that one row does not matter.
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
      bool waited = FUSED;
      for (int kt = kt0; kt < kt1; ++kt) {
        if (!waited && (kt - kt0 >= NST || kt >= n_tiles - 1)) {   // ring full, or the tile holding row `pos`
          pdl_wait();
          waited = true;
        }
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sK = stages + stage * Cfg::STAGE_BYTES;
        att_issue_tile<HD>(&tmK, &tmV, sK, sK + Cfg::TILE_BYTES, &full_bar[stage], pt, m.max_pages, kt, kvh, KV, m.page_tokens);
        if (++stage == NST) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    pdl_wait();   // compute warps: q (or the projection's partial planes), and the newest K/V row

    constexpr int KS = HD / 16, DT = HD / 8;
    if (FUSED) {
      // ---- finish q of this head group (every split needs it) and, where the newest tile lives, the k / v row
      const int heads = H + 2 * KV, half = HD >> 1;
      const long long base = (long long)t * heads * HD;
      const float* rope_row = f.rope + (long long)pos * half * 2;
      const int iph = f.rope_kind == 0 ? HD / 4 : HD / 8;
      const bool owns_last = kt1 == n_tiles && kt1 > kt0;
      const int n_items = (G + (owns_last ? 2 : 0)) * iph;
      const int page = pt[pos / m.page_tokens], off = pos % m.page_tokens;
      const long long grow = (((long long)page * KV + kvh) * m.page_tokens + off) * HD;
      for (int it = threadIdx.x; it < n_items; it += 128) {
        const int hsel = it / iph, c = (it % iph) * 4;
        int head;
        bf16 *sdst, *gdst = nullptr;
        if (hsel < G) { head = kvh * G + hsel; sdst = sQ + hsel * HD; }
        else if (hsel == G) { head = H + kvh; sdst = sKrow; gdst = f.k_pool + grow; }
        else { head = H + KV + kvh; sdst = sVrow; gdst = f.v_pool + grow; }
        float x[4], y[4];
        qkv_item(f.part, f.si, f.split_stride, base, head, c, HD, head < H + KV, f.bias, rope_row, f.rope_kind, x, y);
        qkv_store4(sdst + c, x);
        if (gdst != nullptr) qkv_store4(gdst + c, x);        // KVCache.update: later steps read it from the pool
        if (f.rope_kind != 0) {
          qkv_store4(sdst + half + c, y);
          if (gdst != nullptr) qkv_store4(gdst + half + c, y);
        }
      }
      named_bar_sync(1, 128);
    }

    // ===================== 4 compute warps: each takes 16 keys of every tile =====================
    uint32_t qf[KS][4];
    {
      // A fragment rows: r0 = lane/4 (a head of the group when < G), r1 = r0 + 8 (always padding)
      const int r0 = lane >> 2;
      const bf16* qrow = FUSED ? sQ + (r0 < G ? r0 : 0) * HD + ((lane & 3) << 1)
                               : q + ((long long)t * H + (kvh * G + (r0 < G ? r0 : 0))) * HD + ((lane & 3) << 1);
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
      if (FUSED && kt == n_tiles - 1 && warp == ((pos & (XB_ATT_TILE - 1)) >> 4)) {
        // the step's own row: only this warp reads it (its 16 keys of the tile) -- overwrite what the TMA staged
        const int r = pos & (XB_ATT_TILE - 1);
        uint8_t* sKt = stages + stage * Cfg::STAGE_BYTES;
        constexpr int CH = HD / 8;   // 16-byte chunks per row
        for (int i = lane; i < 2 * CH; i += 32) {
          const int which = i / CH, c16 = i % CH;
          const uint4 v = *reinterpret_cast<const uint4*>((which ? sVrow : sKrow) + c16 * 8);
          *reinterpret_cast<uint4*>(sKt + which * Cfg::TILE_BYTES + att_tile_off(r, c16)) = v;
        }
        __syncwarp();
      }
      if (key_abs0 + warp * 16 <= pos) {
        const uint32_t sK = smem_u32(stages + stage * Cfg::STAGE_BYTES);
        att_warp_block<HD, 1>(qf, sK, sK + Cfg::TILE_BYTES, warp * 16, key_abs0, pos, pos, scale_log2, mrow, lrow, o);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_bar[stage]);
      if (++stage == NST) { stage = 0; phase ^= 1; }
    }

    // ---- merge the 4 warps (each saw a disjoint key subset); only rows < G are real
    named_bar_sync(1, 128);   // every warp is done reading the ring: its memory now holds the merge buffers
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
      } else if (cluster_merge) {
        xc_o[idx] = A;
        if (d == 0) { xc_ml[g * 2] = M; xc_ml[g * 2 + 1] = L; }       // M in log2 units
      } else {
        const long long oo = ((long long)t * H + hh) * n_splits + sp;
        o_part[oo * HD + d] = A;
        if (d == 0) { ml_part[oo * 2] = M; ml_part[oo * 2 + 1] = L; }   // M in log2 units
      }
    }
  }

  if (cluster_merge && n_splits > 1) {
    // ---- the kv splits of one (token, kv head) form a thread-block cluster (rank = split): the first block
    // merges all partials straight out of its peers' shared memory -- the arithmetic, in the order, of
    // xb_attn_combine2_kernel, without that kernel's launch or the round trip through global memory
    cg::cluster_group cluster = cg::this_cluster();
    cluster.sync();
    if (sp == 0 && warp < 4) {
      for (int g = warp; g < G; g += 4) {
        float ms = -INFINITY, ls = 0.f;
        if (lane < n_splits) {
          const float* pml = cluster.map_shared_rank(xc_ml, lane);
          ms = pml[g * 2];
          ls = pml[g * 2 + 1];
        }
        const float M = warp_max(ms);
        const float w = (ms == -INFINITY) ? 0.f : exp2f(ms - M);
        const float L = warp_sum(ls * w);
        constexpr int VPL = HD / 32;
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        for (int s0 = 0; s0 < n_splits; s0 += 8) {
          float4 v[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (s0 + u < n_splits) {
              const float* src = cluster.map_shared_rank(xc_o, s0 + u) + g * HD + lane * VPL;
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
        bf16* dst = out + ((long long)t * H + kvh * G + g) * HD + lane * VPL;
#pragma unroll
        for (int j = 0; j < VPL; j += 2)
          *reinterpret_cast<uint32_t*>(dst + j) = f2_to_bf2(acc[j] * invL, acc[j + 1] * invL);
      }
    }
    cluster.sync();   // the peers' shared memory must outlive the first block's reads
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

// ============================================================================= tcgen05 prefill attention
// Work item = (query tile, kv head): 128 rows = tokens x group heads of one kv head.  Both GEMMs
// run on the 5th-gen tensor cores:
//   S[128 x 64]  = Q[128 x hd] K_tile[64 x hd]^T        tcgen05.mma, A/B K-major from smem -> TMEM
//   O[128 x hd] += P[128 x 64] V_tile[64 x hd]          A = P (bf16, written to swizzled smem by the
//                                                       softmax threads), B = V as an MN-major operand
// One TMEM lane = one query row, so the softmax is thread-private (no shuffles): pass 1 row max,
// pass 2 p = exp2(s - m) -> bf16 -> smem.  O is only rescaled (TMEM ld/st) when some row's max moved
// by more than 2^8 (lazy rescale).  Two blocks per SM: one block's softmax overlaps the other's MMAs.
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t* v) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
        "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// MN-major, 128B-swizzled operand: 64-element (128 B) atoms along MN every `lbo` bytes, 8-row groups
// along K every 1024 B
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t smem_addr, uint32_t lbo_bytes) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) | (64ull << 32) |
         (1ull << 46) | (2ull << 61);
}

// D[tmem] (+)= A[tmem] * B[smem desc]: the A operand (128 rows = lanes, K bf16 elements packed two per
// 32-bit column) is read from tensor memory -- P of the attention never visits shared memory
__device__ __forceinline__ void tc_mma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc,
                                               uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
#ifndef XB_ATT_P_TMEM
#define XB_ATT_P_TMEM 1
#endif

#define XB_ATT_PG_CACHE 128   // page ids prefetched per block (8k keys at 64-token pages; beyond: on demand)

template <int HD>
struct AttTcCfg {
  static constexpr int Q_BYTES = 128 * HD * 2;
  static constexpr int TILE_BYTES = XB_ATT_TILE * HD * 2;   // one K or V tile
  static constexpr int P_BYTES = XB_ATT_P_TMEM ? 0 : 128 * XB_ATT_TILE * 2;   // P lives in tensor memory
  static constexpr int NST = 2;                             // V ring depth
  static constexpr int NSTK = XB_ATT_P_TMEM ? 3 : 2;        // K ring depth: the 16 KB the P buffer used to take hold a third K tile
  // TWO blocks per SM (measured: 240 us vs 328 us for one block with double-buffered P and 4-deep
  // rings at 4 x 2048 tokens -- the four softmax warps of a block are latency bound, a second
  // block doubles the warps per scheduler).  No alignment slack (with the shared-memory P buffer,
  // XB_ATT_P_TMEM=0, 2 x (112.75 KB + 1 KB reserved) only just fits 228 KB); the kernel traps if the
  // dynamic shared memory base is not 1024-byte aligned.
  // P in tensor memory (XB_ATT_P_TMEM=1, default): the softmax warps store bf16 P over the first 32
  // columns of the tile's S buffer with tcgen05.st and O += P V takes its A operand from there -- no
  // shared-memory round trip, no proxy fence, and the softmax of tile kt no longer waits for PV of
  // tile kt-1 to release a single P buffer: 212 us against 238 us per layer at 4 x 2048 tokens.
  static constexpr int SMEM_BYTES = Q_BYTES + (NSTK + NST) * TILE_BYTES + P_BYTES + 256 + XB_ATT_PG_CACHE * 4;
  static constexpr int TM_COLS = 256;                       // S: 64 columns, O: HD columns
};

template <int HD, int G>
__global__ void __launch_bounds__(192, 2) xb_attn_prefill_tc_kernel(
    const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV, const bf16* __restrict__ q,
    BatchMeta m, const int* __restrict__ tile_start, const int* __restrict__ tile_len, bf16* __restrict__ out, int H,
    int KV, float scale_log2) {
  using Cfg = AttTcCfg<HD>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem) & 1023u) != 0u) {
    if (threadIdx.x == 0) printf("xb200: attention smem base not 1024-aligned\n");
    __trap();
  }
  uint8_t* sQ = smem;
  uint8_t* sKV = sQ + Cfg::Q_BYTES;
  uint8_t* sVr = sKV + Cfg::NSTK * Cfg::TILE_BYTES;          // K ring first, then the V ring
  uint8_t* sP = sVr + Cfg::NST * Cfg::TILE_BYTES;
  // K and V have separate rings: a K slot is free as soon as QK^T of its tile retires, a V slot
  // after PV -- the next K load starts a whole tile earlier than with a shared ring
  uint64_t* k_full = reinterpret_cast<uint64_t*>(sP + Cfg::P_BYTES);
  uint64_t* k_empty = k_full + Cfg::NSTK;
  uint64_t* v_full = k_empty + Cfg::NSTK;
  uint64_t* v_empty = v_full + Cfg::NST;
  // the per-tile barriers come in pairs indexed by (tile & 1): each completes every other tile, so a
  // parity wait can never be overrun by the next completion
  uint64_t* s_full = v_empty + Cfg::NST;      // [2] S buffer written (QK^T retired)
  uint64_t* p_full = s_full + 2;              // [2] P buffer written by the softmax warps
  uint64_t* o_done = p_full + 2;              // [2] PV of the tile retired (P buffer free, O updated)
  uint64_t* q_ready = o_done + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(q_ready + 1);
  int* s_pages = reinterpret_cast<int*>(tmem_slot + 2);   // [XB_ATT_PG_CACHE] prefetched page ids

  const int tile = blockIdx.x, kvh = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_launch_dependents();
  pdl_wait();
  const int t0 = tile_start[tile];
  const int nt = tile_len[tile];
  const int pos0 = m.tok_pos[t0];
  const int slot = m.seq_slot[m.tok_seq[t0]];
  const int* pt = m.page_table + (long long)slot * m.max_pages;
  const int n_kt = (pos0 + nt - 1) / XB_ATT_TILE + 1;

  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < Cfg::NSTK; ++i) {
        mbar_init(&k_full[i], 1);
        mbar_init(&k_empty[i], 1);
      }
      for (int i = 0; i < Cfg::NST; ++i) {
        mbar_init(&v_full[i], 1);
        mbar_init(&v_empty[i], 1);
      }
      mbar_init(&s_full[0], 1);
      mbar_init(&s_full[1], 1);
      mbar_init(&p_full[0], 4);
      mbar_init(&p_full[1], 4);
      mbar_init(&o_done[0], 1);
      mbar_init(&o_done[1], 1);
      mbar_init(q_ready, 4);
      fence_mbar_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, Cfg::TM_COLS);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // S is double buffered (columns 0..63 and 64..127): QK^T of tile j+1 overlaps the softmax of tile j
  const uint32_t tm_S = tmem_base, tm_O = tmem_base + 2 * XB_ATT_TILE;

  if (warp == 0) {
    // ===================== TMA producer =====================
    // page ids of the block's key tiles are fetched up front by the whole warp, so the TMA
    // issue loop carries no dependent global load
    const int P = m.page_tokens;
    const int BR = P < XB_ATT_TILE ? P : XB_ATT_TILE;       // rows per TMA box
    const int rgs = XB_ATT_TILE / BR;                        // boxes (row groups) per tile
    const int n_rg = n_kt * rgs;
    for (int i = lane; i < n_rg && i < XB_ATT_PG_CACHE; i += 32) {
      const int pi = (i * BR) / P;
      s_pages[i] = pt[pi < m.max_pages ? pi : m.max_pages - 1];
    }
    __syncwarp();
    if (lane == 0) {
      tma_prefetch_desc(&tmK);
      tma_prefetch_desc(&tmV);
      auto issue = [&](const CUtensorMap* tm, uint8_t* dst, uint64_t* bar, int kt) {
        mbar_expect_tx(bar, Cfg::TILE_BYTES);
        for (int rg = 0; rg < rgs; ++rg) {
          const int i = kt * rgs + rg;
          const int key = kt * XB_ATT_TILE + rg * BR;
          int page;
          if (i < XB_ATT_PG_CACHE) page = s_pages[i];
          else { const int pi = key / P; page = pt[pi < m.max_pages ? pi : m.max_pages - 1]; }
          const int row = (page * KV + kvh) * P + (key % P);
#pragma unroll
          for (int db = 0; db < HD / 64; ++db) tma_load_2d(dst + db * (XB_ATT_TILE * 128) + rg * BR * 128, tm, bar, db * 64, row);
        }
      };
      auto issue_k = [&](int kt) {
        const int ks = kt % Cfg::NSTK;
        mbar_wait(&k_empty[ks], ((kt / Cfg::NSTK) & 1) ^ 1);
        issue(&tmK, sKV + ks * Cfg::TILE_BYTES, &k_full[ks], kt);
      };
      // K runs one tile ahead of V (QK^T of tile kt+1 is issued before PV of tile kt)
      issue_k(0);
      for (int kt = 0; kt < n_kt; ++kt) {
        if (kt + 1 < n_kt) issue_k(kt + 1);
        const int vs = kt % Cfg::NST;
        mbar_wait(&v_empty[vs], ((kt / Cfg::NST) & 1) ^ 1);
        issue(&tmV, sVr + vs * Cfg::TILE_BYTES, &v_full[vs], kt);
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc_s = umma_idesc_bf16(XB_ATT_TILE);            // M=128, N=64, A/B K-major
      constexpr uint32_t idesc_o = umma_idesc_bf16(HD) | (1u << 16);        // M=128, N=HD, B MN-major
      mbar_wait(q_ready, 0);
      tc_fence_after();
      const uint32_t q_addr = smem_u32(sQ), p_addr = smem_u32(sP);
      auto issue_qk = [&](int kt) {
        const int stage = kt % Cfg::NSTK;
        mbar_wait(&k_full[stage], (kt / Cfg::NSTK) & 1);
        tc_fence_after();
        const uint32_t k_addr = smem_u32(sKV + stage * Cfg::TILE_BYTES);
        const uint32_t d = tm_S + (uint32_t)((kt & 1) * XB_ATT_TILE);
        // S = Q K^T over hd in steps of 16; dims 64.. live in the second 128-byte-wide region
#pragma unroll
        for (int ks = 0; ks < HD / 16; ++ks) {
          const uint32_t a = q_addr + (ks >> 2) * (128 * 128) + (ks & 3) * 32;
          const uint32_t b = k_addr + (ks >> 2) * (XB_ATT_TILE * 128) + (ks & 3) * 32;
          tc_mma_bf16(d, umma_desc_sw128(a), umma_desc_sw128(b), idesc_s, ks > 0 ? 1u : 0u);
        }
        tc_commit(&k_empty[stage]);      // the K slot may be refilled as soon as these MMAs retire
        tc_commit(&s_full[kt & 1]);
      };
      issue_qk(0);
      for (int kt = 0; kt < n_kt; ++kt) {
        // S buffer (kt+1)&1 was last read by the softmax of tile kt-1, which finished before p_full(kt-1)
        if (kt + 1 < n_kt) issue_qk(kt + 1);
        const int stage = kt % Cfg::NST;
        const uint32_t v_addr = smem_u32(sVr + stage * Cfg::TILE_BYTES);
        // O += P V over the 64 keys in steps of 16
        mbar_wait(&v_full[stage], (kt / Cfg::NST) & 1);
        mbar_wait(&p_full[kt & 1], (kt >> 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int ks = 0; ks < XB_ATT_TILE / 16; ++ks) {
          const uint32_t b = v_addr + ks * 2048;
#if XB_ATT_P_TMEM
          // P(kt) sits in the first 32 columns of S buffer kt&1 (16 keys = 8 packed columns per step)
          tc_mma_bf16_ts(tm_O, tm_S + (uint32_t)((kt & 1) * XB_ATT_TILE) + ks * 8, umma_desc_mn_sw128(b, XB_ATT_TILE * 128), idesc_o,
                         (kt > 0 || ks > 0) ? 1u : 0u);
#else
          const uint32_t a = p_addr + ks * 32;
          tc_mma_bf16(tm_O, umma_desc_sw128(a), umma_desc_mn_sw128(b, XB_ATT_TILE * 128), idesc_o, (kt > 0 || ks > 0) ? 1u : 0u);
#endif
        }
        tc_commit(&v_empty[stage]);
        tc_commit(&o_done[kt & 1]);
      }
    }
  } else {
    // ===================== softmax / correction / epilogue: warps 2..5, one row per thread =====
    const int quad = warp & 3;
    const int r = quad * 32 + lane;                 // query row = TMEM lane
    const int tq = r / G, gq = r % G;
    const bool valid = tq < nt;
    const int lim = valid ? pos0 + tq : -1;
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    // ---- stage Q into swizzled smem (K-major A operand), zero rows beyond the tile
    {
      const bf16* qrow = q + ((long long)(t0 + (valid ? tq : 0)) * H + (kvh * G + gq)) * HD;
#pragma unroll
      for (int c = 0; c < HD / 8; ++c) {
        uint4 v = make_uint4(0, 0, 0, 0);
        if (valid) v = *reinterpret_cast<const uint4*>(qrow + c * 8);
        *reinterpret_cast<uint4*>(sQ + (c >> 3) * (128 * 128) + r * 128 + (((c & 7) ^ (r & 7)) << 4)) = v;
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(q_ready);
    }
    float m_used = -INFINITY;     // reference max (scaled, log2 units) that O and l are expressed against
    float l_sum = 0.f;
    for (int kt = 0; kt < n_kt; ++kt) {
      mbar_wait(&s_full[kt & 1], (kt >> 1) & 1);
      tc_fence_after();
      const int key0 = kt * XB_ATT_TILE;
      const uint32_t tm_Sk = tm_S + (uint32_t)((kt & 1) * XB_ATT_TILE);
      // ---- the 64 scores of this row: ONE TMEM read, kept in registers for both steps below
      uint32_t v[XB_ATT_TILE];
#pragma unroll
      for (int c = 0; c < XB_ATT_TILE; c += 16) tmem_ld16(tm_Sk + lane_base + c, v + c);
      tmem_ld_wait();
      // only tiles that reach past the first row's position need per-element causal masks
      const bool masked = key0 + XB_ATT_TILE - 1 > pos0;
      float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};   // independent chains (ILP)
      if (masked) {
#pragma unroll
        for (int j = 0; j < XB_ATT_TILE; ++j)
          if (key0 + j <= lim) mx4[j & 3] = fmaxf(mx4[j & 3], __uint_as_float(v[j]));
      } else {
#pragma unroll
        for (int j = 0; j < XB_ATT_TILE; ++j) mx4[j & 3] = fmaxf(mx4[j & 3], __uint_as_float(v[j]));
      }
      const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3])) * scale_log2;   // scale > 0
#if !XB_ATT_P_TMEM
      // ---- the (single) P buffer was last read by PV of tile kt-1
      if (kt > 0) {
        mbar_wait(&o_done[(kt - 1) & 1], ((kt - 1) >> 1) & 1);
        tc_fence_after();
      }
#endif
      // ---- lazy rescale: only when some row of the warp moved by more than 2^8
      const bool need = mx > m_used + 8.0f;
      if (__any_sync(0xffffffffu, need)) {
#if XB_ATT_P_TMEM
        if (kt > 0) {   // O must be quiescent: PV of tile kt-1 retired
          mbar_wait(&o_done[(kt - 1) & 1], ((kt - 1) >> 1) & 1);
          tc_fence_after();
        }
#endif
        const float m_new = need ? mx : m_used;
        const float corr = (m_used == -INFINITY) ? 0.f : exp2f(m_used - m_new);
        l_sum *= corr;
        if (kt > 0) {   // O is quiescent: PV of tile kt-1 retired (waited above)
#pragma unroll 1
          for (int c = 0; c < HD; c += 16) {
            uint32_t o[16];
            tmem_ld16(tm_O + lane_base + c, o);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 16; ++j) o[j] = __float_as_uint(__uint_as_float(o[j]) * corr);
            tmem_st16(tm_O + lane_base + c, o);
          }
          tmem_st_wait();
        }
        m_used = m_new;
      }
      // ---- p = 2^(s*scale - m_used) -> bf16 -> swizzled smem (A operand of the PV MMA)
      const float nmref = (m_used == -INFINITY) ? 0.f : -m_used;
      float ps[4] = {0.f, 0.f, 0.f, 0.f};
#if XB_ATT_P_TMEM
      uint32_t pw[XB_ATT_TILE / 2];
#endif
#pragma unroll
      for (int c = 0; c < XB_ATT_TILE; c += 8) {
        float p[8];
#pragma unroll
        for (int j = 0; j < 8; ++j)
          asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(p[j]) : "f"(fmaf(__uint_as_float(v[c + j]), scale_log2, nmref)));
        if (masked) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            if (key0 + c + j > lim) p[j] = 0.f;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) ps[j & 3] += p[j];
#if XB_ATT_P_TMEM
        pw[(c >> 1)] = f2_to_bf2(p[0], p[1]); pw[(c >> 1) + 1] = f2_to_bf2(p[2], p[3]);
        pw[(c >> 1) + 2] = f2_to_bf2(p[4], p[5]); pw[(c >> 1) + 3] = f2_to_bf2(p[6], p[7]);
      }
      // P -> tensor memory, over the first 32 columns of this tile's S buffer (this row's scores are
      // already in registers; the MMA warp reads them as the A operand of O += P V)
      tmem_st16(tm_Sk + lane_base, pw);
      tmem_st16(tm_Sk + lane_base + 16, pw + 16);
      tmem_st_wait();
#else
        uint4 w;
        w.x = f2_to_bf2(p[0], p[1]); w.y = f2_to_bf2(p[2], p[3]); w.z = f2_to_bf2(p[4], p[5]); w.w = f2_to_bf2(p[6], p[7]);
        *reinterpret_cast<uint4*>(sP + r * 128 + ((((c >> 3)) ^ (r & 7)) << 4)) = w;
      }
      fence_proxy_async();
#endif
      l_sum += (ps[0] + ps[1]) + (ps[2] + ps[3]);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full[kt & 1]);
    }
    // ---- epilogue: O / l -> bf16 -> attention output row
    mbar_wait(&o_done[(n_kt - 1) & 1], ((n_kt - 1) >> 1) & 1);
    tc_fence_after();
    const float inv = l_sum > 0.f ? 1.0f / l_sum : 0.f;
    bf16* dst = out + ((long long)(t0 + (valid ? tq : 0)) * H + (kvh * G + gq)) * HD;
#pragma unroll 1
    for (int c = 0; c < HD; c += 16) {
      uint32_t v[16];
      tmem_ld16(tm_O + lane_base + c, v);
      tmem_ld_wait();
      if (valid) {
        float x[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) x[j] = __uint_as_float(v[j]) * inv;
        uint4 w0, w1;
        w0.x = f2_to_bf2(x[0], x[1]);   w0.y = f2_to_bf2(x[2], x[3]);   w0.z = f2_to_bf2(x[4], x[5]);   w0.w = f2_to_bf2(x[6], x[7]);
        w1.x = f2_to_bf2(x[8], x[9]);   w1.y = f2_to_bf2(x[10], x[11]); w1.z = f2_to_bf2(x[12], x[13]); w1.w = f2_to_bf2(x[14], x[15]);
        reinterpret_cast<uint4*>(dst + c)[0] = w0;
        reinterpret_cast<uint4*>(dst + c)[1] = w1;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::TM_COLS);
  }
}
