// Linear layers of the layer-range forward as ONE warp-specialised tcgen05 kernel.
//
// Replaces the per-projection cuBLAS calls behind torchtune's nn.Linear modules that the
// reference builds at xotorch/inference/torch/models/general_mha.py:83-102 (q/k/v/o) and
// xotorch/inference/torch/models/llm_utils.py:513-522 (gate/up/down), and the LM head at
// llm_utils.py:425-435.
//
//   D[128 x BN] (fp32, TMEM) += A[128 x 64] (bf16, smem) * B[BN x 64]^T (bf16, smem)
//
// Both operands are K-major and arrive by TMA into 128B-swizzled tiles; one elected thread
// issues tcgen05.mma; four epilogue warps drain TMEM with tcgen05.ld.  The kernel is
// persistent over work units (a_tile, b_tile, k_split) and double-buffers the accumulator
// so the epilogue of unit i overlaps the mainloop of unit i+1.
//
// Two ways of mapping a Linear  y[T,N] = x[T,K] * W[N,K]^T  onto it:
//   MODE_SKINNY  (decode, T <= 128): A = W tile (128 output features), B = x (all T tokens).
//       HBM-bound weight streaming; K is split across CTAs so every SM pulls weights;
//       output = fp32 partial sums  out[split][t][n]  reduced by the consumer kernel.
//   MODE_WIDE    (prefill): A = x tile (128 tokens), B = W tile (BN features);
//       tensor-bound; output = bf16  out[t][n] (+ bias, rounded once like nn.Linear).
//
// The engine stores weights TILE-MAJOR in HBM ([N/128][K/64][128][64] bf16): every TMA box is
// one contiguous 16 KB read instead of 128 rows x 128 B at a K*2-byte pitch.
//
// Programmatic dependent launch: weights never depend on the preceding kernel, so the producer
// issues the weight boxes of the first pipeline stages BEFORE griddepcontrol.wait and only the
// activation boxes / the epilogue stores wait for the upstream grid -- the HBM stream keeps
// going across kernel boundaries.  (Tried and dropped: also TMA-prefetching the next 8..64 weight
// boxes into L2 during that wait -- the prefetches compete with the real loads: Linear kernel 0.81 ->
// 0.74 of HBM peak at depth 16, bs=1 decode 297 -> 275 tok/s.)
#pragma once
#include "xb_common.cuh"
#include <stdlib.h>

enum { XB_GEMM_SKINNY = 0, XB_GEMM_WIDE = 1 };

struct GemmArgs {
  int a_tiles, b_tiles, k_blocks, splits;
  int a_rows, b_rows;      // valid rows of the A / B global matrices
  int mode;
  int round_bf16;          // SKINNY: round the fp32 result to bf16 precision (splits == 1 only)
  int w_tiled;             // weight operand is stored tile-major: [N/128][K/64][128][64]
  int sk_q;                // > 0: stream-K -- block b owns items [b*sk_q, (b+1)*sk_q) of the
  int sk_items;            //      a_tiles*k_blocks (tile-major, k-block-minor) item space
  void* out;
  long long ld_out;        // SKINNY: N (floats per token row); WIDE: elements per token row
  const bf16* bias;        // WIDE: per-feature bias or nullptr
  // WIDE fused epilogues
  int epi;                 // XB_EPI_*
  // XB_EPI_QKV: bias + bf16 round + RoPE at the row's position + q store / paged-KV append
  const float* rope;       // [pos][hd/2][cos,sin]
  bf16 *q_out, *k_pool, *v_pool;
  const int *tok_seq, *tok_pos, *seq_slot, *page_table;
  int max_pages, page_tokens, n_heads, n_kv_heads, head_dim, rope_kind;
  int bulk_epi;            // SKINNY: the block's LAST unit leaves through shared memory + bulk stores (see the epilogue)
  float* out2;             // SKINNY: second copy of the result (same layout as `out`), e.g. pinned host memory
  unsigned long long* ts;  // timeline instrumentation (XB_GEMM_TS=1): [gridDim.x][XB_TS_SLOTS] globaltimer stamps, else nullptr
};
// stamps: 0 block entry, 1 producer: first stages issued (before the dependency wait), 2 producer: dependency resolved,
// 3 first MMA issued, 4 last MMA committed, 5 (not a time) %smid of the block, 6 epilogue: last unit stored, 7 block exit,
// 8 epilogue: accumulator of the last unit complete, 9 producer: last load issued, 10 MMA thread: last stage landed
#define XB_TS(g, slot)                                                                   \
  do {                                                                                   \
    if ((g).ts != nullptr) (g).ts[(size_t)blockIdx.x * XB_TS_SLOTS + (slot)] = xb_gtimer(); \
  } while (0)
enum { XB_EPI_PLAIN = 0, XB_EPI_QKV = 1, XB_EPI_SILU = 2 };

// TMA coordinates (innermost, row) of the [128 x 64] weight box holding rows 128*wt.. of k-block kb
__device__ __forceinline__ void gemm_w_coord(const GemmArgs& g, int wt, int kb, int& c0, int& c1) {
  if (g.w_tiled) { c0 = 0; c1 = (wt * g.k_blocks + kb) * 128; }
  else { c0 = kb * 64; c1 = wt * 128; }
}

struct GemmUnit {
  int ta, tb, kb0, kb1, slot;
};
// Walks the work units of one block; all three warp roles walk the same sequence.
struct UnitWalker {
  int u, it, end;
  __device__ __forceinline__ void init(const GemmArgs& g) {
    if (g.sk_q > 0) {
      it = blockIdx.x * g.sk_q;
      end = it + g.sk_q < g.sk_items ? it + g.sk_q : g.sk_items;
    } else {
      u = blockIdx.x;
    }
  }
  __device__ __forceinline__ bool next(const GemmArgs& g, GemmUnit& w) {
    if (g.sk_q > 0) {
      if (it >= end) return false;
      w.ta = it / g.k_blocks;
      w.tb = 0;
      w.kb0 = it - w.ta * g.k_blocks;
      const int left = end - it;
      w.kb1 = w.kb0 + left < g.k_blocks ? w.kb0 + left : g.k_blocks;
      w.slot = blockIdx.x - (w.ta * g.k_blocks) / g.sk_q;
      it += w.kb1 - w.kb0;
      return true;
    }
    if (u >= g.a_tiles * g.b_tiles * g.splits) return false;
    const int s = u % g.splits;
    const int tile = u / g.splits;
    w.ta = tile % g.a_tiles;
    w.tb = tile / g.a_tiles;
    w.kb0 = (int)(((long long)s * g.k_blocks) / g.splits);
    w.kb1 = (int)(((long long)(s + 1) * g.k_blocks) / g.splits);
    w.slot = s;
    u += gridDim.x;
    return true;
  }
};

template <int BN, int STAGES>
struct GemmCfg {
  static constexpr int A_BYTES = 128 * 64 * 2;
  static constexpr int B_BYTES = BN * 64 * 2;
  static constexpr int ACC_STAGES = (2 * BN <= 512) ? 2 : 1;
  static constexpr int TM_COLS_RAW = ACC_STAGES * BN;
  static constexpr int TM_COLS = TM_COLS_RAW <= 32 ? 32 : TM_COLS_RAW <= 64 ? 64 : TM_COLS_RAW <= 128 ? 128 : TM_COLS_RAW <= 256 ? 256 : 512;
  static constexpr int SMEM_BYTES = STAGES * (A_BYTES + B_BYTES) + 1024 /*align slack*/ + 256 /*barriers*/;
};


// ---- WIDE epilogues: one thread = one token row `t`, TMEM chunks of 16 consecutive columns ----
__device__ __forceinline__ void epi_store16(bf16* dst, const float* x) {
  uint4 p0, p1;
  p0.x = f2_to_bf2(x[0], x[1]);   p0.y = f2_to_bf2(x[2], x[3]);
  p0.z = f2_to_bf2(x[4], x[5]);   p0.w = f2_to_bf2(x[6], x[7]);
  p1.x = f2_to_bf2(x[8], x[9]);   p1.y = f2_to_bf2(x[10], x[11]);
  p1.z = f2_to_bf2(x[12], x[13]); p1.w = f2_to_bf2(x[14], x[15]);
  reinterpret_cast<uint4*>(dst)[0] = p0;
  reinterpret_cast<uint4*>(dst)[1] = p1;
}
// nn.Linear rounding point: bf16(acc + bias)
__device__ __forceinline__ void epi_linear16(const GemmArgs& g, const uint32_t* v, int n0, float* x) {
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    float a = __uint_as_float(v[j]);
    if (g.bias != nullptr) a += bf2f(g.bias[n0 + j]);
    x[j] = rbf(a);
  }
}
// where 16 elements [e, e+16) of fused-QKV head `head` of token row t go: q buffer or the paged cache
__device__ __forceinline__ bf16* epi_qkv_dst(const GemmArgs& g, int t, int pos, int head, int e) {
  const int H = g.n_heads, KV = g.n_kv_heads, hd = g.head_dim;
  if (head < H) return g.q_out + ((long long)t * H + head) * hd + e;
  const int slot = g.seq_slot[g.tok_seq[t]];
  const int page = g.page_table[(long long)slot * g.max_pages + pos / g.page_tokens];
  const int off = pos % g.page_tokens;
  const int kvh = (head - H) % KV;
  bf16* pool = head < H + KV ? g.k_pool : g.v_pool;
  return pool + (((long long)page * KV + kvh) * g.page_tokens + off) * hd + e;
}

template <int BN>
__device__ __forceinline__ void epi_wide_qkv(const GemmArgs& g, uint32_t t_addr, int t, int tb) {
  const int hd = g.head_dim, half = hd >> 1, H = g.n_heads, KV = g.n_kv_heads;
  const bool row_ok = t < g.a_rows;
  const int pos = row_ok ? g.tok_pos[t] : 0;
  const float* rope_row = g.rope + (long long)pos * half * 2;
  if (g.rope_kind == 0) {
    // interleaved pairs (2i, 2i+1): both members sit in one 16-column chunk
#pragma unroll 1
    for (int c = 0; c < BN; c += 16) {
      uint32_t v[16];
      tmem_ld16(t_addr + c, v);
      tmem_ld_wait();
      const int n0 = tb * BN + c;
      if (!row_ok || n0 >= g.b_rows) continue;
      const int head = n0 / hd, e = n0 % hd;
      float x[16];
      epi_linear16(g, v, n0, x);
      if (head < H + KV) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float2 cs = *reinterpret_cast<const float2*>(rope_row + (e / 2 + j) * 2);
          const float a = x[2 * j], b = x[2 * j + 1];
          x[2 * j] = __fsub_rn(__fmul_rn(a, cs.x), __fmul_rn(b, cs.y));
          x[2 * j + 1] = __fadd_rn(__fmul_rn(b, cs.x), __fmul_rn(a, cs.y));
        }
      }
      epi_store16(epi_qkv_dst(g, t, pos, head, e), x);
    }
  } else {
    // half-split pairs (i, i + hd/2): take the two chunks of a head together
#pragma unroll 1
    for (int hb = 0; hb < BN; hb += hd) {
#pragma unroll 1
      for (int c = 0; c < half; c += 16) {
        uint32_t vl[16], vh[16];
        tmem_ld16(t_addr + hb + c, vl);
        tmem_ld16(t_addr + hb + half + c, vh);
        tmem_ld_wait();
        const int n0 = tb * BN + hb + c;
        if (!row_ok || n0 >= g.b_rows) continue;
        const int head = n0 / hd;
        float lo[16], hi[16];
        epi_linear16(g, vl, n0, lo);
        epi_linear16(g, vh, n0 + half, hi);
        if (head < H + KV) {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const float2 cs = *reinterpret_cast<const float2*>(rope_row + (c + j) * 2);
            const float cc = rbf(cs.x), ss = rbf(cs.y);
            const float a = lo[j], b = hi[j];
            lo[j] = __fadd_rn(rbf(__fmul_rn(a, cc)), rbf(__fmul_rn(-b, ss)));
            hi[j] = __fadd_rn(rbf(__fmul_rn(b, cc)), rbf(__fmul_rn(a, ss)));
          }
        }
        bf16* dst = epi_qkv_dst(g, t, pos, head, c);
        epi_store16(dst, lo);
        epi_store16(dst + half, hi);
      }
    }
  }
}

// gate/up rows are interleaved in blocks of 16 features: columns [32b, 32b+16) = gate of features
// [16b, 16b+16), columns [32b+16, 32b+32) = up of the same features.  act = bf16(bf16(silu(g)) * u).
template <int BN>
__device__ __forceinline__ void epi_wide_silu(const GemmArgs& g, uint32_t t_addr, int t, int tb) {
  bf16* out = reinterpret_cast<bf16*>(g.out);
#pragma unroll 1
  for (int c = 0; c < BN; c += 32) {
    uint32_t vg[16], vu[16];
    tmem_ld16(t_addr + c, vg);
    tmem_ld16(t_addr + c + 16, vu);
    tmem_ld_wait();
    const int n0 = tb * BN + c;
    if (t >= g.a_rows || n0 >= g.b_rows) continue;
    float x[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const float gg = rbf(__uint_as_float(vg[j])), uu = rbf(__uint_as_float(vu[j]));
      x[j] = rbf(gg / (1.0f + expf(-gg))) * uu;
    }
    epi_store16(out + (long long)t * g.ld_out + (n0 >> 1), x);
  }
}

template <int BN, int STAGES>
__global__ void __launch_bounds__(192, 1)
xb_gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                       const __grid_constant__ CUtensorMap tmC, GemmArgs g) {
  using Cfg = GemmCfg<BN, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * Cfg::A_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(sB + STAGES * Cfg::B_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    XB_TS(g, 0);
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (g.bulk_epi) tma_prefetch_desc(&tmC);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < STAGES; ++i) {
        mbar_init(&full_bar[i], 1);
        mbar_init(&empty_bar[i], 1);
      }
      for (int i = 0; i < 2; ++i) {
        mbar_init(&tfull_bar[i], 1);
        mbar_init(&tempty_bar[i], 4);
      }
      fence_mbar_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, Cfg::TM_COLS);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();   // the next kernel may start its prologue / weight prefetch now

  UnitWalker walk;
  walk.init(g);
  GemmUnit wu;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      bool waited = false;
      int pre = 0;   // stages whose weight box was issued ahead of the dependency wait
      while (walk.next(g, wu)) {
        const int ta = wu.ta, tb = wu.tb, kb0 = wu.kb0, kb1 = wu.kb1;
        if (!waited) {
          // first unit: weight boxes of up to STAGES k-blocks go out before the upstream grid is done
          const int npre = (kb1 - kb0) < STAGES ? (kb1 - kb0) : STAGES;
          for (int i = 0; i < npre; ++i) {
            mbar_expect_tx(&full_bar[i], Cfg::A_BYTES + Cfg::B_BYTES);
            int c0, c1;
            if (g.mode == XB_GEMM_SKINNY) {
              gemm_w_coord(g, ta, kb0 + i, c0, c1);
              tma_load_2d(sA + i * Cfg::A_BYTES, &tmA, &full_bar[i], c0, c1);
            } else {
#pragma unroll
              for (int j = 0; j < BN / 128; ++j) {
                if (g.w_tiled) {
                  gemm_w_coord(g, tb * (BN / 128) + j, kb0 + i, c0, c1);
                  tma_load_2d(sB + i * Cfg::B_BYTES + j * Cfg::A_BYTES, &tmB, &full_bar[i], c0, c1);
                } else if (j == 0) {
                  tma_load_2d(sB + i * Cfg::B_BYTES, &tmB, &full_bar[i], (kb0 + i) * 64, tb * BN);
                }
              }
            }
          }
          XB_TS(g, 1);
          pdl_wait();
          XB_TS(g, 2);
          waited = true;
          pre = npre;
        }
        for (int kb = kb0; kb < kb1; ++kb) {
          const bool w_issued = pre > 0;
          if (!w_issued) {
            mbar_wait(&empty_bar[stage], phase ^ 1);
            mbar_expect_tx(&full_bar[stage], Cfg::A_BYTES + Cfg::B_BYTES);
          } else {
            --pre;
          }
          int c0, c1;
          if (g.mode == XB_GEMM_SKINNY) {
            if (!w_issued) {
              gemm_w_coord(g, ta, kb, c0, c1);
              tma_load_2d(sA + stage * Cfg::A_BYTES, &tmA, &full_bar[stage], c0, c1);
            }
            tma_load_2d(sB + stage * Cfg::B_BYTES, &tmB, &full_bar[stage], kb * 64, tb * BN);
          } else {
            tma_load_2d(sA + stage * Cfg::A_BYTES, &tmA, &full_bar[stage], kb * 64, ta * 128);
            if (!w_issued) {
#pragma unroll
              for (int j = 0; j < BN / 128; ++j) {
                if (g.w_tiled) {
                  gemm_w_coord(g, tb * (BN / 128) + j, kb, c0, c1);
                  tma_load_2d(sB + stage * Cfg::B_BYTES + j * Cfg::A_BYTES, &tmB, &full_bar[stage], c0, c1);
                } else if (j == 0) {
                  tma_load_2d(sB + stage * Cfg::B_BYTES, &tmB, &full_bar[stage], kb * 64, tb * BN);
                }
              }
            }
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
      XB_TS(g, 9);
      if (!waited) pdl_wait();
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (one thread) =====================
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(BN);
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      bool first_mma = true;
      while (walk.next(g, wu)) {
        const int kb0 = wu.kb0, kb1 = wu.kb1;
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);   // epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase);         // TMA bytes have landed
          if (first_mma) { XB_TS(g, 3); first_mma = false; }
          if (kb + 1 == kb1) XB_TS(g, 10);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(sA + stage * Cfg::A_BYTES);
          const uint32_t b_addr = smem_u32(sB + stage * Cfg::B_BYTES);
#pragma unroll
          for (int k = 0; k < 4; ++k) {               // 4 x (K = 16) per 64-wide k-block
            tc_mma_bf16(d_tmem, umma_desc_sw128(a_addr + k * 32), umma_desc_sw128(b_addr + k * 32), idesc,
                        (kb > kb0 || k > 0) ? 1u : 0u);
          }
          tc_commit(&empty_bar[stage]);               // frees the smem slot when the MMAs retire
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        tc_commit(&tfull_bar[acc]);                   // accumulator complete -> epilogue
        if (++acc == Cfg::ACC_STAGES) { acc = 0; acc_phase ^= 1; }
      }
      XB_TS(g, 4);
    }
  } else {
    // ===================== epilogue warps 2..5: TMEM -> registers -> global =====================
    const int quad = warp & 3;                        // TMEM lane quadrant this warp may read
    const int row_in_tile = quad * 32 + lane;
    pdl_wait();                                       // `out` may still be read by the upstream grid
    if (warp == 2 && lane == 0 && g.ts != nullptr) {   // slot 5: the SM this block runs on
      unsigned int smid;
      asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
      g.ts[(size_t)blockIdx.x * XB_TS_SLOTS + 5] = smid;
    }
    int acc = 0;
    uint32_t acc_phase = 0;
    while (walk.next(g, wu)) {
      const int s = wu.slot, ta = wu.ta, tb = wu.tb;
      mbar_wait(&tfull_bar[acc], acc_phase);
      if (warp == 2 && lane == 0) XB_TS(g, 8);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * BN);
      const int arow = ta * 128 + row_in_tile;
      bool last_unit = false;
      if (BN <= 128 && g.mode == XB_GEMM_SKINNY && g.bulk_epi && g.out2 == nullptr) {
        UnitWalker peek = walk;
        GemmUnit nu;
        last_unit = !peek.next(g, nu);
      }
      if (BN <= 128 && last_unit) {
        // The block's LAST unit is pure tail: nothing overlaps its epilogue (earlier units drain while the next
        // unit's MMAs run).  Measured on B200 with per-block timestamps (profiles/gemm_timeline.py): the per-token
        // strided fp32 stores of the general path take 2.4 us per unit at 32 token rows -- more than the whole gap
        // the weight prefetch of the next Linear can cover (32 separate 512-byte bulk copies, one per lane, were no
        // faster: ~64 ns of issue time each).  All loads of the block have landed and all its MMAs have retired by
        // now, so the operand ring is free: the [tokens x 128 features] block is transposed through it and leaves
        // as ONE 2-D TMA store (a [T x 128] fp32 box of the partial plane; columns past N are clipped).
        float* stg = reinterpret_cast<float*>(sA);          // [BN][128] fp32 <= STAGES * 16 KB
#pragma unroll 1
        for (int c = 0; c < BN; c += 16) {
          uint32_t v[16];
          tmem_ld16(t_addr + c, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            float x = __uint_as_float(v[j]);
            if (g.round_bf16) x = rbf(x);
            stg[(c + j) * 128 + row_in_tile] = x;
          }
        }
        if (warp == 2 && lane == 0) XB_TS(g, 11);
        tc_fence_before();
        fence_proxy_async();                                // generic-proxy writes -> visible to the TMA engine
        asm volatile("bar.sync 2, 128;" ::: "memory");
        if (warp == 2 && lane == 0) {
          XB_TS(g, 12);
          asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                       ::"l"(&tmC), "r"(smem_u32(stg)), "r"(ta * 128), "r"(s * g.b_rows) : "memory");
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          XB_TS(g, 13);
          asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // shared memory may be released at exit
          XB_TS(g, 14);
        }
      } else if (BN == 256 && g.mode == XB_GEMM_WIDE && g.epi == XB_EPI_QKV) {
        epi_wide_qkv<BN>(g, t_addr, arow, tb);
      } else if (BN == 256 && g.mode == XB_GEMM_WIDE && g.epi == XB_EPI_SILU) {
        epi_wide_silu<BN>(g, t_addr, arow, tb);
      } else
#pragma unroll 1
      for (int c = 0; c < BN; c += 16) {
        uint32_t v[16];
        tmem_ld16(t_addr + c, v);
        tmem_ld_wait();
        if (g.mode == XB_GEMM_SKINNY) {
          // arow = output feature n, column = token t; lanes write consecutive n (coalesced)
          float* out = reinterpret_cast<float*>(g.out);
          if (arow < g.a_rows) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const int t = tb * BN + c + j;
              if (t < g.b_rows) {
                float x = __uint_as_float(v[j]);
                if (g.round_bf16) x = rbf(x);
                out[((long long)s * g.b_rows + t) * g.ld_out + arow] = x;
                if (g.out2 != nullptr) g.out2[((long long)s * g.b_rows + t) * g.ld_out + arow] = x;
              }
            }
          }
        } else {
          // arow = token t, columns = 16 consecutive features
          bf16* out = reinterpret_cast<bf16*>(g.out);
          const int n0 = tb * BN + c;
          if (arow < g.a_rows && n0 < g.b_rows) {
            float x[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) x[j] = __uint_as_float(v[j]);
            if (g.bias != nullptr) {
#pragma unroll
              for (int j = 0; j < 16; ++j)
                if (n0 + j < g.b_rows) x[j] += bf2f(g.bias[n0 + j]);
            }
            bf16* dst = out + (long long)arow * g.ld_out + n0;
            if (n0 + 16 <= g.b_rows && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
              uint4 p0, p1;
              p0.x = f2_to_bf2(x[0], x[1]);   p0.y = f2_to_bf2(x[2], x[3]);
              p0.z = f2_to_bf2(x[4], x[5]);   p0.w = f2_to_bf2(x[6], x[7]);
              p1.x = f2_to_bf2(x[8], x[9]);   p1.y = f2_to_bf2(x[10], x[11]);
              p1.z = f2_to_bf2(x[12], x[13]); p1.w = f2_to_bf2(x[14], x[15]);
              reinterpret_cast<uint4*>(dst)[0] = p0;
              reinterpret_cast<uint4*>(dst)[1] = p1;
            } else {
              for (int j = 0; j < 16; ++j)
                if (n0 + j < g.b_rows) dst[j] = f2bf(x[j]);
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[acc]);
      if (++acc == Cfg::ACC_STAGES) { acc = 0; acc_phase ^= 1; }
      // the accumulator is released: the next unit's MMAs run while the tile is reduced
    }
    if (warp == 2 && lane == 0) XB_TS(g, 6);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::TM_COLS);
  }
  if (threadIdx.x == 0) XB_TS(g, 7);
}

// ----------------------------------------------------------------------------- CUDA-core checker
// Plain fp32-accumulate Linear used by the kernel-level tests (xb_debug_linear) to check the
// tcgen05 path on the GPU itself.  out[t][n] = sum_k x[t][k] * w[n][k]; one warp per (n, 8 tokens).
__global__ void xb_linear_check_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w, float* __restrict__ out,
                                       int T, int N, int K) {
  const int n = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int t0 = blockIdx.y * 8;
  const int lane = threadIdx.x & 31;
  if (n >= N) return;
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  for (int k = lane * 2; k < K; k += 64) {
    const float2 wv = bf2_to_f2(*reinterpret_cast<const uint32_t*>(w + (long long)n * K + k));
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (t0 + i < T) {
        const float2 xv = bf2_to_f2(*reinterpret_cast<const uint32_t*>(x + (long long)(t0 + i) * K + k));
        acc[i] = fmaf(xv.x, wv.x, acc[i]);
        acc[i] = fmaf(xv.y, wv.y, acc[i]);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float v = warp_sum(acc[i]);
    if (lane == 0 && t0 + i < T) out[(long long)(t0 + i) * N + n] = v;
  }
}

// row-major rows [row0, row0+rows) x K of a weight matrix -> tile-major storage
// dst[((r/128)*KB + k/64)*8192 + (r%128)*64 + k%64]; 16 bytes per thread
// interleave > 0: source row j lands at fused row (j/16)*32 + interleave_off + j%16 (gate/up blocks)
__global__ void xb_tile_weight_kernel(const bf16* __restrict__ src, bf16* __restrict__ dst, long long row0, long long rows,
                                      int K, int KB, int interleave, int interleave_off) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int k8 = K / 8;
  if (i >= rows * k8) return;
  const long long j = i / k8;
  const long long r = interleave ? (j / 16) * 32 + interleave_off + (j % 16) : row0 + j;
  const int k = (int)(i % k8) * 8;
  const uint4 v = *reinterpret_cast<const uint4*>(src + (i / k8) * K + k);
  *reinterpret_cast<uint4*>(dst + ((r >> 7) * KB + (k >> 6)) * 8192 + (r & 127) * 64 + (k & 63)) = v;
}

// ----------------------------------------------------------------------------- host side
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled xb_get_encode_fn() {
  static PFN_encodeTiled fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(p);
  }
  return fn;
}

// 2-D bf16 row-major [rows, cols] tensor, box = [box_rows, 64 cols], 128B swizzle, zero fill out of bounds
static int xb_make_tmap(CUtensorMap* m, const void* ptr, long long rows, long long cols, long long ld_elems, int box_rows) {
  PFN_encodeTiled fn = xb_get_encode_fn();
  XB_CHECK(fn != nullptr, "cuTensorMapEncodeTiled is not available from the driver");
  XB_CHECK((reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && (ld_elems * 2) % 16 == 0,
           "TMA operand must be 16-byte aligned with a 16-byte multiple row pitch (ptr=%p ld=%lld)", ptr, ld_elems);
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)ld_elems * 2};
  cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  XB_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d (rows=%lld cols=%lld box_rows=%d)", (int)r,
           rows, cols, box_rows);
  return 0;
}

// 2-D fp32 row-major [rows, cols] tensor, box = [box_rows, 128 cols], no swizzle (TMA store of partial planes)
static int xb_make_tmap_f32(CUtensorMap* m, const void* ptr, long long rows, long long cols, long long ld_elems, int box_rows) {
  PFN_encodeTiled fn = xb_get_encode_fn();
  XB_CHECK(fn != nullptr, "cuTensorMapEncodeTiled is not available from the driver");
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)ld_elems * 4};
  cuuint32_t box[2] = {128, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(ptr), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  XB_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (fp32 store map) failed with CUresult %d (rows=%lld cols=%lld box_rows=%d)",
           (int)r, rows, cols, box_rows);
  return 0;
}

template <int BN, int STAGES>
static int xb_gemm_launch_t(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC, const GemmArgs& g, int grid,
                            cudaStream_t st, int pdl) {
  using Cfg = GemmCfg<BN, STAGES>;
  {
    // per device: the attribute does not carry over to another GPU of the same process
    static bool done[XB_MAX_DEVICES] = {false};
    int dev = 0;
    XB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= XB_MAX_DEVICES || !done[dev]) {
      XB_CUDA(cudaFuncSetAttribute(xb_gemm_tcgen05_kernel<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   Cfg::SMEM_BYTES));
      if (dev >= 0 && dev < XB_MAX_DEVICES) done[dev] = true;
    }
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid, 1, 1);
  cfg.blockDim = dim3(192, 1, 1);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  XB_CUDA(cudaLaunchKernelEx(&cfg, xb_gemm_tcgen05_kernel<BN, STAGES>, tmA, tmB, tmC, g));
  return 0;
}

// A planned Linear: tensor maps + unit decomposition, built once per (site, token count).
struct GemmPlan {
  CUtensorMap tmA, tmB;
  CUtensorMap tmC;   // SKINNY with bulk_epi: fp32 view [splits * T, N] of the partial planes (TMA store); else a copy of tmA
  GemmArgs args;
  SplitInfo split;   // how a consumer sums the fp32 partial planes (SKINNY)
  int bn;
  int grid;
  int stages;        // SKINNY: depth of the smem pipeline (xb_skinny_stages)
};

// Depth of the shared-memory pipeline of the skinny (decode) Linear by token-column width.  Every stage is a 16 KB box
// of weights plus the token tile; the stages a block can fill before the upstream grid is done are what keeps HBM
// busy across the kernel boundary, while a block too large to share the SM with its neighbours in the stream loses
// that overlap again.  Measured on the graph-replayed 8B step (ms per step; profiles/r02_decode_ab_stages.txt):
//   16 columns (bs=1):  5: 3.294  6: 3.249  7: 3.227  8: 3.280  9: 3.299  10: 3.335
//   32 columns (bs=32): 5: 4.85   6: 4.67   7: 4.68   8: 4.70   9: 4.72   10: 4.76
//   64 columns (bs=64): 4: 6.85   5: 6.60   6: 6.59   7: 6.59   8: 6.57
// XB_GEMM_STAGES=r1 selects the round-1 depths (5 / 5 / 4 / 6) for an A/B; a number selects that depth for the widths
// that have it instantiated (xb_gemm_run) and leaves the others at their default.
static int xb_skinny_stages(int bn) {
  const int w = bn <= 16 ? 0 : bn <= 32 ? 1 : bn <= 64 ? 2 : 3;
  static const int dflt[4] = {7, 6, 8, 6}, r1[4] = {5, 5, 4, 6};
  static const int have[4][4] = {{5, 7, 0, 0}, {5, 6, 7, 8}, {4, 8, 0, 0}, {6, 7, 0, 0}};
  const char* ev = getenv("XB_GEMM_STAGES");
  if (ev == nullptr) return dflt[w];
  if (ev[0] == 'r') return r1[w];
  const int n = atoi(ev);
  for (int i = 0; i < 4; ++i)
    if (n > 0 && have[w][i] == n) return n;
  return dflt[w];
}

static int xb_pick_bn_skinny(int T) { return T <= 16 ? 16 : T <= 32 ? 32 : T <= 64 ? 64 : 128; }

// y[T,N] = x[T,K] W[N,K]^T.  T <= 128 -> SKINNY (fp32 partials [splits][T][N] into `out`);
// otherwise WIDE (bf16 [T, ld_out] into `out`).
static int xb_gemm_plan(GemmPlan* p, const bf16* x, long long ldx, const bf16* w, int w_tiled, int T, int N, int K,
                        void* out, long long ld_out, const bf16* bias, int round_bf16, int num_sms, int force_splits) {
  XB_CHECK(K % 8 == 0, "K=%d must be a multiple of 8", K);
  GemmArgs& g = p->args;
  g.k_blocks = (K + 63) / 64;
  g.out = out;
  g.ld_out = ld_out;
  g.bias = bias;
  g.round_bf16 = round_bf16;
  g.w_tiled = w_tiled;
  g.epi = XB_EPI_PLAIN;
  g.bulk_epi = 0;
  g.out2 = nullptr;
  g.ts = nullptr;
  const long long w_tiles = (N + 127) / 128;
  if (T <= 128) {
    const int bn = xb_pick_bn_skinny(T);
    p->bn = bn;
    g.mode = XB_GEMM_SKINNY;
    g.a_tiles = (N + 127) / 128;
    g.b_tiles = 1;
    g.a_rows = N;
    g.b_rows = T;
    g.sk_q = 0;
    g.sk_items = g.a_tiles * g.k_blocks;
    int splits = force_splits > 0 ? force_splits : 1;
    if (round_bf16) splits = 1;          // the result is final: whole tiles, no partial planes
    if (splits > g.k_blocks) splits = g.k_blocks;
    g.splits = splits;
    if (force_splits <= 0 && !round_bf16) {
      // stream-K: one block per SM, each owning an equal contiguous range of (tile, k-block) items
      int grid = num_sms;
      if (g.sk_items < grid * 2) grid = (g.sk_items + 1) / 2;
      if (grid < 1) grid = 1;
      g.sk_q = (g.sk_items + grid - 1) / grid;
      g.splits = (g.k_blocks + g.sk_q - 1) / g.sk_q + 1;   // most planes any tile can have
    }
    if (w_tiled) XB_TRY(xb_make_tmap(&p->tmA, w, w_tiles * g.k_blocks * 128, 64, 64, 128));
    else XB_TRY(xb_make_tmap(&p->tmA, w, N, K, K, 128));
    XB_TRY(xb_make_tmap(&p->tmB, x, T, K, ldx, bn));
    p->tmC = p->tmA;
  } else {
    p->bn = 256;
    g.mode = XB_GEMM_WIDE;
    g.a_tiles = (T + 127) / 128;
    g.b_tiles = (N + 255) / 256;
    g.a_rows = T;
    g.b_rows = N;
    g.splits = 1;
    g.sk_q = 0;
    g.sk_items = 0;
    XB_TRY(xb_make_tmap(&p->tmA, x, T, K, ldx, 128));
    if (w_tiled) XB_TRY(xb_make_tmap(&p->tmB, w, w_tiles * g.k_blocks * 128, 64, 64, 128));
    else XB_TRY(xb_make_tmap(&p->tmB, w, N, K, K, 256));
  }
  // one block per SM: under programmatic dependent launch the NEXT kernel's block co-resides
  // and prefetches its first weight stages while this one drains
  if (g.mode == XB_GEMM_SKINNY && g.sk_q > 0) {
    p->grid = (g.sk_items + g.sk_q - 1) / g.sk_q;
  } else {
    const int units = g.a_tiles * g.b_tiles * g.splits;
    p->grid = units < num_sms ? units : num_sms;
  }
  if (g.mode == XB_GEMM_SKINNY) {
    // last-unit epilogue through a TMA store: 16-byte aligned base and row pitch (XB_BULK_EPI=0: per-token stores only)
    const char* bv = getenv("XB_BULK_EPI");
    if (!(bv != nullptr && bv[0] == '0') && ld_out % 4 == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0) {
      XB_TRY(xb_make_tmap_f32(&p->tmC, out, (long long)g.splits * T, N, ld_out, T));
      g.bulk_epi = 1;
    }
  } else {
    p->tmC = p->tmA;
  }
  p->stages = g.mode == XB_GEMM_SKINNY ? xb_skinny_stages(p->bn) : 0;
  p->split.k_blocks = g.k_blocks;
  p->split.q = g.mode == XB_GEMM_SKINNY ? g.sk_q : 0;
  p->split.splits = g.mode == XB_GEMM_SKINNY ? g.splits : 1;
  return 0;
}

static int xb_gemm_run(const GemmPlan& p, cudaStream_t st, int pdl) {
  if (p.stages > 0) {
#define XB_STAGE_CASE(BN_, N_) if (p.bn == BN_ && p.stages == N_) return xb_gemm_launch_t<BN_, N_>(p.tmA, p.tmB, p.tmC, p.args, p.grid, st, pdl);
    XB_STAGE_CASE(16, 7) XB_STAGE_CASE(32, 6) XB_STAGE_CASE(64, 8) XB_STAGE_CASE(128, 6)          // defaults
    XB_STAGE_CASE(16, 5) XB_STAGE_CASE(32, 5) XB_STAGE_CASE(64, 4) XB_STAGE_CASE(128, 7)          // A/B (XB_GEMM_STAGES)
    XB_STAGE_CASE(32, 7) XB_STAGE_CASE(32, 8)
#undef XB_STAGE_CASE
    return xb_fail(1, "no skinny Linear instantiated for %d token columns with %d stages", p.bn, p.stages);
  }
  switch (p.bn) {
    case 128: return xb_gemm_launch_t<128, 6>(p.tmA, p.tmB, p.tmC, p.args, p.grid, st, pdl);
    case 256: return xb_gemm_launch_t<256, 4>(p.tmA, p.tmB, p.tmC, p.args, p.grid, st, pdl);
  }
  return xb_fail(1, "unsupported BN %d", p.bn);
}
