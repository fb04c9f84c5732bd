// Sampler with the row resident on chip: a cluster of CL blocks per row, each block keeps its
// 1/CL slice of the SCALED logits in shared memory, so the row is read from L2 exactly once and
// the radix-select / softmax / argmax passes run out of shared memory; per-pass results
// (histograms, sums, best candidates) are merged across the cluster through distributed shared
// memory.  Same arithmetic as xb_sample_kernel (xb_kernels.cuh), which restates
// torchtune.generation.sample as the reference calls it (sharded_inference_engine.py:208-228);
// that kernel remains the fallback for rows too long for CL x 200 KB of shared memory.
//
// One block per row spends ~14 us per pass over a 128256-entry row (a single SM pulls ~40 GB/s
// out of L2): 7 passes = 100 us for 32 rows on 32 of the 148 SMs.
#pragma once
#include "xb_kernels.cuh"
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

// exchange area of one block, read by its cluster peers; two copies used alternately so that one
// cluster barrier per exchange is enough (a copy is rewritten only two exchanges later)
struct SampleXchg {
  unsigned int hist[256];
  float fhist[256];
  float f0, f1;
  int i0;
  int pad;
};

template <int NT, int CL>
__global__ void __launch_bounds__(NT) xb_sample_cluster_kernel(const float* __restrict__ logits, int V, int chunk, float temp,
                                                               int top_k, float top_p, const float* __restrict__ q_noise,
                                                               unsigned long long seed, int* __restrict__ out) {
  extern __shared__ float vals[];                 // this block's slice of the row, already * 1/temp
  __shared__ SampleXchg xc[2];
  __shared__ unsigned int mh[256];
  __shared__ float mfh[256];
  __shared__ uint32_t s_prefix;
  __shared__ int s_need;
  __shared__ float s_needf;
  __shared__ float red[32];
  __shared__ float sv[NT / 32];
  __shared__ int si[NT / 32];
  cg::cluster_group cluster = cg::this_cluster();
  const int rank = (int)cluster.block_rank();
  const int rowi = blockIdx.x / CL;
  const float* row = logits + (long long)rowi * V;
  const float inv_t = 1.0f / fmaxf(temp, 1e-5f);
  const int i0 = rank * chunk;
  const int n = V - i0 < 0 ? 0 : (V - i0 < chunk ? V - i0 : chunk);
  int xi = 0;                                     // which exchange copy the next merge uses

  // ---- load the slice once (float4 when the row allows it), scale, local max
  float mx = -INFINITY;
  if ((V & 3) == 0 && (chunk & 3) == 0 && (reinterpret_cast<uintptr_t>(row) & 15) == 0) {
    const float4* r4 = reinterpret_cast<const float4*>(row + i0);
    float4* v4 = reinterpret_cast<float4*>(vals);
    const int n4 = n >> 2;
    auto put = [&](int i, float4 a) {
      a.x *= inv_t; a.y *= inv_t; a.z *= inv_t; a.w *= inv_t;
      v4[i] = a;
      mx = fmaxf(fmaxf(mx, fmaxf(a.x, a.y)), fmaxf(a.z, a.w));
    };
    int i = threadIdx.x;
    for (; i + 3 * NT < n4; i += 4 * NT) {   // four loads in flight per thread: the slice comes out of L2 once
      const float4 a = r4[i], b = r4[i + NT], c = r4[i + 2 * NT], d = r4[i + 3 * NT];
      put(i, a); put(i + NT, b); put(i + 2 * NT, c); put(i + 3 * NT, d);
    }
    for (; i < n4; i += NT) put(i, r4[i]);
  } else {
    for (int i = threadIdx.x; i < n; i += NT) {
      const float v = row[i0 + i] * inv_t;
      vals[i] = v;
      mx = fmaxf(mx, v);
    }
  }
  mx = warp_max(mx);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
  __syncthreads();
  if (threadIdx.x == 0) {
    float m2 = red[0];
    for (int w = 1; w < NT / 32; ++w) m2 = fmaxf(m2, red[w]);
    xc[0].f0 = m2;                                // travels with the first exchange
  }
  __syncthreads();

  // ---- 1. pivot = k-th largest scaled logit: 4 radix passes over shared memory
  float mxv = -INFINITY;                          // global max = max of the kept set
  bool have_max = false;
  uint32_t pivot_key = 0;
  const bool use_k = top_k > 0 && top_k < V;
  if (use_k) {
    uint32_t prefix = 0;
    int need = top_k;
    for (int pass = 0; pass < 4; ++pass) {
      const int shift = 24 - 8 * pass;
      SampleXchg& x = xc[xi];
      for (int i = threadIdx.x; i < 256; i += NT) x.hist[i] = 0;
      __syncthreads();
      const uint32_t mask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
      for (int i = threadIdx.x; i < n; i += NT) {
        const uint32_t key = float_order_key(vals[i]);
        if ((key & mask) == (prefix & mask)) atomicAdd(&x.hist[(key >> shift) & 255], 1u);
      }
      cluster.sync();
      if (threadIdx.x < 256) {
        unsigned int c = 0;
#pragma unroll
        for (int r = 0; r < CL; ++r) c += cluster.map_shared_rank(&xc[xi], r)->hist[threadIdx.x];
        mh[threadIdx.x] = c;
      }
      if (!have_max && threadIdx.x == 256 % NT) {
        float m2 = -INFINITY;
        for (int r = 0; r < CL; ++r) m2 = fmaxf(m2, cluster.map_shared_rank(&xc[0], r)->f0);
        s_needf = m2;
      }
      __syncthreads();
      if (!have_max) { mxv = s_needf; have_max = true; }
      if (threadIdx.x == 0) {
        int nn = need, b = 255;
        for (; b > 0; --b) {
          if ((int)mh[b] >= nn) break;
          nn -= (int)mh[b];
        }
        s_prefix = prefix | ((uint32_t)b << shift);
        s_need = nn;
      }
      __syncthreads();
      prefix = s_prefix;
      need = s_need;
      xi ^= 1;
    }
    pivot_key = prefix;
  }
  const float pivot = use_k ? key_to_float(pivot_key) : -INFINITY;

  // ---- 2. softmax statistics over the kept values (>= pivot)
  if (!have_max) {                                // no top-k: the max has not been exchanged yet
    cluster.sync();
    if (threadIdx.x == 0) {
      float m2 = -INFINITY;
      for (int r = 0; r < CL; ++r) m2 = fmaxf(m2, cluster.map_shared_rank(&xc[0], r)->f0);
      s_needf = m2;
    }
    __syncthreads();
    mxv = s_needf;
    xi = 1;                                       // copy 0 was just read by the peers
    __syncthreads();
  }
  float sum = 0.f;
  for (int i = threadIdx.x; i < n; i += NT) {
    const float v = vals[i];
    if (v >= pivot) sum += expf(v - mxv);
  }
  sum = block_sum<NT>(sum, red);
  if (threadIdx.x == 0) xc[xi].f1 = sum;
  cluster.sync();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int r = 0; r < CL; ++r) t += cluster.map_shared_rank(&xc[xi], r)->f1;   // fixed order: deterministic
    s_needf = t;
  }
  __syncthreads();
  sum = s_needf;
  xi ^= 1;
  __syncthreads();

  // ---- 2b. nucleus: largest probability threshold whose upper set has mass >= top_p
  float p_cut = 0.f;
  if (top_p > 0.f && top_p < 1.f) {
    uint32_t prefix = 0;
    float needf = top_p * sum;                    // in un-normalised mass
    for (int pass = 0; pass < 4; ++pass) {
      const int shift = 24 - 8 * pass;
      SampleXchg& x = xc[xi];
      for (int i = threadIdx.x; i < 256; i += NT) x.fhist[i] = 0.f;
      __syncthreads();
      const uint32_t mask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
      for (int i = threadIdx.x; i < n; i += NT) {
        const float v = vals[i];
        if (v >= pivot) {
          const float e = expf(v - mxv);
          const uint32_t key = float_order_key(e);
          if ((key & mask) == (prefix & mask)) atomicAdd(&x.fhist[(key >> shift) & 255], e);
        }
      }
      cluster.sync();
      if (threadIdx.x < 256) {
        float c = 0.f;
#pragma unroll
        for (int r = 0; r < CL; ++r) c += cluster.map_shared_rank(&xc[xi], r)->fhist[threadIdx.x];
        mfh[threadIdx.x] = c;
      }
      __syncthreads();
      if (threadIdx.x == 0) {
        float nn = needf;
        int b = 255;
        for (; b > 0; --b) {
          if (mfh[b] >= nn) break;
          nn -= mfh[b];
        }
        s_prefix = prefix | ((uint32_t)b << shift);
        s_needf = nn;
      }
      __syncthreads();
      prefix = s_prefix;
      needf = s_needf;
      xi ^= 1;
      __syncthreads();
    }
    p_cut = key_to_float(prefix);
  }

  // ---- 3. argmax(p / q) over the kept set; ties -> lowest index
  float best = -INFINITY;
  int bi = 0x7fffffff;
  for (int i = threadIdx.x; i < n; i += NT) {
    const float v = vals[i];
    if (v < pivot) continue;
    const float e = expf(v - mxv);
    if (e < p_cut) continue;
    const int gi = i0 + i;
    float qv;
    if (q_noise != nullptr) {
      qv = q_noise[(long long)rowi * V + gi];
    } else {
      const uint32_t r = mix32(seed * 0x9E3779B97F4A7C15ULL + ((unsigned long long)rowi << 32) + (unsigned)gi);
      const float u = ((float)(r & 0xffffff) + 0.5f) * (1.0f / 16777216.0f);
      qv = -logf(u);
    }
    const float sc = (e / sum) / qv;
    if (sc > best || (sc == best && gi < bi)) { best = sc; bi = gi; }
  }
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
    xc[xi].f0 = best;
    xc[xi].i0 = bi;
  }
  cluster.sync();
  if (rank == 0 && threadIdx.x == 0) {
    for (int r = 1; r < CL; ++r) {
      const SampleXchg* px = cluster.map_shared_rank(&xc[xi], r);
      const float ov = px->f0;
      const int oi = px->i0;
      if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    out[rowi] = bi == 0x7fffffff ? 0 : bi;
  }
  cluster.sync();                                 // peers' shared memory must outlive rank 0's reads
}
