// Native checkpoint loader: one .safetensors file -> this shard's weights in HBM.
//
// Replaces, for the layer range of a stage, what load_model_weights_torchtune does through the
// safetensors Python package (xotorch/inference/torch/models/llm_utils.py:136-284): open every
// file, read EVERY tensor into host memory, rename, permute q/k on the CPU, build a state dict,
// load_state_dict, cast.  Here the file is mapped, only the tensors of this shard's layers are
// touched, and each of them streams  page cache -> pinned ring -> device staging -> final layout
// in row chunks: the q/k permutation (llm_utils.py:126-134), the fused q|k|v and gate|up row
// placement, the tile-major repack and the dtype cast all happen in ONE device kernel per chunk,
// overlapped with the next chunk's host copy.
//
// File format (safetensors): u64 little-endian header length N, N bytes of JSON
//   {"tensor.name": {"dtype": "BF16", "shape": [r, c], "data_offsets": [begin, end]}, ...,
//    "__metadata__": {...}}
// then the raw little-endian tensor bytes; offsets are relative to the end of the header.
#pragma once
#include "xb_common.cuh"
#include <cuda_fp16.h>
#include <errno.h>
#include <fcntl.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <string>
#include <thread>
#include <vector>

struct StEntry {
  std::string name, dtype;
  std::vector<long long> shape;
  size_t begin = 0, end = 0;
};

// ----------------------------------------------------------------------------- minimal JSON reader
// Enough for a safetensors header: objects, arrays, strings (with escapes), integers; other
// scalars are skipped.  Returns false on malformed input.
struct StJson {
  const char* p;
  const char* end;
  void ws() {
    while (p < end && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) ++p;
  }
  bool lit(char c) {
    ws();
    if (p < end && *p == c) { ++p; return true; }
    return false;
  }
  bool str(std::string* out) {
    ws();
    if (p >= end || *p != '"') return false;
    ++p;
    std::string s;
    while (p < end && *p != '"') {
      if (*p == '\\' && p + 1 < end) {
        ++p;
        switch (*p) {
          case 'n': s += '\n'; break;
          case 't': s += '\t'; break;
          case 'u': p += 4; s += '?'; break;   // names in HF checkpoints are ASCII
          default: s += *p;
        }
        ++p;
      } else {
        s += *p++;
      }
    }
    if (p >= end) return false;
    ++p;
    if (out) *out = s;
    return true;
  }
  bool integer(long long* v) {
    ws();
    const char* q = p;
    if (q < end && (*q == '-' || *q == '+')) ++q;
    if (q >= end || *q < '0' || *q > '9') return false;
    long long x = 0;
    bool neg = *p == '-';
    while (q < end && *q >= '0' && *q <= '9') x = x * 10 + (*q++ - '0');
    p = q;
    *v = neg ? -x : x;
    return true;
  }
  bool skip_value() {   // any JSON value
    ws();
    if (p >= end) return false;
    if (*p == '"') return str(nullptr);
    if (*p == '{' || *p == '[') {
      const char open = *p, close = open == '{' ? '}' : ']';
      ++p;
      if (lit(close)) return true;
      while (true) {
        if (open == '{') {
          if (!str(nullptr) || !lit(':')) return false;
        }
        if (!skip_value()) return false;
        if (lit(',')) continue;
        return lit(close);
      }
    }
    while (p < end && *p != ',' && *p != '}' && *p != ']') ++p;   // number / true / false / null
    return true;
  }
};

static bool st_parse_header(const char* json, size_t n, std::vector<StEntry>* out) {
  StJson j{json, json + n};
  if (!j.lit('{')) return false;
  if (j.lit('}')) return true;
  while (true) {
    StEntry e;
    if (!j.str(&e.name) || !j.lit(':')) return false;
    if (e.name == "__metadata__") {
      if (!j.skip_value()) return false;
    } else {
      if (!j.lit('{')) return false;
      bool have_off = false;
      if (!j.lit('}')) {
        while (true) {
          std::string key;
          if (!j.str(&key) || !j.lit(':')) return false;
          if (key == "dtype") {
            if (!j.str(&e.dtype)) return false;
          } else if (key == "shape" || key == "data_offsets") {
            std::vector<long long> v;
            if (!j.lit('[')) return false;
            if (!j.lit(']')) {
              while (true) {
                long long x;
                if (!j.integer(&x)) return false;
                v.push_back(x);
                if (j.lit(',')) continue;
                if (!j.lit(']')) return false;
                break;
              }
            }
            if (key == "shape") {
              e.shape = v;
            } else {
              if (v.size() != 2 || v[0] < 0 || v[1] < v[0]) return false;
              e.begin = (size_t)v[0];
              e.end = (size_t)v[1];
              have_off = true;
            }
          } else if (!j.skip_value()) {
            return false;
          }
          if (j.lit(',')) continue;
          if (!j.lit('}')) return false;
          break;
        }
      }
      if (!have_off || e.dtype.empty()) return false;
      out->push_back(e);
    }
    if (j.lit(',')) continue;
    return j.lit('}');
  }
}

// ----------------------------------------------------------------------------- device side
enum { ST_BF16 = 0, ST_F16 = 1, ST_F32 = 2 };
static int st_dtype_code(const std::string& s) { return s == "BF16" ? ST_BF16 : s == "F16" ? ST_F16 : s == "F32" ? ST_F32 : -1; }
static size_t st_dtype_size(int dt) { return dt == ST_F32 ? 4 : 2; }

// 8 consecutive source elements -> 8 bf16 (round to nearest even, what Tensor.to(bfloat16) does)
__device__ __forceinline__ uint4 st_load8(const void* src, long long idx, int dt) {
  uint4 o;
  if (dt == ST_BF16) {
    o = *reinterpret_cast<const uint4*>(reinterpret_cast<const bf16*>(src) + idx);
  } else if (dt == ST_F16) {
    const uint4 v = *reinterpret_cast<const uint4*>(reinterpret_cast<const __half*>(src) + idx);
    const __half2* h = reinterpret_cast<const __half2*>(&v);
    uint32_t* w = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = __half22float2(h[i]);
      w[i] = f2_to_bf2(f.x, f.y);
    }
  } else {
    const float4 a = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(src) + idx);
    const float4 b = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(src) + idx + 4);
    o.x = f2_to_bf2(a.x, a.y); o.y = f2_to_bf2(a.z, a.w);
    o.z = f2_to_bf2(b.x, b.y); o.w = f2_to_bf2(b.z, b.w);
  }
  return o;
}

// A chunk of `rows` source rows (first one = row j0 of the checkpoint tensor) -> tile-major storage
//   dst[((r/128)*KB + k/64)*8192 + (r%128)*64 + k%64],  r = destination row of the fused matrix:
//   perm_hd > 0 : the reference's q/k permutation, new row h*hd + 2i + j  <-  old row h*hd + j*hd/2 + i
//                 (llm_utils.py:126-134: view(n_heads, 2, hd/2, dim).transpose(1, 2))
//   interleave  : gate/up rows interleaved in 16-feature blocks, (row/16)*32 + interleave_off + row%16
//   else        : row0 + row (q | k | v stacked)
__global__ void xb_load_rows_tiled_kernel(const void* __restrict__ src, int dt, long long j0, long long rows, int K,
                                          bf16* __restrict__ dst, int KB, long long row0, int interleave, int interleave_off,
                                          int perm_hd) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int k8 = K / 8;
  if (i >= rows * k8) return;
  const long long j = i / k8;
  const int k = (int)(i % k8) * 8;
  long long js = j0 + j;
  if (perm_hd > 0) {
    const long long h = js / perm_hd;
    const int rem = (int)(js % perm_hd), half = perm_hd >> 1;
    js = h * perm_hd + 2 * (rem % half) + rem / half;
  }
  const long long r = interleave ? (js / 16) * 32 + interleave_off + (js % 16) : row0 + js;
  const uint4 v = st_load8(src, j * K + k, dt);
  *reinterpret_cast<uint4*>(dst + ((r >> 7) * KB + (k >> 6)) * 8192 + (r & 127) * 64 + (k & 63)) = v;
}

// n elements (multiple of 8), converted, stored contiguously from dst
__global__ void xb_load_copy_kernel(const void* __restrict__ src, int dt, long long n8, bf16* __restrict__ dst) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n8) return;
  *reinterpret_cast<uint4*>(dst + i * 8) = st_load8(src, i * 8, dt);
}

// ----------------------------------------------------------------------------- host -> device ring
// NBUF pinned buffers + NBUF device staging buffers.  stage() copies a byte range of the mapped file
// into the next pinned buffer with a few threads (page-cache reads are per-thread memcpy bound),
// starts the H2D copy on the loader stream and hands back the device pointer; the caller launches
// its layout kernel on the same stream and calls done().  The host copy of chunk i+1 overlaps the
// DMA + kernel of chunk i.
struct StRing {
  static constexpr int NBUF = 3;
  size_t chunk = 0;
  void* pin[NBUF] = {nullptr, nullptr, nullptr};
  void* dev[NBUF] = {nullptr, nullptr, nullptr};
  cudaEvent_t ev[NBUF] = {nullptr, nullptr, nullptr};
  cudaStream_t st = nullptr;
  int next = 0, threads = 8;

  int init(size_t chunk_bytes) {
    chunk = chunk_bytes;
    XB_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    for (int i = 0; i < NBUF; ++i) {
      XB_CUDA(cudaMallocHost(&pin[i], chunk));
      XB_CUDA(cudaMalloc(&dev[i], chunk));
      XB_CUDA(cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming));
    }
    const char* t = getenv("XB_LOAD_THREADS");
    if (t != nullptr && atoi(t) > 0) threads = atoi(t);
    return 0;
  }
  void destroy() {
    if (st) cudaStreamSynchronize(st);
    for (int i = 0; i < NBUF; ++i) {
      if (pin[i]) cudaFreeHost(pin[i]);
      if (dev[i]) cudaFree(dev[i]);
      if (ev[i]) cudaEventDestroy(ev[i]);
    }
    if (st) cudaStreamDestroy(st);
  }
  int stage(const uint8_t* src, size_t bytes, void** dptr, int* idx) {
    XB_CHECK(bytes <= chunk, "loader chunk overflow: %zu > %zu", bytes, chunk);
    const int i = next;
    next = (next + 1) % NBUF;
    XB_CUDA(cudaEventSynchronize(ev[i]));   // the previous user of this pair is finished
    const int nt = bytes >= ((size_t)4 << 20) ? threads : 1;
    if (nt <= 1) {
      memcpy(pin[i], src, bytes);
    } else {
      std::vector<std::thread> pool;
      const size_t part = ((bytes / nt) + 4095) & ~(size_t)4095;
      for (int t = 1; t < nt; ++t) {
        const size_t o = (size_t)t * part;
        if (o >= bytes) break;
        const size_t len = o + part < bytes ? part : bytes - o;
        pool.emplace_back([=] { memcpy((uint8_t*)pin[i] + o, src + o, len); });
      }
      memcpy(pin[i], src, part < bytes ? part : bytes);
      for (auto& th : pool) th.join();
    }
    XB_CUDA(cudaMemcpyAsync(dev[i], pin[i], bytes, cudaMemcpyHostToDevice, st));
    *dptr = dev[i];
    *idx = i;
    return 0;
  }
  int done(int idx) {
    XB_CUDA(cudaGetLastError());
    XB_CUDA(cudaEventRecord(ev[idx], st));
    return 0;
  }
};

// A mapped read-only file; unmapped on destruction.
struct StFile {
  int fd = -1;
  const uint8_t* base = nullptr;
  size_t size = 0;
  ~StFile() {
    if (base) munmap((void*)base, size);
    if (fd >= 0) close(fd);
  }
  int open_map(const char* path) {
    fd = open(path, O_RDONLY);
    XB_CHECK(fd >= 0, "cannot open %s: %s", path, strerror(errno));
    struct stat sb;
    XB_CHECK(fstat(fd, &sb) == 0 && sb.st_size >= 8, "%s: not a safetensors file (too short)", path);
    size = (size_t)sb.st_size;
    void* p = mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
    XB_CHECK(p != MAP_FAILED, "mmap(%s) failed: %s", path, strerror(errno));
    base = (const uint8_t*)p;
    madvise(p, size, MADV_SEQUENTIAL);
    return 0;
  }
};
