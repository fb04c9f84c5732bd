// libxb200: the C ABI declared in include/xb200.h.
// Engine object = weights of one layer range + paged KV cache + workspaces + launch plans.
#include "../../include/xb200.h"
#include "xb_common.cuh"
#include "xb_gemm.cuh"
#include "xb_loader.cuh"
#include "xb_kernels.cuh"
#include "xb_attn_mma.cuh"
#include "xb_sample_cluster.cuh"

#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include <map>
#include <vector>

thread_local std::string g_xb_err;

int xb_fail(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_xb_err = buf;
  return code;
}

#define XB_DEC_ROWS 128

struct LayerW {
  bf16 *wqkv = nullptr, *bqkv = nullptr, *wo = nullptr, *wgu = nullptr, *wdown = nullptr;
  bf16 *sa_norm = nullptr, *mlp_norm = nullptr;
  bf16 *k_pool = nullptr, *v_pool = nullptr;
  CUtensorMap tmK, tmV;   // [n_pages*KV*page_tokens, head_dim] views of the pools, boxes of min(page,64) x 64
};

enum Site { SITE_QKV = 0, SITE_O, SITE_GU, SITE_DOWN, SITE_HEAD };

struct xb_engine {
  xb_model_desc d;
  int device = 0;
  int num_sms = 148;
  int reserved_sms = 1;   // SMs kept free of the decode Linears (xb_engine_create, xb_reserve_sms)
  int nl = 0;  // local layers
  int NQ = 0, NKV = 0, NQKV = 0;
  std::vector<LayerW> L;
  bf16 *embed = nullptr, *final_norm = nullptr, *lm_head = nullptr;
  float* rope = nullptr;
  // kv paging
  int n_slots = 0, max_pages = 0, page_tokens = 0, n_pages = 0;
  std::vector<int> free_pages;
  std::vector<std::vector<int>> slot_pages;
  std::vector<int> slot_len;
  std::vector<int> h_page_table;
  int* d_page_table = nullptr;
  // batch meta: one device block [tok_seq | tok_pos | seq_slot | seq_last], refreshed by ONE
  // H2D copy per forward from a ring of pinned host buffers
  int* d_meta = nullptr;
  int *d_tok_seq = nullptr, *d_tok_pos = nullptr, *d_seq_slot = nullptr, *d_seq_last = nullptr;
  int *d_tile_start = nullptr, *d_tile_len = nullptr;
  int max_tiles = 0, tile_tokens = 0;   // prefill query tiles: tile_tokens = 128 / group size
  static constexpr int META_RING = 8;
  int* h_meta_ring[META_RING] = {nullptr};
  cudaEvent_t meta_ev[META_RING] = {nullptr};
  int meta_next = 0;
  size_t meta_ints = 0;
  // decode steps of up to XB_DEC_ROWS sequences describe themselves in a compact block of their own:
  // [ctl 4 | tok_pos R | seq_slot R | tok_seq R = 0..R-1 | seq_last R = 0..R-1]; the last two never change, so a step
  // uploads 4 + 2R ints in ONE copy (it used to take three or four small copies per pipeline tick)
  int* d_dec = nullptr;
  bool meta_is_dec = false;   // which block the last forward filled (measurement hooks replay that batch)
  // engine-owned input/output staging so captured graphs see fixed addresses
  int* in_tokens = nullptr;
  float* logits = nullptr;
  // decode-step graphs keyed by (n_seqs, n_splits)
  struct StepGraph { cudaGraphExec_t exec = nullptr; int warmed = 0; uint64_t launches = 0; };
  std::map<long long, StepGraph> graphs;
  int use_graphs = 1;
  int use_pdl = 1;
  // peer hand-off (pipeline decode): landing buffers of THIS stage (written by the previous GPU) and
  // the peer-mapped landing buffers of the NEXT stage
  bf16* landing = nullptr;
  uint32_t* flags = nullptr;
  unsigned int* done_cnt = nullptr;
  bf16* peer_landing = nullptr;
  uint32_t* peer_flags = nullptr;
  int ho_slots = 0, ho_rows = 0;
  bool peer_is_local = false;       // peer_landing / peer_flags are plain device pointers of an engine in this process
  long long ho_timeout_cycles = 0;  // consumer-side wait budget (XB_HANDOFF_TIMEOUT_S)
  int* ho_err = nullptr;            // mapped host memory {slot + 1, wait value}: written by the waiting kernel on time-out
  int* ho_err_dev = nullptr;        // its device address
  HandoffCtl* d_ctl = nullptr;
  // workspaces
  bf16 *h = nullptr, *hn = nullptr, *q = nullptr, *attn = nullptr, *act = nullptr, *last_rows = nullptr;
  void* gemm_out = nullptr;
  size_t gemm_out_bytes = 0;
  float *o_part = nullptr, *ml_part = nullptr;
  int max_attn_splits = 32;
  int fuse_qkv = 1;       // XB_FUSE_QKV=0: q/k/v finalize as its own kernel instead of the decode attention's prologue
  int attn_cluster = 1;   // XB_ATTN_CLUSTER=0: kv splits merged by xb_attn_combine2_kernel instead of a cluster + DSMEM; 16: clusters of up to 16
  // timeline instrumentation (XB_GEMM_TS=1; eager launches only): globaltimer stamps of every Linear launch
  unsigned long long* ts_buf = nullptr;
  int ts_cap = 0, ts_next = 0, ts_blocks = 160;
  int meta_T = 0;         // token rows the device metadata currently describes
  std::map<long long, GemmPlan> plans;
  xb_counters c{};
  std::map<long long, int> loaded;   // (layer + 1) * 64 + kind -> seen by xb_load_weight / xb_load_safetensors
  StRing* ring = nullptr;            // loader staging, alive from the first file until the shard is complete
};

#define LAUNCH_COUNT(e) ((e)->c.launches++)

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is per DEVICE: one flag per device and call site, so a
// second engine on another GPU of the same process sets it again.
#define XB_SMEM_ATTR_ONCE(kernel, bytes)                                                                  \
  do {                                                                                                    \
    static bool _done[XB_MAX_DEVICES] = {false};                                                          \
    int _dev = 0;                                                                                         \
    XB_CUDA(cudaGetDevice(&_dev));                                                                        \
    if (_dev < 0 || _dev >= XB_MAX_DEVICES || !_done[_dev]) {                                             \
      XB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes)));   \
      if (_dev >= 0 && _dev < XB_MAX_DEVICES) _done[_dev] = true;                                         \
    }                                                                                                     \
  } while (0)

// a stage that begins at the MLP block of layer 0 does not hold the embedding; one that stops after the
// attention block of the last layer does not hold the final norm / LM head
static inline bool shard_is_first(const xb_model_desc& d) { return d.start_layer == 0 && !d.first_from_mlp; }
static inline bool shard_is_last(const xb_model_desc& d) { return d.end_layer == d.n_layers - 1 && !d.last_to_attn; }

// Launch one kernel of the forward: programmatic dependent launch (the kernel calls
// griddepcontrol.wait before touching upstream data) and an optional cluster dimension.
template <typename... KArgs, typename... Args>
static int launch_kc(xb_engine* e, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, dim3 cluster,
                     Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (e->use_pdl) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  if (cluster.x * cluster.y * cluster.z > 1) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = cluster.x;
    attr[na].val.clusterDim.y = cluster.y;
    attr[na].val.clusterDim.z = cluster.z;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  XB_CUDA(cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...));
  e->c.launches++;
  return 0;
}

template <typename... KArgs, typename... Args>
static int launch_k(xb_engine* e, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, int cluster_x,
                    Args... args) {
  return launch_kc(e, kern, grid, block, smem, st, dim3(cluster_x > 1 ? cluster_x : 1, 1, 1), args...);
}

static int dmalloc(void** p, size_t bytes) {
  XB_CUDA(cudaMalloc(p, bytes ? bytes : 16));
  return 0;
}

// tile-major weight storage: rows padded to 128, K padded to 64, zero filled
static size_t tiled_elems(size_t rows, size_t K) { return ((rows + 127) / 128) * ((K + 63) / 64) * 8192; }
static int alloc_tiled(bf16** p, size_t rows, size_t K) {
  XB_TRY(dmalloc((void**)p, tiled_elems(rows, K) * 2));
  XB_CUDA(cudaMemset(*p, 0, tiled_elems(rows, K) * 2));
  return 0;
}
// copy `rows` x K row-major bf16 (host or device memory) into rows [row0, row0+rows) of a tiled matrix
static int store_tiled(bf16* dst, const void* src, size_t row0, size_t rows, size_t K, int interleave = 0, int interleave_off = 0) {
  XB_CHECK(K % 8 == 0, "K=%zu must be a multiple of 8", K);
  cudaPointerAttributes pa;
  const bool on_device = cudaPointerGetAttributes(&pa, src) == cudaSuccess && pa.type == cudaMemoryTypeDevice;
  cudaGetLastError();
  const bf16* dsrc = (const bf16*)src;
  void* staging = nullptr;
  if (!on_device) {
    XB_TRY(dmalloc(&staging, rows * K * 2));
    XB_CUDA(cudaMemcpy(staging, src, rows * K * 2, cudaMemcpyHostToDevice));
    dsrc = (const bf16*)staging;
  }
  const long long n = (long long)rows * (K / 8);
  xb_tile_weight_kernel<<<(unsigned)((n + 255) / 256), 256>>>(dsrc, dst, (long long)row0, (long long)rows, (int)K,
                                                              (int)((K + 63) / 64), interleave, interleave_off);
  XB_CUDA(cudaGetLastError());
  XB_CUDA(cudaDeviceSynchronize());
  if (staging) cudaFree(staging);
  return 0;
}

extern "C" const char* xb_last_error(void) { return g_xb_err.c_str(); }
extern "C" int xb_abi_version(void) { return XB_ABI_VERSION; }

extern "C" int xb_engine_create(const xb_model_desc* desc, int device, xb_engine** out) {
  XB_CHECK(desc != nullptr && out != nullptr, "null argument");
  int ndev = 0;
  cudaError_t ce = cudaGetDeviceCount(&ndev);
  XB_CHECK(ce == cudaSuccess && ndev > 0, "no CUDA device available (%s): libxb200 has no CPU fallback",
           cudaGetErrorString(ce));
  XB_CHECK(device >= 0 && device < ndev, "device %d out of range (%d devices)", device, ndev);
  XB_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  XB_CUDA(cudaGetDeviceProperties(&prop, device));
  XB_CHECK(prop.major == 10, "libxb200 is built for sm_100a only; device %d is sm_%d%d", device, prop.major, prop.minor);
  XB_CHECK(desc->struct_size == sizeof(xb_model_desc),
           "xb_model_desc.struct_size is %u but this library (ABI v%d) expects %zu: the caller was built against another "
           "xb200.h", desc->struct_size, XB_ABI_VERSION, sizeof(xb_model_desc));
  const xb_model_desc& d = *desc;
  XB_CHECK(d.start_layer >= 0 && d.end_layer >= d.start_layer && d.end_layer < d.n_layers, "bad layer range [%d,%d] of %d",
           d.start_layer, d.end_layer, d.n_layers);
  XB_CHECK(d.head_dim == 64 || d.head_dim == 128, "head_dim %d not supported (64 or 128)", d.head_dim);
  XB_CHECK(!(d.first_from_mlp && d.last_to_attn && d.start_layer == d.end_layer), "a stage holding neither block of its only layer is empty");
  XB_CHECK(d.n_heads % d.n_kv_heads == 0 && d.n_heads / d.n_kv_heads <= 8, "n_heads %d must be 1..8 x n_kv_heads %d",
           d.n_heads, d.n_kv_heads);
  XB_CHECK(d.dim % 8 == 0 && d.inter_dim % 16 == 0, "dim must be a multiple of 8 and inter_dim of 16");
  XB_CHECK(d.max_batch_tokens > 0 && d.max_seqs > 0 && d.max_positions > 0, "max_batch_tokens/max_seqs/max_positions must be > 0");
  xb_engine* e = new xb_engine();
  e->d = d;
  e->device = device;
  e->num_sms = prop.multiProcessorCount;
  {
    // One SM is left out of the decode Linears' grids on EVERY engine, not only where a communication kernel is
    // resident: the stream-K decomposition (and with it the fp32 summation order) follows the block count, and the
    // stages of a ring must produce the bits a single-GPU engine produces.
    const char* rv = getenv("XB_RESERVE_SMS");
    e->reserved_sms = rv != nullptr ? atoi(rv) : 1;
    if (e->reserved_sms < 0 || e->reserved_sms >= e->num_sms) e->reserved_sms = 1;
  }
  e->nl = d.end_layer - d.start_layer + 1;
  e->NQ = d.n_heads * d.head_dim;
  e->NKV = d.n_kv_heads * d.head_dim;
  e->NQKV = e->NQ + 2 * e->NKV;
  e->L.resize(e->nl);
  const size_t D = d.dim, I = d.inter_dim;
  uint64_t wbytes = 0;
  for (int l = 0; l < e->nl; ++l) {
    LayerW& w = e->L[l];
    // a boundary layer of a half-layer stage holds only the block that runs here (436 MB per cut at 8B)
    const bool has_attn = !(l == 0 && d.first_from_mlp), has_mlp = !(l + 1 == e->nl && d.last_to_attn);
    if (has_attn) {
      XB_TRY(alloc_tiled(&w.wqkv, e->NQKV, D));
      XB_TRY(alloc_tiled(&w.wo, D, e->NQ));
      XB_TRY(dmalloc((void**)&w.sa_norm, D * 2));
      if (d.qkv_bias) {
        XB_TRY(dmalloc((void**)&w.bqkv, (size_t)e->NQKV * 2));
        XB_CUDA(cudaMemset(w.bqkv, 0, (size_t)e->NQKV * 2));
      }
      wbytes += ((size_t)e->NQKV * D + D * e->NQ) * 2;
    }
    if (has_mlp) {
      XB_TRY(alloc_tiled(&w.wgu, 2 * I, D));
      XB_TRY(alloc_tiled(&w.wdown, D, I));
      XB_TRY(dmalloc((void**)&w.mlp_norm, D * 2));
      wbytes += (size_t)3 * I * D * 2;
    }
  }
  const bool first = shard_is_first(d), last = shard_is_last(d);
  // the embedding table stays row-major (row gather); the LM head is tile-major like every
  // other matrix, so a tied head on a last shard is a second, tiled copy of the table
  if (first) XB_TRY(dmalloc((void**)&e->embed, (size_t)d.vocab * D * 2));
  if (last) {
    XB_TRY(dmalloc((void**)&e->final_norm, D * 2));
    XB_TRY(alloc_tiled(&e->lm_head, d.vocab, D));
    wbytes += (size_t)d.vocab * D * 2;
  }
  XB_TRY(dmalloc((void**)&e->rope, (size_t)d.max_positions * (d.head_dim / 2) * 2 * sizeof(float)));
  e->c.weight_bytes = wbytes;
  e->c.kv_bytes_per_pos = (uint64_t)2 * (e->nl - (d.first_from_mlp ? 1 : 0)) * e->NKV * 2;   // layers whose attention block runs here
  // workspaces
  const size_t T = d.max_batch_tokens;
  XB_TRY(dmalloc((void**)&e->h, T * D * 2));
  XB_TRY(dmalloc((void**)&e->hn, T * D * 2));
  XB_TRY(dmalloc((void**)&e->q, T * e->NQ * 2));
  XB_TRY(dmalloc((void**)&e->attn, T * e->NQ * 2));
  XB_TRY(dmalloc((void**)&e->act, T * I * 2));
  XB_TRY(dmalloc((void**)&e->last_rows, (size_t)d.max_seqs * D * 2));
  size_t maxN = e->NQKV;
  if (2 * I > maxN) maxN = 2 * I;
  if (D > maxN) maxN = D;
  e->gemm_out_bytes = T * maxN * 2;
  if (e->gemm_out_bytes < ((size_t)64 << 20)) e->gemm_out_bytes = (size_t)64 << 20;
  XB_TRY(dmalloc(&e->gemm_out, e->gemm_out_bytes));
  // kv splits > 1 are only used while T * n_kv_heads < 2 * num_sms (see xb_forward)
  const size_t Tsplit = T < (size_t)2 * e->num_sms ? T : (size_t)2 * e->num_sms;
  XB_TRY(dmalloc((void**)&e->o_part, Tsplit * d.n_heads * e->max_attn_splits * d.head_dim * sizeof(float)));
  XB_TRY(dmalloc((void**)&e->ml_part, Tsplit * d.n_heads * e->max_attn_splits * 2 * sizeof(float)));
  e->tile_tokens = 128 / (d.n_heads / d.n_kv_heads);
  e->max_tiles = (int)(T / e->tile_tokens) + d.max_seqs + 1;
  e->meta_ints = 2 * T + 2 * (size_t)d.max_seqs + 2 * (size_t)e->max_tiles + 4;
  XB_TRY(dmalloc((void**)&e->d_meta, e->meta_ints * 4));
  e->d_tok_seq = e->d_meta;
  e->d_tok_pos = e->d_meta + T;
  e->d_seq_slot = e->d_meta + 2 * T;
  e->d_seq_last = e->d_meta + 2 * T + d.max_seqs;
  e->d_tile_start = e->d_meta + 2 * T + 2 * d.max_seqs;
  e->d_tile_len = e->d_tile_start + e->max_tiles;
  {
    XB_TRY(dmalloc((void**)&e->d_dec, (4 + 4 * XB_DEC_ROWS) * 4));
    int ident[4 + 4 * XB_DEC_ROWS] = {0};
    for (int i = 0; i < XB_DEC_ROWS; ++i) ident[4 + 2 * XB_DEC_ROWS + i] = ident[4 + 3 * XB_DEC_ROWS + i] = i;
    XB_CUDA(cudaMemcpy(e->d_dec, ident, sizeof(ident), cudaMemcpyHostToDevice));
  }
  e->d_ctl = reinterpret_cast<HandoffCtl*>(e->d_dec);   // the peer hand-off carries decode steps only
  for (int i = 0; i < xb_engine::META_RING; ++i) {
    XB_CUDA(cudaMallocHost((void**)&e->h_meta_ring[i], (e->meta_ints + 4 + 2 * XB_DEC_ROWS) * 4));
    XB_CUDA(cudaEventCreateWithFlags(&e->meta_ev[i], cudaEventDisableTiming));
  }
  XB_TRY(dmalloc((void**)&e->in_tokens, T * 4));
  if (last) XB_TRY(dmalloc((void**)&e->logits, (size_t)d.max_seqs * d.vocab * sizeof(float)));
  {
    const char* g = getenv("XB_GRAPHS");
    if (g != nullptr && g[0] == '0') e->use_graphs = 0;
    g = getenv("XB_PDL");
    if (g != nullptr && g[0] == '0') e->use_pdl = 0;
    g = getenv("XB_GEMM_TS");
    if (g != nullptr && g[0] == '1') {
      e->ts_cap = 1024;
      e->use_graphs = 0;
      XB_TRY(dmalloc((void**)&e->ts_buf, (size_t)e->ts_cap * e->ts_blocks * XB_TS_SLOTS * 8));
      XB_CUDA(cudaMemset(e->ts_buf, 0, (size_t)e->ts_cap * e->ts_blocks * XB_TS_SLOTS * 8));
    }
    g = getenv("XB_FUSE_QKV");
    if (g != nullptr) e->fuse_qkv = g[0] != '0';
    g = getenv("XB_ATTN_CLUSTER");
    if (g != nullptr) e->attn_cluster = atoi(g);
  }
  *out = e;
  return 0;
}

extern "C" int xb_engine_destroy(xb_engine* e) {
  if (e == nullptr) return 0;
  if (e->ring != nullptr) {
    e->ring->destroy();
    delete e->ring;
    e->ring = nullptr;
  }
  cudaSetDevice(e->device);
  cudaDeviceSynchronize();
  for (LayerW& w : e->L) {
    cudaFree(w.wqkv); cudaFree(w.bqkv); cudaFree(w.wo); cudaFree(w.wgu); cudaFree(w.wdown);
    cudaFree(w.sa_norm); cudaFree(w.mlp_norm); cudaFree(w.k_pool); cudaFree(w.v_pool);
  }
  cudaFree(e->embed); cudaFree(e->final_norm); cudaFree(e->lm_head); cudaFree(e->rope);
  cudaFree(e->d_page_table); cudaFree(e->d_meta); cudaFree(e->d_dec); cudaFree(e->in_tokens); cudaFree(e->logits);
  if (e->peer_landing && !e->peer_is_local) cudaIpcCloseMemHandle(e->peer_landing);
  if (e->peer_flags && !e->peer_is_local) cudaIpcCloseMemHandle(e->peer_flags);
  cudaFree(e->landing); cudaFree(e->flags); cudaFree(e->done_cnt);
  if (e->ho_err) cudaFreeHost(e->ho_err);
  for (int i = 0; i < xb_engine::META_RING; ++i) {
    if (e->h_meta_ring[i]) cudaFreeHost(e->h_meta_ring[i]);
    if (e->meta_ev[i]) cudaEventDestroy(e->meta_ev[i]);
  }
  for (auto& kv : e->graphs)
    if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
  cudaFree(e->h); cudaFree(e->hn); cudaFree(e->q); cudaFree(e->attn); cudaFree(e->act);
  cudaFree(e->ts_buf);
  cudaFree(e->last_rows); cudaFree(e->gemm_out); cudaFree(e->o_part); cudaFree(e->ml_part);
  delete e;
  return 0;
}

// Where a checkpoint tensor goes: a row range of a tile-major matrix, or a plain bf16 vector/table.
struct WTarget {
  bf16* tdst = nullptr;         // tiled matrix target
  size_t row0 = 0, rows = 0, K = 0;
  int il = 0, il_off = 0;       // gate/up rows are interleaved in 16-feature blocks
  void* dst = nullptr;          // plain copy target (vectors, embedding, RoPE table)
  size_t want = 0;              // bytes of a plain target
  bool embed = false;           // embedding: row-major copy and/or the tied head's tiled copy
};

static int resolve_target(xb_engine* e, int layer, int kind, WTarget* t) {
  const xb_model_desc& d = e->d;
  const size_t D = d.dim, I = d.inter_dim;
  if (layer >= 0) {
    XB_CHECK(layer >= d.start_layer && layer <= d.end_layer, "layer %d outside this shard [%d,%d]", layer, d.start_layer,
             d.end_layer);
    LayerW& w = e->L[layer - d.start_layer];
    {
      const bool attn_kind = kind == XB_W_Q || kind == XB_W_K || kind == XB_W_V || kind == XB_W_O || kind == XB_W_SA_NORM ||
                             kind == XB_W_Q_BIAS || kind == XB_W_K_BIAS || kind == XB_W_V_BIAS;
      XB_CHECK(!(attn_kind && w.wqkv == nullptr), "layer %d: the attention block of this layer runs on the previous stage", layer);
      XB_CHECK(!(!attn_kind && kind <= XB_W_MLP_NORM && w.wgu == nullptr), "layer %d: the MLP block of this layer runs on the next stage", layer);
    }
    switch (kind) {
      case XB_W_Q: t->tdst = w.wqkv; t->row0 = 0; t->rows = e->NQ; t->K = D; break;
      case XB_W_K: t->tdst = w.wqkv; t->row0 = e->NQ; t->rows = e->NKV; t->K = D; break;
      case XB_W_V: t->tdst = w.wqkv; t->row0 = e->NQ + e->NKV; t->rows = e->NKV; t->K = D; break;
      case XB_W_O: t->tdst = w.wo; t->rows = D; t->K = e->NQ; break;
      case XB_W_GATE: t->tdst = w.wgu; t->rows = I; t->K = D; t->il = 1; t->il_off = 0; break;
      case XB_W_UP: t->tdst = w.wgu; t->rows = I; t->K = D; t->il = 1; t->il_off = 16; break;
      case XB_W_DOWN: t->tdst = w.wdown; t->rows = D; t->K = I; break;
      case XB_W_SA_NORM: t->dst = w.sa_norm; t->want = D * 2; break;
      case XB_W_MLP_NORM: t->dst = w.mlp_norm; t->want = D * 2; break;
      case XB_W_Q_BIAS: XB_CHECK(w.bqkv, "model has no qkv bias"); t->dst = w.bqkv; t->want = (size_t)e->NQ * 2; break;
      case XB_W_K_BIAS: XB_CHECK(w.bqkv, "model has no qkv bias"); t->dst = w.bqkv + e->NQ; t->want = (size_t)e->NKV * 2; break;
      case XB_W_V_BIAS: XB_CHECK(w.bqkv, "model has no qkv bias"); t->dst = w.bqkv + e->NQ + e->NKV; t->want = (size_t)e->NKV * 2; break;
      default: return xb_fail(1, "weight kind %d is not a per-layer kind", kind);
    }
  } else {
    switch (kind) {
      case XB_W_EMBED:
        XB_CHECK(e->embed || (e->lm_head && d.tied_head), "this shard holds no embedding");
        t->embed = true; t->rows = d.vocab; t->K = D; t->want = (size_t)d.vocab * D * 2;
        break;
      case XB_W_FINAL_NORM: XB_CHECK(e->final_norm, "this shard holds no final norm"); t->dst = e->final_norm; t->want = D * 2; break;
      case XB_W_LM_HEAD:
        XB_CHECK(e->lm_head && !d.tied_head, "this shard holds no separate LM head");
        t->tdst = e->lm_head; t->rows = d.vocab; t->K = D;
        break;
      case XB_W_ROPE_TABLE: t->dst = e->rope; t->want = (size_t)d.max_positions * (d.head_dim / 2) * 2 * sizeof(float); break;
      default: return xb_fail(1, "weight kind %d is not a global kind", kind);
    }
  }
  if (t->tdst != nullptr) t->want = t->rows * t->K * 2;
  return 0;
}

extern "C" int xb_load_weight(xb_engine* e, int layer, int kind, const void* src, size_t nbytes) {
  XB_CHECK(e != nullptr && src != nullptr, "null argument");
  XB_CUDA(cudaSetDevice(e->device));
  WTarget t;
  XB_TRY(resolve_target(e, layer, kind, &t));
  XB_CHECK(nbytes == t.want, "weight kind %d layer %d: got %zu bytes, expected %zu", kind, layer, nbytes, t.want);
  if (t.embed) {
    if (e->embed) XB_CUDA(cudaMemcpy(e->embed, src, nbytes, cudaMemcpyDefault));
    if (e->lm_head && e->d.tied_head) XB_TRY(store_tiled(e->lm_head, src, 0, e->d.vocab, e->d.dim));
    e->plans.clear();
  } else if (t.tdst != nullptr) {
    XB_TRY(store_tiled(t.tdst, src, t.row0, t.rows, t.K, t.il, t.il_off));
  } else {
    XB_CUDA(cudaMemcpy(t.dst, src, nbytes, cudaMemcpyDefault));
  }
  e->loaded[(long long)(layer + 1) * 64 + kind] = 1;
  return 0;
}

// HF tensor name -> (layer, kind) of this ABI; false for names the path does not use (rotary inv_freq ...)
static bool hf_name_to_kind(const std::string& name, int* layer, int* kind) {
  static const struct { const char* suffix; int kind; } per_layer[] = {
    {"self_attn.q_proj.weight", XB_W_Q}, {"self_attn.k_proj.weight", XB_W_K}, {"self_attn.v_proj.weight", XB_W_V},
    {"self_attn.o_proj.weight", XB_W_O}, {"mlp.gate_proj.weight", XB_W_GATE}, {"mlp.up_proj.weight", XB_W_UP},
    {"mlp.down_proj.weight", XB_W_DOWN}, {"input_layernorm.weight", XB_W_SA_NORM},
    {"post_attention_layernorm.weight", XB_W_MLP_NORM}, {"self_attn.q_proj.bias", XB_W_Q_BIAS},
    {"self_attn.k_proj.bias", XB_W_K_BIAS}, {"self_attn.v_proj.bias", XB_W_V_BIAS}};
  const std::string pre = "model.layers.";
  if (name.compare(0, pre.size(), pre) == 0) {
    size_t i = pre.size();
    int l = 0;
    if (i >= name.size() || name[i] < '0' || name[i] > '9') return false;
    while (i < name.size() && name[i] >= '0' && name[i] <= '9') l = l * 10 + (name[i++] - '0');
    if (i >= name.size() || name[i] != '.') return false;
    const std::string rest = name.substr(i + 1);
    for (const auto& pl : per_layer)
      if (rest == pl.suffix) { *layer = l; *kind = pl.kind; return true; }
    return false;
  }
  *layer = -1;
  if (name == "model.embed_tokens.weight") { *kind = XB_W_EMBED; return true; }
  if (name == "model.norm.weight") { *kind = XB_W_FINAL_NORM; return true; }
  if (name == "lm_head.weight") { *kind = XB_W_LM_HEAD; return true; }
  return false;
}

static bool shard_wants(const xb_engine* e, int layer, int kind) {
  const xb_model_desc& d = e->d;
  if (layer >= 0) {
    if (layer < d.start_layer || layer > d.end_layer) return false;
    const bool attn_kind = kind == XB_W_Q || kind == XB_W_K || kind == XB_W_V || kind == XB_W_O || kind == XB_W_SA_NORM ||
                           kind == XB_W_Q_BIAS || kind == XB_W_K_BIAS || kind == XB_W_V_BIAS;
    if (attn_kind && layer == d.start_layer && d.first_from_mlp) return false;   // that block lives on the previous stage
    if (!attn_kind && layer == d.end_layer && d.last_to_attn) return false;      // ... on the next stage
    return (kind != XB_W_Q_BIAS && kind != XB_W_K_BIAS && kind != XB_W_V_BIAS) || d.qkv_bias;
  }
  const bool first = shard_is_first(d), last = shard_is_last(d);
  if (kind == XB_W_EMBED) return first || (last && d.tied_head);
  if (kind == XB_W_FINAL_NORM) return last;
  if (kind == XB_W_LM_HEAD) return last && !d.tied_head;
  return false;
}

// One .safetensors file -> every tensor of this shard found in it (see xb_loader.cuh).
extern "C" int xb_load_safetensors(xb_engine* e, const char* path, int permute_qk, int* n_tensors, uint64_t* n_bytes) {
  XB_CHECK(e != nullptr && path != nullptr, "null argument");
  XB_CUDA(cudaSetDevice(e->device));
  StFile f;
  XB_TRY(f.open_map(path));
  uint64_t hlen = 0;
  memcpy(&hlen, f.base, 8);
  XB_CHECK(hlen > 0 && hlen <= f.size - 8, "%s: bad safetensors header length %llu", path, (unsigned long long)hlen);
  std::vector<StEntry> entries;
  XB_CHECK(st_parse_header((const char*)f.base + 8, (size_t)hlen, &entries), "%s: cannot parse the safetensors header", path);
  const uint8_t* data = f.base + 8 + hlen;
  const size_t data_size = f.size - 8 - (size_t)hlen;
  int rc = 0;
  if (e->ring == nullptr) {
    e->ring = new StRing();
    rc = e->ring->init((size_t)32 << 20);
  }
  StRing& ring = *e->ring;
  int count = 0;
  uint64_t total = 0;
  const xb_model_desc& d = e->d;
  for (size_t ei = 0; rc == 0 && ei < entries.size(); ++ei) {
    const StEntry& en = entries[ei];
    int layer = 0, kind = 0;
    if (!hf_name_to_kind(en.name, &layer, &kind) || !shard_wants(e, layer, kind)) continue;
    const int dt = st_dtype_code(en.dtype);
    if (dt < 0) { rc = xb_fail(1, "%s: tensor %s has unsupported dtype %s (BF16, F16, F32)", path, en.name.c_str(), en.dtype.c_str()); break; }
    WTarget t;
    if ((rc = resolve_target(e, layer, kind, &t)) != 0) break;
    const size_t es = st_dtype_size(dt);
    long long numel = 1;
    for (long long x : en.shape) numel *= x;
    const size_t want_elems = t.want / 2;
    if ((size_t)numel != want_elems || en.end - en.begin != (size_t)numel * es || en.end > data_size) {
      rc = xb_fail(1, "%s: tensor %s has %lld elements / %zu bytes, this model wants %zu elements", path, en.name.c_str(), numel,
                   en.end - en.begin, want_elems);
      break;
    }
    if ((t.tdst != nullptr || t.embed) && (en.shape.size() != 2 || (size_t)en.shape[0] != t.rows || (size_t)en.shape[1] != t.K)) {
      rc = xb_fail(1, "%s: tensor %s is not [%zu, %zu]", path, en.name.c_str(), t.rows, t.K);
      break;
    }
    const uint8_t* src = data + en.begin;
    if (t.tdst != nullptr || t.embed) {
      // matrix: row chunks through the ring; Llama q/k rows are permuted on the way (llm_utils.py:213-231)
      const int perm = (permute_qk && (kind == XB_W_Q || kind == XB_W_K)) ? d.head_dim : 0;
      const size_t row_bytes = t.K * es;
      size_t rows_per = ring.chunk / row_bytes;
      if (rows_per > 128) rows_per &= ~(size_t)127;
      if (rows_per == 0) { rc = xb_fail(1, "row of %zu bytes exceeds the loader chunk", row_bytes); break; }
      if (t.K % 8 != 0) { rc = xb_fail(1, "K=%zu must be a multiple of 8", t.K); break; }
      for (size_t j0 = 0; rc == 0 && j0 < t.rows; j0 += rows_per) {
        const size_t nr = j0 + rows_per <= t.rows ? rows_per : t.rows - j0;
        void* dsrc = nullptr;
        int bi = 0;
        if ((rc = ring.stage(src + j0 * row_bytes, nr * row_bytes, &dsrc, &bi)) != 0) break;
        const long long n = (long long)nr * (long long)(t.K / 8);
        const unsigned blocks = (unsigned)((n + 255) / 256);
        if (t.embed) {
          if (e->embed) xb_load_copy_kernel<<<blocks, 256, 0, ring.st>>>(dsrc, dt, n, e->embed + j0 * t.K);
          if (e->lm_head && d.tied_head)
            xb_load_rows_tiled_kernel<<<blocks, 256, 0, ring.st>>>(dsrc, dt, (long long)j0, (long long)nr, (int)t.K, e->lm_head,
                                                                   (int)((t.K + 63) / 64), 0, 0, 0, 0);
        } else {
          xb_load_rows_tiled_kernel<<<blocks, 256, 0, ring.st>>>(dsrc, dt, (long long)j0, (long long)nr, (int)t.K, t.tdst,
                                                                 (int)((t.K + 63) / 64), (long long)t.row0, t.il, t.il_off, perm);
        }
        rc = ring.done(bi);
      }
    } else {
      // vector (norm scale, bias): one small chunk
      void* dsrc = nullptr;
      int bi = 0;
      if ((size_t)numel % 8 != 0) { rc = xb_fail(1, "%s: length of %s must be a multiple of 8", path, en.name.c_str()); break; }
      if ((rc = ring.stage(src, (size_t)numel * es, &dsrc, &bi)) != 0) break;
      xb_load_copy_kernel<<<(unsigned)((numel / 8 + 255) / 256), 256, 0, ring.st>>>(dsrc, dt, numel / 8, (bf16*)t.dst);
      rc = ring.done(bi);
    }
    if (rc != 0) break;
    e->loaded[(long long)(layer + 1) * 64 + kind] = 1;
    ++count;
    total += en.end - en.begin;
  }
  if (rc == 0) {
    cudaError_t ce = cudaStreamSynchronize(ring.st);
    if (ce != cudaSuccess) rc = xb_fail(2, "%s: loader stream failed: %s", path, cudaGetErrorString(ce));
  }
  if (rc != 0) {
    ring.destroy();
    delete e->ring;
    e->ring = nullptr;
    return rc;
  }
  e->plans.clear();
  if (n_tensors) *n_tensors = count;
  if (n_bytes) *n_bytes = total;
  return 0;
}

// Header of a .safetensors file as text, one tensor per line: "name dtype begin end dim0 dim1 ...".
// Pure host code (no CUDA call): what xb_load_safetensors would see, for tools and CPU-side tests.
extern "C" int xb_safetensors_list(const char* path, char* buf, size_t buflen, int* n_tensors) {
  XB_CHECK(path != nullptr && n_tensors != nullptr, "null argument");
  StFile f;
  XB_TRY(f.open_map(path));
  uint64_t hlen = 0;
  memcpy(&hlen, f.base, 8);
  XB_CHECK(hlen > 0 && hlen <= f.size - 8, "%s: bad safetensors header length %llu", path, (unsigned long long)hlen);
  std::vector<StEntry> entries;
  XB_CHECK(st_parse_header((const char*)f.base + 8, (size_t)hlen, &entries), "%s: cannot parse the safetensors header", path);
  std::string out;
  for (const StEntry& en : entries) {
    XB_CHECK(en.end <= f.size - 8 - (size_t)hlen, "%s: tensor %s reaches past the end of the file", path, en.name.c_str());
    out += en.name + " " + en.dtype + " " + std::to_string(en.begin) + " " + std::to_string(en.end);
    for (long long x : en.shape) out += " " + std::to_string(x);
    out += "\n";
  }
  *n_tensors = (int)entries.size();
  if (buf != nullptr && buflen > 0) {
    XB_CHECK(out.size() < buflen, "listing needs %zu bytes, buffer has %zu", out.size() + 1, buflen);
    memcpy(buf, out.c_str(), out.size() + 1);
  }
  return 0;
}

// Names (ABI kinds) this shard still lacks; 0 in *n_missing = ready.  `buf` receives a readable list.
extern "C" int xb_weights_missing(xb_engine* e, char* buf, size_t buflen, int* n_missing) {
  XB_CHECK(e != nullptr && n_missing != nullptr, "null argument");
  static const char* names[] = {"q_proj", "k_proj", "v_proj", "o_proj", "gate_proj", "up_proj", "down_proj", "input_layernorm",
                                "post_attention_layernorm", "q_proj.bias", "k_proj.bias", "v_proj.bias", "embed_tokens",
                                "final norm", "lm_head", "rope table"};
  std::string out;
  int n = 0;
  auto need = [&](int layer, int kind) {
    if (e->loaded.count((long long)(layer + 1) * 64 + kind)) return;
    ++n;
    if (out.size() < 900) out += (layer >= 0 ? "layer " + std::to_string(layer) + " " : std::string()) + names[kind] + "; ";
  };
  for (int l = e->d.start_layer; l <= e->d.end_layer; ++l)
    for (int k = XB_W_Q; k <= XB_W_V_BIAS; ++k)
      if (shard_wants(e, l, k)) need(l, k);
  for (int k = XB_W_EMBED; k <= XB_W_LM_HEAD; ++k)
    if (shard_wants(e, -1, k)) need(-1, k);
  need(-1, XB_W_ROPE_TABLE);
  *n_missing = n;
  if (n == 0 && e->ring != nullptr) {   // complete: the staging buffers are no longer needed
    e->ring->destroy();
    delete e->ring;
    e->ring = nullptr;
  }
  if (buf != nullptr && buflen > 0) {
    strncpy(buf, out.c_str(), buflen - 1);
    buf[buflen - 1] = 0;
  }
  return 0;
}

// ----------------------------------------------------------------------------- paged KV
extern "C" int xb_kv_reserve(xb_engine* e, int n_slots, int max_tokens_per_slot, int page_tokens) {
  XB_CHECK(e != nullptr, "null engine");
  XB_CHECK(n_slots > 0 && max_tokens_per_slot > 0 && page_tokens > 0, "bad kv_reserve arguments");
  XB_CHECK(page_tokens % 8 == 0 && (page_tokens >= 64 ? page_tokens % 64 == 0 : 64 % page_tokens == 0),
           "page_tokens must be 8, 16, 32 or a multiple of 64 (got %d)", page_tokens);
  XB_CUDA(cudaSetDevice(e->device));
  XB_CUDA(cudaDeviceSynchronize());
  for (LayerW& w : e->L) {
    cudaFree(w.k_pool); cudaFree(w.v_pool);
    w.k_pool = w.v_pool = nullptr;
  }
  cudaFree(e->d_page_table);
  e->d_page_table = nullptr;
  // captured decode graphs and launch plans hold the old pool / page-table addresses
  for (auto& kv : e->graphs)
    if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
  e->graphs.clear();
  e->plans.clear();
  e->n_slots = n_slots;
  e->page_tokens = page_tokens;
  e->max_pages = (max_tokens_per_slot + page_tokens - 1) / page_tokens;
  e->n_pages = n_slots * e->max_pages;
  const size_t pool_bytes = (size_t)e->n_pages * e->NKV * page_tokens * 2;
  const long long pool_rows = (long long)e->n_pages * e->d.n_kv_heads * page_tokens;
  for (LayerW& w : e->L) {
    if (w.wqkv == nullptr) continue;   // boundary layer whose attention block runs on the previous stage: no cache here
    XB_TRY(dmalloc((void**)&w.k_pool, pool_bytes));
    XB_TRY(dmalloc((void**)&w.v_pool, pool_bytes));
    // zeroed ONCE here (never per prompt): rows of a page beyond a sequence's length are
    // read by the attention tiles and masked; they must hold finite values
    XB_CUDA(cudaMemset(w.k_pool, 0, pool_bytes));
    XB_CUDA(cudaMemset(w.v_pool, 0, pool_bytes));
    const int box_rows = page_tokens < 64 ? page_tokens : 64;
    XB_TRY(xb_make_tmap(&w.tmK, w.k_pool, pool_rows, e->d.head_dim, e->d.head_dim, box_rows));
    XB_TRY(xb_make_tmap(&w.tmV, w.v_pool, pool_rows, e->d.head_dim, e->d.head_dim, box_rows));
  }
  e->free_pages.clear();
  for (int p = e->n_pages - 1; p >= 0; --p) e->free_pages.push_back(p);
  e->slot_pages.assign(n_slots, std::vector<int>());
  e->slot_len.assign(n_slots, 0);
  e->h_page_table.assign((size_t)n_slots * e->max_pages, 0);
  XB_TRY(dmalloc((void**)&e->d_page_table, e->h_page_table.size() * 4));
  XB_CUDA(cudaMemset(e->d_page_table, 0, e->h_page_table.size() * 4));
  e->c.kv_pages_total = e->n_pages;
  e->c.kv_pages_free = e->free_pages.size();
  return 0;
}

extern "C" int xb_kv_release_slot(xb_engine* e, int slot) {
  XB_CHECK(e != nullptr && slot >= 0 && slot < e->n_slots, "bad slot %d", slot);
  for (int p : e->slot_pages[slot]) e->free_pages.push_back(p);
  e->slot_pages[slot].clear();
  e->slot_len[slot] = 0;
  e->c.kv_pages_free = e->free_pages.size();
  return 0;
}

extern "C" int xb_kv_slot_len(xb_engine* e, int slot, int* len) {
  XB_CHECK(e != nullptr && len != nullptr && slot >= 0 && slot < e->n_slots, "bad slot %d", slot);
  *len = e->slot_len[slot];
  return 0;
}

// ----------------------------------------------------------------------------- forward
// every Linear of the engine goes through here (timeline instrumentation hooks in when enabled)
static int run_gemm(xb_engine* e, const GemmPlan& gp, cudaStream_t st) {
  if (e->ts_buf != nullptr && e->ts_next < e->ts_cap && gp.grid <= e->ts_blocks) {
    GemmPlan p = gp;
    p.args.ts = e->ts_buf + (size_t)e->ts_next * e->ts_blocks * XB_TS_SLOTS;
    ++e->ts_next;
    return xb_gemm_run(p, st, e->use_pdl);
  }
  return xb_gemm_run(gp, st, e->use_pdl);
}

static int get_plan(xb_engine* e, int site, int layer, const bf16* x, long long ldx, const bf16* w, int T, int N, int K,
                    void* out, long long ld_out, const bf16* bias, int round_bf16, const GemmPlan** pp) {
  const long long key = (((long long)site * 4096 + layer) << 32) | (unsigned)T;
  auto it = e->plans.find(key);
  if (it == e->plans.end()) {
    GemmPlan p;
    // (a decode Linear is one block per SM; SMs set aside with xb_reserve_sms stay free for a resident communication kernel)
    const int sms = (T <= 128 && e->reserved_sms > 0 && e->reserved_sms < e->num_sms) ? e->num_sms - e->reserved_sms : e->num_sms;
    XB_TRY(xb_gemm_plan(&p, x, ldx, w, 1, T, N, K, out, ld_out, bias, round_bf16, sms, 0));
    if (p.args.mode == XB_GEMM_WIDE && site == SITE_QKV) {
      // prefill: bias + RoPE + q store + paged-KV append happen in the GEMM epilogue
      const xb_model_desc& d = e->d;
      GemmArgs& g = p.args;
      g.epi = XB_EPI_QKV;
      g.rope = e->rope;
      g.q_out = e->q;
      g.k_pool = e->L[layer].k_pool;
      g.v_pool = e->L[layer].v_pool;
      g.tok_seq = e->d_tok_seq;
      g.tok_pos = e->d_tok_pos;
      g.seq_slot = e->d_seq_slot;
      g.page_table = e->d_page_table;
      g.max_pages = e->max_pages;
      g.page_tokens = e->page_tokens;
      g.n_heads = d.n_heads;
      g.n_kv_heads = d.n_kv_heads;
      g.head_dim = d.head_dim;
      g.rope_kind = d.rope_kind;
    } else if (p.args.mode == XB_GEMM_WIDE && site == SITE_GU) {
      // prefill: SiLU(gate) * up in the GEMM epilogue, straight into the down projection's input
      p.args.epi = XB_EPI_SILU;
      p.args.out = e->act;
      p.args.ld_out = e->d.inter_dim;
    }
    if (p.args.mode == XB_GEMM_SKINNY && out == e->gemm_out) {
      const size_t need = (size_t)p.split.splits * T * N * 4;
      XB_CHECK(need <= e->gemm_out_bytes, "gemm workspace too small: need %zu have %zu", need, e->gemm_out_bytes);
    } else if (p.args.mode == XB_GEMM_WIDE && site != SITE_QKV && site != SITE_GU && out == e->gemm_out) {
      const size_t need = p.args.mode == XB_GEMM_SKINNY ? (size_t)p.split.splits * T * N * 4 : (size_t)T * N * 2;
      XB_CHECK(need <= e->gemm_out_bytes, "gemm workspace too small: need %zu have %zu", need, e->gemm_out_bytes);
    }
    it = e->plans.emplace(key, p).first;
  }
  *pp = &it->second;
  return 0;
}

template <int HD, int G, bool FUSED>
static int launch_attn_decode(xb_engine* e, const LayerW& w, const BatchMeta& m, int T, int n_splits, cudaStream_t st,
                              const QkvFin& fin) {
  const xb_model_desc& d = e->d;
  const float scale_log2 = (1.0f / sqrtf((float)d.head_dim)) * 1.4426950408889634f;
#ifndef XB_ATT_DEC_NST
#define XB_ATT_DEC_NST 3
#endif
  constexpr int NST = XB_ATT_DEC_NST;
  // decode: one block per (token, kv head, kv split); the splits of one (token, kv head) are a cluster
  using Cfg = AttDecCfg<HD, G, NST>;
  auto kern = xb_attn_decode_mma_kernel<HD, G, NST, FUSED>;
  XB_SMEM_ATTR_ONCE(kern, Cfg::SMEM_BYTES);
  const int cl = (e->attn_cluster && n_splits > 1) ? 1 : 0;
  if (cl && n_splits > 8) {
    static bool np_done[XB_MAX_DEVICES] = {false};
    int dev = 0;
    XB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= XB_MAX_DEVICES || !np_done[dev]) {
      XB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
      if (dev >= 0 && dev < XB_MAX_DEVICES) np_done[dev] = true;
    }
  }
  dim3 grid(T, d.n_kv_heads, n_splits);
  XB_TRY(launch_kc(e, kern, grid, dim3(160), Cfg::SMEM_BYTES, st, dim3(1, 1, cl ? n_splits : 1), w.tmK, w.tmV, e->q, m, e->o_part,
                   e->ml_part, e->attn, d.n_heads, d.n_kv_heads, n_splits, scale_log2, fin, cl));
  if (n_splits > 1 && !cl) {
    const int rows = T * d.n_heads;
    XB_TRY(launch_k(e, xb_attn_combine2_kernel<HD>, dim3((rows + 3) / 4), dim3(128), 0, st, 1, e->o_part, e->ml_part, e->attn, n_splits, rows));
  }
  return 0;
}

// fin != nullptr: decode with the q/k/v finalize folded into the attention kernel (see QkvFin)
template <int HD, int G>
static int launch_attn_g(xb_engine* e, const LayerW& w, const BatchMeta& m, int T, int n_splits, int n_tiles,
                         cudaStream_t st, const QkvFin* fin) {
  const xb_model_desc& d = e->d;
  const float scale_log2 = (1.0f / sqrtf((float)d.head_dim)) * 1.4426950408889634f;
  if (n_tiles == 0) {
    if (fin != nullptr) return launch_attn_decode<HD, G, true>(e, w, m, T, n_splits, st, *fin);
    return launch_attn_decode<HD, G, false>(e, w, m, T, n_splits, st, QkvFin{});
  } else {
    // prefill: one block per (query tile, kv head), tcgen05 flash attention
    using Cfg = AttTcCfg<HD>;
    XB_SMEM_ATTR_ONCE((xb_attn_prefill_tc_kernel<HD, G>), Cfg::SMEM_BYTES);
    XB_TRY(launch_k(e, xb_attn_prefill_tc_kernel<HD, G>, dim3(n_tiles, d.n_kv_heads), dim3(192), Cfg::SMEM_BYTES, st, 1, w.tmK,
                    w.tmV, e->q, m, m.tile_start, m.tile_len, e->attn, d.n_heads, d.n_kv_heads, scale_log2));
  }
  return 0;
}

template <int HD>
static int launch_attn(xb_engine* e, const LayerW& w, const BatchMeta& m, int T, int n_splits, int n_tiles, cudaStream_t st,
                       const QkvFin* fin = nullptr) {
  switch (e->d.n_heads / e->d.n_kv_heads) {
    case 1: return launch_attn_g<HD, 1>(e, w, m, T, n_splits, n_tiles, st, fin);
    case 2: return launch_attn_g<HD, 2>(e, w, m, T, n_splits, n_tiles, st, fin);
    case 3: return launch_attn_g<HD, 3>(e, w, m, T, n_splits, n_tiles, st, fin);
    case 4: return launch_attn_g<HD, 4>(e, w, m, T, n_splits, n_tiles, st, fin);
    case 5: return launch_attn_g<HD, 5>(e, w, m, T, n_splits, n_tiles, st, fin);
    case 6: return launch_attn_g<HD, 6>(e, w, m, T, n_splits, n_tiles, st, fin);
    case 7: return launch_attn_g<HD, 7>(e, w, m, T, n_splits, n_tiles, st, fin);
    case 8: return launch_attn_g<HD, 8>(e, w, m, T, n_splits, n_tiles, st, fin);
  }
  return xb_fail(1, "GQA group size %d not supported (1..8)", e->d.n_heads / e->d.n_kv_heads);
}

static int launch_resid_norm(xb_engine* e, const GemmPlan* gp, int T, const bf16* nw, cudaStream_t st) {
  const xb_model_desc& d = e->d;
  const size_t smem = (size_t)d.dim * sizeof(float);
  if (gp->args.mode == XB_GEMM_SKINNY) {
    // decode-sized batches: one 1024-thread block per row
    if (d.dim % 4 == 0 && d.dim <= 1024 * 4 * 2) {
      if (d.dim <= 1024 * 4)
        return launch_k(e, xb_resid_norm_row_kernel<1024, 1>, dim3(T), dim3(1024), 0, st, 1, (const float*)e->gemm_out, gp->split,
                        (long long)T * d.dim, e->h, nw, e->hn, d.dim, d.norm_eps, d.rmsnorm_two_round);
      return launch_k(e, xb_resid_norm_row_kernel<1024, 2>, dim3(T), dim3(1024), 0, st, 1, (const float*)e->gemm_out, gp->split,
                      (long long)T * d.dim, e->h, nw, e->hn, d.dim, d.norm_eps, d.rmsnorm_two_round);
    }
    XB_TRY(launch_k(e, xb_resid_norm_kernel<256, 0>, dim3(T), dim3(256), smem, st, 1, (const float*)e->gemm_out, gp->split, (long long)T * d.dim,
                                                        nullptr, e->h, nw, e->hn, d.dim, d.norm_eps, d.rmsnorm_two_round));
  } else {
    XB_TRY(launch_k(e, xb_resid_norm_kernel<256, 1>, dim3(T), dim3(256), smem, st, 1, nullptr, SplitInfo{1, 0, 1}, 0, (const bf16*)e->gemm_out, e->h, nw, e->hn, d.dim,
                                                        d.norm_eps, d.rmsnorm_two_round));
  }
  return 0;
}

// kv splits of a decode batch.  Measured on B200 (profiles/attn_sweep.py, Llama-3.1-8B geometry,
// ctx 2048): the kernel is fastest when the whole grid is ONE wave of about two blocks per SM --
// 32 seqs x 8 kv heads unsplit (45 us, 5.96 TB/s) beats any split; 8 seqs want 4 splits, 1 seq 16.
// More splits only add partial traffic and combine work.
static int choose_kv_splits(xb_engine* e, int T, int max_ctx) {
  const char* ov = getenv("XB_ATTN_SPLITS");
  const int blocks = T * e->d.n_kv_heads;
  const int tiles = (max_ctx + XB_ATT_TILE - 1) / XB_ATT_TILE;
  int n = (2 * e->num_sms) / blocks;          // largest split count that still fits one wave
  if (n > 16) n = 16;
  if (ov != nullptr && atoi(ov) > 0) n = atoi(ov);
  if (n > tiles) n = tiles;
  // the splits of a (token, kv head) are merged inside one thread-block cluster: 8 blocks portable, 16 with XB_ATTN_CLUSTER=16
  if (e->attn_cluster && n > (e->attn_cluster >= 16 ? 16 : 8)) n = e->attn_cluster >= 16 ? 16 : 8;
  if (n > e->max_attn_splits) n = e->max_attn_splits;
  if (n < 1) n = 1;
  return n;
}

// All kernel launches of one forward over T token rows.  Reads e->in_tokens (first shard) or
// e->h (other shards); leaves the residual stream in e->h and, on a last shard, the logits
// of each sequence's last position in e->logits.  Pure stream work on fixed addresses, so it
// can be captured into a CUDA graph.
// LM head on the last position of each sequence (the final norm is fused into the last layer's
// residual kernel).  host_logits != nullptr: the Linear's epilogue ALSO stores the fp32 logits
// straight into that pinned host buffer -- for the plugin API, which has to hand them to the host
// every step: stores over PCIe (~51 GB/s measured) while the weights stream, instead of a D2H
// copy (~54 GB/s) that can only start when the kernel has finished.
static int run_head(xb_engine* e, int n_seqs, cudaStream_t st, float* host_logits) {
  const xb_model_desc& d = e->d;
  const int D = d.dim;
  const int* seq_last = e->meta_is_dec ? e->d_dec + 4 + 3 * XB_DEC_ROWS : e->d_seq_last;
  XB_TRY(launch_k(e, xb_gather_rows_kernel, dim3(n_seqs), dim3(256), 0, st, 1, e->hn, seq_last, e->last_rows, D));
  const GemmPlan* gp;
  XB_TRY(get_plan(e, SITE_HEAD, 0, e->last_rows, D, e->lm_head, n_seqs, d.vocab, D, e->logits, d.vocab, nullptr, 1, &gp));
  if (host_logits != nullptr) {
    GemmPlan p = *gp;
    p.args.out2 = host_logits;
    XB_TRY(run_gemm(e, p, st));
  } else {
    XB_TRY(run_gemm(e, *gp, st));
  }
  LAUNCH_COUNT(e);
  return 0;
}

static void fill_batch_meta(const xb_engine* e, BatchMeta& m) {
  if (e->meta_is_dec) {
    m.tok_pos = e->d_dec + 4;
    m.seq_slot = e->d_dec + 4 + XB_DEC_ROWS;
    m.tok_seq = e->d_dec + 4 + 2 * XB_DEC_ROWS;
    m.seq_last = e->d_dec + 4 + 3 * XB_DEC_ROWS;
  } else {
    m.tok_seq = e->d_tok_seq;
    m.tok_pos = e->d_tok_pos;
    m.seq_slot = e->d_seq_slot;
    m.seq_last = e->d_seq_last;
  }
}

static int run_body(xb_engine* e, int T, int n_seqs, int n_splits, int n_tiles, void* const* taps, cudaStream_t st,
                    int ho_in = 0, int ho_out = 0, int with_head = 1) {
  const xb_model_desc& d = e->d;
  const bool first = shard_is_first(d), last = shard_is_last(d);
  BatchMeta m;
  fill_batch_meta(e, m);
  m.page_table = e->d_page_table;
  m.tile_start = e->d_tile_start;
  m.tile_len = e->d_tile_len;
  m.max_pages = e->max_pages;
  m.page_tokens = e->page_tokens;
  const int D = d.dim, I = d.inter_dim, hd = d.head_dim;
  const size_t norm_smem = (size_t)D * sizeof(float);
  // half-layer shards: a stage may start at the MLP block of its first layer (the attention block ran
  // on the previous GPU) and/or stop after the attention block of its last layer; the residual stream
  // crosses the boundary exactly as it does between whole layers
  const bf16* norm0 = d.first_from_mlp ? e->L[0].mlp_norm : e->L[0].sa_norm;
  // batches of up to 128 rows run the row kernels after every Linear (launch_resid_norm: T <= 128, D % 4 == 0, D <= 8192)
  const bool row_norm = T <= 128 && D % 4 == 0 && D <= 8192;
  if (first) {
    XB_TRY(launch_k(e, xb_embed_norm_kernel<256>, dim3(T), dim3(256), 0, st, 1, e->in_tokens, e->embed, e->L[0].sa_norm, e->h, e->hn, D, d.vocab, d.norm_eps,
                                                  d.rmsnorm_two_round));
  } else if (ho_in) {
    // stage input arrives in the landing buffer, stored there by the previous GPU's last kernel.  Decode rows take the
    // row kernels' layout: the norm's sum of squares is then accumulated in the order the unsplit model uses (row_ss4)
    if (row_norm && D <= 4096)
      XB_TRY(launch_k(e, xb_recv_norm_row_kernel<1024, 1>, dim3(T), dim3(1024), 0, st, 1, (const bf16*)e->landing,
                      (const uint32_t*)e->flags, (const HandoffCtl*)e->d_ctl, (long long)e->ho_rows * D, e->h, norm0, e->hn, D,
                      d.norm_eps, d.rmsnorm_two_round, e->ho_timeout_cycles, e->ho_err_dev));
    else if (row_norm)
      XB_TRY(launch_k(e, xb_recv_norm_row_kernel<1024, 2>, dim3(T), dim3(1024), 0, st, 1, (const bf16*)e->landing,
                      (const uint32_t*)e->flags, (const HandoffCtl*)e->d_ctl, (long long)e->ho_rows * D, e->h, norm0, e->hn, D,
                      d.norm_eps, d.rmsnorm_two_round, e->ho_timeout_cycles, e->ho_err_dev));
    else
    XB_TRY(launch_k(e, xb_recv_norm_kernel<256>, dim3(T), dim3(256), norm_smem, st, 1, (const bf16*)e->landing,
                    (const uint32_t*)e->flags, (const HandoffCtl*)e->d_ctl, (long long)e->ho_rows * D, e->h, norm0, e->hn, D,
                    d.norm_eps, d.rmsnorm_two_round, e->ho_timeout_cycles, e->ho_err_dev));
  } else if (row_norm) {
    if (D <= 4096)
      XB_TRY(launch_k(e, xb_norm_row_kernel<1024, 1>, dim3(T), dim3(1024), 0, st, 1, (const bf16*)e->h, norm0, e->hn, D, d.norm_eps, d.rmsnorm_two_round));
    else
      XB_TRY(launch_k(e, xb_norm_row_kernel<1024, 2>, dim3(T), dim3(1024), 0, st, 1, (const bf16*)e->h, norm0, e->hn, D, d.norm_eps, d.rmsnorm_two_round));
  } else {
    XB_TRY(launch_k(e, xb_resid_norm_kernel<256, 2>, dim3(T), dim3(256), norm_smem, st, 1, nullptr, SplitInfo{1, 0, 1}, 0, nullptr, e->h, norm0, e->hn, D,
                                                             d.norm_eps, d.rmsnorm_two_round));
  }

  for (int l = 0; l < e->nl; ++l) {
    const LayerW& w = e->L[l];
    const GemmPlan* gp;
    const bool do_attn = !(l == 0 && d.first_from_mlp);
    const bool do_mlp = !(l + 1 == e->nl && d.last_to_attn);
    if (do_attn) {
    // q/k/v projections as one GEMM (general_mha.py:83-97)
    XB_TRY(get_plan(e, SITE_QKV, l, e->hn, D, w.wqkv, T, e->NQKV, D, e->gemm_out, e->NQKV, w.bqkv, 0, &gp));
    XB_TRY(run_gemm(e, *gp, st));
    LAUNCH_COUNT(e);
    QkvFin fin{};
    const QkvFin* finp = nullptr;
    if (gp->args.mode == XB_GEMM_SKINNY && gp->args.epi == XB_EPI_PLAIN) {
      if (e->fuse_qkv && n_tiles == 0) {
        // decode: the attention blocks reduce the planes, apply bias + RoPE and append the k / v row themselves
        fin = QkvFin{(const float*)e->gemm_out, gp->split, (long long)T * e->NQKV, w.bqkv, e->rope, w.k_pool, w.v_pool, d.rope_kind};
        finp = &fin;
      } else {
        const int qkv_items = (d.n_heads + 2 * d.n_kv_heads) * (d.rope_kind == 0 ? hd / 4 : hd / 8);
        XB_TRY(launch_k(e, xb_qkv_finalize_kernel, dim3(T, (qkv_items + 255) / 256), dim3(256), 0, st, 1, (const float*)e->gemm_out, gp->split,
                        (long long)T * e->NQKV, w.bqkv, e->rope, e->q, w.k_pool, w.v_pool, m, d.n_heads, d.n_kv_heads, hd,
                        d.rope_kind));
      }
    }
    if (hd == 128) XB_TRY(launch_attn<128>(e, w, m, T, n_splits, n_tiles, st, finp));
    else XB_TRY(launch_attn<64>(e, w, m, T, n_splits, n_tiles, st, finp));
    // output projection + residual + mlp_norm
    XB_TRY(get_plan(e, SITE_O, l, e->attn, e->NQ, w.wo, T, D, e->NQ, e->gemm_out, D, nullptr, 0, &gp));
    XB_TRY(run_gemm(e, *gp, st));
    LAUNCH_COUNT(e);
    if (do_mlp) XB_TRY(launch_resid_norm(e, gp, T, w.mlp_norm, st));
    }
    if (do_mlp) {
    // gate/up as one GEMM, SiLU*up, down projection (llm_utils.py:513-522)
    XB_TRY(get_plan(e, SITE_GU, l, e->hn, D, w.wgu, T, 2 * I, D, e->gemm_out, 2 * I, nullptr, 0, &gp));
    XB_TRY(run_gemm(e, *gp, st));
    LAUNCH_COUNT(e);
    if (gp->args.mode == XB_GEMM_SKINNY && gp->args.epi == XB_EPI_PLAIN) {
      const long long n = (long long)T * (I / 4);
      XB_TRY(launch_k(e, xb_silu_mul_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, st, 1, (const float*)e->gemm_out,
                      gp->split, (long long)T * 2 * I, e->act, T, I));
    }
    XB_TRY(get_plan(e, SITE_DOWN, l, e->act, I, w.wdown, T, D, I, e->gemm_out, D, nullptr, 0, &gp));
    XB_TRY(run_gemm(e, *gp, st));
    LAUNCH_COUNT(e);
    }
    // `gp` is the block's last Linear (down, or the output projection of a stage that stops after attention)
    const bf16* next_norm = (l + 1 < e->nl) ? e->L[l + 1].sa_norm : (last ? e->final_norm : nullptr);
    if (ho_out && l + 1 == e->nl && gp->args.mode == XB_GEMM_SKINNY) {
      // stage output: reduce + residual, stored straight into the next GPU's landing buffer
      XB_TRY(launch_k(e, xb_resid_send_kernel<256>, dim3(T), dim3(256), 0, st, 1, (const float*)e->gemm_out, gp->split,
                      (long long)T * D, e->h, e->peer_landing, e->peer_flags, e->done_cnt, (const HandoffCtl*)e->d_ctl,
                      (long long)e->ho_rows * D, D));
    } else {
      XB_TRY(launch_resid_norm(e, gp, T, next_norm, st));
    }
    if (taps != nullptr && taps[l] != nullptr)
      XB_CUDA(cudaMemcpyAsync(taps[l], e->h, (size_t)T * D * 2, cudaMemcpyDeviceToDevice, st));
  }

  if (last && with_head) XB_TRY(run_head(e, n_seqs, st, nullptr));
  return 0;
}

static int forward_impl(xb_engine* e, const xb_batch* b, void* out, void* const* taps, void* stream, const int* ctl,
                        float* host_logits = nullptr) {
  XB_CHECK(e != nullptr && b != nullptr, "null argument");
  const int ho_in = ctl != nullptr && ctl[0] >= 0, ho_out = ctl != nullptr && ctl[2] >= 0;
  const xb_model_desc& d = e->d;
  cudaStream_t st = (cudaStream_t)stream;
  XB_CUDA(cudaSetDevice(e->device));
  XB_CHECK(e->n_slots > 0, "xb_kv_reserve must be called before xb_forward");
  XB_CHECK(b->n_seqs > 0 && b->n_seqs <= d.max_seqs, "n_seqs %d out of range (max %d)", b->n_seqs, d.max_seqs);
  const bool first = shard_is_first(d), last = shard_is_last(d);
  XB_CHECK(first ? (b->tokens != nullptr) : (ho_in || b->hidden_in != nullptr),
           first ? "first shard needs token ids" : "non-first shard needs a hidden state");
  XB_CHECK(!last || b->n_seqs <= 128, "LM head path supports up to 128 sequences per call");
  // ---- host-side batch bookkeeping
  int T = 0, max_ctx = 0;
  bool decode_only = true;
  for (int i = 0; i < b->n_seqs; ++i) {
    const int slot = b->slot_ids[i], ql = b->q_lens[i], pos = b->positions[i];
    XB_CHECK(slot >= 0 && slot < e->n_slots, "seq %d: slot %d out of range", i, slot);
    XB_CHECK(ql > 0 && pos >= 0, "seq %d: bad q_len %d / position %d", i, ql, pos);
    XB_CHECK(pos <= e->slot_len[slot], "seq %d: position %d leaves a gap after cached length %d", i, pos, e->slot_len[slot]);
    XB_CHECK(pos + ql <= e->max_pages * e->page_tokens, "seq %d: %d tokens exceed the reserved %d per slot", i, pos + ql,
             e->max_pages * e->page_tokens);
    XB_CHECK(pos + ql <= d.max_positions, "seq %d: position %d beyond the RoPE table (%d rows)", i, pos + ql, d.max_positions);
    for (int j = 0; j < i; ++j) XB_CHECK(b->slot_ids[j] != slot, "slot %d appears twice in one batch", slot);
    T += ql;
    if (ql != 1) decode_only = false;
    if (pos + ql > max_ctx) max_ctx = pos + ql;
  }
  XB_CHECK(T <= d.max_batch_tokens, "batch carries %d token rows, engine was created for %d", T, d.max_batch_tokens);
  if (e->ho_err != nullptr && *reinterpret_cast<volatile int*>(e->ho_err) != 0)
    return xb_fail(3, "an earlier hand-off step timed out waiting for landing slot %d to reach %d (producer stage absent or stuck); "
                      "its result is invalid -- xb_handoff_clear_error() to go on", e->ho_err[0] - 1, e->ho_err[1]);
  {
    // every check comes before the first change of slot_len / the page lists: a batch that cannot be
    // served leaves the engine exactly as it found it
    size_t need_new = 0;
    for (int i = 0; i < b->n_seqs; ++i) {
      const int need_pages = (b->positions[i] + b->q_lens[i] + e->page_tokens - 1) / e->page_tokens;
      const size_t have = e->slot_pages[b->slot_ids[i]].size();
      if ((size_t)need_pages > have) need_new += (size_t)need_pages - have;
    }
    XB_CHECK(need_new <= e->free_pages.size(), "KV page pool exhausted: the batch needs %zu more pages, %zu are free", need_new,
             e->free_pages.size());
  }
  if (ctl != nullptr) {
    XB_CHECK(decode_only, "the peer hand-off carries decode steps only");
    XB_CHECK(!ho_in || (!first && e->landing != nullptr && ctl[0] < e->ho_slots && T <= e->ho_rows),
             "hand-off input: needs a non-first shard with xb_handoff_create done, slot < %d, rows <= %d", e->ho_slots, e->ho_rows);
    XB_CHECK(!ho_out || (!last && e->peer_landing != nullptr && e->done_cnt != nullptr && ctl[2] < e->ho_slots && T <= e->ho_rows),
             "hand-off output: needs a non-last shard with xb_handoff_create + connect done, slot < %d, rows <= %d", e->ho_slots, e->ho_rows);
  }
  // pinned staging buffer from the ring (wait until its previous upload has been consumed)
  const int ring = e->meta_next;
  e->meta_next = (e->meta_next + 1) % xb_engine::META_RING;
  XB_CUDA(cudaEventSynchronize(e->meta_ev[ring]));
  int* hm = e->h_meta_ring[ring];
  const size_t Tmax = d.max_batch_tokens;
  int* tok_seq = hm;
  int* tok_pos = hm + Tmax;
  int* seq_slot = hm + 2 * Tmax;
  int* seq_last = hm + 2 * Tmax + d.max_seqs;
  int* tile_start = hm + 2 * Tmax + 2 * d.max_seqs;
  int* tile_len = tile_start + e->max_tiles;
  int n_tiles = 0;
  int r = 0;
  for (int i = 0; i < b->n_seqs; ++i) {
    const int slot = b->slot_ids[i], ql = b->q_lens[i], pos = b->positions[i];
    // grow the slot's page list (pages come from the shared free list: non-contiguous)
    const int need_pages = (pos + ql + e->page_tokens - 1) / e->page_tokens;
    std::vector<int>& pages = e->slot_pages[slot];
    if ((int)pages.size() < need_pages) {
      const int old = (int)pages.size();
      while ((int)pages.size() < need_pages) {
        const int p = e->free_pages.back();
        e->free_pages.pop_back();
        e->h_page_table[(size_t)slot * e->max_pages + pages.size()] = p;
        pages.push_back(p);
      }
      XB_CUDA(cudaMemcpyAsync(e->d_page_table + (size_t)slot * e->max_pages + old,
                              e->h_page_table.data() + (size_t)slot * e->max_pages + old, (need_pages - old) * 4,
                              cudaMemcpyHostToDevice, st));
    }
    if (!decode_only) {
      for (int j = 0; j < ql; j += e->tile_tokens) {
        tile_start[n_tiles] = r + j;
        tile_len[n_tiles] = ql - j < e->tile_tokens ? ql - j : e->tile_tokens;
        ++n_tiles;
      }
    }
    for (int j = 0; j < ql; ++j) {
      tok_seq[r] = i;
      tok_pos[r] = pos + j;
      ++r;
    }
    seq_slot[i] = slot;
    seq_last[i] = r - 1;
    e->slot_len[slot] = pos + ql;
  }
  e->c.kv_pages_free = e->free_pages.size();
  XB_CHECK(ctl == nullptr || T <= XB_DEC_ROWS, "the peer hand-off carries at most %d rows", XB_DEC_ROWS);
  e->meta_is_dec = decode_only && T <= XB_DEC_ROWS;
  if (e->meta_is_dec) {
    // decode step: positions, slots and the hand-off control words in one small copy (tok_seq / seq_last are the identity)
    int* hd = hm + e->meta_ints;
    hd[0] = hd[1] = hd[2] = hd[3] = -1;
    if (ctl != nullptr) { hd[0] = ctl[0]; hd[1] = ctl[1]; hd[2] = ctl[2]; hd[3] = ctl[3]; }
    for (int i = 0; i < T; ++i) {
      hd[4 + i] = tok_pos[i];
      hd[4 + XB_DEC_ROWS + i] = seq_slot[i];
    }
    XB_CUDA(cudaMemcpyAsync(e->d_dec, hd, (4 + XB_DEC_ROWS + T) * 4, cudaMemcpyHostToDevice, st));
  } else {
    XB_CUDA(cudaMemcpyAsync(e->d_meta, hm, e->meta_ints * 4, cudaMemcpyHostToDevice, st));
  }
  e->meta_T = T;
  XB_CUDA(cudaEventRecord(e->meta_ev[ring], st));

  // ---- inputs to fixed addresses
  if (first) XB_CUDA(cudaMemcpyAsync(e->in_tokens, b->tokens, (size_t)T * 4, cudaMemcpyDeviceToDevice, st));
  else if (!ho_in) XB_CUDA(cudaMemcpyAsync(e->h, b->hidden_in, (size_t)T * d.dim * 2, cudaMemcpyDeviceToDevice, st));

  // decode: split every context across blocks (split-KV); prefill: one block per (query tile,
  // kv head), no kv split
  int n_splits = 1;
  if (decode_only) {
    n_splits = choose_kv_splits(e, T, max_ctx);   // 1 once T * kv heads exceeds one wave (no split workspace needed)
  }

  if (decode_only && e->use_graphs && taps == nullptr && st != nullptr && st != cudaStreamLegacy) {
    // decode step: replay a captured graph (first call of a shape runs eagerly to build the
    // launch plans and set kernel attributes, second call captures)
    // host_logits: the head runs after the graph (its host pointer changes from call to call)
    const int headless = host_logits != nullptr && last;
    const long long key = ((long long)b->n_seqs << 16) | (n_splits << 3) | (headless << 2) | (ho_in << 1) | ho_out;
    xb_engine::StepGraph& sg = e->graphs[key];
    if (sg.exec == nullptr && sg.warmed) {
      cudaGraph_t graph = nullptr;
      const uint64_t before = e->c.launches;
      XB_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
      const int rc = run_body(e, T, b->n_seqs, n_splits, 0, nullptr, st, ho_in, ho_out, !headless);
      cudaError_t ce = cudaStreamEndCapture(st, &graph);
      if (rc != 0) { if (graph) cudaGraphDestroy(graph); return rc; }
      XB_CUDA(ce);
      sg.launches = e->c.launches - before;
      e->c.launches = before;
      XB_CUDA(cudaGraphInstantiate(&sg.exec, graph, 0));
      XB_CUDA(cudaGraphDestroy(graph));
    }
    if (sg.exec != nullptr) {
      XB_CUDA(cudaGraphLaunch(sg.exec, st));
      e->c.launches += sg.launches;
    } else {
      XB_TRY(run_body(e, T, b->n_seqs, n_splits, 0, nullptr, st, ho_in, ho_out, !headless));
      sg.warmed = 1;
    }
    if (headless) XB_TRY(run_head(e, b->n_seqs, st, host_logits));
  } else {
    const int headless = host_logits != nullptr && last;
    XB_TRY(run_body(e, T, b->n_seqs, n_splits, n_tiles, taps, st, ho_in, ho_out, !headless));
    if (headless) XB_TRY(run_head(e, b->n_seqs, st, host_logits));
  }

  if (out != nullptr && !(ho_out && !last)) {
    if (last) XB_CUDA(cudaMemcpyAsync(out, e->logits, (size_t)b->n_seqs * d.vocab * sizeof(float), cudaMemcpyDeviceToDevice, st));
    else XB_CUDA(cudaMemcpyAsync(out, e->h, (size_t)T * d.dim * 2, cudaMemcpyDeviceToDevice, st));
  }
  e->c.forwards++;
  return 0;
}

extern "C" int xb_forward(xb_engine* e, const xb_batch* b, void* out, void* const* taps, void* stream) {
  return forward_impl(e, b, out, taps, stream, nullptr);
}

extern "C" int xb_forward_host_logits(xb_engine* e, const xb_batch* b, void* out, float* host_logits, void* stream) {
  XB_CHECK(e != nullptr && host_logits != nullptr, "null argument");
  XB_CHECK(shard_is_last(e->d), "xb_forward_host_logits needs a last shard (it produces logits)");
  cudaPointerAttributes pa;
  XB_CHECK(cudaPointerGetAttributes(&pa, host_logits) == cudaSuccess && pa.type == cudaMemoryTypeHost && pa.devicePointer != nullptr,
           "host_logits must be page-locked host memory the device can address (cudaHostAlloc / torch pin_memory)");
  return forward_impl(e, b, out, nullptr, stream, nullptr, (float*)pa.devicePointer);
}

extern "C" int xb_forward_handoff(xb_engine* e, const xb_batch* b, int slot_in, int wait_value, int slot_out, int signal_value,
                                  void* out, void* stream) {
  const int ctl[4] = {slot_in, wait_value, slot_out, signal_value};
  return forward_impl(e, b, out, nullptr, stream, ctl);
}

extern "C" int xb_handoff_create(xb_engine* e, int n_slots, int rows, void* landing_handle, void* flag_handle) {
  XB_CHECK(e != nullptr && n_slots > 0 && rows > 0 && rows <= 128, "bad argument");
  XB_CHECK((landing_handle == nullptr) == (flag_handle == nullptr), "pass both IPC handle buffers, or neither (in-process ring)");
  XB_CUDA(cudaSetDevice(e->device));
  XB_CHECK(e->landing == nullptr, "hand-off buffers already created");
  const size_t bytes = (size_t)n_slots * rows * e->d.dim * 2;
  XB_TRY(dmalloc((void**)&e->landing, bytes));
  XB_TRY(dmalloc((void**)&e->flags, (size_t)n_slots * 4));
  XB_TRY(dmalloc((void**)&e->done_cnt, (size_t)n_slots * 4));
  XB_CUDA(cudaMemset(e->landing, 0, bytes));
  XB_CUDA(cudaMemset(e->flags, 0, (size_t)n_slots * 4));
  XB_CUDA(cudaMemset(e->done_cnt, 0, (size_t)n_slots * 4));
  e->ho_slots = n_slots;
  e->ho_rows = rows;
  if (landing_handle != nullptr) {
    XB_CUDA(cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(landing_handle), e->landing));
    XB_CUDA(cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(flag_handle), e->flags));
  }
  // consumer-side wait budget and the host-visible word a timed-out wait reports through
  {
    const char* tv = getenv("XB_HANDOFF_TIMEOUT_S");
    double secs = tv != nullptr ? atof(tv) : 30.0;
    if (!(secs > 0.0)) secs = 30.0;
    int khz = 0;
    XB_CUDA(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, e->device));
    e->ho_timeout_cycles = (long long)(secs * (double)khz * 1000.0);
    XB_CUDA(cudaHostAlloc((void**)&e->ho_err, 2 * sizeof(int), cudaHostAllocMapped));
    e->ho_err[0] = e->ho_err[1] = 0;
    XB_CUDA(cudaHostGetDevicePointer((void**)&e->ho_err_dev, e->ho_err, 0));
  }
  XB_CUDA(cudaDeviceSynchronize());
  return 0;
}

extern "C" int xb_reserve_sms(xb_engine* e, int n) {
  XB_CHECK(e != nullptr && n >= 0 && n < e->num_sms, "bad argument");
  XB_CUDA(cudaSetDevice(e->device));
  XB_CUDA(cudaDeviceSynchronize());
  for (auto& kv : e->graphs)
    if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
  e->graphs.clear();     // captured steps and launch plans were sized for the old SM count
  e->plans.clear();
  e->reserved_sms = n;
  return 0;
}

extern "C" int xb_handoff_clear_error(xb_engine* e) {
  XB_CHECK(e != nullptr, "null engine");
  if (e->ho_err != nullptr) {
    XB_CUDA(cudaSetDevice(e->device));
    XB_CUDA(cudaDeviceSynchronize());
    e->ho_err[0] = e->ho_err[1] = 0;
  }
  return 0;
}

extern "C" int xb_handoff_connect_local(xb_engine* e, xb_engine* next) {
  XB_CHECK(e != nullptr && next != nullptr && e != next, "bad argument");
  XB_CHECK(e->ho_slots > 0 && next->landing != nullptr, "both stages must call xb_handoff_create first");
  XB_CHECK(e->ho_slots == next->ho_slots && e->ho_rows == next->ho_rows && e->d.dim == next->d.dim,
           "hand-off geometry differs between the stages (%d x %d x %d vs %d x %d x %d)", e->ho_slots, e->ho_rows, e->d.dim,
           next->ho_slots, next->ho_rows, next->d.dim);
  XB_CHECK(e->peer_landing == nullptr, "this stage is already connected");
  XB_CUDA(cudaSetDevice(e->device));
  if (e->device != next->device) {
    int can = 0;
    XB_CUDA(cudaDeviceCanAccessPeer(&can, e->device, next->device));
    XB_CHECK(can, "device %d cannot address the memory of device %d", e->device, next->device);
    cudaError_t ce = cudaDeviceEnablePeerAccess(next->device, 0);
    if (ce == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
    else XB_CUDA(ce);
  }
  e->peer_landing = next->landing;
  e->peer_flags = next->flags;
  e->peer_is_local = true;
  return 0;
}

extern "C" int xb_handoff_connect(xb_engine* e, const void* next_landing_handle, const void* next_flag_handle) {
  XB_CHECK(e != nullptr && next_landing_handle && next_flag_handle, "bad argument");
  XB_CHECK(e->ho_slots > 0, "call xb_handoff_create first (slot geometry is shared by all stages)");
  XB_CUDA(cudaSetDevice(e->device));
  cudaIpcMemHandle_t hl, hf;
  memcpy(&hl, next_landing_handle, sizeof(hl));
  memcpy(&hf, next_flag_handle, sizeof(hf));
  void *pl = nullptr, *pf = nullptr;
  XB_CUDA(cudaIpcOpenMemHandle(&pl, hl, cudaIpcMemLazyEnablePeerAccess));
  XB_CUDA(cudaIpcOpenMemHandle(&pf, hf, cudaIpcMemLazyEnablePeerAccess));
  e->peer_landing = (bf16*)pl;
  e->peer_flags = (uint32_t*)pf;
  return 0;
}

// Timing hook for the roofline measurement: launches the Linear kernels of one forward over T
// token rows (QKV, O, gate/up, down per layer, + LM head on a last shard) `iters` times,
// back to back on `stream`, on the engine's own weights and workspaces.  The caller brackets
// the call with CUDA events; the launch count is returned in *n_launches.
extern "C" int xb_bench_linears(xb_engine* e, int T, int iters, int* n_launches, void* stream) {
  XB_CHECK(e != nullptr && T > 0 && T <= e->d.max_batch_tokens && iters > 0, "bad argument");
  const xb_model_desc& d = e->d;
  cudaStream_t st = (cudaStream_t)stream;
  XB_CUDA(cudaSetDevice(e->device));
  const bool last = shard_is_last(d);
  const int D = d.dim, I = d.inter_dim;
  // XB_BENCH_SITE=0..4 restricts the pass to one site (qkv, o, gate/up, down, head) for per-site numbers
  const char* sv = getenv("XB_BENCH_SITE");
  const int only = sv != nullptr ? atoi(sv) : -1;
  int n = 0;
  for (int it = 0; it < iters; ++it) {
    for (int l = 0; l < e->nl; ++l) {
      const LayerW& w = e->L[l];
      const GemmPlan* gp;
      const bool has_attn = w.wqkv != nullptr, has_mlp = w.wgu != nullptr;
      if (has_attn && (only < 0 || only == SITE_QKV)) {
        XB_TRY(get_plan(e, SITE_QKV, l, e->hn, D, w.wqkv, T, e->NQKV, D, e->gemm_out, e->NQKV, w.bqkv, 0, &gp));
        XB_TRY(run_gemm(e, *gp, st));
        ++n;
      }
      if (has_attn && (only < 0 || only == SITE_O)) {
        XB_TRY(get_plan(e, SITE_O, l, e->attn, e->NQ, w.wo, T, D, e->NQ, e->gemm_out, D, nullptr, 0, &gp));
        XB_TRY(run_gemm(e, *gp, st));
        ++n;
      }
      if (has_mlp && (only < 0 || only == SITE_GU)) {
        XB_TRY(get_plan(e, SITE_GU, l, e->hn, D, w.wgu, T, 2 * I, D, e->gemm_out, 2 * I, nullptr, 0, &gp));
        XB_TRY(run_gemm(e, *gp, st));
        ++n;
      }
      if (has_mlp && (only < 0 || only == SITE_DOWN)) {
        XB_TRY(get_plan(e, SITE_DOWN, l, e->act, I, w.wdown, T, D, I, e->gemm_out, D, nullptr, 0, &gp));
        XB_TRY(run_gemm(e, *gp, st));
        ++n;
      }
    }
    if ((only < 0 || only == SITE_HEAD) && last && T <= d.max_seqs && T <= 128) {
      const bf16* head = e->lm_head;
      const GemmPlan* gp;
      XB_TRY(get_plan(e, SITE_HEAD, 0, e->last_rows, D, head, T, d.vocab, D, e->logits, d.vocab, nullptr, 1, &gp));
      XB_TRY(run_gemm(e, *gp, st));
      n += 1;
    }
  }
  e->c.launches += n;
  if (n_launches) *n_launches = n;
  return 0;
}

// Measurement hook: the decode attention of local layer 0 for the batch of the LAST xb_forward
// call (its metadata is still on the device), `iters` times back to back with the given kv-split
// count (0 = the engine's own choice for that batch).  The caller brackets it with CUDA events.
extern "C" int xb_bench_attention(xb_engine* e, int n_seqs, int max_ctx, int n_splits, int iters, int* splits_used,
                                  void* stream) {
  XB_CHECK(e != nullptr && n_seqs > 0 && n_seqs <= e->d.max_seqs && iters > 0, "bad argument");
  const xb_model_desc& d = e->d;
  cudaStream_t st = (cudaStream_t)stream;
  XB_CUDA(cudaSetDevice(e->device));
  BatchMeta m;
  fill_batch_meta(e, m);
  m.page_table = e->d_page_table; m.tile_start = e->d_tile_start; m.tile_len = e->d_tile_len;
  m.max_pages = e->max_pages; m.page_tokens = e->page_tokens;
  if (n_splits <= 0) n_splits = choose_kv_splits(e, n_seqs, max_ctx);
  XB_CHECK(n_splits <= e->max_attn_splits, "n_splits %d above the workspace limit %d", n_splits, e->max_attn_splits);
  for (int it = 0; it < iters; ++it) {
    XB_CHECK(e->L[0].wqkv != nullptr || e->nl > 1, "this stage runs no attention block");
    const LayerW& w0 = e->L[e->L[0].wqkv != nullptr ? 0 : 1];   // first layer whose attention block runs here
    if (d.head_dim == 128) XB_TRY(launch_attn<128>(e, w0, m, n_seqs, n_splits, 0, st));
    else XB_TRY(launch_attn<64>(e, w0, m, n_seqs, n_splits, 0, st));
  }
  if (splits_used) *splits_used = n_splits;
  return 0;
}

// ----------------------------------------------------------------------------- sampling
extern "C" int xb_sample(xb_engine* e, const float* logits, int n, float temp, int top_k, float top_p, const float* q,
                         uint64_t seed, int32_t* tokens_out, void* stream) {
  XB_CHECK(e != nullptr && logits != nullptr && tokens_out != nullptr && n > 0, "bad argument");
  XB_CUDA(cudaSetDevice(e->device));
  cudaStream_t st = (cudaStream_t)stream;
  if (temp <= 0.f) {
    xb_argmax_kernel<1024><<<n, 1024, 0, st>>>(logits, e->d.vocab, tokens_out);
  } else {
    // a cluster of 4 blocks per row keeps the row in shared memory (xb_sample_cluster.cuh); rows too long
    // for 4 x 200 KB, or XB_SAMPLE_CLUSTER=0, take the one-block-per-row kernel
    constexpr int CL = 4;
    const int V = e->d.vocab;
    const int chunk = (((V + CL - 1) / CL) + 3) & ~3;
    const char* cv = getenv("XB_SAMPLE_CLUSTER");
    const bool use_cluster = !(cv != nullptr && cv[0] == '0');
    if (use_cluster && (size_t)chunk * 4 <= (size_t)200 * 1024) {
      XB_SMEM_ATTR_ONCE((xb_sample_cluster_kernel<1024, CL>), 200 * 1024);
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(n * CL);
      cfg.blockDim = dim3(1024);
      cfg.dynamicSmemBytes = (size_t)chunk * 4;
      cfg.stream = st;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeClusterDimension;
      attr[0].val.clusterDim.x = CL;
      attr[0].val.clusterDim.y = 1;
      attr[0].val.clusterDim.z = 1;
      cfg.attrs = attr;
      cfg.numAttrs = 1;
      XB_CUDA(cudaLaunchKernelEx(&cfg, xb_sample_cluster_kernel<1024, CL>, logits, V, chunk, temp, top_k, top_p, q,
                                 (unsigned long long)seed, (int*)tokens_out));
    } else {
      xb_sample_kernel<1024><<<n, 1024, 0, st>>>(logits, V, temp, top_k, top_p, q, (unsigned long long)seed, tokens_out);
    }
  }
  XB_CUDA(cudaGetLastError());
  LAUNCH_COUNT(e);
  return 0;
}

// Timeline instrumentation read-out (engine created with XB_GEMM_TS=1): copies the stamps of the Linear launches
// recorded since the last call -- [n_launches][blocks][XB_TS_SLOTS] nanosecond globaltimer values, 0 = not reached -- and
// starts a new recording.  Synchronises the device.
extern "C" int xb_debug_timeline(xb_engine* e, unsigned long long* host_out, size_t cap_u64, int* n_launches, int* blocks) {
  XB_CHECK(e != nullptr && n_launches != nullptr && blocks != nullptr, "null argument");
  XB_CHECK(e->ts_buf != nullptr, "create the engine with XB_GEMM_TS=1 to record a timeline");
  XB_CUDA(cudaSetDevice(e->device));
  XB_CUDA(cudaDeviceSynchronize());
  const size_t per = (size_t)e->ts_blocks * XB_TS_SLOTS;
  int n = e->ts_next;
  if (host_out != nullptr) {
    if ((size_t)n * per > cap_u64) n = (int)(cap_u64 / per);
    XB_CUDA(cudaMemcpy(host_out, e->ts_buf, (size_t)n * per * 8, cudaMemcpyDeviceToHost));
  }
  XB_CUDA(cudaMemset(e->ts_buf, 0, (size_t)e->ts_cap * per * 8));
  e->ts_next = 0;
  *n_launches = n;
  *blocks = e->ts_blocks;
  return 0;
}

extern "C" int xb_get_counters(xb_engine* e, xb_counters* out) {
  XB_CHECK(e != nullptr && out != nullptr, "null argument");
  *out = e->c;
  return 0;
}

// ----------------------------------------------------------------------------- kernel-level test hook
__global__ void xb_reduce_partials_kernel(const float* __restrict__ part, SplitInfo si, int N, long long n,
                                          float* __restrict__ y, int round_bf16) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float a = 0.f;
  const int splits = xb_n_segments(si, (int)(i % N));
  for (int s = 0; s < splits; ++s) a += part[s * n + i];
  y[i] = round_bf16 ? rbf(a) : a;
}
__global__ void xb_bf16_to_f32_kernel(const bf16* __restrict__ x, long long n, float* __restrict__ y) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] = bf2f(x[i]);
}

extern "C" int xb_debug_linear(const void* x, const void* w, int T, int N, int K, float* y, float* scratch,
                               int use_tensor_cores, int force_splits, int round_bf16, void* stream) {
  XB_CHECK(x && w && y && scratch && T > 0 && N > 0 && K > 0, "bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  int dev = 0, sms = 148;
  XB_CUDA(cudaGetDevice(&dev));
  XB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  if (!use_tensor_cores) {
    dim3 grid((N + 3) / 4, (T + 7) / 8);
    xb_linear_check_kernel<<<grid, 128, 0, st>>>((const bf16*)x, (const bf16*)w, y, T, N, K);
    XB_CUDA(cudaGetLastError());
    return 0;
  }
  GemmPlan p;
  XB_TRY(xb_gemm_plan(&p, (const bf16*)x, K, (const bf16*)w, 0, T, N, K, scratch, N, nullptr, 0, sms, force_splits));
  XB_TRY(xb_gemm_run(p, st, 0));
  const long long n = (long long)T * N;
  const int blocks = (int)((n + 255) / 256);
  if (p.args.mode == XB_GEMM_SKINNY)
    xb_reduce_partials_kernel<<<blocks, 256, 0, st>>>(scratch, p.split, N, n, y, round_bf16);
  else
    xb_bf16_to_f32_kernel<<<blocks, 256, 0, st>>>((const bf16*)scratch, n, y);
  XB_CUDA(cudaGetLastError());
  return 0;
}

// ----------------------------------------------------------------------------- allocation-free Linear
// Built-in scratch for xb_linear_noalloc: lives in the library image, so every GPU that loads
// libxb200 carries these 8 MiB for the life of the process.
#define XB_NOALLOC_SCRATCH_FLOATS (2u * 1024u * 1024u)
static __device__ __align__(256) float g_xb_noalloc_scratch[XB_NOALLOC_SCRATCH_FLOATS];
// Per-device DEVICE address of the array above.  In host code `g_xb_noalloc_scratch` is only a
// host-side handle for the symbol -- handing it to a kernel would make the kernel write through
// a host address -- so the real address is looked up once per device and cached here.
#define XB_NOALLOC_MAX_DEVICES XB_MAX_DEVICES
static float* g_xb_noalloc_ptr[XB_NOALLOC_MAX_DEVICES] = {nullptr};

// y[T,N] = x[T,K] * w[N,K]^T with the engine's split-K tcgen05 Linear, for callers that must
// not allocate.  A serving loop that calls cudaMalloc/cudaFree (or lets a framework allocator do
// it) between decode steps stalls every stream of the device on the allocator's implicit
// synchronisation, and a process that has already handed all of its HBM to weights and KV pages
// can be refused the few MiB of split-K workspace in the middle of a request.  Here the
// workspace is the static array above: the call enqueues two kernels and touches no allocator
// (no cudaMalloc, no cudaFree, no torch tensor is created).
//
// * T <= 128 (the weight-streaming "skinny" shape); splits*T*N floats of workspace must fit
//   XB_NOALLOC_SCRATCH_FLOATS, otherwise the call fails with the required and available sizes.
// * The first call on each device looks up the array's device address
//   (cudaGetSymbolAddress, with that device current); make that first call outside any stream
//   capture.
// * All calls on a device share the one array and it keeps whatever the previous call left in
//   it.  Nothing has to be cleared: every workspace element this call reads is written earlier in
//   the same call (each (split, token, feature) partial is stored by exactly one work unit before
//   the reduction kernel runs).  Two calls whose kernels can overlap on the GPU (different
//   streams, or one stream per thread) race on the array -- serialise them.
extern "C" int xb_linear_noalloc(const void* x, const void* w, int T, int N, int K, float* y, int round_bf16,
                                 void* stream) {
  XB_CHECK(x && w && y && T > 0 && N > 0 && K > 0, "bad argument");
  XB_CHECK(T <= 128, "xb_linear_noalloc handles up to 128 token rows per call (got %d)", T);
  cudaStream_t st = (cudaStream_t)stream;
  int dev = 0, sms = 148;
  XB_CUDA(cudaGetDevice(&dev));
  XB_CHECK(dev >= 0 && dev < XB_NOALLOC_MAX_DEVICES, "device index %d not supported", dev);
  {
    // the scratch array used is the CURRENT device's copy: the operands must live there too
    cudaPointerAttributes pa;
    XB_CUDA(cudaPointerGetAttributes(&pa, x));
    XB_CHECK(pa.type == cudaMemoryTypeDevice || pa.type == cudaMemoryTypeManaged, "x must be device memory");
    XB_CHECK(pa.type == cudaMemoryTypeManaged || pa.device == dev,
             "xb_linear_noalloc: x lives on device %d but device %d is current; make the operands' device current "
             "(cudaSetDevice) before the call", pa.device, dev);
  }
  if (g_xb_noalloc_ptr[dev] == nullptr) {
    float* p = nullptr;
    XB_CUDA(cudaGetSymbolAddress((void**)&p, g_xb_noalloc_scratch));
    XB_CHECK(p != nullptr, "cudaGetSymbolAddress returned a null address for the built-in scratch");
    g_xb_noalloc_ptr[dev] = p;
  }
  float* scratch = g_xb_noalloc_ptr[dev];
  XB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  GemmPlan p;
  XB_TRY(xb_gemm_plan(&p, (const bf16*)x, K, (const bf16*)w, 0, T, N, K, scratch, N, nullptr, 0, sms, 0));
  const unsigned long long need = (unsigned long long)p.split.splits * T * N;
  XB_CHECK(need <= XB_NOALLOC_SCRATCH_FLOATS,
           "xb_linear_noalloc: this shape needs %llu floats of split-K workspace (splits=%d x T=%d x N=%d); the built-in "
           "buffer holds %u floats (%u MiB)",
           need, p.split.splits, T, N, XB_NOALLOC_SCRATCH_FLOATS, (unsigned)(XB_NOALLOC_SCRATCH_FLOATS * 4u >> 20));
  XB_TRY(xb_gemm_run(p, st, 0));
  const long long n = (long long)T * N;
  xb_reduce_partials_kernel<<<(int)((n + 255) / 256), 256, 0, st>>>(scratch, p.split, N, n, y, round_bf16);
  XB_CUDA(cudaGetLastError());
  return 0;
}
