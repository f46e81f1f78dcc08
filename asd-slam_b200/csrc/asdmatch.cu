// asdmatch.cu -- C ABI of libasdmatch_cuda.so (include/asdmatch.h). Host-side orchestration only: every
// distance, selection, ratio test and histogram runs in the sm_100a kernels of kernels_*.cuh / tc_shortlist.cu.
// There is no CPU fallback: without a CUDA device asdm_create fails.
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <new>

#include "internal.cuh"
#include "kernels_dense.cuh"
#include "kernels_mappoint.cuh"
#include "kernels_proj.cuh"
#include "kernels_set.cuh"
#include "kernels_vocab.cuh"
#include "kernels_bowdb.cuh"

// ------------------------------------------------------------------------------------------------ utilities
static int fail(asdm_ctx* ctx, int code, const char* what, cudaError_t e = cudaSuccess) {
  if (ctx) {
    char buf[512];
    if (e != cudaSuccess)
      snprintf(buf, sizeof buf, "%s: %s", what, cudaGetErrorString(e));
    else
      snprintf(buf, sizeof buf, "%s", what);
    ctx->err = buf;
  }
  if (e != cudaSuccess) cudaGetLastError();  // clear the sticky-less error state
  return code;
}

#define CK(call)                                                        \
  do {                                                                  \
    cudaError_t e__ = (call);                                           \
    if (e__ != cudaSuccess) return fail(ctx, e__ == cudaErrorMemoryAllocation ? ASDM_ERR_OOM : ASDM_ERR_CUDA, #call, e__); \
  } while (0)

#define LAUNCHED(ctx)           \
  do {                          \
    (ctx)->stat_launches++;     \
    (ctx)->stat_total_launches++; \
  } while (0)

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// slots of asdm_ctx::scratch
enum { SC_QUERY = 0, SC_QDESC = 1, SC_SEG = 2, SC_ENT = 3, SC_STATE = 4, SC_BLOCK = 5, SC_MISC = 6, SC_MP_ARGS = 7, SC_MP_DIST = 8, SC_VOC = 9, SC_INIT = 10 };
static int flatten_featvec(asdm_ctx* ctx, const asdm_featvec* fv, int n, int n_pad, uint32_t* node, int32_t* pos);

// Device memory comes from the device's stream-ordered pool (cudaMallocAsync) with the release threshold lifted in
// asdm_create: allocation and free are ordered on the context stream, never synchronise the device, and the memory
// of released sets / jobs is recycled by the next ones (keyframes come and go at frame rate in the SLAM system).
static cudaError_t dev_alloc(asdm_ctx* ctx, void** p, size_t bytes) { return cudaMallocAsync(p, bytes, ctx->stream); }
static void dev_free(asdm_ctx* ctx, void* p) {
  if (p && cudaFreeAsync(p, ctx->stream) != cudaSuccess) cudaGetLastError();
}

// size classes of the block cache: 4 per octave (waste <= 25 %), at least 4 KB
static size_t block_class(size_t bytes) {
  size_t c = 4096;
  while (c < bytes) {
    const size_t q = c / 4;
    for (int k = 1; k <= 4; k++)
      if (c + q * k >= bytes) return c + q * k;
    c *= 2;
  }
  return c;
}
static const size_t BLOCK_CACHE_LIMIT = (size_t)8 << 30;

static void block_put(asdm_ctx* ctx, void* p, size_t cap) {
  if (!p) return;
  if (ctx->block_cache_bytes + cap > BLOCK_CACHE_LIMIT) {
    dev_free(ctx, p);
    return;
  }
  ctx->block_cache[cap].push_back(p);
  ctx->block_cache_bytes += cap;
}

static void block_cache_flush(asdm_ctx* ctx) {
  for (auto& kv : ctx->block_cache)
    for (void* p : kv.second) dev_free(ctx, p);
  ctx->block_cache.clear();
  ctx->block_cache_bytes = 0;
}

static int ensure(asdm_ctx* ctx, DevBuf& b, size_t bytes) {
  if (bytes <= b.cap) return ASDM_OK;
  block_put(ctx, b.p, b.cap);  // stream order: kernels already queued on the old block finish before any reuse
  b.p = nullptr;
  b.cap = 0;
  const size_t want = block_class(std::max(bytes, (size_t)4096));
  auto it = ctx->block_cache.find(want);
  if (it != ctx->block_cache.end() && !it->second.empty()) {
    b.p = it->second.back();
    it->second.pop_back();
    ctx->block_cache_bytes -= want;
    b.cap = want;
    return ASDM_OK;
  }
  cudaError_t e = dev_alloc(ctx, &b.p, want);
  if (e == cudaErrorMemoryAllocation) {  // give the cached blocks back and try once more
    cudaGetLastError();
    block_cache_flush(ctx);
    e = dev_alloc(ctx, &b.p, want);
  }
  CK(e);
  b.cap = want;
  return ASDM_OK;
}

static void release(asdm_ctx* ctx, DevBuf& b) {
  if (b.cap) block_put(ctx, b.p, b.cap);  // cap == 0: a view into another block (job input / zero arenas)
  b.p = nullptr;
  b.cap = 0;
}

// the library synchronised the stream: everything staged so far has been consumed
static inline void synced(asdm_ctx* ctx) { ctx->stage_off = 0; }

static int stage_alloc(asdm_ctx* ctx, size_t bytes, void** out) {
  bytes = align_up(std::max(bytes, (size_t)1), 256);
  if (ctx->stage_off + bytes > ctx->stage_cap) {
    CK(cudaStreamSynchronize(ctx->stream));
    synced(ctx);
    if (bytes > ctx->stage_cap) {
      if (ctx->stage) cudaFreeHost(ctx->stage);
      ctx->stage = nullptr;
      ctx->stage_cap = 0;
      const size_t want = std::max(bytes * 2, (size_t)16 << 20);
      CK(cudaHostAlloc((void**)&ctx->stage, want, cudaHostAllocMapped));
      ctx->stage_cap = want;
    }
  }
  *out = ctx->stage + ctx->stage_off;
  ctx->stage_off += bytes;
  return ASDM_OK;
}

// Small inputs of a search reach the device through a COPY KERNEL that reads the page-locked staging arena over PCIe
// (zero-copy loads), not through the copy engine: the engine serves a stream's backlog in long slices, so a 4 KB
// work list queued behind a burst of 2 MB keyframe uploads (another stream, same engine) would only arrive when the
// burst has drained -- measured: the first wave of a streamed sweep finished together with the LAST upload.
__global__ void __launch_bounds__(256) k_pull16(uint4* __restrict__ dst, const uint4* __restrict__ src, size_t n16) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i];
}
__global__ void __launch_bounds__(256) k_pull4(uint32_t* __restrict__ dst, const uint32_t* __restrict__ src, size_t n4) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i];
}

// zero fill by kernel, for the same reason (cudaMemsetAsync may be served by the copy engine)
__global__ void __launch_bounds__(256) k_zero16(uint4* __restrict__ dst, size_t n16) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x)
    dst[i] = make_uint4(0u, 0u, 0u, 0u);
}
static int zero_device(asdm_ctx* ctx, void* dst, size_t bytes) {  // dst 16-byte aligned; rounds bytes up to 16
  const size_t n = (bytes + 15) / 16;
  if (!n) return ASDM_OK;
  k_zero16<<<(int)std::min<size_t>((n + 255) / 256, 1184), 256, 0, ctx->stream>>>((uint4*)dst, n);
  LAUNCHED(ctx);
  CK(cudaGetLastError());
  return ASDM_OK;
}

// `h` lies in the staging arena (page-locked, device-visible under UVA); queued on the main stream
static int pull_to_device(asdm_ctx* ctx, void* dst, const void* h, size_t bytes) {
  if (!bytes) return ASDM_OK;
  const bool a16 = (((uintptr_t)dst | (uintptr_t)h | bytes) & 15) == 0, a4 = (((uintptr_t)dst | (uintptr_t)h | bytes) & 3) == 0;
  if (!a4 || bytes > ((size_t)8 << 20)) {  // odd sizes / big blocks: the copy engine
    CK(cudaMemcpyAsync(dst, h, bytes, cudaMemcpyHostToDevice, ctx->stream));
    return ASDM_OK;
  }
  const size_t n = a16 ? bytes / 16 : bytes / 4;
  const int blocks = (int)std::min<size_t>((n + 255) / 256, 64);
  if (a16) k_pull16<<<blocks, 256, 0, ctx->stream>>>((uint4*)dst, (const uint4*)h, n);
  else k_pull4<<<blocks, 256, 0, ctx->stream>>>((uint32_t*)dst, (const uint32_t*)h, n);
  LAUNCHED(ctx);
  CK(cudaGetLastError());
  return ASDM_OK;
}

// host -> device through the staging arena: returns at once, the caller's buffer is free on return
static int stage_upload(asdm_ctx* ctx, void* dst, const void* src, size_t bytes, cudaStream_t stream = nullptr) {
  if (!bytes) return ASDM_OK;
  void* h = nullptr;
  int rc = stage_alloc(ctx, bytes, &h);
  if (rc) return rc;
  memcpy(h, src, bytes);
  if (!stream || stream == ctx->stream) return pull_to_device(ctx, dst, h, bytes);
  CK(cudaMemcpyAsync(dst, h, bytes, cudaMemcpyHostToDevice, stream));
  return ASDM_OK;
}

enum { SLOT_BYTES = 64, SLOT_COUNT = 4096 };
static int32_t* slot_get(asdm_ctx* ctx) {
  if (!ctx->slot_slab) {
    if (cudaHostAlloc((void**)&ctx->slot_slab, (size_t)SLOT_BYTES * SLOT_COUNT, cudaHostAllocMapped) != cudaSuccess) {
      cudaGetLastError();
      ctx->slot_slab = nullptr;
      return nullptr;
    }
    ctx->slot_free.reserve(SLOT_COUNT);
    for (int i = SLOT_COUNT - 1; i >= 0; i--) ctx->slot_free.push_back(i);
  }
  if (ctx->slot_free.empty()) return nullptr;
  const int i = ctx->slot_free.back();
  ctx->slot_free.pop_back();
  int32_t* p = (int32_t*)(ctx->slot_slab + (size_t)i * SLOT_BYTES);
  memset(p, 0, SLOT_BYTES);
  return p;
}
static void slot_put(asdm_ctx* ctx, int32_t* p) {
  if (p) ctx->slot_free.push_back((int)(((char*)p - ctx->slot_slab) / SLOT_BYTES));
}

// Smallest keep threshold t such that a candidate farther than t can neither be accepted (d > TH_LOW) nor make
// the ratio test fail as second best: for every fp32 d > t, fl(nnratio*d) > th_accept >= best1.
static float bow_keep_threshold(float nnratio, float th_accept) {
  if (!(nnratio > 0.f)) return ASDM_BEST_INIT_BOW;
  float t = th_accept / nnratio;
  if (!(t < ASDM_BEST_INIT_BOW)) return ASDM_BEST_INIT_BOW;
  for (int it = 0; it < 64 && !(nnratio * t > th_accept); it++) t = nextafterf(t, INFINITY);
  t = nextafterf(nextafterf(t, INFINITY), INFINITY);
  return std::min(std::max(t, th_accept), ASDM_BEST_INIT_BOW);
}

// ------------------------------------------------------------------------------------------------- context
extern "C" int asdm_create(int device, asdm_ctx** out) {
  if (!out) return ASDM_ERR_INVALID;
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev <= 0 || device < 0 || device >= ndev) {
    cudaGetLastError();
    return ASDM_ERR_CUDA;  // no device: there is no CPU path behind this library
  }
  asdm_ctx* ctx = new (std::nothrow) asdm_ctx();
  if (!ctx) return ASDM_ERR_OOM;
  ctx->device = device;
  if (cudaSetDevice(device) != cudaSuccess || cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaStreamCreateWithFlags(&ctx->upload_stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaStreamCreateWithFlags(&ctx->prep_stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaStreamCreateWithFlags(&ctx->download_stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreateWithFlags(&ctx->run_done, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&ctx->upload_done, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&ctx->prep_done, cudaEventDisableTiming) != cudaSuccess) {
    cudaGetLastError();
    delete ctx;
    return ASDM_ERR_CUDA;
  }
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) {
    ctx->sm_count = prop.multiProcessorCount;
    if (prop.major != 10) {
      cudaStreamDestroy(ctx->stream);
      delete ctx;
      return ASDM_ERR_CUDA;  // kernels are built for sm_100a only
    }
  }
  cudaFuncSetAttribute(k_cand_exact, cudaFuncAttributeMaxDynamicSharedMemorySize,
                       (int)(ASDM_DESC_DIM * (CE_TM + CE_TN) * sizeof(float)));
  {  // keep freed blocks in the pool instead of returning them to the driver at every synchronisation
    cudaMemPool_t pool = nullptr;
    uint64_t keep = UINT64_MAX;
    if (cudaDeviceGetDefaultMemPool(&pool, device) != cudaSuccess ||
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep) != cudaSuccess) {
      cudaGetLastError();
      cudaStreamDestroy(ctx->stream);
      delete ctx;
      return ASDM_ERR_CUDA;
    }
  }
  *out = ctx;
  return ASDM_OK;
}

extern "C" int asdm_destroy(asdm_ctx* ctx) {
  if (!ctx) return ASDM_ERR_INVALID;
  cudaSetDevice(ctx->device);
  for (auto& b : ctx->scratch) release(ctx, b);
  block_cache_flush(ctx);
  cudaStreamSynchronize(ctx->stream);
  for (cudaEvent_t e : ctx->ev) cudaEventDestroy(e);
  if (ctx->stage) cudaFreeHost(ctx->stage);
  if (ctx->slot_slab) cudaFreeHost(ctx->slot_slab);
  if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
  if (ctx->upload_stream) cudaStreamDestroy(ctx->upload_stream);
  if (ctx->upload_done) cudaEventDestroy(ctx->upload_done);
  if (ctx->prep_stream) cudaStreamDestroy(ctx->prep_stream);
  if (ctx->prep_done) cudaEventDestroy(ctx->prep_done);
  if (ctx->download_stream) cudaStreamDestroy(ctx->download_stream);
  if (ctx->run_done) cudaEventDestroy(ctx->run_done);
  delete ctx;
  return ASDM_OK;
}

extern "C" const char* asdm_last_error(const asdm_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }
extern "C" void* asdm_stream(asdm_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

extern "C" int asdm_set_stream(asdm_ctx* ctx, void* stream) {
  if (!ctx) return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  synced(ctx);
  if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
  if (stream) {
    ctx->stream = (cudaStream_t)stream;
    ctx->own_stream = false;
  } else {  // back to a stream of our own
    ctx->stream = nullptr;
    ctx->own_stream = true;
    CK(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
  }
  return ASDM_OK;
}

extern "C" int asdm_synchronize(asdm_ctx* ctx) {
  if (!ctx) return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  CK(cudaStreamSynchronize(ctx->stream));
  synced(ctx);
  return ASDM_OK;
}

extern "C" int asdm_set_option(asdm_ctx* ctx, const char* key, int value) {
  if (!ctx || !key) return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  if (!strcmp(key, "force_exact")) ctx->opt_force_exact = value;
  else if (!strcmp(key, "init_serial")) ctx->opt_init_serial = value;
  else if (!strcmp(key, "voc_direct")) ctx->opt_voc_direct = value;
  else if (!strcmp(key, "voc_exact")) ctx->opt_voc_exact = value;
  else if (!strcmp(key, "bowdb_search")) ctx->opt_bowdb_search = value;
  else if (!strcmp(key, "bowdb_filter_bits")) ctx->opt_bowdb_filter_bits = std::min(std::max(value, 8), 1024);
  else if (!strcmp(key, "cand_cap")) ctx->opt_cand_cap = std::max(1, value);
  else if (!strcmp(key, "list_cap_per_query")) ctx->opt_list_cap_per_query = std::max(1, value);
  else if (!strcmp(key, "time_kernels")) ctx->opt_time_kernels = value;
  else if (!strcmp(key, "tc_epi_warps")) ctx->opt_tc_epi_warps = value == 8 ? 8 : 4;
  else if (!strcmp(key, "tc_variant")) ctx->opt_tc_variant = value;
  else return fail(ctx, ASDM_ERR_INVALID, "unknown option");
  return ASDM_OK;
}

extern "C" int asdm_get_stat(asdm_ctx* ctx, const char* key, int64_t* value) {
  if (!ctx || !key || !value) return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  if (!strcmp(key, "gpu_launches")) *value = ctx->stat_launches;
  else if (!strcmp(key, "total_launches")) *value = ctx->stat_total_launches;
  else if (!strcmp(key, "cand_cap")) *value = ctx->stat_cand_cap;
  else if (!strcmp(key, "retries")) *value = ctx->stat_retries;
  else if (!strcmp(key, "init_sweeps")) *value = ctx->stat_init_sweeps;
  else if (!strcmp(key, "voc_exact_levels")) *value = ctx->stat_voc_exact;
  else if (!strcmp(key, "tc_used")) *value = ctx->stat_tc_used;
  else if (!strcmp(key, "tc_pairs")) *value = ctx->stat_tc_pairs;
  else if (!strcmp(key, "sm_count")) *value = ctx->sm_count;
  else if (!strcmp(key, "dist_stage_us") || !strcmp(key, "dist_stage_runs")) {
    // total device time of the distance stage (tensor-core shortlist, or the exact kernel) over the runs recorded
    // since the last drain; the caller must have synchronised
    double us = 0;
    const int64_t runs = (int64_t)ctx->ev.size() / 2;
    if (!strcmp(key, "dist_stage_runs")) {
      *value = runs;
      return ASDM_OK;
    }
    for (size_t i = 0; i + 1 < ctx->ev.size(); i += 2) {
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, ctx->ev[i], ctx->ev[i + 1]) == cudaSuccess) us += (double)ms * 1000.0;
      cudaEventDestroy(ctx->ev[i]);
      cudaEventDestroy(ctx->ev[i + 1]);
    }
    cudaGetLastError();
    ctx->ev.clear();
    *value = (int64_t)(us + 0.5);
  }
  else return fail(ctx, ASDM_ERR_INVALID, "unknown stat");
  return ASDM_OK;
}

extern "C" int asdm_host_alloc(void** out, size_t bytes) {
  if (!out) return ASDM_ERR_INVALID;
  if (cudaHostAlloc(out, bytes, cudaHostAllocDefault) != cudaSuccess) {
    cudaGetLastError();
    return ASDM_ERR_OOM;
  }
  return ASDM_OK;
}
extern "C" int asdm_host_free(void* p) {
  if (p && cudaFreeHost(p) != cudaSuccess) {
    cudaGetLastError();
    return ASDM_ERR_CUDA;
  }
  return ASDM_OK;
}

// --------------------------------------------------------------------------------------- descriptor sets
static void set_free(asdm_set* s) {
  if (!s) return;
  asdm_ctx* ctx = s->ctx;
  cudaSetDevice(ctx->device);
  dev_free(ctx, s->desc);  // one arena: every other device pointer lives inside it (or in the batch's side block)
  dev_free(ctx, s->fv_mem);
  if (s->side && --s->side->ref == 0) {
    dev_free(ctx, s->side->p);
    delete s->side;
  }
  slot_put(ctx, s->h_flag);
  delete s;
}

// Frame::AssignFeaturesToGrid (Frame.cc:123-138) on the device, on first use: only the projection-gated searches read
// the bucket, so keyframes that only ever meet BoW-gated searches (loop-closure sweeps) never pay for it.
static int ensure_grid(asdm_ctx* ctx, asdm_set* s) {
  if (!s->has_grid || s->grid_built) return ASDM_OK;
  const int nc = s->grid.cols * s->grid.rows;
  const size_t smem = (size_t)(nc + 1) * sizeof(int32_t);
  if (smem > 48 * 1024) cudaFuncSetAttribute(k_build_grid, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  k_build_grid<<<1, 1024, smem, ctx->stream>>>(s->kps, s->n, s->grid, s->cell_tmp, s->cell_start, s->cell_items);
  LAUNCHED(ctx);
  CK(cudaGetLastError());
  s->grid_built = true;
  return ASDM_OK;
}

// reads the tensor-core range flags of sets whose flag copy is still in flight (one stream wait for all of them)
static int resolve_tc_flags(asdm_ctx* ctx, asdm_set* const* sets, int n) {
  bool pending = false;
  for (int k = 0; k < n; k++) pending |= sets[k]->tc_ok < 0;
  if (!pending) return ASDM_OK;
  CK(cudaStreamSynchronize(ctx->stream));
  synced(ctx);
  for (int k = 0; k < n; k++) {
    asdm_set* s = sets[k];
    if (s->tc_ok >= 0) continue;
    s->tc_ok = (s->h_flag && *s->h_flag == 0) ? 1 : 0;
    slot_put(ctx, s->h_flag);
    s->h_flag = nullptr;
  }
  return ASDM_OK;
}

static int set_create_impl(asdm_ctx* ctx, const float* desc, int n, int row_stride_bytes, const asdm_kp* kps,
                           const asdm_grid* grid, asdm_set** out, bool wait) {
  if (!ctx || !out || n < 0 || (n > 0 && !desc)) return ASDM_ERR_INVALID;
  *out = nullptr;
  if (row_stride_bytes == 0) row_stride_bytes = ASDM_DESC_DIM * sizeof(float);
  if (row_stride_bytes < (int)(ASDM_DESC_DIM * sizeof(float))) return fail(ctx, ASDM_ERR_INVALID, "row stride < 512 B");
  if (grid && (!kps || grid->cols <= 0 || grid->rows <= 0 || (int64_t)grid->cols * grid->rows > (1 << 20)))
    return fail(ctx, ASDM_ERR_INVALID, "bad grid");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  asdm_set* s = new (std::nothrow) asdm_set();
  if (!s) return fail(ctx, ASDM_ERR_OOM, "host allocation");
  s->ctx = ctx;
  s->n = n;
  s->n_pad = (int)align_up((size_t)std::max(n, 1), 256);
  const int nc = grid ? grid->cols * grid->rows : 0;
  // one arena per set
  size_t off = 0;
  const size_t o_desc = off; off += align_up((size_t)s->n_pad * ASDM_DESC_DIM * sizeof(float), 1024);
  const size_t o_bf = off;   off += align_up((size_t)s->n_pad * ASDM_DESC_DIM * sizeof(__nv_bfloat16), 1024);
  const size_t o_norm = off; off += align_up((size_t)s->n_pad * sizeof(float), 256);
  const size_t o_flags = off; off += 256;
  const size_t o_kps = off;  off += kps ? align_up((size_t)s->n_pad * sizeof(asdm_kp), 256) : 0;
  const size_t o_cs = off;   off += grid ? align_up((size_t)(nc + 1) * sizeof(int32_t), 256) : 0;
  const size_t o_ci = off;   off += grid ? align_up((size_t)s->n_pad * sizeof(int32_t), 256) : 0;
  const size_t o_tmp = off;  off += grid ? align_up((size_t)s->n_pad * sizeof(int32_t), 256) : 0;
  // allocated on the upload stream (the pool orders the reuse of blocks freed on the main stream by itself)
  char* base = nullptr;
  cudaError_t e = cudaMallocAsync((void**)&base, off, ctx->upload_stream);
  if (e != cudaSuccess) {
    delete s;
    return fail(ctx, ASDM_ERR_OOM, "cudaMallocAsync(set)", e);
  }
  s->desc = (float*)(base + o_desc);
  s->desc_bf = (__nv_bfloat16*)(base + o_bf);
  s->norm2 = (float*)(base + o_norm);
  s->d_flags = (int32_t*)(base + o_flags);
  auto bail = [&](int code) {
    set_free(s);
    return code;
  };
#define CKS(call)                                              \
  do {                                                         \
    cudaError_t e__ = (call);                                  \
    if (e__ != cudaSuccess) return bail(fail(ctx, ASDM_ERR_CUDA, #call, e__)); \
  } while (0)
  // All host->device traffic of a set rides the upload stream: a small copy queued on the main stream would sit in
  // the copy engine's FIFO behind its own stream dependencies (pack kernel <- this upload) and stall the next
  // keyframe's descriptor DMA behind it (measured: 26 GB/s instead of 48 GB/s over 64 keyframes).
  if (kps) {
    s->kps = (asdm_kp*)(base + o_kps);
    s->has_kps = true;
    if (n > 0) {
      const int rc = stage_upload(ctx, s->kps, kps, (size_t)n * sizeof(asdm_kp), ctx->upload_stream);
      if (rc) return bail(rc);
    }
  }
  // `desc` may also be device memory (e.g. keyframes replicated over NVLink by an all-gather that the caller queued on
  // the context stream): that copy is ordered on the context stream instead.
  bool desc_on_device = false;
  if (n > 0) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, desc) == cudaSuccess) desc_on_device = at.type == cudaMemoryTypeDevice;
    else cudaGetLastError();
  }
  if (n > 0 && !desc_on_device) {
    // descriptors are DMA'd straight from the caller's buffer (true async when it is page-locked)
    if (row_stride_bytes == (int)(ASDM_DESC_DIM * sizeof(float)))
      CKS(cudaMemcpyAsync(s->desc, desc, (size_t)n * ASDM_DESC_DIM * sizeof(float), cudaMemcpyHostToDevice,
                          ctx->upload_stream));
    else
      CKS(cudaMemcpy2DAsync(s->desc, ASDM_DESC_DIM * sizeof(float), desc, row_stride_bytes,
                            ASDM_DESC_DIM * sizeof(float), n, cudaMemcpyHostToDevice, ctx->upload_stream));
  }
  // Preparation (bf16 copy, norms, range flag) runs on its own stream behind the upload, so the copy engine moves on
  // to the next keyframe at once and searches already queued on the main stream are not in its way.
  CKS(cudaEventRecord(ctx->upload_done, ctx->upload_stream));
  CKS(cudaStreamWaitEvent(ctx->prep_stream, ctx->upload_done, 0));
  if (n > 0 && desc_on_device) {  // the producer of device-resident rows was queued on the main stream
    CKS(cudaEventRecord(ctx->prep_done, ctx->stream));
    CKS(cudaStreamWaitEvent(ctx->prep_stream, ctx->prep_done, 0));
    CKS(cudaMemcpy2DAsync(s->desc, ASDM_DESC_DIM * sizeof(float), desc, row_stride_bytes, ASDM_DESC_DIM * sizeof(float), n,
                          cudaMemcpyDeviceToDevice, ctx->prep_stream));
  }
  CKS(cudaMemsetAsync(s->d_flags, 0, 256, ctx->prep_stream));
  if (s->n_pad > n)
    CKS(cudaMemsetAsync(s->desc + (size_t)n * ASDM_DESC_DIM, 0, (size_t)(s->n_pad - n) * ASDM_DESC_DIM * sizeof(float),
                        ctx->prep_stream));
  ctx->stat_launches = 0;
  k_pack_desc<<<(s->n_pad * 32 + 255) / 256, 256, 0, ctx->prep_stream>>>(s->desc, n, s->n_pad, s->desc_bf, s->norm2, s->d_flags);
  LAUNCHED(ctx);
  s->h_flag = slot_get(ctx);  // the range flag travels back behind the kernel; read when a job first needs it
  if (s->h_flag) CKS(cudaMemcpyAsync(s->h_flag, s->d_flags, sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->prep_stream));
  CKS(cudaEventRecord(ctx->prep_done, ctx->prep_stream));
  CKS(cudaStreamWaitEvent(ctx->stream, ctx->prep_done, 0));  // whatever is queued on the main stream from here on sees the set
  if (grid) {  // the grid bucket itself is built by the first projection-gated search (ensure_grid)
    s->grid = *grid;
    s->has_grid = true;
    s->cell_start = (int32_t*)(base + o_cs);
    s->cell_items = (int32_t*)(base + o_ci);
    s->cell_tmp = (int32_t*)(base + o_tmp);
  }
  CKS(cudaGetLastError());
  if (wait || !s->h_flag) {  // the caller's descriptor buffer may be reused as soon as we return
    CKS(cudaStreamSynchronize(ctx->stream));
    synced(ctx);
    if (!s->h_flag) {
      int32_t f = 1;
      CKS(cudaMemcpy(&f, s->d_flags, sizeof f, cudaMemcpyDeviceToHost));
      s->tc_ok = f ? 0 : 1;
    }
  }
#undef CKS
  *out = s;
  return ASDM_OK;
}

extern "C" int asdm_set_create(asdm_ctx* ctx, const float* desc, int n, int row_stride_bytes, const asdm_kp* kps,
                               const asdm_grid* grid, asdm_set** out) {
  return set_create_impl(ctx, desc, n, row_stride_bytes, kps, grid, out, true);
}

extern "C" int asdm_set_create_async(asdm_ctx* ctx, const float* desc, int n, int row_stride_bytes, const asdm_kp* kps,
                                     const asdm_grid* grid, asdm_set** out) {
  return set_create_impl(ctx, desc, n, row_stride_bytes, kps, grid, out, false);
}

// A batch of sets (asdmatch.h). Copy-engine work per batch: the pack table, ONE copy with the small arrays of every
// set (range flag | keypoints | FeatureVector, into a side block the sets share), then the descriptor DMAs back to
// back; one pack launch on the preparation stream.
extern "C" int asdm_set_create_batch(asdm_ctx* ctx, const asdm_set_desc* descs, int n_sets, asdm_set** out) {
  if (!ctx || n_sets < 0 || (n_sets > 0 && (!descs || !out))) return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  for (int k = 0; k < n_sets; k++) out[k] = nullptr;
  size_t side_bytes = 0;
  std::vector<size_t> side_off((size_t)n_sets);
  for (int k = 0; k < n_sets; k++) {
    const asdm_set_desc& d = descs[k];
    if (d.n < 0 || (d.n > 0 && !d.desc)) return fail(ctx, ASDM_ERR_INVALID, "bad set description");
    if (d.row_stride_bytes != 0 && d.row_stride_bytes < (int)(ASDM_DESC_DIM * sizeof(float)))
      return fail(ctx, ASDM_ERR_INVALID, "row stride < 512 B");
    if (d.grid && (!d.kps || d.grid->cols <= 0 || d.grid->rows <= 0 || (int64_t)d.grid->cols * d.grid->rows > (1 << 20)))
      return fail(ctx, ASDM_ERR_INVALID, "bad grid");
    if (d.featvec && d.featvec->n_nodes < 0) return fail(ctx, ASDM_ERR_INVALID, "bad FeatureVector");
    const size_t n_pad = align_up((size_t)std::max(d.n, 1), 256);
    side_off[k] = side_bytes;
    side_bytes += 256 + (d.kps ? n_pad * sizeof(asdm_kp) : 0) + (d.featvec ? n_pad * 8 : 0);
  }
  if (n_sets == 0) return ASDM_OK;
  // rows that already live on this GPU (keyframes replicated over NVLink by an all-gather queued on the main stream):
  // the "upload" is then a device-to-device copy, ordered behind whatever the main stream holds so far
  bool any_on_device = false;
  for (int k = 0; k < n_sets && !any_on_device; k++) {
    cudaPointerAttributes at;
    if (descs[k].n > 0 && cudaPointerGetAttributes(&at, descs[k].desc) == cudaSuccess) any_on_device = at.type == cudaMemoryTypeDevice;
    else cudaGetLastError();
  }
  if (any_on_device) {
    CK(cudaEventRecord(ctx->prep_done, ctx->stream));
    CK(cudaStreamWaitEvent(ctx->upload_stream, ctx->prep_done, 0));
  }
  SideBlock* side = new (std::nothrow) SideBlock;
  if (!side) return fail(ctx, ASDM_ERR_OOM, "host allocation");
  auto bail = [&](int code) {
    for (int k = 0; k < n_sets; k++) {
      if (out[k]) set_free(out[k]);
      out[k] = nullptr;
    }
    if (side->ref == 0) {  // no set took a share yet
      dev_free(ctx, side->p);
      delete side;
    }
    return code;
  };
  const size_t table_bytes = align_up((size_t)n_sets * sizeof(PackItem), 256);
  void* h_stage = nullptr;
  int rc = stage_alloc(ctx, table_bytes + side_bytes, &h_stage);
  if (rc) return bail(rc);
  PackItem* items = (PackItem*)h_stage;
  char* h_side = (char*)h_stage + table_bytes;
  // the table sits in front of the side block on the device as well: one copy brings both
  cudaError_t e = cudaMallocAsync(&side->p, table_bytes + side_bytes, ctx->upload_stream);
  if (e != cudaSuccess) return bail(fail(ctx, ASDM_ERR_OOM, "cudaMallocAsync(side block)", e));
  char* d_side = (char*)side->p + table_bytes;
  int max_pad = 0;
  for (int k = 0; k < n_sets; k++) {
    const asdm_set_desc& d = descs[k];
    asdm_set* s = new (std::nothrow) asdm_set();
    if (!s) return bail(fail(ctx, ASDM_ERR_OOM, "host allocation"));
    out[k] = s;
    s->ctx = ctx;
    s->side = side;
    side->ref++;
    s->n = d.n;
    s->n_pad = (int)align_up((size_t)std::max(d.n, 1), 256);
    max_pad = std::max(max_pad, s->n_pad);
    const int nc = d.grid ? d.grid->cols * d.grid->rows : 0;
    size_t off = 0;
    const size_t o_desc = off; off += align_up((size_t)s->n_pad * ASDM_DESC_DIM * sizeof(float), 1024);
    const size_t o_bf = off;   off += align_up((size_t)s->n_pad * ASDM_DESC_DIM * sizeof(__nv_bfloat16), 1024);
    const size_t o_norm = off; off += align_up((size_t)s->n_pad * sizeof(float), 256);
    const size_t o_cs = off;   off += d.grid ? align_up((size_t)(nc + 1) * sizeof(int32_t), 256) : 0;
    const size_t o_ci = off;   off += d.grid ? align_up((size_t)s->n_pad * sizeof(int32_t), 256) : 0;
    const size_t o_tmp = off;  off += d.grid ? align_up((size_t)s->n_pad * sizeof(int32_t), 256) : 0;
    char* base = nullptr;
    e = cudaMallocAsync((void**)&base, off, ctx->upload_stream);
    if (e != cudaSuccess) return bail(fail(ctx, ASDM_ERR_OOM, "cudaMallocAsync(set)", e));
    s->desc = (float*)(base + o_desc);
    s->desc_bf = (__nv_bfloat16*)(base + o_bf);
    s->norm2 = (float*)(base + o_norm);
    if (d.grid) {
      s->grid = *d.grid;
      s->has_grid = true;
      s->cell_start = (int32_t*)(base + o_cs);
      s->cell_items = (int32_t*)(base + o_ci);
      s->cell_tmp = (int32_t*)(base + o_tmp);
    }
    // the descriptors start crossing PCIe at once; the small arrays are staged below, while the copy engine works
    if (d.n > 0) {
      const int stride = d.row_stride_bytes ? d.row_stride_bytes : (int)(ASDM_DESC_DIM * sizeof(float));
      if (stride == (int)(ASDM_DESC_DIM * sizeof(float)))
        e = cudaMemcpyAsync(s->desc, d.desc, (size_t)d.n * ASDM_DESC_DIM * sizeof(float), cudaMemcpyDefault, ctx->upload_stream);
      else
        e = cudaMemcpy2DAsync(s->desc, ASDM_DESC_DIM * sizeof(float), d.desc, stride, ASDM_DESC_DIM * sizeof(float), d.n,
                              cudaMemcpyDefault, ctx->upload_stream);
      if (e != cudaSuccess) return bail(fail(ctx, ASDM_ERR_CUDA, "descriptor upload", e));
    }
  }
  for (int k = 0; k < n_sets; k++) {
    const asdm_set_desc& d = descs[k];
    asdm_set* s = out[k];
    // side block share: flags | keypoints | node ids | FeatureVector positions
    size_t so = side_off[k];
    s->d_flags = (int32_t*)(d_side + so);
    memset(h_side + so, 0, 256);
    so += 256;
    if (d.kps) {
      s->kps = (asdm_kp*)(d_side + so);
      s->has_kps = true;
      memcpy(h_side + so, d.kps, (size_t)d.n * sizeof(asdm_kp));
      memset(h_side + so + (size_t)d.n * sizeof(asdm_kp), 0, (size_t)(s->n_pad - d.n) * sizeof(asdm_kp));
      so += (size_t)s->n_pad * sizeof(asdm_kp);
    }
    if (d.featvec) {
      s->node_id = (uint32_t*)(d_side + so);
      s->fv_pos = (int32_t*)(d_side + so + (size_t)s->n_pad * 4);
      uint32_t* hn = (uint32_t*)(h_side + so);
      if ((rc = flatten_featvec(ctx, d.featvec, d.n, s->n_pad, hn, (int32_t*)(hn + s->n_pad)))) return bail(rc);
      s->has_fv = true;
    }
    s->h_flag = slot_get(ctx);
    items[k] = PackItem{s->desc, s->desc_bf, s->norm2, s->d_flags, s->n, s->n_pad, s->h_flag};
  }
#define CKB(call)                                              \
  do {                                                         \
    cudaError_t e__ = (call);                                  \
    if (e__ != cudaSuccess) return bail(fail(ctx, ASDM_ERR_CUDA, #call, e__)); \
  } while (0)
  CKB(cudaMemcpyAsync(side->p, h_stage, table_bytes + side_bytes, cudaMemcpyHostToDevice, ctx->upload_stream));
  CKB(cudaEventRecord(ctx->upload_done, ctx->upload_stream));
  CKB(cudaStreamWaitEvent(ctx->prep_stream, ctx->upload_done, 0));
  ctx->stat_launches = 0;
  k_pack_batch<<<dim3(std::min((max_pad * 32 + 255) / 256, 64), n_sets), 256, 0, ctx->prep_stream>>>((const PackItem*)side->p);
  LAUNCHED(ctx);
  CKB(cudaEventRecord(ctx->prep_done, ctx->prep_stream));
  CKB(cudaStreamWaitEvent(ctx->stream, ctx->prep_done, 0));
  CKB(cudaGetLastError());
  bool need_sync = false;
  for (int k = 0; k < n_sets; k++) need_sync |= !out[k]->h_flag;
  if (need_sync) {  // pinned flag slots exhausted: read the flags now
    CKB(cudaStreamSynchronize(ctx->stream));
    synced(ctx);
    for (int k = 0; k < n_sets; k++)
      if (!out[k]->h_flag) {
        int32_t f = 1;
        CKB(cudaMemcpy(&f, out[k]->d_flags, sizeof f, cudaMemcpyDeviceToHost));
        out[k]->tc_ok = f ? 0 : 1;
      }
  }
#undef CKB
  return ASDM_OK;
}

extern "C" int asdm_set_retain(asdm_set* s) {
  if (!s) return ASDM_ERR_INVALID;
  s->ref.fetch_add(1);
  return ASDM_OK;
}

extern "C" int asdm_set_release(asdm_set* s) {
  if (!s) return ASDM_ERR_INVALID;
  if (s->ref.fetch_sub(1) == 1) {
    asdm_ctx* ctx = s->ctx;
    std::lock_guard<std::recursive_mutex> lk(ctx->mu);
    set_free(s);  // stream-ordered free: searches already queued on this set finish first
  }
  return ASDM_OK;
}

extern "C" int asdm_set_size(const asdm_set* s) { return s ? s->n : ASDM_ERR_INVALID; }

// DBoW2::FeatureVector -> per-feature node id + CSR position (-1 = not in the vector)
static int flatten_featvec(asdm_ctx* ctx, const asdm_featvec* fv, int n, int n_pad, uint32_t* node, int32_t* pos) {
  memset(node, 0, (size_t)n_pad * 4);
  memset(pos, 0xff, (size_t)n_pad * 4);
  for (int k = 0; k < fv->n_nodes; k++) {
    if (k > 0 && !(fv->node_ids[k] > fv->node_ids[k - 1])) return fail(ctx, ASDM_ERR_INVALID, "node ids not ascending");
    for (int e = fv->off[k]; e < fv->off[k + 1]; e++) {
      const int i = fv->idx[e];
      if (i < 0 || i >= n) return fail(ctx, ASDM_ERR_INVALID, "feature index out of range");
      if (pos[i] >= 0) return fail(ctx, ASDM_ERR_INVALID, "feature listed twice in the FeatureVector");
      node[i] = fv->node_ids[k];
      pos[i] = e;
    }
  }
  return ASDM_OK;
}

extern "C" int asdm_set_featvec(asdm_set* s, const asdm_featvec* fv) {
  if (!s || !fv || fv->n_nodes < 0) return ASDM_ERR_INVALID;
  asdm_ctx* ctx = s->ctx;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  // per-feature node id and scan position: the flattening of DBoW2::FeatureVector the kernels gate on, written
  // straight into the pinned staging arena
  void* h = nullptr;
  int rc = stage_alloc(ctx, (size_t)s->n_pad * 8, &h);
  if (rc) return rc;
  if ((rc = flatten_featvec(ctx, fv, s->n, s->n_pad, (uint32_t*)h, (int32_t*)((uint32_t*)h + s->n_pad)))) return rc;
  if (!s->node_id) {
    CK(cudaMallocAsync(&s->fv_mem, (size_t)s->n_pad * 8, ctx->upload_stream));
    s->node_id = (uint32_t*)s->fv_mem;
    s->fv_pos = (int32_t*)((char*)s->fv_mem + (size_t)s->n_pad * 4);
  } else {
    // replacing a FeatureVector: searches already queued on the main stream must have read the old one
    CK(cudaEventRecord(ctx->upload_done, ctx->stream));
    CK(cudaStreamWaitEvent(ctx->upload_stream, ctx->upload_done, 0));
  }
  CK(cudaMemcpyAsync(s->node_id, h, (size_t)s->n_pad * 8, cudaMemcpyHostToDevice, ctx->upload_stream));
  CK(cudaEventRecord(ctx->upload_done, ctx->upload_stream));
  CK(cudaStreamWaitEvent(ctx->stream, ctx->upload_done, 0));
  s->has_fv = true;
  return ASDM_OK;
}

extern "C" int asdm_set_read_desc(asdm_set* s, int row0, int n_rows, float* out) {
  if (!s || !out || row0 < 0 || n_rows < 0 || row0 + n_rows > s->n) return ASDM_ERR_INVALID;
  asdm_ctx* ctx = s->ctx;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  CK(cudaMemcpyAsync(out, s->desc + (size_t)row0 * ASDM_DESC_DIM, (size_t)n_rows * ASDM_DESC_DIM * sizeof(float),
                     cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  synced(ctx);
  return ASDM_OK;
}

// chamo.map descriptor blocks (visual_map_seri.cc:134-146 / :272-285): [int32 128][int32 n][n x 128 fp32]
extern "C" size_t asdm_desc_block_bytes(int n) {
  return n < 0 ? 0 : 8 + (size_t)n * ASDM_DESC_DIM * sizeof(float);
}

extern "C" int asdm_set_save_blocks(asdm_ctx* ctx, asdm_set* const* sets, int n_sets, void* out, size_t cap,
                                    size_t* offsets) {
  if (!ctx || n_sets < 0 || (n_sets > 0 && (!sets || !out))) return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  size_t off = 0;
  for (int k = 0; k < n_sets; k++) {
    if (!sets[k] || sets[k]->ctx != ctx) return fail(ctx, ASDM_ERR_INVALID, "set of another context");
    off += asdm_desc_block_bytes(sets[k]->n);
  }
  if (off > cap) return fail(ctx, ASDM_ERR_INVALID, "output buffer too small for the descriptor blocks");
  off = 0;
  for (int k = 0; k < n_sets; k++) {
    const asdm_set* s = sets[k];
    if (offsets) offsets[k] = off;
    const int32_t hdr[2] = {ASDM_DESC_DIM, s->n};
    memcpy((char*)out + off, hdr, 8);
    if (s->n)
      CK(cudaMemcpyAsync((char*)out + off + 8, s->desc, (size_t)s->n * ASDM_DESC_DIM * sizeof(float), cudaMemcpyDeviceToHost,
                         ctx->stream));
    off += asdm_desc_block_bytes(s->n);
  }
  if (offsets) offsets[n_sets] = off;
  CK(cudaStreamSynchronize(ctx->stream));
  synced(ctx);
  return ASDM_OK;
}

extern "C" int asdm_set_load_block(asdm_ctx* ctx, const void* block, size_t bytes, const asdm_kp* kps,
                                   const asdm_grid* grid, asdm_set** out, size_t* consumed) {
  if (!ctx || !block || !out) return ASDM_ERR_INVALID;
  if (bytes < 8) return fail(ctx, ASDM_ERR_INVALID, "descriptor block truncated (no header)");
  int32_t hdr[2];
  memcpy(hdr, block, 8);
  if (hdr[0] != ASDM_DESC_DIM) return fail(ctx, ASDM_ERR_INVALID, "descriptor block is not 128 floats wide (not an ASD map)");
  if (hdr[1] < 0 || asdm_desc_block_bytes(hdr[1]) > bytes) return fail(ctx, ASDM_ERR_INVALID, "descriptor block truncated");
  // the payload starts 8 bytes into the block: rows are only 4-byte aligned in the file, which the copy engine accepts
  const int rc = set_create_impl(ctx, (const float*)((const char*)block + 8), hdr[1], 0, kps, grid, out, false);
  if (rc == ASDM_OK && consumed) *consumed = asdm_desc_block_bytes(hdr[1]);
  return rc;
}

// ------------------------------------------------------------------------------- batched keyframe-pair jobs
struct asdm_job {
  asdm_ctx* ctx = nullptr;
  int kind = 0;
  int n_pairs = 0;
  int n_stride = 0;    // max nA over pairs (query slots per pair)
  int m_stride = 0;    // max nB over pairs
  int out_stride = 0;  // entries per pair in match12 / dist12
  PublishArgs pub{};   // gathered arrays of every rank the resolve kernel also stores into (asdm_job_set_publish)
  BowParams P{};
  std::vector<asdm_set*> sets;  // retained
  std::vector<PairDesc> h_pairs;
  DevBuf d_pairs, d_masks, d_cnt, d_cand, d_owner, d_acc, d_flags, d_match, d_dist, d_nm;
  DevBuf d_work, d_tmaps, d_list, d_seg_count, d_rb, d_cmin, d_setdesc;
  // d_masks / d_pairs / d_work / d_tmaps / d_setdesc are views into d_in (every input of the job, brought over by ONE
  // pull); d_flags / d_seg_count / d_cnt are views into d_zero (everything a run has to clear, ONE fill)
  DevBuf d_in, d_zero;
  size_t zero_bytes = 0;
  int cap = 8;
  int use_tc = 0;
  int n_work = 0;
  int list_cap = 0;  // total hit slots; split evenly into n_seg segments of seg_cap
  int n_seg = 0, seg_cap = 0;
  float t_keep_tc = 0.f;
  bool ran = false;
  void* prefetched[3] = {nullptr, nullptr, nullptr};  // destinations of asdm_job_fetch_async
  bool prefetch_valid = false;
  bool tc_pending = false;     // tensor-core route taken before every set's range flag was back (validated in settle)
  int32_t* h_flags = nullptr;  // pinned, 8 ints: [0] cand overflow, [1] list overflow, [2] list count
  bool h_flags_owned = false;  // true when the context's slot slab was exhausted and the job allocated its own
};

static void job_free(asdm_job* j) {
  if (!j) return;
  asdm_ctx* ctx = j->ctx;
  cudaSetDevice(ctx->device);
  if (j->prefetch_valid && ctx->downloads_pending) {  // destroyed between fetch_async and fetch
    cudaStreamSynchronize(ctx->download_stream);
    ctx->downloads_pending = false;
  }
  DevBuf* bufs[] = {&j->d_pairs, &j->d_masks, &j->d_cnt, &j->d_cand, &j->d_owner, &j->d_acc, &j->d_flags,
                    &j->d_match, &j->d_dist, &j->d_nm, &j->d_work, &j->d_tmaps, &j->d_list, &j->d_seg_count,
                    &j->d_rb, &j->d_cmin, &j->d_setdesc, &j->d_in, &j->d_zero};
  for (DevBuf* b : bufs) release(ctx, *b);  // stream-ordered: a run still in flight completes first
  if (j->h_flags_owned) {
    cudaStreamSynchronize(ctx->stream);
    synced(ctx);
    cudaFreeHost(j->h_flags);
  } else {
    slot_put(ctx, j->h_flags);
  }
  for (asdm_set* s : j->sets) asdm_set_release(s);
  delete j;
}

static int job_alloc_scratch(asdm_job* j) {
  asdm_ctx* ctx = j->ctx;
  const size_t P = (size_t)j->n_pairs;
  int rc;
  if (j->use_tc) j->n_seg = ctx->sm_count * ASDM_TC_MAX_EPI_WARPS;
  {
    const size_t o_seg = 256, o_cnt = o_seg + align_up((size_t)(j->use_tc ? j->n_seg : 0) * sizeof(int32_t), 256);
    j->zero_bytes = o_cnt + align_up(P * j->n_stride * sizeof(int32_t), 256);
    if ((rc = ensure(ctx, j->d_zero, j->zero_bytes))) return rc;
    char* z = (char*)j->d_zero.p;
    j->d_flags = DevBuf{z, 0};
    j->d_seg_count = DevBuf{z + o_seg, 0};
    j->d_cnt = DevBuf{z + o_cnt, 0};
  }
  if ((rc = ensure(ctx, j->d_cand, P * j->n_stride * (size_t)j->cap * sizeof(Cand)))) return rc;
  if ((rc = ensure(ctx, j->d_owner, P * 2 * j->m_stride * sizeof(int32_t)))) return rc;
  if ((rc = ensure(ctx, j->d_acc, P * j->n_stride * sizeof(int32_t)))) return rc;
  if ((rc = ensure(ctx, j->d_match, P * j->out_stride * sizeof(int32_t)))) return rc;
  if ((rc = ensure(ctx, j->d_dist, P * j->out_stride * sizeof(float)))) return rc;
  if ((rc = ensure(ctx, j->d_nm, P * sizeof(int32_t)))) return rc;
  if (j->use_tc) {
    {  // only the CTAs that get work fill their segments
      const int active = std::max(1, std::min(j->n_work, ctx->sm_count)) * ASDM_TC_MAX_EPI_WARPS;
      j->seg_cap = std::max(1024, (j->list_cap + active - 1) / active);
    }
    if ((rc = ensure(ctx, j->d_list, (size_t)j->seg_cap * j->n_seg * sizeof(TcHit)))) return rc;
  }
  return ASDM_OK;
}

static int job_prepare(asdm_ctx* ctx, const asdm_job_desc* d, int kind_internal, asdm_job** out) {
  if (!ctx || !d || !out) return ASDM_ERR_INVALID;
  *out = nullptr;
  if (d->n_kf <= 0 || d->n_pairs < 0 || !d->sets || !d->masks || (d->n_pairs > 0 && (!d->ia || !d->ib)))
    return fail(ctx, ASDM_ERR_INVALID, "bad job description");
  if (kind_internal == KIND_TRIANGULATION &&
      (!d->F12 || !d->epipoles || !d->scale_factors || !d->level_sigma2 || d->n_levels <= 0 || d->n_levels > ASDM_MAX_LEVELS))
    return fail(ctx, ASDM_ERR_INVALID, "triangulation job needs F12, epipoles and level tables");
  if ((d->row_begin || d->row_end) && (kind_internal != KIND_TRIANGULATION || d->check_ori || !d->row_begin || !d->row_end))
    return fail(ctx, ASDM_ERR_INVALID, "row ranges need both arrays and a triangulation job without rotation check");
  for (int k = 0; k < d->n_kf; k++) {
    const asdm_set* s = d->sets[k];
    if (!s || s->ctx != ctx || !d->masks[k]) return fail(ctx, ASDM_ERR_INVALID, "bad keyframe set / mask");
    if (!s->has_fv) return fail(ctx, ASDM_ERR_INVALID, "keyframe set has no FeatureVector (asdm_set_featvec)");
    if (!s->has_kps) return fail(ctx, ASDM_ERR_INVALID, "keyframe set has no keypoints");
  }
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  asdm_job* j = new (std::nothrow) asdm_job();
  if (!j) return fail(ctx, ASDM_ERR_OOM, "host allocation");
  j->ctx = ctx;
  j->kind = kind_internal;
  j->n_pairs = d->n_pairs;
  auto bail = [&](int code) {
    job_free(j);
    return code;
  };
  for (int k = 0; k < d->n_kf; k++) {
    asdm_set_retain(d->sets[k]);
    j->sets.push_back(d->sets[k]);
  }
  // masks of all keyframes, packed
  std::vector<size_t> mask_off(d->n_kf);
  size_t mtot = 0;
  for (int k = 0; k < d->n_kf; k++) {
    mask_off[k] = mtot;
    mtot += align_up((size_t)std::max(d->sets[k]->n, 1), 16);
  }
  int rc;
  // Every input of the job lives in ONE device block (d_in) that is filled by ONE pull from the staging arena:
  // masks | pair table | tensor maps | per-keyframe bound descriptors | chunk -> keyframe map | work list.
  // Sizes are known after a pre-pass over the pairs (the work list is the only data-dependent one).
  const long target_units = (long)ctx->sm_count * 4;
  long tc_tiles = 0;
  auto m_range_of = [&](int p, int& m_lo, int& m_hi) {  // 128-row tiles of set A that hold this unit's query rows
    const int nA = d->sets[d->ia[p]]->n;
    const int r0 = d->row_begin ? d->row_begin[p] : 0, r1 = d->row_end ? d->row_end[p] : nA;
    m_lo = r0 / 128;
    m_hi = r1 > r0 ? (r1 + 127) / 128 : m_lo;
  };
  for (int p = 0; p < d->n_pairs; p++) {
    const int a = d->ia[p], b = d->ib[p];
    if (a < 0 || a >= d->n_kf || b < 0 || b >= d->n_kf) return bail(fail(ctx, ASDM_ERR_INVALID, "pair index out of range"));
    if (d->row_begin && (d->row_begin[p] < 0 || d->row_end[p] < d->row_begin[p] || d->row_end[p] > d->sets[a]->n))
      return bail(fail(ctx, ASDM_ERR_INVALID, "row range out of bounds"));
    int m_lo, m_hi;
    m_range_of(p, m_lo, m_hi);
    tc_tiles += (long)((d->sets[b]->n + 255) / 256) * (m_hi - m_lo);
  }
  const int tc_run = (int)std::max<long>(1, std::min<long>(32, tc_tiles / std::max<long>(1, target_units)));
  size_t work_bound = 0, n_chunks = 0;
  for (int p = 0; p < d->n_pairs; p++) {
    int m_lo, m_hi;
    m_range_of(p, m_lo, m_hi);
    work_bound += (size_t)((d->sets[d->ib[p]]->n + 255) / 256) * (size_t)((m_hi - m_lo + tc_run - 1) / tc_run);
  }
  for (int k = 0; k < d->n_kf; k++) n_chunks += (size_t)d->sets[k]->n_pad / 128;
  size_t in_off = 0;
  auto carve = [&](size_t bytes) { const size_t o = in_off; in_off += align_up(std::max(bytes, (size_t)1), 256); return o; };
  const size_t o_masks = carve(mtot), o_pairs = carve((size_t)d->n_pairs * sizeof(PairDesc)),
               o_tmaps = carve((size_t)d->n_kf * 128), o_sd = carve((size_t)d->n_kf * sizeof(TcSetDesc)),
               o_chunk = carve(n_chunks * sizeof(int32_t)), o_work = carve(work_bound * sizeof(TcWork));
  if ((rc = ensure(ctx, j->d_in, in_off))) return bail(rc);
  void* h_in_v = nullptr;
  if ((rc = stage_alloc(ctx, in_off, &h_in_v))) return bail(rc);
  char* h_in = (char*)h_in_v;
  char* d_in = (char*)j->d_in.p;
  j->d_masks = DevBuf{d_in + o_masks, 0};
  j->d_pairs = DevBuf{d_in + o_pairs, 0};
  j->d_tmaps = DevBuf{d_in + o_tmaps, 0};
  j->d_setdesc = DevBuf{d_in + o_sd, 0};
  j->d_work = DevBuf{d_in + o_work, 0};
  for (int k = 0; k < d->n_kf; k++) memcpy(h_in + o_masks + mask_off[k], d->masks[k], (size_t)d->sets[k]->n);
  size_t in_used = o_tmaps;  // masks + pairs always travel; the tensor-core tables only when that route is taken
  size_t row_bound_chunks = 0;

  BowParams& P = j->P;
  memset(&P, 0, sizeof P);
  P.kind = kind_internal;
  P.nnratio = d->nnratio;
  P.check_ori = d->check_ori ? 1 : 0;
  P.th_accept = ASDM_TH_LOW;
  if (kind_internal == KIND_TRIANGULATION) {
    P.mask_sense = 0;          // usable features are those WITHOUT a map point (ORBmatcher.cc:708-712, :730)
    P.t_keep = ASDM_TH_LOW;    // :737
    P.n_levels = d->n_levels;
    for (int l = 0; l < d->n_levels; l++) {
      P.scale_factors[l] = d->scale_factors[l];
      P.level_sigma2[l] = d->level_sigma2[l];
    }
  } else {
    P.mask_sense = 1;
    P.strict = kind_internal == KIND_BOW_KF_KF ? 1 : 0;  // :609 vs :231
    P.t_keep = bow_keep_threshold(d->nnratio, ASDM_TH_LOW);
  }

  j->h_pairs.resize((size_t)d->n_pairs);
  int all_tc = 1;
  for (int p = 0; p < d->n_pairs; p++) {
    const int a = d->ia[p], b = d->ib[p];
    if (a < 0 || a >= d->n_kf || b < 0 || b >= d->n_kf) return bail(fail(ctx, ASDM_ERR_INVALID, "pair index out of range"));
    const asdm_set *A = d->sets[a], *B = d->sets[b];
    PairDesc& pd = j->h_pairs[p];
    memset(&pd, 0, sizeof pd);
    pd.descA = A->desc; pd.descB = B->desc;
    pd.bfA = A->desc_bf; pd.bfB = B->desc_bf;
    pd.normA = A->norm2; pd.normB = B->norm2;
    pd.nodeA = A->node_id; pd.nodeB = B->node_id;
    pd.posA = A->fv_pos; pd.posB = B->fv_pos;
    pd.maskA = (const uint8_t*)j->d_masks.p + mask_off[a];
    pd.maskB = (const uint8_t*)j->d_masks.p + mask_off[b];
    pd.kpA = A->kps; pd.kpB = B->kps;
    pd.nA = A->n; pd.nB = B->n;
    pd.rowA0 = 0; pd.rowA1 = A->n;
    if (d->row_begin) {
      if (d->row_begin[p] < 0 || d->row_end[p] < d->row_begin[p] || d->row_end[p] > A->n)
        return bail(fail(ctx, ASDM_ERR_INVALID, "row range out of bounds"));
      pd.rowA0 = d->row_begin[p];
      pd.rowA1 = d->row_end[p];
    }
    pd.setA = a; pd.setB = b;
    if (kind_internal == KIND_TRIANGULATION) {
      memcpy(pd.F12, d->F12 + 9 * (size_t)p, 9 * sizeof(float));
      pd.ex = d->epipoles[2 * p];
      pd.ey = d->epipoles[2 * p + 1];
    }
    j->n_stride = std::max(j->n_stride, A->n);
    j->m_stride = std::max(j->m_stride, B->n);
  }
  j->n_stride = std::max(j->n_stride, 1);
  j->m_stride = std::max(j->m_stride, 1);
  j->out_stride = kind_internal == KIND_BOW_KF_F ? j->m_stride : j->n_stride;
  j->cap = ctx->opt_cand_cap;

  // tensor-core shortlist only when every set passed the range check of k_pack_desc
  if (!ctx->opt_force_exact && d->n_pairs > 0) {
    // The range flag of a set travels back behind its k_pack_desc. When some flags are still in flight (keyframes
    // uploaded asynchronously just before this call) the job does not wait for them: it takes the tensor-core route
    // speculatively and job_settle() checks the flags once the run has finished anyway (a raised flag = rerun on the
    // exact route; results are only ever taken from a validated run). So preparing and launching a job never blocks
    // on uploads that are still crossing PCIe -- later keyframes keep streaming in while this job computes.
    bool pending = false;
    for (int k = 0; k < d->n_kf; k++) pending |= d->sets[k]->tc_ok < 0;
    if (pending && cudaStreamQuery(ctx->stream) == cudaSuccess) {  // nothing in flight: the flags are already here
      if ((rc = resolve_tc_flags(ctx, d->sets, d->n_kf))) return bail(rc);
      pending = false;
    }
    cudaGetLastError();  // cudaErrorNotReady from the query is not an error
    for (int k = 0; k < d->n_kf; k++) all_tc &= d->sets[k]->tc_ok != 0;
    j->tc_pending = pending && all_tc;
    if (all_tc) {
      // work units: one resident 256-row B tile x a run of 128-row A tiles; split runs until the grid is full
      TcWork* work = (TcWork*)(h_in + o_work);
      size_t n_work = 0;
      for (int p = 0; p < d->n_pairs; p++) {
        const int nt_n = (j->h_pairs[p].nB + 255) / 256;
        int m_lo, m_hi;
        m_range_of(p, m_lo, m_hi);
        for (int tn = 0; tn < nt_n; tn++)
          for (int m0 = m_lo; m0 < m_hi; m0 += tc_run) work[n_work++] = TcWork{p, tn, m0, std::min(tc_run, m_hi - m0)};
      }
      bool maps_ok = true;
      for (int k = 0; k < d->n_kf && maps_ok; k++) maps_ok = asdm_tc_build_tmap(ctx, d->sets[k]) == ASDM_OK;
      if (maps_ok && n_work > 0) {
        j->use_tc = 1;
        j->n_work = (int)n_work;
        for (int k = 0; k < d->n_kf; k++) memcpy(h_in + o_tmaps + (size_t)k * 128, d->sets[k]->tmap_bf.data(), 128);
        // per-keyframe filter bounds (k_tc_row_bounds): one launch over all 128-row chunks of all keyframes
        {
          TcSetDesc* sd = (TcSetDesc*)(h_in + o_sd);
          int32_t* chunk_set = (int32_t*)(h_in + o_chunk);
          size_t nch = 0;
          for (int k = 0; k < d->n_kf; k++) {
            const asdm_set* S = d->sets[k];
            sd[k] = TcSetDesc{S->norm2, (const uint8_t*)j->d_masks.p + mask_off[k], S->fv_pos, S->n, (int32_t)nch};
            for (size_t c = 0; c < (size_t)S->n_pad / 128; c++) chunk_set[nch++] = k;
          }
          if ((rc = ensure(ctx, j->d_rb, nch * 128 * sizeof(float))) || (rc = ensure(ctx, j->d_cmin, nch * sizeof(float))))
            return bail(rc);
          row_bound_chunks = nch;  // launched below, once the inputs are on the device
          for (int p = 0; p < d->n_pairs; p++) {
            PairDesc& pd = j->h_pairs[p];
            pd.rbA = (const float*)j->d_rb.p + (size_t)sd[pd.setA].chunk0 * 128;
            pd.cminB = (const float*)j->d_cmin.p + sd[pd.setB].chunk0;
          }
        }
        in_used = o_work + align_up(n_work * sizeof(TcWork), 256);
        long q = 0;
        for (int p = 0; p < d->n_pairs; p++) q += j->h_pairs[p].nA;
        j->list_cap = (int)std::min<long>((long)1 << 30, std::max<long>(4096, q * ctx->opt_list_cap_per_query));
      }
    }
  }
  if (!j->h_pairs.empty()) memcpy(h_in + o_pairs, j->h_pairs.data(), j->h_pairs.size() * sizeof(PairDesc));
  if ((rc = pull_to_device(ctx, d_in, h_in, in_used))) return bail(rc);
  if (row_bound_chunks) {
    if (asdm_tc_row_bounds_launch(ctx, (const TcSetDesc*)j->d_setdesc.p, (const int32_t*)(d_in + o_chunk), (int)row_bound_chunks,
                                  P.mask_sense, (float*)j->d_rb.p, (float*)j->d_cmin.p) != cudaSuccess)
      return bail(fail(ctx, ASDM_ERR_CUDA, "row-bound launch"));
    ctx->stat_total_launches++;
  }
  if ((rc = job_alloc_scratch(j))) return bail(rc);
  j->h_flags = slot_get(ctx);
  if (!j->h_flags) {
    if (cudaHostAlloc((void**)&j->h_flags, 64, cudaHostAllocDefault) != cudaSuccess)
      return bail(fail(ctx, ASDM_ERR_OOM, "pinned flags"));
    j->h_flags_owned = true;
  }
  *out = j;
  return ASDM_OK;
}

static int job_run(asdm_job* j) {
  asdm_ctx* ctx = j->ctx;
  CK(cudaSetDevice(ctx->device));
  ctx->stat_launches = 0;
  ctx->stat_tc_used = j->use_tc;
  ctx->stat_cand_cap = j->cap;
  j->ran = true;
  if (j->n_pairs == 0) return ASDM_OK;
  const PairDesc* pairs = (const PairDesc*)j->d_pairs.p;
  int32_t* flags = (int32_t*)j->d_flags.p;
  {
    const int rc = zero_device(ctx, j->d_zero.p, j->zero_bytes);  // flags | hit-list segment counts | candidate counts
    if (rc) return rc;
  }
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  if (ctx->opt_time_kernels) {
    CK(cudaEventCreate(&ev0));
    CK(cudaEventCreate(&ev1));
    CK(cudaEventRecord(ev0, ctx->stream));
  }
  if (j->use_tc) {
    CK(asdm_tc_shortlist_launch(ctx, pairs, j->d_tmaps.p, (const TcWork*)j->d_work.p, j->n_work, j->P.t_keep,
                                (TcHit*)j->d_list.p, j->seg_cap, (int32_t*)j->d_seg_count.p, flags + 2));
    LAUNCHED(ctx);
    if (ev1) CK(cudaEventRecord(ev1, ctx->stream));
    k_rerank<<<j->n_seg, 256, 0, ctx->stream>>>(pairs, j->P, j->cap, j->n_stride, (const TcHit*)j->d_list.p,
                                                (const int32_t*)j->d_seg_count.p, j->seg_cap, (int32_t*)j->d_cnt.p,
                                                (Cand*)j->d_cand.p, flags);
    LAUNCHED(ctx);
  } else {
    dim3 grid((j->n_stride + CE_TM - 1) / CE_TM, (j->m_stride + CE_TN - 1) / CE_TN, j->n_pairs);
    if (grid.y > 65535 || grid.z > 65535) return fail(ctx, ASDM_ERR_INVALID, "job too large for the exact kernel grid");
    k_cand_exact<<<grid, 256, ASDM_DESC_DIM * (CE_TM + CE_TN) * sizeof(float), ctx->stream>>>(
        pairs, j->P, j->cap, j->n_stride, (int32_t*)j->d_cnt.p, (Cand*)j->d_cand.p, flags);
    LAUNCHED(ctx);
    if (ev1) CK(cudaEventRecord(ev1, ctx->stream));
  }
  if (ev0) {
    ctx->ev.push_back(ev0);
    ctx->ev.push_back(ev1);
  }
  if (j->kind == KIND_TRIANGULATION)
    k_resolve_tri<<<j->n_pairs, 512, 0, ctx->stream>>>(pairs, j->P, j->cap, j->n_stride, j->out_stride,
                                                       (const int32_t*)j->d_cnt.p, (const Cand*)j->d_cand.p,
                                                       (int32_t*)j->d_acc.p, (int32_t*)j->d_match.p,
                                                       (float*)j->d_dist.p, (int32_t*)j->d_nm.p, j->pub);
  else
    k_resolve_bow<<<j->n_pairs, 512, 0, ctx->stream>>>(pairs, j->P, j->cap, j->n_stride, j->m_stride, j->out_stride,
                                                       (const int32_t*)j->d_cnt.p, (Cand*)j->d_cand.p,
                                                       (int32_t*)j->d_owner.p, (int32_t*)j->d_acc.p,
                                                       (int32_t*)j->d_match.p, (float*)j->d_dist.p,
                                                       (int32_t*)j->d_nm.p, j->pub);
  LAUNCHED(ctx);
  CK(cudaGetLastError());
  return ASDM_OK;
}

// Waits for the run; when a candidate pool overflowed, grows it and runs again (results are only ever taken
// from a run without overflow).
static int job_settle(asdm_job* j) {
  asdm_ctx* ctx = j->ctx;
  if (!j->ran) {
    int rc = job_run(j);
    if (rc) return rc;
  }
  ctx->stat_retries = 0;
  for (int attempt = 0; attempt < 24; attempt++) {
    CK(cudaMemcpyAsync(j->h_flags, j->d_flags.p, 16, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    synced(ctx);
    if (j->tc_pending) {  // the stream is idle now, so every set's flag copy has landed
      j->tc_pending = false;
      int rc = resolve_tc_flags(ctx, j->sets.data(), (int)j->sets.size());
      if (rc) return rc;
      bool all_tc = true;
      for (asdm_set* s : j->sets) all_tc &= s->tc_ok != 0;
      if (!all_tc && j->use_tc) {  // speculation failed: some descriptors are outside the shortlist's proven range
        j->use_tc = 0;
        ctx->stat_retries++;
        if ((rc = job_alloc_scratch(j)) || (rc = job_run(j))) return rc;
        continue;
      }
    }
    const bool cand_over = j->h_flags[0] != 0, list_over = j->h_flags[1] != 0;
    ctx->stat_tc_pairs = j->use_tc ? j->h_flags[2] : 0;
    if (!cand_over && !list_over) return ASDM_OK;
    ctx->stat_retries++;
    if (cand_over) j->cap = std::min(std::max(j->cap * 2, 2), std::max(j->m_stride, 2));
    if (list_over) {
      const long need = std::max<long>(2 * (long)j->h_flags[2] + 4096, (long)j->list_cap * 2);
      if (need > ((long)1 << 30)) j->use_tc = 0;  // absurdly dense: take the exact kernel
      else j->list_cap = (int)need;
    }
    int rc = job_alloc_scratch(j);
    if (rc) return rc;
    if ((rc = job_run(j))) return rc;
  }
  return fail(ctx, ASDM_ERR_INTERNAL, "candidate pool kept overflowing");
}

extern "C" int asdm_job_prepare(asdm_ctx* ctx, const asdm_job_desc* desc, asdm_job** out) {
  if (!desc) return ASDM_ERR_INVALID;
  if (desc->kind != ASDM_JOB_BOW_KF_KF && desc->kind != ASDM_JOB_TRIANGULATION)
    return fail(ctx, ASDM_ERR_INVALID, "unknown job kind");
  return job_prepare(ctx, desc, desc->kind == ASDM_JOB_BOW_KF_KF ? KIND_BOW_KF_KF : KIND_TRIANGULATION, out);
}

extern "C" int asdm_job_run(asdm_job* j) {
  if (!j) return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(j->ctx->mu);
  return job_run(j);
}

extern "C" int asdm_job_out_stride(const asdm_job* j) { return j ? j->out_stride : ASDM_ERR_INVALID; }

extern "C" int asdm_job_fetch(asdm_job* j, int32_t* match12, float* dist12, int32_t* nmatches) {
  if (!j) return ASDM_ERR_INVALID;
  asdm_ctx* ctx = j->ctx;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  const bool had_prefetch = j->prefetch_valid && j->prefetched[0] == match12 && j->prefetched[1] == dist12 &&
                            j->prefetched[2] == nmatches;
  j->prefetch_valid = false;
  int rc = job_settle(j);
  if (rc) return rc;
  // results queued by asdm_job_fetch_async stand when the run they were copied from was the validated one
  if (ctx->downloads_pending) {  // also before copying again: the two must not race on the destination
    CK(cudaStreamSynchronize(ctx->download_stream));
    ctx->downloads_pending = false;
  }
  if (had_prefetch && ctx->stat_retries == 0) return ASDM_OK;
  const size_t n = (size_t)j->n_pairs * j->out_stride;
  if (match12 && n) CK(cudaMemcpyAsync(match12, j->d_match.p, n * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
  if (dist12 && n) CK(cudaMemcpyAsync(dist12, j->d_dist.p, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
  if (nmatches && j->n_pairs)
    CK(cudaMemcpyAsync(nmatches, j->d_nm.p, (size_t)j->n_pairs * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  synced(ctx);
  return ASDM_OK;
}

extern "C" int asdm_job_fetch_async(asdm_job* j, int32_t* match12, float* dist12, int32_t* nmatches) {
  if (!j) return ASDM_ERR_INVALID;
  asdm_ctx* ctx = j->ctx;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  if (!j->ran) {
    const int rc = job_run(j);
    if (rc) return rc;
  }
  const size_t n = (size_t)j->n_pairs * j->out_stride;
  cudaStream_t ds = ctx->download_stream;
  CK(cudaEventRecord(ctx->run_done, ctx->stream));
  CK(cudaStreamWaitEvent(ds, ctx->run_done, 0));
  if (match12 && n) CK(cudaMemcpyAsync(match12, j->d_match.p, n * sizeof(int32_t), cudaMemcpyDeviceToHost, ds));
  if (dist12 && n) CK(cudaMemcpyAsync(dist12, j->d_dist.p, n * sizeof(float), cudaMemcpyDeviceToHost, ds));
  if (nmatches && j->n_pairs)
    CK(cudaMemcpyAsync(nmatches, j->d_nm.p, (size_t)j->n_pairs * sizeof(int32_t), cudaMemcpyDeviceToHost, ds));
  ctx->downloads_pending = true;
  j->prefetched[0] = match12; j->prefetched[1] = dist12; j->prefetched[2] = nmatches;
  j->prefetch_valid = true;
  return ASDM_OK;
}

extern "C" int asdm_job_settle(asdm_job* j) {
  if (!j) return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(j->ctx->mu);
  return job_settle(j);
}

extern "C" int asdm_job_device_outputs(asdm_job* j, void** match12, void** dist12, void** nmatches) {
  if (!j) return ASDM_ERR_INVALID;
  if (match12) *match12 = j->d_match.p;
  if (dist12) *dist12 = j->d_dist.p;
  if (nmatches) *nmatches = j->d_nm.p;
  return ASDM_OK;
}

// --------------------------------------------------------------- results published to every rank
extern "C" int asdm_shared_alloc(asdm_ctx* ctx, size_t bytes, void** dptr, uint8_t handle[ASDM_SHARED_HANDLE_BYTES]) {
  if (!ctx || !dptr || !handle || bytes == 0) return ASDM_ERR_INVALID;
  static_assert(sizeof(cudaIpcMemHandle_t) == ASDM_SHARED_HANDLE_BYTES, "handle size");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, bytes);  // not from the context's block cache: IPC handles name whole allocations
  if (e != cudaSuccess) return fail(ctx, ASDM_ERR_OOM, "shared allocation", e);
  cudaIpcMemHandle_t h;
  e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    return fail(ctx, ASDM_ERR_CUDA, "cudaIpcGetMemHandle", e);
  }
  memcpy(handle, &h, sizeof h);
  *dptr = p;
  return ASDM_OK;
}

extern "C" int asdm_shared_open(asdm_ctx* ctx, const uint8_t handle[ASDM_SHARED_HANDLE_BYTES], void** dptr) {
  if (!ctx || !dptr || !handle) return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof h);
  void* p = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) return fail(ctx, ASDM_ERR_CUDA, "cudaIpcOpenMemHandle (no peer access between these GPUs?)", e);
  *dptr = p;
  return ASDM_OK;
}

extern "C" int asdm_shared_close(asdm_ctx* ctx, void* dptr) {
  if (!ctx || !dptr) return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  CK(cudaStreamSynchronize(ctx->stream));
  CK(cudaIpcCloseMemHandle(dptr));
  return ASDM_OK;
}

extern "C" int asdm_shared_free(asdm_ctx* ctx, void* dptr) {
  if (!ctx || !dptr) return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  CK(cudaStreamSynchronize(ctx->stream));
  CK(cudaFree(dptr));
  return ASDM_OK;
}

extern "C" int asdm_job_set_publish(asdm_job* j, int n_ranks, void* const* match, void* const* dist, void* const* nmatches,
                                    int row0, int stride, int flags) {
  if (!j) return ASDM_ERR_INVALID;
  asdm_ctx* ctx = j->ctx;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  if (n_ranks == 0) {
    j->pub.n = 0;
    return ASDM_OK;
  }
  if (n_ranks < 0 || n_ranks > ASDM_MAX_PEERS || !match || !dist || !nmatches || row0 < 0)
    return fail(ctx, ASDM_ERR_INVALID, "publish: 1..16 destinations");
  if (stride < j->out_stride) return fail(ctx, ASDM_ERR_INVALID, "publish: gathered stride below the job's row length");
  for (int r = 0; r < n_ranks; r++)
    if (!match[r] || !dist[r] || !nmatches[r]) return fail(ctx, ASDM_ERR_INVALID, "publish: null destination");
  j->pub.n = n_ranks;
  j->pub.row0 = row0;
  j->pub.stride = stride;
  j->pub.multicast = (flags & ASDM_PUBLISH_MULTICAST) ? 1 : 0;
  j->pub.no_dist = (flags & ASDM_PUBLISH_NO_DIST) ? 1 : 0;
  for (int r = 0; r < n_ranks; r++) {
    j->pub.match[r] = (int32_t*)match[r];
    j->pub.dist[r] = (float*)dist[r];
    j->pub.nm[r] = (int32_t*)nmatches[r];
  }
  return ASDM_OK;
}

extern "C" int asdm_job_destroy(asdm_job* j) {
  if (!j) return ASDM_ERR_INVALID;
  asdm_ctx* ctx = j->ctx;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  job_free(j);
  return ASDM_OK;
}

// ------------------------------------------------------------------- single keyframe-pair entry points
static int single_pair(asdm_ctx* ctx, int kind, asdm_set* a, const uint8_t* mask_a, asdm_set* b, const uint8_t* mask_b,
                       float nnratio, int check_ori, const float* F12, float ex, float ey, const float* sf,
                       const float* ls, int n_levels, int32_t* match, float* dist, int32_t* nmatches) {
  if (!ctx || !a || !b || !mask_a || !match) return ASDM_ERR_INVALID;
  std::vector<uint8_t> ones;
  if (!mask_b) {
    ones.assign((size_t)std::max(b->n, 1), 1);
    mask_b = ones.data();
  }
  asdm_set* sets[2] = {a, b};
  const uint8_t* masks[2] = {mask_a, mask_b};
  const int32_t ia = 0, ib = 1;
  const float ep[2] = {ex, ey};
  asdm_job_desc d;
  memset(&d, 0, sizeof d);
  d.n_kf = 2;
  d.sets = sets;
  d.masks = masks;
  d.n_pairs = 1;
  d.ia = &ia;
  d.ib = &ib;
  d.nnratio = nnratio;
  d.check_ori = check_ori;
  d.F12 = F12;
  d.epipoles = ep;
  d.scale_factors = sf;
  d.level_sigma2 = ls;
  d.n_levels = n_levels;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  asdm_job* j = nullptr;
  int rc = job_prepare(ctx, &d, kind, &j);
  if (rc) return rc;
  rc = job_run(j);
  if (!rc) rc = job_settle(j);
  const int64_t launches = ctx->stat_launches;
  if (!rc) {
    int32_t nm = 0;
    const size_t n = (size_t)j->out_stride;
    cudaError_t e = cudaMemcpyAsync(match, j->d_match.p, n * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess && dist)
      e = cudaMemcpyAsync(dist, j->d_dist.p, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(&nm, j->d_nm.p, sizeof nm, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) rc = fail(ctx, ASDM_ERR_CUDA, "result download", e);
    if (nmatches) *nmatches = nm;
  }
  job_free(j);
  ctx->stat_launches = launches;
  return rc;
}

extern "C" int asdm_search_by_bow_kf_f(asdm_ctx* ctx, asdm_set* kf, const uint8_t* valid_kf, asdm_set* f, float nnratio,
                                       int check_ori, int32_t* match_f, float* dist_f, int32_t* nmatches) {
  return single_pair(ctx, KIND_BOW_KF_F, kf, valid_kf, f, nullptr, nnratio, check_ori, nullptr, 0.f, 0.f, nullptr,
                     nullptr, 0, match_f, dist_f, nmatches);
}

extern "C" int asdm_search_by_bow_kf_kf(asdm_ctx* ctx, asdm_set* kf1, const uint8_t* valid1, asdm_set* kf2,
                                        const uint8_t* valid2, float nnratio, int check_ori, int32_t* match12,
                                        float* dist12, int32_t* nmatches) {
  if (!valid2) return ASDM_ERR_INVALID;
  return single_pair(ctx, KIND_BOW_KF_KF, kf1, valid1, kf2, valid2, nnratio, check_ori, nullptr, 0.f, 0.f, nullptr,
                     nullptr, 0, match12, dist12, nmatches);
}

extern "C" int asdm_search_for_triangulation(asdm_ctx* ctx, asdm_set* kf1, const uint8_t* has_mp1, asdm_set* kf2,
                                             const uint8_t* has_mp2, const float* F12, float ex, float ey,
                                             const float* scale_factors2, const float* level_sigma2_2, int n_levels,
                                             int check_ori, int32_t* match12, float* dist12, int32_t* nmatches) {
  if (!has_mp2 || !F12) return ASDM_ERR_INVALID;
  return single_pair(ctx, KIND_TRIANGULATION, kf1, has_mp1, kf2, has_mp2, 0.f, check_ori, F12, ex, ey, scale_factors2,
                     level_sigma2_2, n_levels, match12, dist12, nmatches);
}

// ------------------------------------------------------------------------- projection-gated entry points
// scratch slots of the context

struct ProjDev {
  asdm_proj_query* q;
  ProjSeg* seg;
  int32_t *ent_j, *ent_q;
  float* ent_d;
  int32_t *total, *overflow;
};

// gather + distance with automatic growth of the entry pool; queries are already on the device
static int proj_candidates(asdm_ctx* ctx, asdm_set* target, int nq, const float* d_desc_q, const int32_t* d_q_rows,
                           const ProjGate& gate, int* ent_cap_io, ProjDev* out, bool large_windows = false,
                           bool defer_check = false) {
  // defer_check: no wait for the entry count -- the distance kernel sizes itself from the device counter and the caller
  // reads (total, overflow) back together with its results, growing the pool and repeating the search if needed
  int ent_cap = std::max(*ent_cap_io, 1 << 14);
  {
    const int rc = ensure_grid(ctx, target);
    if (rc) return rc;
  }
  for (int attempt = 0; attempt < 20; attempt++) {
    int rc;
    if ((rc = ensure(ctx, ctx->scratch[SC_SEG], (size_t)std::max(nq, 1) * sizeof(ProjSeg)))) return rc;
    if ((rc = ensure(ctx, ctx->scratch[SC_ENT], (size_t)ent_cap * 12))) return rc;
    if ((rc = ensure(ctx, ctx->scratch[SC_MISC], 256))) return rc;
    ProjDev pd;
    pd.q = (asdm_proj_query*)ctx->scratch[SC_QUERY].p;
    pd.seg = (ProjSeg*)ctx->scratch[SC_SEG].p;
    pd.ent_j = (int32_t*)ctx->scratch[SC_ENT].p;
    pd.ent_q = pd.ent_j + ent_cap;
    pd.ent_d = (float*)(pd.ent_q + ent_cap);
    pd.total = (int32_t*)ctx->scratch[SC_MISC].p;
    pd.overflow = pd.total + 1;
    CK(cudaMemsetAsync(pd.total, 0, 8, ctx->stream));
    if (nq > 0) {
      if (large_windows)
        k_proj_gather_warp<<<(nq * 32 + 255) / 256, 256, 0, ctx->stream>>>(pd.q, nq, target->kps, target->grid, gate,
                                                                          target->cell_start, target->cell_items, pd.seg,
                                                                          pd.ent_j, pd.ent_q, ent_cap, pd.total, pd.overflow);
      else
        k_proj_gather<<<(nq + 127) / 128, 128, 0, ctx->stream>>>(pd.q, nq, target->kps, target->grid, gate, target->cell_start,
                                                                 target->cell_items, pd.seg, pd.ent_j, pd.ent_q, ent_cap,
                                                                 pd.total, pd.overflow);
      LAUNCHED(ctx);
    }
    if (defer_check) {
      if (nq > 0) {
        k_proj_dist<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(d_desc_q, d_q_rows, target->desc, pd.ent_j, pd.ent_q, ent_cap,
                                                               pd.total, pd.ent_d);
        LAUNCHED(ctx);
      }
      *ent_cap_io = ent_cap;
      *out = pd;
      return ASDM_OK;
    }
    int32_t h[2] = {0, 0};
    CK(cudaMemcpyAsync(h, pd.total, 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    synced(ctx);
    if (h[1]) {  // pool too small: h[0] is the exact demand
      ent_cap = std::max(ent_cap * 2, h[0] + 1024);
      ctx->stat_retries++;
      continue;
    }
    if (h[0] > 0) {
      const int blocks = std::min((h[0] + 255) / 256, ctx->sm_count * 8);
      k_proj_dist<<<blocks, 256, 0, ctx->stream>>>(d_desc_q, d_q_rows, target->desc, pd.ent_j, pd.ent_q, ent_cap, pd.total,
                                                   pd.ent_d);
      LAUNCHED(ctx);
    }
    *ent_cap_io = ent_cap;
    *out = pd;
    return ASDM_OK;
  }
  return fail(ctx, ASDM_ERR_INTERNAL, "projection entry pool kept overflowing");
}

static int search_by_projection_impl(asdm_ctx* ctx, asdm_set* target, const asdm_proj_query* q, int nq,
                                     const float* desc_q, asdm_set* q_set, const int32_t* q_rows,
                                     const uint8_t* blocked0, const asdm_proj_params* p, int32_t* assign, float* dist_t,
                                     int32_t* nmatches, int32_t* query_best, float* query_dist) {
  if (!ctx || !target || nq < 0 || (nq > 0 && !q) || !p || (!assign && !query_best)) return ASDM_ERR_INVALID;
  if (!target->has_grid) return fail(ctx, ASDM_ERR_INVALID, "target set has no grid");
  if (!desc_q && !q_set) return fail(ctx, ASDM_ERR_INVALID, "query descriptors missing");
  if (p->mode != ASDM_PROJ_RATIO_LEVEL && p->mode != ASDM_PROJ_BEST_ONLY) return fail(ctx, ASDM_ERR_INVALID, "bad mode");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  ctx->stat_launches = 0;
  ctx->stat_retries = 0;
  const int n_t = target->n;
  int rc;
  if ((rc = ensure(ctx, ctx->scratch[SC_QUERY], (size_t)std::max(nq, 1) * sizeof(asdm_proj_query)))) return rc;
  // queries, masks and row maps are caller memory (pageable): staged through the pinned arena so nothing blocks
  if (nq > 0 && (rc = stage_upload(ctx, ctx->scratch[SC_QUERY].p, q, (size_t)nq * sizeof(asdm_proj_query)))) return rc;
  const float* d_desc_q = nullptr;
  const int32_t* d_rows = nullptr;
  if (desc_q) {
    // only the descriptors of queries that survived the pre-descriptor gates travel (C2: 2000 of 8000 local map
    // points are in view): compacted into the pinned staging arena, one DMA, and a row map for the kernels
    int nv = 0;
    for (int i = 0; i < nq; i++) nv += q[i].valid ? 1 : 0;
    const size_t row_bytes = ASDM_DESC_DIM * sizeof(float);
    const size_t map_bytes = align_up((size_t)std::max(nq, 1) * sizeof(int32_t), 256);
    if ((rc = ensure(ctx, ctx->scratch[SC_QDESC], map_bytes + (size_t)std::max(nv, 1) * row_bytes))) return rc;
    void* h = nullptr;
    if ((rc = stage_alloc(ctx, map_bytes + (size_t)nv * row_bytes, &h))) return rc;
    int32_t* h_map = (int32_t*)h;
    char* h_rows = (char*)h + map_bytes;
    int c = 0;
    for (int i = 0; i < nq; i++) {
      h_map[i] = q[i].valid ? c : 0;
      if (q[i].valid) memcpy(h_rows + (size_t)c++ * row_bytes, desc_q + (size_t)i * ASDM_DESC_DIM, row_bytes);
    }
    CK(cudaMemcpyAsync(ctx->scratch[SC_QDESC].p, h, map_bytes + (size_t)nv * row_bytes, cudaMemcpyHostToDevice, ctx->stream));
    d_rows = (const int32_t*)ctx->scratch[SC_QDESC].p;
    d_desc_q = (const float*)((const char*)ctx->scratch[SC_QDESC].p + map_bytes);
  } else {
    if (q_set->ctx != ctx) return fail(ctx, ASDM_ERR_INVALID, "query set belongs to another context");
    d_desc_q = q_set->desc;
    if (q_rows) {
      for (int i = 0; i < nq; i++)
        if (q[i].valid && (q_rows[i] < 0 || q_rows[i] >= q_set->n)) return fail(ctx, ASDM_ERR_INVALID, "query row out of range");
      if ((rc = ensure(ctx, ctx->scratch[SC_QDESC], (size_t)std::max(nq, 1) * sizeof(int32_t)))) return rc;
      if (nq > 0 && (rc = stage_upload(ctx, ctx->scratch[SC_QDESC].p, q_rows, (size_t)nq * sizeof(int32_t)))) return rc;
      d_rows = (const int32_t*)ctx->scratch[SC_QDESC].p;
    } else if (nq > q_set->n) {
      return fail(ctx, ASDM_ERR_INVALID, "more queries than rows in the query set");
    }
  }
  const uint8_t* d_blocked = nullptr;
  if (blocked0) {
    if ((rc = ensure(ctx, ctx->scratch[SC_BLOCK], (size_t)std::max(n_t, 1)))) return rc;
    if (n_t > 0 && (rc = stage_upload(ctx, ctx->scratch[SC_BLOCK].p, blocked0, (size_t)n_t))) return rc;
    d_blocked = (const uint8_t*)ctx->scratch[SC_BLOCK].p;
  }
  ProjDev pd;
  ProjGate gate;
  memset(&gate, 0, sizeof gate);
  if (p->chi2_max > 0.0) {
    if (p->n_levels <= 0 || p->n_levels > ASDM_MAX_LEVELS) return fail(ctx, ASDM_ERR_INVALID, "chi2 gate needs 1..32 levels");
    gate.chi2_max = p->chi2_max;
    gate.n_levels = p->n_levels;
    memcpy(gate.inv_level_sigma2, p->inv_level_sigma2, (size_t)p->n_levels * sizeof(float));
  }
  // state: acc[nq] own[2*n_t] assign[n_t] dist[n_t] nm[4] qbest[nq] qdist[nq]
  const size_t NQ = (size_t)std::max(nq, 1), NT = (size_t)std::max(n_t, 1);
  const size_t state = (3 * NQ + 4 * NT + 4) * 4;
  if ((rc = ensure(ctx, ctx->scratch[SC_STATE], state))) return rc;
  int32_t* acc = (int32_t*)ctx->scratch[SC_STATE].p;
  int32_t* own = acc + NQ;
  int32_t* d_assign = own + 2 * NT;
  float* d_dist = (float*)(d_assign + NT);
  int32_t* d_nm = (int32_t*)(d_dist + NT);
  int32_t* d_qbest = d_nm + 4;
  float* d_qdist = (float*)(d_qbest + NQ);
  ProjParamsDev P{p->mode, p->th_accept, p->best_init, p->nnratio, p->check_ori ? 1 : 0, p->count_twice ? 1 : 0};
  int32_t nm = 0;
  for (int attempt = 0;; attempt++) {
    if (attempt >= 20) return fail(ctx, ASDM_ERR_INTERNAL, "projection entry pool kept overflowing");
    // gather -> distances -> resolve back to back; the entry pool is checked with the results (one wait per call)
    if ((rc = proj_candidates(ctx, target, nq, d_desc_q, d_rows, gate, &ctx->proj_ent_cap, &pd, false, true))) return rc;
    k_proj_resolve<<<1, 1024, 0, ctx->stream>>>(pd.q, nq, target->kps, n_t, d_blocked, pd.seg, pd.ent_j, pd.ent_d, P, acc,
                                                own, d_assign, d_dist, d_nm, query_best ? d_qbest : nullptr, d_qdist);
    LAUNCHED(ctx);
    CK(cudaGetLastError());
    // results land in the pinned arena (the caller's arrays are pageable) and are handed over after the wait
    void* ho = nullptr;
    if ((rc = stage_alloc(ctx, (2 * NQ + 2 * NT + 8) * 4, &ho))) return rc;
    int32_t* h_pool = (int32_t*)ho;  // total, overflow, nm
    int32_t* h_assign = h_pool + 8;
    float* h_dist = (float*)(h_assign + NT);
    int32_t* h_qbest = (int32_t*)(h_dist + NT);
    float* h_qdist = (float*)(h_qbest + NQ);
    CK(cudaMemcpyAsync(h_pool, pd.total, 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(h_pool + 2, d_nm, 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (query_best && nq > 0) {
      CK(cudaMemcpyAsync(h_qbest, d_qbest, (size_t)nq * 4, cudaMemcpyDeviceToHost, ctx->stream));
      if (query_dist) CK(cudaMemcpyAsync(h_qdist, d_qdist, (size_t)nq * 4, cudaMemcpyDeviceToHost, ctx->stream));
    }
    if ((assign || dist_t) && n_t > 0)  // assign and dist are adjacent on both sides: one copy
      CK(cudaMemcpyAsync(h_assign, d_assign, (size_t)n_t * 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (h_pool[1]) {  // pool too small: h_pool[0] is the exact demand. Inputs are still resident: search again
      ctx->proj_ent_cap = std::max(ctx->proj_ent_cap * 2, h_pool[0] + 1024);
      ctx->stat_retries++;
      continue;
    }
    nm = h_pool[2];
    if (query_best && nq > 0) {
      memcpy(query_best, h_qbest, (size_t)nq * 4);
      if (query_dist) memcpy(query_dist, h_qdist, (size_t)nq * 4);
    }
    if (assign && n_t > 0) memcpy(assign, h_assign, (size_t)n_t * 4);
    if (dist_t && n_t > 0) memcpy(dist_t, h_dist, (size_t)n_t * 4);
    break;
  }
  synced(ctx);
  if (nmatches) *nmatches = nm;
  return ASDM_OK;
}

extern "C" int asdm_search_by_projection(asdm_ctx* ctx, asdm_set* target, const asdm_proj_query* q, int nq,
                                         const float* desc_q, asdm_set* q_set, const int32_t* q_rows,
                                         const uint8_t* blocked0, const asdm_proj_params* p, int32_t* assign,
                                         float* dist_t, int32_t* nmatches) {
  if (!assign) return ASDM_ERR_INVALID;
  return search_by_projection_impl(ctx, target, q, nq, desc_q, q_set, q_rows, blocked0, p, assign, dist_t, nmatches, nullptr,
                                   nullptr);
}

extern "C" int asdm_search_by_projection_q(asdm_ctx* ctx, asdm_set* target, const asdm_proj_query* q, int nq,
                                           const float* desc_q, asdm_set* q_set, const int32_t* q_rows,
                                           const uint8_t* blocked0, const asdm_proj_params* p, int32_t* query_best,
                                           float* query_dist, int32_t* nmatches) {
  if (!query_best) return ASDM_ERR_INVALID;
  return search_by_projection_impl(ctx, target, q, nq, desc_q, q_set, q_rows, blocked0, p, nullptr, nullptr, nmatches,
                                   query_best, query_dist);
}

extern "C" int asdm_search_for_initialization(asdm_ctx* ctx, asdm_set* f1, asdm_set* f2, float* prev_matched,
                                              int window_size, float nnratio, int check_ori, int32_t* matches12,
                                              float* dist12, int32_t* nmatches) {
  if (!ctx || !f1 || !f2 || !prev_matched || !matches12) return ASDM_ERR_INVALID;
  if (!f1->has_kps || !f2->has_grid) return fail(ctx, ASDM_ERR_INVALID, "F1 needs keypoints and F2 a grid");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  ctx->stat_launches = 0;
  ctx->stat_retries = 0;
  const int n1 = f1->n, n2 = f2->n;
  int rc;
  if ((rc = ensure(ctx, ctx->scratch[SC_QUERY], (size_t)std::max(n1, 1) * sizeof(asdm_proj_query)))) return rc;
  // state: prev[2*n1] matched_dist[n2] matches21[n2] matches12[n1] dist12[n1] pushed[n1] nm
  const size_t N1 = (size_t)std::max(n1, 1), N2 = (size_t)std::max(n2, 1);
  if ((rc = ensure(ctx, ctx->scratch[SC_STATE], (5 * N1 + 2 * N2 + 4) * 4))) return rc;
  float* d_prev = (float*)ctx->scratch[SC_STATE].p;
  float* d_md = d_prev + 2 * N1;
  int32_t* d_m21 = (int32_t*)(d_md + N2);
  int32_t* d_m12 = d_m21 + N2;
  float* d_d12 = (float*)(d_m12 + N1);
  int32_t* d_push = (int32_t*)(d_d12 + N1);
  int32_t* d_nm = d_push + N1;
  if (n1 > 0) {
    CK(cudaMemcpyAsync(d_prev, prev_matched, (size_t)n1 * 8, cudaMemcpyHostToDevice, ctx->stream));
    k_init_queries<<<(n1 + 255) / 256, 256, 0, ctx->stream>>>(f1->kps, n1, d_prev, (float)window_size,
                                                              (asdm_proj_query*)ctx->scratch[SC_QUERY].p);
    LAUNCHED(ctx);
  }
  ProjDev pd;
  ProjGate gate;
  memset(&gate, 0, sizeof gate);
  if ((rc = proj_candidates(ctx, f2, n1, f1->desc, nullptr, gate, &ctx->proj_ent_cap, &pd, window_size >= 30))) return rc;
  // Jacobi sweeps to the fixed point of the serial replay (k_init_sweep); the serial kernel remains as the fallback
  // for pathological inputs (more than INIT_K accepts of one keypoint, or no convergence within the sweep budget)
  bool resolved = false;
  ctx->stat_init_sweeps = 0;
  if (n1 > 0 && !ctx->opt_init_serial) {
    const int SWEEPS_PER_ROUND = 8, MAX_SWEEPS = 96;
    const size_t per_buf = align_up(N1 * 8 + N2 * 4 + N2 * INIT_K * 8, 256);
    const size_t state_bytes = align_up((size_t)(INIT_STATE_CHG + MAX_SWEEPS + 2) * 4, 256);
    if ((rc = ensure(ctx, ctx->scratch[SC_INIT], state_bytes + 3 * per_buf))) return rc;
    char* base = (char*)ctx->scratch[SC_INIT].p;
    int32_t* d_state = (int32_t*)base;
    InitBuf B[3];
    for (int b = 0; b < 3; b++) {
      char* p = base + state_bytes + (size_t)b * per_buf;
      B[b].dec_j = (int32_t*)p;
      B[b].dec_d = (float*)(p + N1 * 4);
      B[b].cnt = (int32_t*)(p + N1 * 8);
      B[b].lst_i = (int32_t*)(p + N1 * 8 + N2 * 4);
      B[b].lst_d = (float*)(p + N1 * 8 + N2 * 4 + N2 * INIT_K * 4);
    }
    if ((rc = zero_device(ctx, base, state_bytes + 3 * per_buf))) return rc;
    CK(cudaMemsetAsync(B[0].dec_j, 0xff, N1 * 4, ctx->stream));  // sweep 0: nobody has accepted anything
    int32_t h_state[4] = {0, 0, 0, 0};
    for (int t0 = 1; t0 + SWEEPS_PER_ROUND <= MAX_SWEEPS && !h_state[0] && !h_state[2]; t0 += SWEEPS_PER_ROUND) {
      for (int t = t0; t < t0 + SWEEPS_PER_ROUND; t++) {
        k_init_sweep<<<(n1 * 32 + 255) / 256, 256, 0, ctx->stream>>>(t, n1, n2, pd.seg, pd.ent_j, pd.ent_d, nnratio,
                                                                     B[(t - 1) % 3], B[t % 3], B[(t + 1) % 3], d_state);
        LAUNCHED(ctx);
      }
      // an empty launch in the next slot notices a last sweep of the round that changed nothing
      k_init_sweep<<<1, 32, 0, ctx->stream>>>(t0 + SWEEPS_PER_ROUND, 0, 0, pd.seg, pd.ent_j, pd.ent_d, nnratio, B[0], B[0], B[0],
                                              d_state);
      k_init_finalize<<<1, 1024, 0, ctx->stream>>>(f1->kps, n1, f2->kps, check_ori ? 1 : 0, B[0], B[1], B[2], d_state, d_m12,
                                                   d_d12, d_prev, d_nm);
      LAUNCHED(ctx);
      CK(cudaGetLastError());
      CK(cudaMemcpyAsync(h_state, d_state, 16, cudaMemcpyDeviceToHost, ctx->stream));
      CK(cudaStreamSynchronize(ctx->stream));
      if (!h_state[0] && !h_state[2]) ctx->stat_retries++;  // another round of sweeps
    }
    resolved = h_state[0] && !h_state[2];
    ctx->stat_init_sweeps = resolved ? 1 : 0;
  }
  if (!resolved) {
    k_init_resolve<<<1, 32, 0, ctx->stream>>>(f1->kps, n1, f2->kps, n2, pd.seg, pd.ent_j, pd.ent_d, nnratio,
                                              check_ori ? 1 : 0, d_md, d_m21, d_m12, d_d12, d_push, d_prev, d_nm);
    LAUNCHED(ctx);
  }
  CK(cudaGetLastError());
  int32_t nm = 0;
  if (n1 > 0) {
    CK(cudaMemcpyAsync(matches12, d_m12, (size_t)n1 * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (dist12) CK(cudaMemcpyAsync(dist12, d_d12, (size_t)n1 * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(prev_matched, d_prev, (size_t)n1 * 8, cudaMemcpyDeviceToHost, ctx->stream));
  }
  CK(cudaMemcpyAsync(&nm, d_nm, 4, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  synced(ctx);
  if (nmatches) *nmatches = nm;
  return ASDM_OK;
}

extern "C" int asdm_stereo_row_match(asdm_ctx* ctx, asdm_set* left, asdm_set* right, const float* scale_factors,
                                     int n_levels, int n_rows, float min_d, float max_d, int32_t* best_r,
                                     float* best_dist) {
  if (!ctx || !left || !right || !scale_factors || n_levels <= 0 || n_levels > ASDM_MAX_LEVELS || !best_r)
    return ASDM_ERR_INVALID;
  if (!left->has_kps || !right->has_kps) return fail(ctx, ASDM_ERR_INVALID, "sets need keypoints");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  ctx->stat_launches = 0;
  const int n_l = left->n;
  int rc;
  const size_t NL = (size_t)std::max(n_l, 1);
  if ((rc = ensure(ctx, ctx->scratch[SC_STATE], (2 * NL + ASDM_MAX_LEVELS) * 4))) return rc;
  int32_t* d_best = (int32_t*)ctx->scratch[SC_STATE].p;
  float* d_dist = (float*)(d_best + NL);
  float* d_sf = d_dist + NL;
  CK(cudaMemcpyAsync(d_sf, scale_factors, (size_t)n_levels * 4, cudaMemcpyHostToDevice, ctx->stream));
  if (n_l > 0) {
    k_stereo_match<<<(n_l * 32 + 255) / 256, 256, 0, ctx->stream>>>(left->desc, left->kps, n_l, right->desc, right->kps,
                                                                    right->n, d_sf, n_levels, n_rows, min_d, max_d, d_best,
                                                                    d_dist);
    LAUNCHED(ctx);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(best_r, d_best, (size_t)n_l * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (best_dist) CK(cudaMemcpyAsync(best_dist, d_dist, (size_t)n_l * 4, cudaMemcpyDeviceToHost, ctx->stream));
  }
  CK(cudaStreamSynchronize(ctx->stream));
  synced(ctx);
  return ASDM_OK;
}


// ------------------------------------------------------------------ MapPoint::ComputeDistinctiveDescriptors
extern "C" int asdm_distinctive_descriptors(asdm_ctx* ctx, asdm_set* const* sets, int n_sets, int n_points,
                                            const int32_t* obs_off, const int32_t* obs_set, const int32_t* obs_row,
                                            int32_t* best, float* best_median) {
  if (!ctx || n_sets < 0 || n_points < 0 || (n_points > 0 && (!obs_off || !best)) || (n_sets > 0 && !sets))
    return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  ctx->stat_launches = 0;
  if (n_points == 0) return ASDM_OK;
  const int n_obs = obs_off[n_points];
  if (obs_off[0] != 0 || n_obs < 0 || (n_obs > 0 && (!obs_set || !obs_row)))
    return fail(ctx, ASDM_ERR_INVALID, "bad observation offsets");
  std::vector<int32_t> small, large;
  std::vector<int64_t> d_off;
  int64_t scratch_floats = 0;
  for (int p = 0; p < n_points; p++) {
    const int N = obs_off[p + 1] - obs_off[p];
    if (N < 0) return fail(ctx, ASDM_ERR_INVALID, "observation offsets not ascending");
    for (int e = obs_off[p]; e < obs_off[p + 1]; e++) {
      if (obs_set[e] < 0 || obs_set[e] >= n_sets || !sets[obs_set[e]] || sets[obs_set[e]]->ctx != ctx)
        return fail(ctx, ASDM_ERR_INVALID, "observation names an unknown keyframe set");
      if (obs_row[e] < 0 || obs_row[e] >= sets[obs_set[e]]->n) return fail(ctx, ASDM_ERR_INVALID, "observation row out of range");
    }
    if (N == 0) continue;
    if (N <= 32) {
      small.push_back(p);
    } else {
      large.push_back(p);
      d_off.push_back(scratch_floats);
      scratch_floats += (int64_t)N * N;
    }
  }
  // one staged block: set pointers | obs_off | obs_set | obs_row | small ids | large ids | large offsets
  size_t off = 0;
  auto take = [&](size_t bytes) { const size_t o = off; off += align_up(std::max(bytes, (size_t)1), 256); return o; };
  const size_t o_ptr = take((size_t)n_sets * sizeof(float*)), o_off = take((size_t)(n_points + 1) * 4),
               o_set = take((size_t)n_obs * 4), o_row = take((size_t)n_obs * 4), o_small = take(small.size() * 4),
               o_large = take(large.size() * 4), o_doff = take(d_off.size() * 8), o_best = take((size_t)n_points * 4),
               o_med = take((size_t)n_points * 4);
  const size_t in_bytes = o_best;
  int rc;
  if ((rc = ensure(ctx, ctx->scratch[SC_MP_ARGS], off))) return rc;
  if (scratch_floats && (rc = ensure(ctx, ctx->scratch[SC_MP_DIST], (size_t)scratch_floats * sizeof(float)))) return rc;
  void* h = nullptr;
  if ((rc = stage_alloc(ctx, in_bytes, &h))) return rc;
  char* hb = (char*)h;
  for (int k = 0; k < n_sets; k++) ((const float**)(hb + o_ptr))[k] = sets[k] ? sets[k]->desc : nullptr;
  memcpy(hb + o_off, obs_off, (size_t)(n_points + 1) * 4);
  if (n_obs) {
    memcpy(hb + o_set, obs_set, (size_t)n_obs * 4);
    memcpy(hb + o_row, obs_row, (size_t)n_obs * 4);
  }
  if (!small.empty()) memcpy(hb + o_small, small.data(), small.size() * 4);
  if (!large.empty()) {
    memcpy(hb + o_large, large.data(), large.size() * 4);
    memcpy(hb + o_doff, d_off.data(), d_off.size() * 8);
  }
  char* db = (char*)ctx->scratch[SC_MP_ARGS].p;
  CK(cudaMemcpyAsync(db, hb, in_bytes, cudaMemcpyHostToDevice, ctx->stream));
  DistinctArgs a;
  a.set_desc = (const float* const*)(db + o_ptr);
  a.obs_off = (const int32_t*)(db + o_off);
  a.obs_set = (const int32_t*)(db + o_set);
  a.obs_row = (const int32_t*)(db + o_row);
  a.best = (int32_t*)(db + o_best);
  a.best_median = (float*)(db + o_med);
  CK(cudaMemsetAsync(a.best, 0xff, (size_t)n_points * 4, ctx->stream));  // -1: no observation (MapPoint.cc:285-286)
  CK(cudaMemsetAsync(a.best_median, 0, (size_t)n_points * 4, ctx->stream));
  if (!small.empty()) {
    k_distinct_small<<<((int)small.size() + DS_WARPS - 1) / DS_WARPS, DS_WARPS * 32, 0, ctx->stream>>>(
        a, (const int32_t*)(db + o_small), (int)small.size());
    LAUNCHED(ctx);
  }
  if (!large.empty()) {
    k_distinct_large<<<(int)large.size(), 256, 0, ctx->stream>>>(a, (const int32_t*)(db + o_large), (const int64_t*)(db + o_doff),
                                                                 (float*)ctx->scratch[SC_MP_DIST].p);
    LAUNCHED(ctx);
  }
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(best, a.best, (size_t)n_points * 4, cudaMemcpyDeviceToHost, ctx->stream));
  if (best_median) CK(cudaMemcpyAsync(best_median, a.best_median, (size_t)n_points * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  synced(ctx);
  return ASDM_OK;
}

// ============================================================================ vocabulary tree
extern "C" int asdm_vocab_create(asdm_ctx* ctx, const asdm_vocab_desc* d, asdm_vocab** out) {
  if (!ctx || !d || !out) return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  *out = nullptr;
  if (d->n_nodes < 2 || !d->child_off || !d->child_idx || !d->desc || !d->weight || !d->word_id)
    return fail(ctx, ASDM_ERR_INVALID, "vocabulary is empty or incomplete");
  if (d->weighting < 0 || d->weighting > 3 || d->scoring < 0 || d->scoring > 5)
    return fail(ctx, ASDM_ERR_INVALID, "unknown weighting / scoring type");
  const int nn = d->n_nodes;
  if (d->child_off[0] != 0 || d->child_off[nn] != nn - 1)
    return fail(ctx, ASDM_ERR_INVALID, "children lists must name every node but the root exactly once");
  if (d->child_off[1] == 0) return fail(ctx, ASDM_ERR_INVALID, "the root has no children");
  std::vector<uint8_t> seen(nn, 0);
  int kmax = 0;
  for (int p = 0; p < nn; p++) {
    const int kp = d->child_off[p + 1] - d->child_off[p];
    if (kp < 0) return fail(ctx, ASDM_ERR_INVALID, "child offsets not ascending");
    kmax = std::max(kmax, kp);
    for (int e = d->child_off[p]; e < d->child_off[p + 1]; e++) {
      const int c = d->child_idx[e];
      if (c <= 0 || c >= nn || seen[c] || c == p) return fail(ctx, ASDM_ERR_INVALID, "children lists do not form a tree");
      seen[c] = 1;
    }
  }
  // a node reachable from itself cannot also be reachable from the root: every node has one parent (checked above),
  // so walking down from the root must visit all n_nodes
  {
    std::vector<int> stack{0};
    int visited = 0;
    while (!stack.empty()) {
      const int p = stack.back();
      stack.pop_back();
      visited++;
      for (int e = d->child_off[p]; e < d->child_off[p + 1]; e++) stack.push_back(d->child_idx[e]);
    }
    if (visited != nn) return fail(ctx, ASDM_ERR_INVALID, "children lists contain a cycle");
  }
  asdm_vocab* v = new (std::nothrow) asdm_vocab;
  if (!v) return fail(ctx, ASDM_ERR_OOM, "host allocation");
  v->ctx = ctx;
  v->k = d->k; v->L = d->L; v->weighting = d->weighting; v->scoring = d->scoring; v->n_nodes = nn;
  v->G = std::min(std::max(kmax, 1), 32);
  v->kmax = kmax;
  v->h_weight.assign(d->weight, d->weight + nn);
  v->h_word.assign(d->word_id, d->word_id + nn);
  // one descriptor row and one record per child slot, built on the host once (vocabulary load time), one upload
  const size_t ns = (size_t)nn - 1, o_meta = 0, o_rows = align_up(ns * sizeof(int4), 512), total = o_rows + ns * ASDM_DESC_DIM * 4;
  std::vector<char> hb(total, 0);
  int4* meta = (int4*)(hb.data() + o_meta);
  float* rows = (float*)(hb.data() + o_rows);
  for (size_t sl = 0; sl < ns; sl++) {
    const int c = d->child_idx[sl];
    meta[sl] = make_int4(c, d->child_off[c], d->child_off[c + 1] - d->child_off[c], d->weight[c] > 0.0 ? 1 : 0);
    memcpy(rows + sl * ASDM_DESC_DIM, d->desc + (size_t)c * ASDM_DESC_DIM, ASDM_DESC_DIM * 4);
  }
  cudaError_t e = dev_alloc(ctx, &v->mem, total);
  if (e == cudaSuccess) e = cudaMemcpyAsync(v->mem, hb.data(), total, cudaMemcpyHostToDevice, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);  // hb is pageable and dies with this scope
  if (e != cudaSuccess) {
    dev_free(ctx, v->mem);
    delete v;
    return fail(ctx, e == cudaErrorMemoryAllocation ? ASDM_ERR_OOM : ASDM_ERR_CUDA, "vocabulary upload", e);
  }
  synced(ctx);
  v->meta = (int4*)((char*)v->mem + o_meta);
  v->rows = (float*)((char*)v->mem + o_rows);
  v->root_k = d->child_off[1];
  *out = v;
  return ASDM_OK;
}

extern "C" int asdm_vocab_release(asdm_vocab* v) {
  if (!v) return ASDM_ERR_INVALID;
  asdm_ctx* ctx = v->ctx;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  cudaSetDevice(ctx->device);
  dev_free(ctx, v->mem);
  delete v;
  return ASDM_OK;
}

extern "C" int asdm_vocab_transform(asdm_ctx* ctx, asdm_vocab* voc, asdm_set* s, int levelsup, int32_t* word_id,
                                    int32_t* node_id, uint8_t* in_fv, int32_t* bow_n, uint32_t* bow_id, double* bow_val) {
  if (!ctx || !voc || !s || voc->ctx != ctx || s->ctx != ctx) return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  ctx->stat_launches = 0;
  if (bow_n) *bow_n = 0;
  const int n = s->n;
  if (!s->node_id) {
    CK(dev_alloc(ctx, &s->fv_mem, (size_t)s->n_pad * 8));
    s->node_id = (uint32_t*)s->fv_mem;
    s->fv_pos = (int32_t*)((char*)s->fv_mem + (size_t)s->n_pad * 4);
  }
  // padding rows: node 0, not in the FeatureVector (as asdm_set_featvec leaves them)
  CK(cudaMemsetAsync(s->node_id, 0, (size_t)s->n_pad * 4, ctx->stream));
  CK(cudaMemsetAsync(s->fv_pos, 0xff, (size_t)s->n_pad * 4, ctx->stream));
  s->has_fv = true;
  if (n == 0) return ASDM_OK;
  int rc;
  const size_t o_key = 0, o_leaf = align_up((size_t)n * 8, 256), o_cnt = o_leaf + align_up((size_t)n * 4, 256), sc_bytes = o_cnt + 256;
  if ((rc = ensure(ctx, ctx->scratch[SC_VOC], sc_bytes))) return rc;
  unsigned long long* d_key = (unsigned long long*)((char*)ctx->scratch[SC_VOC].p + o_key);
  int32_t* d_leaf = (int32_t*)((char*)ctx->scratch[SC_VOC].p + o_leaf);
  int32_t* d_cnt = (int32_t*)((char*)ctx->scratch[SC_VOC].p + o_cnt);
  VocDev vd{voc->rows, voc->meta, voc->root_k, voc->G};
  const int fpw = 32 / voc->G;
  bool staged = false;
  if (voc->kmax <= 32 && !ctx->opt_voc_direct) {  // child blocks staged through shared memory (cp.async)
    const int per_cta = VOC_WARPS * fpw;
    const size_t smem = (size_t)VOC_WARPS * voc_warp_words(voc->G) * sizeof(float);
    if (smem > 48 * 1024) CK(cudaFuncSetAttribute(k_voc_descend_staged, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // several CTAs per SM must fit (the descent is latency bound): ask for the largest shared-memory carve-out
    CK(cudaFuncSetAttribute(k_voc_descend_staged, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    CK(cudaMemsetAsync(d_cnt, 0, 4, ctx->stream));
    k_voc_descend_staged<<<(n + per_cta - 1) / per_cta, VOC_WARPS * 32, smem, ctx->stream>>>(
        vd, s->desc, n, voc->L - levelsup, d_leaf, s->node_id, s->fv_pos, d_key, ctx->opt_voc_exact, d_cnt);
    staged = true;
  } else {
    const int per_cta = VOC_WARPS_DIRECT * fpw;
    k_voc_descend<<<(n + per_cta - 1) / per_cta, VOC_WARPS_DIRECT * 32, (size_t)per_cta * ASDM_DESC_DIM * 4, ctx->stream>>>(
        vd, s->desc, n, voc->L - levelsup, d_leaf, s->node_id, s->fv_pos, d_key);
  }
  LAUNCHED(ctx);
  k_fv_rank<<<dim3((n + 127) / 128, (n + FV_JT - 1) / FV_JT), 128, 0, ctx->stream>>>(d_key, n, s->fv_pos);
  LAUNCHED(ctx);
  CK(cudaGetLastError());
  const bool want_host = word_id || node_id || in_fv || bow_n || bow_id || bow_val;
  if (!want_host) return ASDM_OK;  // stays asynchronous: the searches that follow are ordered on the same stream
  void* h = nullptr;
  if ((rc = stage_alloc(ctx, (size_t)n * 8 + 4, &h))) return rc;
  int32_t* h_leaf = (int32_t*)h;
  uint32_t* h_nid = (uint32_t*)h + n;
  int32_t* h_cnt = (int32_t*)h + 2 * (size_t)n;
  *h_cnt = 0;
  CK(cudaMemcpyAsync(h_leaf, d_leaf, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaMemcpyAsync(h_nid, s->node_id, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  if (staged) CK(cudaMemcpyAsync(h_cnt, d_cnt, 4, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  ctx->stat_voc_exact = *h_cnt;
  // marshalling for the caller's std::map containers: per-feature ids, then the BowVector in map (ascending word) order
  std::vector<std::pair<uint32_t, double>> words;  // (word, weight) of the features that count, in feature order
  words.reserve(n);
  for (int i = 0; i < n; i++) {
    const int lf = h_leaf[i];
    const double w = voc->h_weight[lf];
    if (word_id) word_id[i] = voc->h_word[lf];
    if (node_id) node_id[i] = (int32_t)h_nid[i];
    if (in_fv) in_fv[i] = w > 0 ? 1 : 0;
    if (w > 0) words.emplace_back((uint32_t)voc->h_word[lf], w);
  }
  synced(ctx);
  if (!(bow_n || bow_id || bow_val)) return ASDM_OK;
  std::stable_sort(words.begin(), words.end(),
                   [](const std::pair<uint32_t, double>& a, const std::pair<uint32_t, double>& b) { return a.first < b.first; });
  std::vector<uint32_t> ids;
  std::vector<double> vals;
  const bool tf = voc->weighting == 0 || voc->weighting == 1;
  for (size_t e = 0; e < words.size(); e++) {
    if (e > 0 && words[e].first == words[e - 1].first) {
      if (tf) vals.back() += words[e].second;  // BowVector::addWeight (BowVector.cpp:35-47); addIfNotExist keeps the first
    } else {
      ids.push_back(words[e].first);
      vals.push_back(words[e].second);
    }
  }
  const bool must = voc->scoring != 5;  // DotProductScoring alone does not normalise (ScoringObject.h:72-89)
  const bool l2 = voc->scoring == 1;
  if (tf && !ids.empty() && !must) {
    const double nd = (double)ids.size();
    for (double& x : vals) x /= nd;  // :1163-1169
  }
  if (must) {  // BowVector::normalize (BowVector.cpp:61-84)
    double norm = 0.0;
    if (!l2) {
      for (double x : vals) norm += fabs(x);
    } else {
      for (double x : vals) norm += x * x;
      norm = sqrt(norm);
    }
    if (norm > 0.0)
      for (double& x : vals) x /= norm;
  }
  if (bow_n) *bow_n = (int32_t)ids.size();
  if (bow_id) memcpy(bow_id, ids.data(), ids.size() * 4);
  if (bow_val) memcpy(bow_val, vals.data(), vals.size() * 8);
  return ASDM_OK;
}


// ============================================================================ keyframe database (BoW candidates)
// KeyFrameDatabase.cc:34-320. Device: the BowVectors of every keyframe (SoA arena) + one slot record each; host: the
// mirrors the list logic needs. See kernels_bowdb.cuh for the kernel and why its sums equal the reference's.
struct asdm_bowdb {
  asdm_ctx* ctx = nullptr;
  uint32_t* d_word = nullptr;
  double* d_val = nullptr;
  BowSlot* d_slot = nullptr;
  size_t cap_entries = 0, n_entries = 0;
  int cap_slots = 0;
  std::vector<BowSlot> slots;       // host mirror of d_slot
  std::vector<float> reloc_score;   // KeyFrame::mRelocScore of every slot (persists between queries, :270 / :294)
  DevBuf query, result;             // query words | values; common | first word | score
};

static int bowdb_reserve(asdm_bowdb* db, size_t entries, int n_slots) {
  asdm_ctx* ctx = db->ctx;
  if (entries > db->cap_entries) {
    // multiple of 64 entries: k_bowdb_query_hash reads whole 16-byte quads, up to 3 words past a keyframe's last entry
    const size_t want = align_up(std::max(std::max(entries, db->cap_entries * 2), (size_t)1 << 20), 64);
    void *w = nullptr, *v = nullptr;
    CK(dev_alloc(ctx, &w, want * sizeof(uint32_t)));
    CK(dev_alloc(ctx, &v, want * sizeof(double)));
    if (db->n_entries) {
      CK(cudaMemcpyAsync(w, db->d_word, db->n_entries * sizeof(uint32_t), cudaMemcpyDeviceToDevice, ctx->stream));
      CK(cudaMemcpyAsync(v, db->d_val, db->n_entries * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
    }
    dev_free(ctx, db->d_word);
    dev_free(ctx, db->d_val);
    db->d_word = (uint32_t*)w;
    db->d_val = (double*)v;
    db->cap_entries = want;
  }
  if (n_slots > db->cap_slots) {
    const int want = std::max(std::max(n_slots, db->cap_slots * 2), 1024);
    void* r = nullptr;
    CK(dev_alloc(ctx, &r, (size_t)want * sizeof(BowSlot)));
    if (!db->slots.empty())
      CK(cudaMemcpyAsync(r, db->d_slot, db->slots.size() * sizeof(BowSlot), cudaMemcpyDeviceToDevice, ctx->stream));
    dev_free(ctx, db->d_slot);
    db->d_slot = (BowSlot*)r;
    db->cap_slots = want;
  }
  return ASDM_OK;
}

extern "C" int asdm_bowdb_create(asdm_ctx* ctx, asdm_bowdb** out) {
  if (!ctx || !out) return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  asdm_bowdb* db = new asdm_bowdb();
  db->ctx = ctx;
  *out = db;
  return ASDM_OK;
}

extern "C" int asdm_bowdb_release(asdm_bowdb* db) {
  if (!db) return ASDM_ERR_INVALID;
  asdm_ctx* ctx = db->ctx;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  cudaSetDevice(ctx->device);
  dev_free(ctx, db->d_word);
  dev_free(ctx, db->d_val);
  dev_free(ctx, db->d_slot);
  release(ctx, db->query);
  release(ctx, db->result);
  delete db;
  return ASDM_OK;
}

extern "C" int asdm_bowdb_size(asdm_bowdb* db) { return db ? (int)db->slots.size() : ASDM_ERR_INVALID; }

extern "C" int asdm_bowdb_add(asdm_bowdb* db, const uint32_t* word_id, const double* value, int n, int32_t* slot) {
  if (!db || n < 0 || (n > 0 && (!word_id || !value))) return ASDM_ERR_INVALID;
  asdm_ctx* ctx = db->ctx;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  for (int e = 1; e < n; e++)
    if (word_id[e] <= word_id[e - 1]) return fail(ctx, ASDM_ERR_INVALID, "BowVector word ids must ascend strictly (std::map order)");
  int rc;
  if ((rc = bowdb_reserve(db, db->n_entries + (size_t)n, (int)db->slots.size() + 1))) return rc;
  const size_t wb = align_up((size_t)n * 4, 256), vb = align_up((size_t)n * 8, 256);
  void* h = nullptr;
  if ((rc = stage_alloc(ctx, wb + vb + 256, &h))) return rc;
  char* hb = (char*)h;
  BowSlot rec;
  rec.off = (long long)db->n_entries;
  rec.len = n;
  rec.alive = 1;
  if (n) {
    memcpy(hb, word_id, (size_t)n * 4);
    memcpy(hb + wb, value, (size_t)n * 8);
    CK(cudaMemcpyAsync(db->d_word + db->n_entries, hb, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(db->d_val + db->n_entries, hb + wb, (size_t)n * 8, cudaMemcpyHostToDevice, ctx->stream));
  }
  memcpy(hb + wb + vb, &rec, sizeof rec);
  CK(cudaMemcpyAsync(db->d_slot + db->slots.size(), hb + wb + vb, sizeof rec, cudaMemcpyHostToDevice, ctx->stream));
  if (slot) *slot = (int32_t)db->slots.size();
  db->slots.push_back(rec);
  db->reloc_score.push_back(0.f);
  db->n_entries += (size_t)n;
  return ASDM_OK;
}

extern "C" int asdm_bowdb_erase(asdm_bowdb* db, int32_t slot) {
  if (!db) return ASDM_ERR_INVALID;
  asdm_ctx* ctx = db->ctx;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  if (slot < 0 || slot >= (int)db->slots.size()) return fail(ctx, ASDM_ERR_INVALID, "unknown keyframe slot");
  if (!db->slots[slot].alive) return ASDM_OK;  // the reference's erase of an absent keyframe finds nothing (:57-67)
  db->slots[slot].alive = 0;
  void* h = nullptr;
  int rc;
  if ((rc = stage_alloc(ctx, sizeof(BowSlot), &h))) return rc;
  memcpy(h, &db->slots[slot], sizeof(BowSlot));
  CK(cudaMemcpyAsync(db->d_slot + slot, h, sizeof(BowSlot), cudaMemcpyHostToDevice, ctx->stream));
  return ASDM_OK;
}

extern "C" int asdm_bowdb_clear(asdm_bowdb* db) {
  if (!db) return ASDM_ERR_INVALID;
  asdm_ctx* ctx = db->ctx;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  // the inverted file is emptied (:72-76); the keyframes keep their slots (and their mRelocScore), none is listed any more
  for (size_t s = 0; s < db->slots.size(); s++) db->slots[s].alive = 0;
  if (!db->slots.empty()) {
    void* h = nullptr;
    int rc;
    if ((rc = stage_alloc(ctx, db->slots.size() * sizeof(BowSlot), &h))) return rc;
    memcpy(h, db->slots.data(), db->slots.size() * sizeof(BowSlot));
    CK(cudaMemcpyAsync(db->d_slot, h, db->slots.size() * sizeof(BowSlot), cudaMemcpyHostToDevice, ctx->stream));
  }
  return ASDM_OK;
}

// queues the device pass and waits; the three result arrays (pinned staging memory, valid until the next staged call)
static int bowdb_run(asdm_bowdb* db, const uint32_t* word_id, const double* value, int n, const int32_t** common,
                     const uint32_t** first_word, const float** score) {
  asdm_ctx* ctx = db->ctx;
  const int ns = (int)db->slots.size();
  for (int e = 1; e < n; e++)
    if (word_id[e] <= word_id[e - 1]) return fail(ctx, ASDM_ERR_INVALID, "BowVector word ids must ascend strictly (std::map order)");
  ctx->stat_launches = 0;
  const size_t wb = align_up((size_t)std::max(n, 1) * 4, 256), vb = align_up((size_t)std::max(n, 1) * 8, 256);
  const size_t rb = align_up((size_t)std::max(ns, 1) * 4, 256);
  int rc;
  if ((rc = ensure(ctx, db->query, wb + vb))) return rc;
  if ((rc = ensure(ctx, db->result, 3 * rb + 256))) return rc;  // common | first word | score | slot counter
  void* h = nullptr;
  if ((rc = stage_alloc(ctx, wb + vb + 3 * rb, &h))) return rc;
  char* hb = (char*)h;
  if (n) {
    memcpy(hb, word_id, (size_t)n * 4);
    memcpy(hb + wb, value, (size_t)n * 8);
    CK(cudaMemcpyAsync(db->query.p, hb, wb + vb, cudaMemcpyHostToDevice, ctx->stream));
  }
  char* dq = (char*)db->query.p;
  char* dr = (char*)db->result.p;
  char* hr = hb + wb + vb;
  if (ns) {
    // the query as a shared-memory hash table + Bloom filter when they fit and no word collides with the empty mark
    int log2cap = 6;
    while ((1 << log2cap) * 2 < 3 * n) log2cap++;  // load factor <= 2/3 (the filter in front keeps probes rare)
    int log2bits = 15;  // Bloom filter in front of the table: >= 128 bits per query word, 4 KB .. 64 KB
    while (log2bits < 19 && (1 << log2bits) < ctx->opt_bowdb_filter_bits * n) log2bits++;
    const size_t tab_bytes = ((size_t)8 << log2cap) + ((size_t)1 << (log2bits - 3));
    if (!ctx->opt_bowdb_search && n > 0 && tab_bytes <= (size_t)200 * 1024 && word_id[n - 1] != 0xffffffffu) {
      static bool attr_set = false;
      if (!attr_set) {
        CK(cudaFuncSetAttribute(k_bowdb_query_hash, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_set = true;
      }
      int per_sm = 1;
      CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_bowdb_query_hash, BOWDB_HASH_WARPS * 32, tab_bytes));
      const int grid = std::min((ns + BOWDB_HASH_WARPS - 1) / BOWDB_HASH_WARPS, std::max(per_sm, 1) * ctx->sm_count);
      CK(cudaMemsetAsync(dr + 3 * rb, 0, 4, ctx->stream));
      k_bowdb_query_hash<<<grid, BOWDB_HASH_WARPS * 32, tab_bytes, ctx->stream>>>(
          db->d_slot, ns, db->d_word, db->d_val, (const uint32_t*)dq, (const double*)(dq + wb), n, log2cap, log2bits,
          (int32_t*)dr, (uint32_t*)(dr + rb), (float*)(dr + 2 * rb), (int*)(dr + 3 * rb));
    } else {
      k_bowdb_query<<<(ns + BOWDB_WARPS - 1) / BOWDB_WARPS, BOWDB_WARPS * 32, 0, ctx->stream>>>(
          db->d_slot, ns, db->d_word, db->d_val, (const uint32_t*)dq, (const double*)(dq + wb), n, (int32_t*)dr,
          (uint32_t*)(dr + rb), (float*)(dr + 2 * rb));
    }
    LAUNCHED(ctx);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(hr, dr, 3 * rb, cudaMemcpyDeviceToHost, ctx->stream));
  }
  CK(cudaStreamSynchronize(ctx->stream));
  // NOT synced(ctx): the results are read from the staging block by the caller of this helper
  *common = (const int32_t*)hr;
  *first_word = (const uint32_t*)(hr + rb);
  *score = (const float*)(hr + 2 * rb);
  return ASDM_OK;
}

extern "C" int asdm_bowdb_query(asdm_bowdb* db, const uint32_t* word_id, const double* value, int n, int32_t* common,
                                uint32_t* first_word, float* score) {
  if (!db || n < 0 || (n > 0 && (!word_id || !value))) return ASDM_ERR_INVALID;
  asdm_ctx* ctx = db->ctx;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  const int32_t* c;
  const uint32_t* f;
  const float* sc;
  int rc;
  if ((rc = bowdb_run(db, word_id, value, n, &c, &f, &sc))) return rc;
  const size_t ns = db->slots.size();
  if (common) memcpy(common, c, ns * 4);
  if (first_word) memcpy(first_word, f, ns * 4);
  if (score) memcpy(score, sc, ns * 4);
  synced(ctx);
  return ASDM_OK;
}

// lKFsSharingWords (:93-113 / :226-243): a keyframe enters the list at the smallest word it shares with the query, and the
// keyframes of one inverted list stand in add order -> sort by (first common word, slot).  Only the members above the
// common-word gate are ever looked at in list order (:137 / :264), and the gate itself (the maximum over the list) does
// not depend on the order, so the gate is found in one pass over all members and only the survivors are sorted.
static void bowdb_sharing_list(const asdm_bowdb* db, const int32_t* common, const uint32_t* first_word, const uint8_t* skip,
                               const uint8_t* need, float gate_factor, std::vector<int32_t>& list, int* min_common) {
  list.clear();
  const int ns = (int)db->slots.size();
  auto member = [&](int s) { return db->slots[s].alive && common[s] > 0 && !(skip && skip[s]) && (!need || need[s]); };
  int maxCommonWords = 0;
  bool any = false;
  for (int s = 0; s < ns; s++)
    if (member(s)) {
      any = true;
      maxCommonWords = std::max(maxCommonWords, common[s]);
    }
  *min_common = (int)(maxCommonWords * gate_factor);  // :128 / :256 (int * float -> float, truncated)
  if (!any) return;
  for (int s = 0; s < ns; s++)
    if (member(s) && common[s] > *min_common) list.push_back(s);
  std::sort(list.begin(), list.end(), [&](int32_t a, int32_t b) {
    return first_word[a] != first_word[b] ? first_word[a] < first_word[b] : a < b;
  });
}

static int bowdb_detect(asdm_bowdb* db, bool reloc, const uint32_t* word_id, const double* value, int n,
                        const asdm_bowdb_gates* g, int32_t* out, int cap, int32_t* n_out, int32_t* common_out = nullptr,
                        uint32_t* first_out = nullptr, float* score_out = nullptr) {
  if (!db || !g || !n_out || n < 0 || (n > 0 && (!word_id || !value)) || cap < 0 || (cap > 0 && !out)) return ASDM_ERR_INVALID;
  asdm_ctx* ctx = db->ctx;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  *n_out = 0;
  const int ns = (int)db->slots.size();
  if (g->only_global_map && !g->global_flag && ns) return fail(ctx, ASDM_ERR_INVALID, "only_global_map needs global_flag");
  if (g->neigh_off && ns) {
    if (g->neigh_off[0] != 0) return fail(ctx, ASDM_ERR_INVALID, "bad neighbour offsets");
    for (int s = 0; s < ns; s++) {
      if (g->neigh_off[s + 1] < g->neigh_off[s]) return fail(ctx, ASDM_ERR_INVALID, "neighbour offsets not ascending");
      for (int e = g->neigh_off[s]; e < g->neigh_off[s + 1]; e++)
        if (g->neigh_slot[e] < 0 || g->neigh_slot[e] >= ns) return fail(ctx, ASDM_ERR_INVALID, "neighbour names an unknown slot");
    }
  }
  const int32_t* common;
  const uint32_t* first_word;
  const float* score;
  int rc;
  if ((rc = bowdb_run(db, word_id, value, n, &common, &first_word, &score))) return rc;
  if (common_out) memcpy(common_out, common, (size_t)ns * 4);
  if (first_out) memcpy(first_out, first_word, (size_t)ns * 4);
  if (score_out) memcpy(score_out, score, (size_t)ns * 4);
  const uint8_t* glob = g->only_global_map ? g->global_flag : nullptr;
  std::vector<int32_t> sharing;  // the members of lKFsSharingWords above the common-word gate, in list order
  int minCommonWords = 0;
  // loop: connected keyframes never enter the list (:104); relocalisation: the global-map gate sits on the list (:236)
  bowdb_sharing_list(db, common, first_word, reloc ? nullptr : g->connected, reloc ? glob : nullptr, reloc ? 0.8f : 0.6f,
                     sharing, &minCommonWords);
  synced(ctx);  // (the staging block stays valid: nothing else is staged before this call returns)
  if (sharing.empty()) return ASDM_OK;
  // "listed": mnLoopQuery == pKF->mnId (the sharing list) / mnRelocQuery == F->mnId (every alive keyframe sharing a word)
  auto listed = [&](int s2) {
    if (!db->slots[s2].alive || common[s2] <= 0) return false;
    return reloc ? true : !(g->connected && g->connected[s2]);
  };
  std::vector<std::pair<float, int32_t> > scoreAndMatch, acc;
  for (int32_t s : sharing) {
    if (common[s] > minCommonWords) {
      const float si = score[s];
      if (reloc) {
        db->reloc_score[s] = si;  // :270
        scoreAndMatch.push_back(std::make_pair(si, s));
      } else if (si >= g->min_score && (!glob || glob[s])) {  // :146-148
        scoreAndMatch.push_back(std::make_pair(si, s));
      }
    }
  }
  if (scoreAndMatch.empty()) return ASDM_OK;
  float bestAccScore = reloc ? 0.f : g->min_score;  // :280 / :155
  for (const std::pair<float, int32_t>& sm : scoreAndMatch) {
    const int32_t s = sm.second;
    float bestScore = sm.first, accScore = sm.first;
    int32_t best = s;
    if (g->neigh_off)
      for (int e = g->neigh_off[s]; e < g->neigh_off[s + 1]; e++) {
        const int32_t s2 = g->neigh_slot[e];
        if (!listed(s2)) continue;
        float s2score;
        if (reloc) {
          s2score = db->reloc_score[s2];  // :294 -- fresh if scored in this query, else what an earlier query left
        } else {
          if (common[s2] <= minCommonWords) continue;  // :168
          s2score = score[s2];                         // mLoopScore, computed above for every keyframe over the gate
        }
        accScore += s2score;
        if (s2score > bestScore) {
          best = s2;
          bestScore = s2score;
        }
      }
    acc.push_back(std::make_pair(accScore, best));
    if (accScore > bestAccScore) bestAccScore = accScore;
  }
  const float minScoreToRetain = (reloc ? 0.75f : 0.55f) * bestAccScore;  // :305 / :184
  std::vector<uint8_t> added((size_t)ns, 0);
  int k = 0;
  for (const std::pair<float, int32_t>& a : acc)
    if (a.first > minScoreToRetain && !added[a.second]) {
      added[a.second] = 1;
      if (k >= cap) return fail(ctx, ASDM_ERR_INVALID, "more candidates than the output holds");
      out[k++] = a.second;
    }
  *n_out = k;
  return ASDM_OK;
}

extern "C" int asdm_bowdb_detect_loop(asdm_bowdb* db, const uint32_t* word_id, const double* value, int n,
                                      const asdm_bowdb_gates* g, int32_t* out, int cap, int32_t* n_out) {
  return bowdb_detect(db, false, word_id, value, n, g, out, cap, n_out);
}

extern "C" int asdm_bowdb_detect_reloc(asdm_bowdb* db, const uint32_t* word_id, const double* value, int n,
                                       const asdm_bowdb_gates* g, int32_t* out, int cap, int32_t* n_out) {
  return bowdb_detect(db, true, word_id, value, n, g, out, cap, n_out);
}

extern "C" int asdm_bowdb_detect_loop_full(asdm_bowdb* db, const uint32_t* word_id, const double* value, int n,
                                           const asdm_bowdb_gates* g, int32_t* out, int cap, int32_t* n_out,
                                           int32_t* common, uint32_t* first_word, float* score) {
  return bowdb_detect(db, false, word_id, value, n, g, out, cap, n_out, common, first_word, score);
}

extern "C" int asdm_bowdb_detect_reloc_full(asdm_bowdb* db, const uint32_t* word_id, const double* value, int n,
                                            const asdm_bowdb_gates* g, int32_t* out, int cap, int32_t* n_out,
                                            int32_t* common, uint32_t* first_word, float* score) {
  return bowdb_detect(db, true, word_id, value, n, g, out, cap, n_out, common, first_word, score);
}

extern "C" int asdm_bowdb_set_reloc_scores(asdm_bowdb* db, const float* score) {
  if (!db || (!score && !db->reloc_score.empty())) return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(db->ctx->mu);
  if (!db->reloc_score.empty()) memcpy(db->reloc_score.data(), score, db->reloc_score.size() * sizeof(float));
  return ASDM_OK;
}

extern "C" int asdm_bowdb_reloc_scores(asdm_bowdb* db, float* score) {
  if (!db || (!score && !db->reloc_score.empty())) return ASDM_ERR_INVALID;
  std::lock_guard<std::recursive_mutex> lk(db->ctx->mu);
  if (!db->reloc_score.empty()) memcpy(score, db->reloc_score.data(), db->reloc_score.size() * sizeof(float));
  return ASDM_OK;
}
