// ctx.cu -- context, stream-ordered pools, pinned staging arena, options and counters of libasdmatch_cuda.so
// (include/asdmatch.h). Host-side orchestration only. There is no CPU fallback: without a CUDA device asdm_create fails.
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <new>

#include "host_util.cuh"

// ------------------------------------------------------------------------------------------------ utilities
int fail(asdm_ctx* ctx, int code, const char* what, cudaError_t e) {
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

// Device memory comes from the device's stream-ordered pool (cudaMallocAsync) with the release threshold lifted in
// asdm_create: allocation and free are ordered on the context stream, never synchronise the device, and the memory
// of released sets / jobs is recycled by the next ones (keyframes come and go at frame rate in the SLAM system).
cudaError_t dev_alloc(asdm_ctx* ctx, void** p, size_t bytes) { return cudaMallocAsync(p, bytes, ctx->stream); }
void dev_free(asdm_ctx* ctx, void* p) {
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

void block_put(asdm_ctx* ctx, void* p, size_t cap) {
  if (!p) return;
  if (ctx->block_cache_bytes + cap > BLOCK_CACHE_LIMIT) {
    dev_free(ctx, p);
    return;
  }
  ctx->block_cache[cap].push_back(p);
  ctx->block_cache_bytes += cap;
}

void block_cache_flush(asdm_ctx* ctx) {
  for (auto& kv : ctx->block_cache)
    for (void* p : kv.second) dev_free(ctx, p);
  ctx->block_cache.clear();
  ctx->block_cache_bytes = 0;
}

int ensure(asdm_ctx* ctx, DevBuf& b, size_t bytes) {
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

void release(asdm_ctx* ctx, DevBuf& b) {
  if (b.cap) block_put(ctx, b.p, b.cap);  // cap == 0: a view into another block (job input / zero arenas)
  b.p = nullptr;
  b.cap = 0;
}

int stage_alloc(asdm_ctx* ctx, size_t bytes, void** out) {
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
int zero_device(asdm_ctx* ctx, void* dst, size_t bytes) {  // dst 16-byte aligned; rounds bytes up to 16
  const size_t n = (bytes + 15) / 16;
  if (!n) return ASDM_OK;
  k_zero16<<<(int)std::min<size_t>((n + 255) / 256, 1184), 256, 0, ctx->stream>>>((uint4*)dst, n);
  LAUNCHED(ctx);
  CK(cudaGetLastError());
  return ASDM_OK;
}

// `h` lies in the staging arena (page-locked, device-visible under UVA); queued on the main stream
int pull_to_device(asdm_ctx* ctx, void* dst, const void* h, size_t bytes) {
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
int stage_upload(asdm_ctx* ctx, void* dst, const void* src, size_t bytes, cudaStream_t stream) {
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
int32_t* slot_get(asdm_ctx* ctx) {
  std::lock_guard<std::mutex> lk(ctx->slot_mu);
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
void slot_put(asdm_ctx* ctx, int32_t* p) {
  std::lock_guard<std::mutex> lk(ctx->slot_mu);
  if (p) ctx->slot_free.push_back((int)(((char*)p - ctx->slot_slab) / SLOT_BYTES));
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
  ctx->family = ctx;
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
  jobs_init_device();
  proj_init_device();
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

// One context per thread of the SLAM system (upstream ORB-SLAM2 runs tracking, local mapping and loop closing
// concurrently): a forked context has its own streams, staging arena, scratch and lock, so a thread never waits for
// another thread's search -- and it may search the sets the other contexts of the family created (Frame / KeyFrame
// descriptors are uploaded once, by whoever builds the object).
extern "C" int asdm_fork(asdm_ctx* parent, asdm_ctx** out) {
  if (!parent || !out) return ASDM_ERR_INVALID;
  const int rc = asdm_create(parent->device, out);
  if (rc) return rc;
  (*out)->family = parent->family;
  (*out)->opt_force_exact = parent->opt_force_exact;
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
  if (ctx->res_host) cudaFreeHost(ctx->res_host);
  if (ctx->proj_ctl) cudaFree(ctx->proj_ctl);
  for (cudaEvent_t e : ctx->stage_ev)
    if (e) cudaEventDestroy(e);
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
  else if (!strcmp(key, "proj_legacy")) ctx->opt_proj_legacy = value;
  else if (!strcmp(key, "resolve_level")) ctx->resolve_level_hint = std::min(std::max(value, 0), 2);
  else if (!strcmp(key, "proj_ent_cap")) ctx->proj_ent_cap = std::max(value, 64);
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
  else if (!strncmp(key, "job_us_", 7)) {  // stage durations of the last job run (time_kernels == 2), after a synchronisation
    static const char* names[4] = {"zero", "distance", "rerank", "resolve"};
    *value = -1;
    for (int k = 0; k < 4; k++)
      if (!strcmp(key + 7, names[k]) && ctx->stage_ev[k] && ctx->stage_ev[k + 1]) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, ctx->stage_ev[k], ctx->stage_ev[k + 1]) == cudaSuccess) *value = (int64_t)(ms * 1000.0f + 0.5f);
        else cudaGetLastError();
      }
  }
  else if (!strncmp(key, "proj_host_ns_", 13) && key[13] >= '0' && key[13] <= '2' && !key[14]) *value = ctx->stat_proj_host_ns[key[13] - '0'];
  else if (!strncmp(key, "proj_ns_", 8) && key[8] >= '0' && key[8] <= '5' && !key[9]) *value = ctx->stat_proj_ns[key[8] - '0'];
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
