// host_util.cuh -- host-side helpers shared by the translation units of libasdmatch_cuda.so (defined in ctx.cu /
// sets.cu): error reporting, the stream-ordered block cache, the pinned staging arena and the set internals.
#pragma once
#include "internal.cuh"

int fail(asdm_ctx* ctx, int code, const char* what, cudaError_t e = cudaSuccess);

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

// ---- ctx.cu
cudaError_t dev_alloc(asdm_ctx* ctx, void** p, size_t bytes);
void dev_free(asdm_ctx* ctx, void* p);
void block_put(asdm_ctx* ctx, void* p, size_t cap);
void block_cache_flush(asdm_ctx* ctx);
int ensure(asdm_ctx* ctx, DevBuf& b, size_t bytes);
void release(asdm_ctx* ctx, DevBuf& b);
// the library synchronised the stream: everything staged so far has been consumed
static inline void synced(asdm_ctx* ctx) { ctx->stage_off = 0; }
int stage_alloc(asdm_ctx* ctx, size_t bytes, void** out);
int zero_device(asdm_ctx* ctx, void* dst, size_t bytes);
int pull_to_device(asdm_ctx* ctx, void* dst, const void* h, size_t bytes);
int stage_upload(asdm_ctx* ctx, void* dst, const void* src, size_t bytes, cudaStream_t stream = nullptr);
int32_t* slot_get(asdm_ctx* ctx);
void slot_put(asdm_ctx* ctx, int32_t* p);

// ---- sets.cu
void set_free(asdm_set* s);
// records s->ready on `stream` (creating the event on first use): call behind everything that writes the set's arrays
int set_mark_ready(asdm_ctx* ctx, asdm_set* s, cudaStream_t stream);
// a search of `ctx` is about to read `s`: same family required; a foreign context's stream is ordered behind s->ready
int use_set(asdm_ctx* ctx, asdm_set* s);
int ensure_grid(asdm_ctx* ctx, asdm_set* s);
int resolve_tc_flags(asdm_ctx* ctx, asdm_set* const* sets, int n);
int set_create_impl(asdm_ctx* ctx, const float* desc, int n, int row_stride_bytes, const asdm_kp* kps,
                    const asdm_grid* grid, asdm_set** out, bool wait);
int flatten_featvec(asdm_ctx* ctx, const asdm_featvec* fv, int n, int n_pad, uint32_t* node, int32_t* pos);

// ---- jobs.cu
float bow_keep_threshold(float nnratio, float th_accept);
void jobs_init_device();
// ---- proj.cu
void proj_init_device();  // per-device kernel attributes of the dense search kernels (called from asdm_create)
