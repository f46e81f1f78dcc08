// sets.cu -- device-resident descriptor sets (Frame.h:147, KeyFrame.h:199): upload, bf16 pack, grid, FeatureVector,
// map (de)serialisation blocks.
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <new>

#include "host_util.cuh"
#include "kernels_set.cuh"

// --------------------------------------------------------------------------------------- descriptor sets
int set_mark_ready(asdm_ctx* ctx, asdm_set* s, cudaStream_t stream) {
  if (!s->ready) CK(cudaEventCreateWithFlags(&s->ready, cudaEventDisableTiming));
  CK(cudaEventRecord(s->ready, stream));
  return ASDM_OK;
}

int use_set(asdm_ctx* ctx, asdm_set* s) {
  if (!s) return fail(ctx, ASDM_ERR_INVALID, "null set");
  if (s->ctx == ctx) return ASDM_OK;  // the owner's main stream already waits for every write at the time it is queued
  if (s->ctx->family != ctx->family) return fail(ctx, ASDM_ERR_INVALID, "set belongs to a context of another family (asdm_fork)");
  if (s->ready) CK(cudaStreamWaitEvent(ctx->stream, s->ready, 0));
  return ASDM_OK;
}

void set_free(asdm_set* s) {
  if (!s) return;
  asdm_ctx* ctx = s->ctx;
  cudaSetDevice(ctx->device);
  if (s->ready) cudaEventDestroy(s->ready);
  if (s->grid_ev) cudaEventDestroy(s->grid_ev);
  if (s->slab) {  // a view into a group window: the slab goes back when its last set does
    if (--s->slab->ref == 0) group_slab_release(s->slab);
    s->h_flag = nullptr;  // lives in the slab's page-locked block
  } else {
    dev_free(ctx, s->desc);  // one arena: every other device pointer lives inside it (or in the batch's side block)
  }
  dev_free(ctx, s->grid_mem);
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
int ensure_grid(asdm_ctx* ctx, asdm_set* s) {
  if (!s->has_grid) return ASDM_OK;
  std::lock_guard<std::mutex> lks(s->mu);
  if (s->grid_built) {  // built on another context's stream: order behind it
    if (s->grid_builder != ctx && s->grid_ev) CK(cudaStreamWaitEvent(ctx->stream, s->grid_ev, 0));
    return ASDM_OK;
  }
  const int nc = s->grid.cols * s->grid.rows;
  if (!s->cell_start) {  // view sets carry no bucket arrays of their own until a projection-gated search asks
    const size_t o_ci = align_up((size_t)(nc + 1) * sizeof(int32_t), 256), o_tmp = o_ci + align_up((size_t)s->n_pad * sizeof(int32_t), 256),
                 o_ckp = o_tmp + align_up((size_t)s->n_pad * sizeof(int32_t), 256), total = o_ckp + (size_t)s->n_pad * sizeof(int4);
    CK(dev_alloc(ctx, &s->grid_mem, total));
    char* b = (char*)s->grid_mem;
    s->cell_start = (int32_t*)b;
    s->cell_items = (int32_t*)(b + o_ci);
    s->cell_tmp = (int32_t*)(b + o_tmp);
    s->cell_kp = (int4*)(b + o_ckp);
  }
  const size_t smem = (size_t)(nc + 1) * sizeof(int32_t);
  if (smem > 48 * 1024) cudaFuncSetAttribute(k_build_grid, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  k_build_grid<<<1, 1024, smem, ctx->stream>>>(s->kps, s->n, s->grid, s->cell_tmp, s->cell_start, s->cell_items, s->cell_kp);
  LAUNCHED(ctx);
  CK(cudaGetLastError());
  if (!s->grid_ev) CK(cudaEventCreateWithFlags(&s->grid_ev, cudaEventDisableTiming));
  CK(cudaEventRecord(s->grid_ev, ctx->stream));
  s->grid_builder = ctx;
  s->grid_built = true;
  return ASDM_OK;
}

// reads the tensor-core range flags of sets whose flag copy is still in flight (one stream wait for all of them)
int resolve_tc_flags(asdm_ctx* ctx, asdm_set* const* sets, int n) {
  bool pending = false;
  for (int k = 0; k < n; k++) pending |= sets[k]->tc_ok < 0;
  if (!pending) return ASDM_OK;
  CK(cudaStreamSynchronize(ctx->stream));
  // (no synced(ctx): job_prepare calls this with its input block staged and not pulled yet -- the arena must not rewind)
  for (int k = 0; k < n; k++) {
    asdm_set* s = sets[k];
    std::lock_guard<std::mutex> lks(s->mu);
    if (s->tc_ok >= 0) continue;
    s->tc_ok = (s->h_flag && *s->h_flag == 0) ? 1 : 0;
    if (!s->slab) slot_put(s->ctx, s->h_flag);  // the owner's slab; a view's flag lives in its slab's page-locked block
    s->h_flag = nullptr;
  }
  return ASDM_OK;
}

int set_create_impl(asdm_ctx* ctx, const float* desc, int n, int row_stride_bytes, const asdm_kp* kps,
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
  const size_t o_ckp = off;  off += grid ? align_up((size_t)s->n_pad * sizeof(int4), 256) : 0;
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
  if (set_mark_ready(ctx, s, ctx->prep_stream)) return bail(ASDM_ERR_CUDA);  // ... and so does any other context of the family
  if (grid) {  // the grid bucket itself is built by the first projection-gated search (ensure_grid)
    s->grid = *grid;
    s->has_grid = true;
    s->cell_start = (int32_t*)(base + o_cs);
    s->cell_items = (int32_t*)(base + o_ci);
    s->cell_tmp = (int32_t*)(base + o_tmp);
    s->cell_kp = (int4*)(base + o_ckp);
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
    const size_t o_ckp = off;  off += d.grid ? align_up((size_t)s->n_pad * sizeof(int4), 256) : 0;
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
      s->cell_kp = (int4*)(base + o_ckp);
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
  for (int k = 0; k < n_sets; k++)
    if (set_mark_ready(ctx, out[k], ctx->prep_stream)) return bail(ASDM_ERR_CUDA);
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
  if (s->ref.fetch_sub(1) == 1)
    set_free(s);  // stream-ordered free on the owner's stream (searches it queued on the set finish first); takes no context
                  // lock: a thread of the family may drop the last reference while the owner waits inside a search
  return ASDM_OK;
}

extern "C" int asdm_set_size(const asdm_set* s) { return s ? s->n : ASDM_ERR_INVALID; }

// DBoW2::FeatureVector -> per-feature node id + CSR position (-1 = not in the vector)
int flatten_featvec(asdm_ctx* ctx, const asdm_featvec* fv, int n, int n_pad, uint32_t* node, int32_t* pos) {
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
  {  // the other contexts of the family wait for the pack (already behind `ready`) AND this upload
    if (s->ready) CK(cudaStreamWaitEvent(ctx->upload_stream, s->ready, 0));
    const int rr = set_mark_ready(ctx, s, ctx->upload_stream);
    if (rr) return rr;
  }
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
    if (!sets[k] || use_set(ctx, sets[k])) return fail(ctx, ASDM_ERR_INVALID, "set of another context family");
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
