// proj.cu -- projection-gated entry points: SearchByProjection x4, SearchForInitialization, stereo row matching.
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <new>

#include "host_util.cuh"
#include <time.h>

#include "kernels_proj_fused.cuh"

static inline int64_t host_ns() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (int64_t)ts.tv_sec * 1000000000 + ts.tv_nsec;
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
  int ent_cap = std::max(*ent_cap_io, 64);
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

static int search_by_projection_legacy(asdm_ctx* ctx, asdm_set* target, const asdm_proj_query* q, int nq,
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
  if ((rc = use_set(ctx, target)) || (q_set && (rc = use_set(ctx, q_set)))) return rc;
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
    if (q_set->ctx->family != ctx->family) return fail(ctx, ASDM_ERR_INVALID, "query set belongs to a context of another family");
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

// ---- the fused path (kernels_proj_fused.cuh): ONE launch and ONE wait per call. Inputs are staged in the page-locked
// arena and read by the kernel over PCIe (zero-copy): the compacted valid queries (32 B each), blocked0, and -- when the
// caller's descriptors are pageable -- the descriptor rows of the valid queries. Descriptors that already sit in
// page-locked host memory (asdm_host_alloc) or on the device are read in place. Results are stored by the kernel into a
// page-locked block and handed to the caller's arrays after the wait.
void proj_init_device() {
  cudaFuncSetAttribute(k_stereo_match, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
  cudaFuncSetAttribute(k_proj_fused<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(k_proj_fused<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
}

static int search_by_projection_impl(asdm_ctx* ctx, asdm_set* target, const asdm_proj_query* q, int nq,
                                     const float* desc_q, asdm_set* q_set, const int32_t* q_rows,
                                     const uint8_t* blocked0, const asdm_proj_params* p, int32_t* assign, float* dist_t,
                                     int32_t* nmatches, int32_t* query_best, float* query_dist) {
  if (!ctx || !target || nq < 0 || (nq > 0 && !q) || !p || (!assign && !query_best)) return ASDM_ERR_INVALID;
  if (ctx->opt_proj_legacy)
    return search_by_projection_legacy(ctx, target, q, nq, desc_q, q_set, q_rows, blocked0, p, assign, dist_t, nmatches,
                                       query_best, query_dist);
  if (!target->has_grid) return fail(ctx, ASDM_ERR_INVALID, "target set has no grid");
  if (!desc_q && !q_set) return fail(ctx, ASDM_ERR_INVALID, "query descriptors missing");
  if (p->mode != ASDM_PROJ_RATIO_LEVEL && p->mode != ASDM_PROJ_BEST_ONLY) return fail(ctx, ASDM_ERR_INVALID, "bad mode");
  if (target->n >= (1 << 24)) return fail(ctx, ASDM_ERR_INVALID, "target set too large for the projection search");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  const int64_t hn0 = host_ns();
  ctx->stat_launches = 0;
  ctx->stat_retries = 0;
  const int n_t = target->n;
  int rc;
  if ((rc = use_set(ctx, target)) || (q_set && (rc = use_set(ctx, q_set)))) return rc;
  ProjFusedArgs A;
  memset(&A, 0, sizeof A);
  if (p->chi2_max > 0.0) {
    if (p->n_levels <= 0 || p->n_levels > ASDM_MAX_LEVELS) return fail(ctx, ASDM_ERR_INVALID, "chi2 gate needs 1..32 levels");
    A.G.chi2_max = p->chi2_max;
    A.G.n_levels = p->n_levels;
    memcpy(A.G.inv_level_sigma2, p->inv_level_sigma2, (size_t)p->n_levels * sizeof(float));
  }
  if (q_set && !desc_q) {
    if (q_set->ctx->family != ctx->family) return fail(ctx, ASDM_ERR_INVALID, "query set belongs to a context of another family");
    if (!q_rows && nq > q_set->n) return fail(ctx, ASDM_ERR_INVALID, "more queries than rows in the query set");
  }
  int nv = 0;
  for (int i = 0; i < nq; i++) nv += q[i].valid ? 1 : 0;
  // where the query rows are read from
  enum { ROWS_SET, ROWS_IN_PLACE, ROWS_STAGED } rows_from = ROWS_SET;
  if (desc_q) {
    rows_from = ROWS_STAGED;
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, desc_q) == cudaSuccess) {
      if (at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeHost || at.type == cudaMemoryTypeManaged) {
        rows_from = ROWS_IN_PLACE;
        A.desc_q = at.type == cudaMemoryTypeHost ? (const float*)at.devicePointer : desc_q;
      }
    } else {
      cudaGetLastError();
    }
  } else {
    A.desc_q = q_set->desc;
  }
  const size_t NV = (size_t)std::max(nv, 1), NQ = (size_t)std::max(nq, 1), NT = (size_t)std::max(n_t, 1);
  const size_t row_bytes = ASDM_DESC_DIM * sizeof(float);
  // staged inputs: queries | blocked0 | rows      results: hdr | assign | dist | qbest | qdist
  const size_t o_blk = align_up(NV * sizeof(ProjQ), 256), o_rows = o_blk + align_up(blocked0 ? NT : 0, 256),
               o_res = o_rows + (rows_from == ROWS_STAGED ? align_up(NV * row_bytes, 256) : 0),
               o_asg = o_res + 64, o_dst = o_asg + align_up(NT * 4, 64), o_qb = o_dst + align_up(NT * 4, 64),
               o_qd = o_qb + (query_best ? align_up(NQ * 4, 64) : 0), stage_bytes = o_qd + (query_best ? align_up(NQ * 4, 64) : 0);
  void* hv = nullptr;
  if ((rc = stage_alloc(ctx, stage_bytes, &hv))) return rc;
  char* h = (char*)hv;
  ProjQ* hq = (ProjQ*)h;
  {
    int c = 0;
    for (int i = 0; i < nq; i++) {
      const asdm_proj_query& s = q[i];
      if (!s.valid) continue;
      uint32_t row = (uint32_t)i;
      if (rows_from == ROWS_STAGED) {
        row = (uint32_t)c;
        memcpy(h + o_rows + (size_t)c * row_bytes, desc_q + (size_t)i * ASDM_DESC_DIM, row_bytes);
      } else if (rows_from == ROWS_SET && q_rows) {
        if (q_rows[i] < 0 || q_rows[i] >= q_set->n) return fail(ctx, ASDM_ERR_INVALID, "query row out of range");
        row = (uint32_t)q_rows[i];
      }
      hq[c++] = ProjQ{s.u, s.v, s.r, s.angle, s.min_level, s.max_level, i, row | (s.blocks ? 0x80000000u : 0u)};
    }
  }
  if (blocked0 && n_t > 0) memcpy(h + o_blk, blocked0, (size_t)n_t);
  if (rows_from == ROWS_STAGED) A.desc_q = (const float*)(h + o_rows);
  if ((rc = ensure_grid(ctx, target))) return rc;
  if (!ctx->proj_ctl) {
    CK(cudaMalloc(&ctx->proj_ctl, 256));
    CK(cudaMemsetAsync(ctx->proj_ctl, 0, 256, ctx->stream));
  }
  const int n_cells = target->grid.cols * target->grid.rows;
  const size_t smem1 = pf_phase1_smem(n_cells), bytes2 = pf_phase2_bytes(n_t, nv);
  if (smem1 > 200 * 1024) return fail(ctx, ASDM_ERR_INVALID, "grid too fine for the projection search");
  A.smem2 = bytes2 <= 200 * 1024 ? 1 : 0;
  const size_t smem = A.smem2 ? std::max(smem1, bytes2) : smem1;
  if (!A.smem2 && (rc = ensure(ctx, ctx->scratch[SC_STATE], bytes2))) return rc;
  if ((rc = ensure(ctx, ctx->scratch[SC_SEG], NV * sizeof(ProjRec)))) return rc;
  if (query_best && (rc = ensure(ctx, ctx->scratch[SC_QDESC], NQ * 8))) return rc;
  if (blocked0 && (rc = ensure(ctx, ctx->scratch[SC_BLOCK], align_up(NT, 16)))) return rc;
  A.q = hq;
  A.nv = nv;
  A.nq = nq;
  A.desc_t = target->desc;
  A.kps_t = target->kps;
  A.cell_kp = target->cell_kp;
  A.cell_start = target->cell_start;
  A.g = target->grid;
  A.n_t = n_t;
  A.blocked0 = blocked0 ? (const uint8_t*)(h + o_blk) : nullptr;
  A.g_blk = blocked0 ? (uint32_t*)ctx->scratch[SC_BLOCK].p : nullptr;
  A.rec = (ProjRec*)ctx->scratch[SC_SEG].p;
  A.g_q = query_best ? (int32_t*)ctx->scratch[SC_QDESC].p : nullptr;
  A.ctl = (ProjCtl*)ctx->proj_ctl;
  A.g_state = A.smem2 ? nullptr : (int32_t*)ctx->scratch[SC_STATE].p;
  A.want_q = query_best ? 1 : 0;
  A.r_hdr = (int32_t*)(h + o_res);
  A.r_assign = (int32_t*)(h + o_asg);
  A.r_dist = (float*)(h + o_dst);
  A.r_qbest = (int32_t*)(h + o_qb);
  A.r_qdist = (float*)(h + o_qd);
  A.P = ProjParamsDev{p->mode, p->th_accept, p->best_init, p->nnratio, p->check_ori ? 1 : 0, p->count_twice ? 1 : 0};
  const int grid_x = std::max(1, std::min((nv + PF_WARPS - 1) / PF_WARPS, 2 * ctx->sm_count));
  int32_t nm = 0;
  for (int attempt = 0;; attempt++) {
    if (attempt >= 20) return fail(ctx, ASDM_ERR_INTERNAL, "projection entry pool kept overflowing");
    if ((rc = ensure(ctx, ctx->scratch[SC_ENT], (size_t)ctx->proj_ent_cap * sizeof(int2)))) return rc;
    A.ent = (int2*)ctx->scratch[SC_ENT].p;
    A.ent_cap = ctx->proj_ent_cap;
    A.r_hdr[0] = A.r_hdr[1] = A.r_hdr[2] = 0;
    const int64_t hn1 = host_ns();
    if (A.smem2) k_proj_fused<true><<<grid_x, PF_WARPS * 32, smem, ctx->stream>>>(A);
    else k_proj_fused<false><<<grid_x, PF_WARPS * 32, smem, ctx->stream>>>(A);
    LAUNCHED(ctx);
    CK(cudaGetLastError());
    const int64_t hn2 = host_ns();
    CK(cudaStreamSynchronize(ctx->stream));
    const int64_t hn3 = host_ns();
    ctx->stat_proj_host_ns[0] = (int32_t)(hn1 - hn0);  // staging
    ctx->stat_proj_host_ns[1] = (int32_t)(hn2 - hn1);  // launch
    ctx->stat_proj_host_ns[2] = (int32_t)(hn3 - hn2);  // wait
    if (A.r_hdr[1]) {  // pool too small: hdr[0] is the exact demand. The staged inputs are still in place: search again
      ctx->proj_ent_cap = std::max(ctx->proj_ent_cap * 2, A.r_hdr[0] + 1024);
      ctx->stat_retries++;
      continue;
    }
    nm = A.r_hdr[2];
    for (int k = 0; k < 6; k++) ctx->stat_proj_ns[k] = A.r_hdr[8 + k];
    break;
  }
  if (query_best && nq > 0) {
    memcpy(query_best, A.r_qbest, (size_t)nq * 4);
    if (query_dist) memcpy(query_dist, A.r_qdist, (size_t)nq * 4);
  }
  if (assign && n_t > 0) memcpy(assign, A.r_assign, (size_t)n_t * 4);
  if (dist_t && n_t > 0) memcpy(dist_t, A.r_dist, (size_t)n_t * 4);
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
  if ((rc = use_set(ctx, f1)) || (rc = use_set(ctx, f2))) return rc;
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
  if (n_l == 0) return ASDM_OK;
  int rc;
  if ((rc = use_set(ctx, left)) || (rc = use_set(ctx, right))) return rc;
  // one launch and one wait: the level table rides in the kernel parameters, the results are stored by the kernel into
  // the page-locked arena (zero-copy) and handed to the caller's arrays after the wait
  const size_t NL = ((size_t)n_l + 3) & ~(size_t)3;  // best_r | best_dist, each padded to 16 bytes
  void* hv = nullptr;
  if ((rc = stage_alloc(ctx, 2 * NL * 4, &hv))) return rc;
  if ((rc = ensure(ctx, ctx->scratch[SC_STATE], 2 * NL * 4))) return rc;
  if (!ctx->proj_ctl) {
    CK(cudaMalloc(&ctx->proj_ctl, 256));
    CK(cudaMemsetAsync(ctx->proj_ctl, 0, 256, ctx->stream));
  }
  int32_t* d_best = (int32_t*)ctx->scratch[SC_STATE].p;
  float* d_dist = (float*)(d_best + NL);
  int32_t* h_best = (int32_t*)hv;
  float* h_dist = (float*)(h_best + NL);
  StereoLevels sf;
  memset(&sf, 0, sizeof sf);
  memcpy(sf.v, scale_factors, (size_t)n_levels * 4);
  const size_t band_bytes = (size_t)std::max(right->n, 1) * sizeof(int4);
  if (band_bytes > 160 * 1024) return fail(ctx, ASDM_ERR_INVALID, "right image has too many keypoints for the stereo search (10240)");
  k_stereo_match<<<(n_l + STEREO_WARPS - 1) / STEREO_WARPS, STEREO_WARPS * 32, band_bytes, ctx->stream>>>(
      left->desc, left->kps, n_l, right->desc, right->kps, right->n, sf, n_levels, n_rows, min_d, max_d, d_best, d_dist,
      (uint32_t*)ctx->proj_ctl + 8, (int4*)hv);
  LAUNCHED(ctx);
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(ctx->stream));
  memcpy(best_r, h_best, (size_t)n_l * 4);
  if (best_dist) memcpy(best_dist, h_dist, (size_t)n_l * 4);
  synced(ctx);
  return ASDM_OK;
}


// ------------------------------------------------------------------------------------------------------------------
// asdm_stereo_row_match_static: the stereo descriptor stage for callers that may not allocate (code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc). No counterpart in the reference.
// Same kernel as asdm_stereo_row_match; what that call takes from the context's pools -- the
// ticket of the last-CTA copy and the device copy of best_r | best_dist -- lives in ONE file-scope device array that every
// GPU loading the library holds until exit:
//   [0, 256)                         control block (ticket at +0), cleared on the stream before every launch
//   [256, 256 + 2 * 16384 * 4)       best_r | best_dist of up to ASDM_STATIC_MAX_LEFT left keypoints
// Every call shares the array and it keeps the last call's contents: calls that overlap on the GPU (two streams, two
// threads) race on it. In host code `g_stereo_static` is a host-side stand-in for the symbol, NOT a device address: the
// address the kernel gets comes from cudaGetSymbolAddress, once per device, with that device current.
#define STEREO_STATIC_CTL_BYTES 256
#define STEREO_STATIC_BYTES (STEREO_STATIC_CTL_BYTES + 2 * ASDM_STATIC_MAX_LEFT * 4)
static __device__ __align__(256) unsigned char g_stereo_static[STEREO_STATIC_BYTES];

enum { STATIC_MAX_DEVICES = 64 };
static std::mutex g_stereo_static_mu;
static unsigned char* g_stereo_static_addr[STATIC_MAX_DEVICES];  // per device, resolved on the first call on that device

static int stereo_static_scratch(asdm_ctx* ctx, unsigned char** out) {
  if (ctx->device < 0 || ctx->device >= STATIC_MAX_DEVICES) return fail(ctx, ASDM_ERR_INVALID, "device index beyond the static-scratch table");
  std::lock_guard<std::mutex> lk(g_stereo_static_mu);
  if (!g_stereo_static_addr[ctx->device]) {  // set-up, once per device (the caller made ctx->device current)
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_stereo_static);
    if (e != cudaSuccess || !p) {
      cudaGetLastError();
      return fail(ctx, ASDM_ERR_CUDA, "cudaGetSymbolAddress(static stereo scratch) failed");
    }
    g_stereo_static_addr[ctx->device] = (unsigned char*)p;
  }
  *out = g_stereo_static_addr[ctx->device];
  return ASDM_OK;
}

extern "C" int asdm_stereo_row_match_static(asdm_ctx* ctx, asdm_set* left, asdm_set* right, const float* scale_factors,
                                            int n_levels, int n_rows, float min_d, float max_d, void* out, void* stream) {
  if (!ctx || !left || !right || !scale_factors || n_levels <= 0 || n_levels > ASDM_MAX_LEVELS || !out) return ASDM_ERR_INVALID;
  if (!left->has_kps || !right->has_kps) return fail(ctx, ASDM_ERR_INVALID, "sets need keypoints");
  if (left->ctx->family != ctx->family || right->ctx->family != ctx->family)
    return fail(ctx, ASDM_ERR_INVALID, "set belongs to a context of another family (asdm_fork)");
  if (left->n > ASDM_STATIC_MAX_LEFT)
    return fail(ctx, ASDM_ERR_INVALID, "asdm_stereo_row_match_static: more left keypoints than the static scratch holds "
                                       "(ASDM_STATIC_MAX_LEFT = 16384); use asdm_stereo_row_match");
  if (reinterpret_cast<uintptr_t>(out) & 15) return fail(ctx, ASDM_ERR_INVALID, "asdm_stereo_row_match_static: `out` must be 16-byte aligned");
  const size_t band_bytes = (size_t)std::max(right->n, 1) * sizeof(int4);
  if (band_bytes > 160 * 1024) return fail(ctx, ASDM_ERR_INVALID, "right image has too many keypoints for the stereo search (10240)");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  CK(cudaSetDevice(ctx->device));
  const int n_l = left->n;
  if (n_l == 0) return ASDM_OK;
  unsigned char* scratch = nullptr;
  int rc = stereo_static_scratch(ctx, &scratch);
  if (rc) return rc;
  cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
  // the sets' arrays are complete behind their `ready` events (a no-op on the stream that wrote them)
  if (left->ready) CK(cudaStreamWaitEvent(st, left->ready, 0));
  if (right->ready) CK(cudaStreamWaitEvent(st, right->ready, 0));
  const size_t NL = ((size_t)n_l + 3) & ~(size_t)3;
  int32_t* d_best = (int32_t*)(scratch + STEREO_STATIC_CTL_BYTES);
  float* d_dist = (float*)(d_best + NL);
  StereoLevels sf;
  memset(&sf, 0, sizeof sf);
  memcpy(sf.v, scale_factors, (size_t)n_levels * 4);
  // what must start at zero: the ticket (a call that was aborted mid-kernel would otherwise leave it counting)
  CK(cudaMemsetAsync(scratch, 0, STEREO_STATIC_CTL_BYTES, st));
  k_stereo_match<<<(n_l + STEREO_WARPS - 1) / STEREO_WARPS, STEREO_WARPS * 32, band_bytes, st>>>(
      left->desc, left->kps, n_l, right->desc, right->kps, right->n, sf, n_levels, n_rows, min_d, max_d, d_best, d_dist,
      (uint32_t*)scratch, (int4*)out);
  ctx->stat_total_launches++;
  CK(cudaGetLastError());
  return ASDM_OK;
}
