// jobs.cu -- batched keyframe-pair jobs (LoopClosing.cc:291-318, LocalMapping.cc:335-370), their publish epilogue and
// the single-pair BoW / triangulation entry points.
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <new>

#include "host_util.cuh"
#include "kernels_dense.cuh"
#include "jobs.cuh"

// Smallest keep threshold t such that a candidate farther than t can neither be accepted (d > TH_LOW) nor make
// the ratio test fail as second best: for every fp32 d > t, fl(nnratio*d) > th_accept >= best1.
float bow_keep_threshold(float nnratio, float th_accept) {
  if (!(nnratio > 0.f)) return ASDM_BEST_INIT_BOW;
  float t = th_accept / nnratio;
  if (!(t < ASDM_BEST_INIT_BOW)) return ASDM_BEST_INIT_BOW;
  for (int it = 0; it < 64 && !(nnratio * t > th_accept); it++) t = nextafterf(t, INFINITY);
  t = nextafterf(nextafterf(t, INFINITY), INFINITY);
  return std::min(std::max(t, th_accept), ASDM_BEST_INIT_BOW);
}

void jobs_init_device() {
  cudaFuncSetAttribute(k_cand_exact, cudaFuncAttributeMaxDynamicSharedMemorySize,
                       (int)(ASDM_DESC_DIM * (CE_TM + CE_TN) * sizeof(float)));
  cudaFuncSetAttribute(k_resolve_bow, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
}

// ------------------------------------------------------------------------------- batched keyframe-pair jobs
void job_free(asdm_job* j) {
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

int job_alloc_scratch(asdm_job* j) {
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

int job_prepare(asdm_ctx* ctx, const asdm_job_desc* d, int kind_internal, asdm_job** out) {
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
    if (!s || !d->masks[k]) return fail(ctx, ASDM_ERR_INVALID, "bad keyframe set / mask");
    if (s->ctx->family != ctx->family) return fail(ctx, ASDM_ERR_INVALID, "keyframe set belongs to a context of another family (asdm_fork)");
    if (s->n >= (1 << 24)) return fail(ctx, ASDM_ERR_INVALID, "keyframe set too large for a batched search (2^24 features)");
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
    const int ru = use_set(ctx, d->sets[k]);  // sets created on another context of the family: order behind their writes
    if (ru) return bail(ru);
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

int job_run(asdm_job* j) {
  asdm_ctx* ctx = j->ctx;
  CK(cudaSetDevice(ctx->device));
  ctx->stat_launches = 0;
  ctx->stat_tc_used = j->use_tc;
  ctx->stat_cand_cap = j->cap;
  j->ran = true;
  if (j->n_pairs == 0) return ASDM_OK;
  const PairDesc* pairs = (const PairDesc*)j->d_pairs.p;
  int32_t* flags = (int32_t*)j->d_flags.p;
  if (ctx->opt_time_kernels == 2) {
    if (!ctx->stage_ev[0]) CK(cudaEventCreate(&ctx->stage_ev[0]));
    CK(cudaEventRecord(ctx->stage_ev[0], ctx->stream));
  }
  {
    const int rc = zero_device(ctx, j->d_zero.p, j->zero_bytes);  // flags | hit-list segment counts | candidate counts
    if (rc) return rc;
  }
  // time_kernels == 2: one event behind every stage of this run (zero | distance | rerank | resolve), read back through
  // asdm_get_stat("job_us_<stage>") after a synchronisation -- the in-step duration of each kernel, warm caches, no profiler
  const bool stages = ctx->opt_time_kernels == 2;
  if (stages) {
    for (int k = 0; k < 5; k++)
      if (!ctx->stage_ev[k]) CK(cudaEventCreate(&ctx->stage_ev[k]));
    CK(cudaEventRecord(ctx->stage_ev[1], ctx->stream));  // [0] sits in front of the fill above
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
    if (stages) CK(cudaEventRecord(ctx->stage_ev[2], ctx->stream));
    if (j->n_seg > RR_MAX_SEG) return fail(ctx, ASDM_ERR_INTERNAL, "more hit-list segments than k_rerank can scan");
    int rr_per_sm = 4;  // persistent: exactly the CTAs that are resident (the batches are dealt round-robin over the grid)
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&rr_per_sm, k_rerank, RR_WARPS * 32, 0) != cudaSuccess || rr_per_sm < 1) {
      cudaGetLastError();
      rr_per_sm = 1;
    }
    k_rerank<<<ctx->sm_count * rr_per_sm, RR_WARPS * 32, 0, ctx->stream>>>(pairs, j->P, j->cap, j->n_stride, (const TcHit*)j->d_list.p,
                                                                   (const int32_t*)j->d_seg_count.p, j->n_seg, j->seg_cap,
                                                                   (int32_t*)j->d_cnt.p, (Cand*)j->d_cand.p, flags);
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
  if (stages) CK(cudaEventRecord(ctx->stage_ev[3], ctx->stream));
  if (j->kind == KIND_TRIANGULATION)
    k_resolve_tri<<<j->n_pairs, 512, 0, ctx->stream>>>(pairs, j->P, j->cap, j->n_stride, j->out_stride,
                                                       (const int32_t*)j->d_cnt.p, (const Cand*)j->d_cand.p,
                                                       (int32_t*)j->d_acc.p, (int32_t*)j->d_match.p,
                                                       (float*)j->d_dist.p, (int32_t*)j->d_nm.p, j->pub);
  else {
    // greedy resolve with its working set in shared memory; the record budget escalates when a pair had more queries
    // with candidates than fit: 100 KB (two CTAs per SM) -> 200 KB -> the global-memory kernel
    int level = std::max(j->resolve_level, ctx->resolve_level_hint);
    size_t smem = 0;
    int max_act = 0;
    for (; level < 2; level++) {
      const size_t budget = level == 0 ? 100 * 1024 : 200 * 1024, fixed = (size_t)j->m_stride * 4 + 64;
      if (fixed + 64 * (sizeof(BowRec) + 8) > budget) continue;
      max_act = (int)std::min<size_t>((size_t)j->n_stride, (budget - fixed) / (sizeof(BowRec) + 8));
      smem = rb_smem_bytes(j->m_stride, max_act);
      break;
    }
    j->resolve_level = level;
    if (level < 2)
      k_resolve_bow<<<j->n_pairs, 512, smem, ctx->stream>>>(pairs, j->P, j->cap, j->n_stride, j->m_stride, j->out_stride,
                                                            (const int32_t*)j->d_cnt.p, (Cand*)j->d_cand.p, (int32_t*)j->d_match.p,
                                                            (float*)j->d_dist.p, (int32_t*)j->d_nm.p, max_act, flags + 3, j->pub);
    else
      k_resolve_bow_global<<<j->n_pairs, 512, 0, ctx->stream>>>(pairs, j->P, j->cap, j->n_stride, j->m_stride, j->out_stride,
                                                                (const int32_t*)j->d_cnt.p, (Cand*)j->d_cand.p,
                                                                (int32_t*)j->d_owner.p, (int32_t*)j->d_acc.p,
                                                                (int32_t*)j->d_match.p, (float*)j->d_dist.p,
                                                                (int32_t*)j->d_nm.p, j->pub);
  }
  LAUNCHED(ctx);
  if (stages) CK(cudaEventRecord(ctx->stage_ev[4], ctx->stream));
  CK(cudaGetLastError());
  return ASDM_OK;
}

// Waits for the run; when a candidate pool overflowed, grows it and runs again (results are only ever taken
// from a run without overflow).
int job_settle(asdm_job* j) {
  asdm_ctx* ctx = j->ctx;
  CK(cudaSetDevice(ctx->device));  // a process may drive several GPUs (asdm_group_create): never rely on the current device
  if (!j->ran) {
    int rc = job_run(j);
    if (rc) return rc;
  }
  ctx->stat_retries = 0;
  if (j->n_pairs == 0) {  // a rank that was dealt no unit: nothing ran, nothing to validate (d_flags was never cleared)
    ctx->stat_tc_pairs = 0;
    return ASDM_OK;
  }
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
    const bool cand_over = j->h_flags[0] != 0, list_over = j->h_flags[1] != 0, rec_over = j->h_flags[3] != 0;
    ctx->stat_tc_pairs = j->use_tc ? j->h_flags[2] : 0;
    if (!cand_over && !list_over && !rec_over) return ASDM_OK;
    ctx->stat_retries++;
    if (rec_over) {  // more queries with candidates than the resolve kernel's shared-memory records: next budget
      j->resolve_level = std::min(j->resolve_level + 1, 2);
      ctx->resolve_level_hint = std::max(ctx->resolve_level_hint, j->resolve_level);
    }
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
  // The resolve kernel stores the finished row straight into the page-locked arena (its publish epilogue, one
  // destination): no copy nodes behind the kernels, one wait per call.
  const size_t n = (size_t)j->out_stride, row_bytes = align_up(n * 4, 64);
  void* hres = nullptr;
  if (!rc && 2 * row_bytes + 64 > ctx->res_cap) {  // (not the staging arena: a rerun inside job_settle may stage again)
    cudaStreamSynchronize(ctx->stream);
    if (ctx->res_host) cudaFreeHost(ctx->res_host);
    ctx->res_host = nullptr;
    ctx->res_cap = 0;
    const size_t want = std::max(2 * (2 * row_bytes + 64), (size_t)256 << 10);
    if (cudaHostAlloc((void**)&ctx->res_host, want, cudaHostAllocMapped) != cudaSuccess) {
      cudaGetLastError();
      rc = fail(ctx, ASDM_ERR_OOM, "page-locked result block");
    } else {
      ctx->res_cap = want;
    }
  }
  hres = ctx->res_host;
  if (!rc) {
    char* hb = (char*)hres;
    j->pub = PublishArgs();
    j->pub.n = 1;
    j->pub.stride = (int)n;
    j->pub.match[0] = (int32_t*)hb;
    j->pub.dist[0] = (float*)(hb + row_bytes);
    j->pub.nm[0] = (int32_t*)(hb + 2 * row_bytes);
    j->pub.no_dist = dist ? 0 : 1;
    memset(hb, 0xff, row_bytes);  // an empty search launches nothing: -1 rows, zero matches
    memset(hb + row_bytes, 0, row_bytes + 64);
    rc = job_run(j);
  }
  if (!rc) rc = job_settle(j);
  const int64_t launches = ctx->stat_launches;
  if (!rc) {
    const char* hb = (const char*)hres;
    memcpy(match, hb, n * sizeof(int32_t));
    if (dist) memcpy(dist, hb + row_bytes, n * sizeof(float));
    if (nmatches) *nmatches = *(const int32_t*)(hb + 2 * row_bytes);
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
