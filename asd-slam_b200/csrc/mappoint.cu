// mappoint.cu -- MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:271-338) for batches of map points.
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <new>

#include "host_util.cuh"
#include "kernels_mappoint.cuh"

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
  for (int k = 0; k < n_sets; k++)
    if (sets[k]) {
      const int ru = use_set(ctx, sets[k]);
      if (ru) return ru;
    }
  std::vector<int32_t> small, large;
  std::vector<int64_t> d_off;
  int64_t scratch_floats = 0;
  for (int p = 0; p < n_points; p++) {
    const int N = obs_off[p + 1] - obs_off[p];
    if (N < 0) return fail(ctx, ASDM_ERR_INVALID, "observation offsets not ascending");
    for (int e = obs_off[p]; e < obs_off[p + 1]; e++) {
      if (obs_set[e] < 0 || obs_set[e] >= n_sets || !sets[obs_set[e]] || sets[obs_set[e]]->ctx->family != ctx->family)
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
