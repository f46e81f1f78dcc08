// kernels_proj.cuh -- projection-gated searches (K3): the descriptor loops of the four SearchByProjection
// overloads (ORBmatcher.cc:44-122, 300-413, 1318-1452, 1455-1582), SearchForInitialization (:416-531) and the
// descriptor stage of Frame::ComputeStereoMatches (Frame.cc:360-444).
//
//  k_proj_gather  : Frame/KeyFrame::GetFeaturesInArea (Frame.cc:219-274, KeyFrame.cc:839-878) per query, written
//                   as CSR segments in the reference's scan order (ix outer, iy inner, insertion order in a cell)
//  k_proj_dist    : exact fp32 distance of every gathered (query, keypoint) entry
//  k_proj_resolve : replays the greedy query order as a parallel fixed point (same argument as k_resolve_bow)
//  k_init_resolve : SearchForInitialization's stealing semantics, replayed in query order by one warp
//  k_stereo_match : row-band stereo candidates, independent per left keypoint
#pragma once
#include "internal.cuh"

struct ProjSeg {
  int32_t off, cnt;
};

__device__ __forceinline__ bool area_cells(const asdm_grid& g, float x, float y, float r, int& x0, int& x1, int& y0,
                                           int& y1) {
  // Frame.cc:226-240. Float->int conversions are clamped first so that huge windows stay defined.
  const float fx0 = floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, g.min_x), r), g.inv_w));
  const float fx1 = ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, g.min_x), r), g.inv_w));
  const float fy0 = floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, g.min_y), r), g.inv_h));
  const float fy1 = ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, g.min_y), r), g.inv_h));
  if (!(fx0 == fx0) || !(fx1 == fx1) || !(fy0 == fy0) || !(fy1 == fy1)) return false;
  const float lim = 1.0e6f;
  x0 = max(0, (int)fminf(fmaxf(fx0, -lim), lim));
  if (x0 >= g.cols) return false;
  x1 = min(g.cols - 1, (int)fminf(fmaxf(fx1, -lim), lim));
  if (x1 < 0) return false;
  y0 = max(0, (int)fminf(fmaxf(fy0, -lim), lim));
  if (y0 >= g.rows) return false;
  y1 = min(g.rows - 1, (int)fminf(fmaxf(fy1, -lim), lim));
  if (y1 < 0) return false;
  return true;
}

// Fuse's per-keypoint reprojection gate (ORBmatcher.cc:919-926); chi2_max <= 0 switches it off
struct ProjGate {
  double chi2_max;
  int32_t n_levels;
  float inv_level_sigma2[ASDM_MAX_LEVELS];
};

__device__ __forceinline__ bool area_keep(const asdm_kp& kp, float x, float y, float r, bool check_levels,
                                          int min_level, int max_level, const ProjGate& G) {
  if (check_levels) {  // Frame.cc:254-261
    if (kp.octave < min_level) return false;
    if (max_level >= 0 && kp.octave > max_level) return false;
  }
  const float distx = __fsub_rn(kp.x, x), disty = __fsub_rn(kp.y, y);
  if (!(fabsf(distx) < r && fabsf(disty) < r)) return false;  // Frame.cc:266
  if (G.chi2_max > 0.0 && kp.octave >= 0 && kp.octave < G.n_levels) {
    const float ex = __fsub_rn(x, kp.x), ey = __fsub_rn(y, kp.y);
    const float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
    if ((double)__fmul_rn(e2, G.inv_level_sigma2[kp.octave]) > G.chi2_max) return false;
  }
  return true;
}

// One thread per query. total[0] = entries allocated so far; overflow[0] set when the entry pool is too small.
__global__ void __launch_bounds__(128) k_proj_gather(const asdm_proj_query* __restrict__ q, int nq,
                                                     const asdm_kp* __restrict__ kps, asdm_grid g, ProjGate G,
                                                     const int32_t* __restrict__ cell_start,
                                                     const int32_t* __restrict__ cell_items,
                                                     ProjSeg* __restrict__ seg, int32_t* __restrict__ ent_j,
                                                     int32_t* __restrict__ ent_q, int ent_cap,
                                                     int32_t* __restrict__ total, int32_t* __restrict__ overflow) {
  const int iq = blockIdx.x * blockDim.x + threadIdx.x;
  if (iq >= nq) return;
  const asdm_proj_query Q = q[iq];
  ProjSeg s{0, 0};
  int x0, x1, y0, y1;
  if (Q.valid && area_cells(g, Q.u, Q.v, Q.r, x0, x1, y0, y1)) {
    const bool check_levels = (Q.min_level > 0) || (Q.max_level >= 0);  // Frame.cc:242
    int cnt = 0;
    for (int ix = x0; ix <= x1; ix++)
      for (int e = cell_start[ix * g.rows + y0]; e < cell_start[ix * g.rows + y1 + 1]; e++)
        cnt += area_keep(kps[cell_items[e]], Q.u, Q.v, Q.r, check_levels, Q.min_level, Q.max_level, G) ? 1 : 0;
    if (cnt > 0) {
      const int off = atomicAdd(total, cnt);
      if (off + cnt > ent_cap) {
        *overflow = 1;
      } else {
        s.off = off;
        s.cnt = cnt;
        int w = off;
        // cells (ix, y0..y1) are contiguous in the CSR (cell = ix*rows + iy), so this is the reference's scan order
        for (int ix = x0; ix <= x1; ix++)
          for (int e = cell_start[ix * g.rows + y0]; e < cell_start[ix * g.rows + y1 + 1]; e++) {
            const int j = cell_items[e];
            if (area_keep(kps[j], Q.u, Q.v, Q.r, check_levels, Q.min_level, Q.max_level, G)) {
              ent_j[w] = j;
              ent_q[w] = iq;
              w++;
            }
          }
      }
    }
  }
  seg[iq] = s;
}

// One thread per entry: 2 x 512 B rows, 384 flops. desc_q rows are indexed by q_row[iq] when given (queries that
// live in a descriptor set, e.g. the last frame), else by iq (map-point descriptors uploaded for this call).
__global__ void __launch_bounds__(256) k_proj_dist(const float* __restrict__ desc_q,
                                                   const int32_t* __restrict__ q_row,
                                                   const float* __restrict__ desc_t,
                                                   const int32_t* __restrict__ ent_j,
                                                   const int32_t* __restrict__ ent_q, int ent_cap,
                                                   const int32_t* __restrict__ total, float* __restrict__ ent_d) {
  // total[1] = overflow: a query that did not fit left a hole of unwritten entries behind it (the gather still advances
  // total[0] so that the host learns the exact demand) -- nothing may be dereferenced, the host grows the pool and
  // searches again
  if (total[1]) return;
  const int n = min(*total, ent_cap);
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x) {
    const int iq = ent_q[e];
    const int row = q_row ? q_row[iq] : iq;
    ent_d[e] = asdm_exact_dist(desc_q + (size_t)row * ASDM_DESC_DIM, desc_t + (size_t)ent_j[e] * ASDM_DESC_DIM);
  }
}

struct ProjParamsDev {
  int32_t mode;  // 0 = best & second with octaves (:44-122), 1 = best only
  float th_accept;
  float best_init;
  float nnratio;
  int32_t check_ori;
  int32_t count_twice;
};

// Single CTA. acc[q] = entry index accepted by query q (or -1); own[2][n_t] = smallest blocking query index that
// currently accepts keypoint j.
__global__ void __launch_bounds__(1024) k_proj_resolve(const asdm_proj_query* __restrict__ q, int nq,
                                                       const asdm_kp* __restrict__ kps_t, int n_t,
                                                       const uint8_t* __restrict__ blocked0,
                                                       const ProjSeg* __restrict__ seg,
                                                       const int32_t* __restrict__ ent_j,
                                                       const float* __restrict__ ent_d, ProjParamsDev P,
                                                       int32_t* __restrict__ acc, int32_t* __restrict__ own,
                                                       int32_t* __restrict__ assign, float* __restrict__ dist_t,
                                                       int32_t* __restrict__ nmatches,
                                                       int32_t* __restrict__ qbest /* [nq] or null */,
                                                       float* __restrict__ qdist /* [nq] or null */) {
  __shared__ int s_changed, s_count;
  __shared__ int s_hist[ASDM_HISTO_LENGTH];
  __shared__ int s_keep[3];
  const int t = threadIdx.x, nt = blockDim.x;
  for (int i = t; i < nq; i += nt) acc[i] = -1;
  for (int j = t; j < n_t; j += nt) {
    own[j] = INT_MAX;
    assign[j] = -1;
    dist_t[j] = -1.f;
  }
  if (t < ASDM_HISTO_LENGTH) s_hist[t] = 0;
  if (t == 0) s_count = 0;
  __syncthreads();

  for (int iter = 0;; iter++) {
    const int32_t* cur = own + (size_t)(iter & 1) * n_t;
    int32_t* nxt = own + (size_t)((iter & 1) ^ 1) * n_t;
    if (t == 0) s_changed = 0;
    __syncthreads();
    for (int i = t; i < nq; i += nt) {
      const ProjSeg s = seg[i];
      if (s.cnt == 0) continue;
      float best = P.best_init, best2 = P.best_init;
      int lvl = -1, lvl2 = -1, be = -1;
      for (int e = s.off; e < s.off + s.cnt; e++) {
        const int j = ent_j[e];
        if ((blocked0 && blocked0[j]) || cur[j] < i) continue;  // :86-88, :385, :1393-1395, :1524
        const float d = ent_d[e];
        if (d < best) {
          best2 = best; best = d;
          lvl2 = lvl; lvl = kps_t[j].octave;
          be = e;
        } else if (P.mode == 0 && d < best2) {  // :102-106
          lvl2 = kps_t[j].octave;
          best2 = d;
        }
      }
      int dec = -1;
      if (be >= 0 && best <= P.th_accept) {  // :110, :404, :1408, :1538
        dec = be;
        if (P.mode == 0 && lvl == lvl2 && best > __fmul_rn(P.nnratio, best2)) dec = -1;  // :112
      }
      if (dec != acc[i]) {
        acc[i] = dec;
        s_changed = 1;
      }
    }
    for (int j = t; j < n_t; j += nt) nxt[j] = INT_MAX;
    __syncthreads();
    for (int i = t; i < nq; i += nt) {
      const int e = acc[i];
      if (e >= 0 && q[i].blocks) atomicMin(nxt + ent_j[e], i);
    }
    __syncthreads();
    if (!s_changed) break;
    __syncthreads();
  }

  // F.mvpMapPoints[bestIdx] = pMP: the last accepting query owns the keypoint
  for (int i = t; i < nq; i += nt) {
    const int e = acc[i];
    if (qbest) {  // per-query view of the same decision (SearchBySim3 / Fuse have no exclusivity)
      qbest[i] = e >= 0 ? ent_j[e] : -1;
      qdist[i] = e >= 0 ? ent_d[e] : -1.f;
    }
    if (e < 0) continue;
    const int j = ent_j[e];
    atomicMax(assign + j, i);
    atomicAdd(&s_count, P.count_twice ? 2 : 1);  // :116-117
    if (P.check_ori) atomicAdd(&s_hist[asdm_rot_bin(q[i].angle, kps_t[j].angle)], 1);  // :1413-1423
  }
  __syncthreads();
  if (t == 0) {
    int pruned = 0;
    if (P.check_ori) {
      asdm_three_maxima(s_hist, ASDM_HISTO_LENGTH, s_keep[0], s_keep[1], s_keep[2]);
      for (int b = 0; b < ASDM_HISTO_LENGTH; b++)
        if (b != s_keep[0] && b != s_keep[1] && b != s_keep[2]) pruned += s_hist[b];
    }
    *nmatches = s_count - pruned;
  }
  for (int i = t; i < nq; i += nt) {
    const int e = acc[i];
    if (e >= 0 && assign[ent_j[e]] == i) dist_t[ent_j[e]] = ent_d[e];
  }
  __syncthreads();
  if (P.check_ori) {  // :1430-1449: every entry pushed into a pruned bin clears its keypoint
    for (int i = t; i < nq; i += nt) {
      const int e = acc[i];
      if (e < 0) continue;
      const int j = ent_j[e];
      const int b = asdm_rot_bin(q[i].angle, kps_t[j].angle);
      if (b != s_keep[0] && b != s_keep[1] && b != s_keep[2]) {
        assign[j] = -2;  // cleared by the rotation check, as opposed to -1 = never assigned
        dist_t[j] = -1.f;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// SearchForInitialization (ORBmatcher.cc:416-531). Later queries may steal a keypoint from an earlier one
// (:474-478) and vMatchedDistance gates candidates (:456), so the queries are replayed in order by ONE warp;
// the 32 lanes share each query's candidate scan (two-minimum reduction with scan-order tie-breaking).
__device__ __forceinline__ bool two_min_before(float d, int p, float d2, int p2) {
  return d < d2 || (d == d2 && p < p2);
}

__global__ void __launch_bounds__(32) k_init_resolve(const asdm_kp* __restrict__ kps1, int n1,
                                                     const asdm_kp* __restrict__ kps2, int n2,
                                                     const ProjSeg* __restrict__ seg,
                                                     const int32_t* __restrict__ ent_j,
                                                     const float* __restrict__ ent_d, float nnratio, int check_ori,
                                                     float* __restrict__ matched_dist /* [n2] */,
                                                     int32_t* __restrict__ matches21 /* [n2] */,
                                                     int32_t* __restrict__ matches12, float* __restrict__ dist12,
                                                     int32_t* __restrict__ pushed /* [n1] bin of the push or -1 */,
                                                     float* __restrict__ prev_matched,
                                                     int32_t* __restrict__ nmatches_out) {
  const int lane = threadIdx.x;
  for (int j = lane; j < n2; j += 32) {
    matched_dist[j] = 100.f;  // :427
    matches21[j] = -1;
  }
  for (int i = lane; i < n1; i += 32) {
    matches12[i] = -1;
    dist12[i] = -1.f;
    pushed[i] = -1;
  }
  __syncwarp();
  int nmatches;
  const int BIG = INT_MAX;
  for (int i1 = 0; i1 < n1; i1++) {
    const ProjSeg s = seg[i1];
    if (s.cnt == 0) continue;  // level1 > 0 (:433-435) or empty window (:439)
    // per-lane two smallest (distance, scan position) among admissible candidates
    float b1 = 100.f, b2 = 100.f;
    int p1 = BIG, p2 = BIG, j1 = -1;
    for (int c = lane; c < s.cnt; c += 32) {
      const int j = ent_j[s.off + c];
      const float d = ent_d[s.off + c];
      if (matched_dist[j] <= d) continue;  // :456
      if (!(d < 100.f)) continue;          // cannot beat the initial best/second
      if (two_min_before(d, c, b1, p1)) {
        b2 = b1; p2 = p1;
        b1 = d; p1 = c; j1 = j;
      } else if (two_min_before(d, c, b2, p2)) {
        b2 = d; p2 = c;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ob1 = __shfl_xor_sync(0xffffffffu, b1, o), ob2 = __shfl_xor_sync(0xffffffffu, b2, o);
      const int op1 = __shfl_xor_sync(0xffffffffu, p1, o), op2 = __shfl_xor_sync(0xffffffffu, p2, o);
      const int oj1 = __shfl_xor_sync(0xffffffffu, j1, o);
      if (two_min_before(ob1, op1, b1, p1)) {
        // other's best wins; second = min(my best, other's second)
        if (two_min_before(b1, p1, ob2, op2)) { b2 = b1; p2 = p1; } else { b2 = ob2; p2 = op2; }
        b1 = ob1; p1 = op1; j1 = oj1;
      } else {
        if (two_min_before(ob1, op1, b2, p2)) { b2 = ob1; p2 = op1; }
      }
    }
    // all lanes hold the same (b1, j1, b2)
    if (j1 >= 0 && b1 <= ASDM_TH_LOW && b1 < __fmul_rn(b2, nnratio)) {  // :470-472
      if (lane == 0) {
        const int prev = matches21[j1];
        if (prev >= 0) {  // :474-478
          matches12[prev] = -1;
          dist12[prev] = -1.f;
        }
        matches12[i1] = j1;
        dist12[i1] = b1;
        matches21[j1] = i1;
        matched_dist[j1] = b1;
        if (check_ori) pushed[i1] = asdm_rot_bin(kps1[i1].angle, kps2[j1].angle);  // :486-496
      }
    }
    __syncwarp();
  }
  // nmatches after the loop = accepts - steals (:477, :482) = queries that still hold a match
  int live = 0;
  for (int i = lane; i < n1; i += 32) live += matches12[i] >= 0 ? 1 : 0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) live += __shfl_xor_sync(0xffffffffu, live, o);
  nmatches = live;

  if (check_ori) {  // :500-523
    __shared__ int s_hist[ASDM_HISTO_LENGTH];
    __shared__ int s_keep[3];
    if (lane < ASDM_HISTO_LENGTH) s_hist[lane] = 0;
    __syncwarp();
    for (int i = lane; i < n1; i += 32)
      if (pushed[i] >= 0) atomicAdd(&s_hist[pushed[i]], 1);  // stale entries still count in the bin sizes
    __syncwarp();
    if (lane == 0) asdm_three_maxima(s_hist, ASDM_HISTO_LENGTH, s_keep[0], s_keep[1], s_keep[2]);
    __syncwarp();
    int removed = 0;
    for (int i = lane; i < n1; i += 32) {
      const int b = pushed[i];
      if (b < 0 || b == s_keep[0] || b == s_keep[1] || b == s_keep[2]) continue;
      if (matches12[i] >= 0) {  // :515
        matches12[i] = -1;
        dist12[i] = -1.f;
        removed++;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) removed += __shfl_xor_sync(0xffffffffu, removed, o);
    nmatches -= removed;
  }
  __syncwarp();
  for (int i = lane; i < n1; i += 32)  // :526-528
    if (matches12[i] >= 0) {
      prev_matched[2 * i] = kps2[matches12[i]].x;
      prev_matched[2 * i + 1] = kps2[matches12[i]].y;
    }
  if (lane == 0) *nmatches_out = nmatches;
}

// ---------------------------------------------------------------------------------------------------------
// Descriptor stage of Frame::ComputeStereoMatches (Frame.cc:360-444). One warp per left keypoint: the lanes test the
// right keypoints in index order (vRowIndices[row] is filled in ascending iR, Frame.cc:378-388) against the row band /
// octave / disparity gates and compact the survivors with ballots; every STEREO_BATCH survivors the warp computes their
// exact distances together -- the left row sits in registers, each right row is read by one coalesced 512-byte load,
// the squared differences go through shared memory and lane c adds candidate c's 128 terms in index order
// (ORBmatcher.cc:1637-1644) -- and keeps the first minimum (strict '<' in ascending iR, Frame.cc:438).
struct StereoLevels {  // mvScaleFactors by value: no upload, no extra launch
  float v[ASDM_MAX_LEVELS];
};
#define STEREO_WARPS 8
#define STEREO_BATCH 8
#define STEREO_ROW 132
__global__ void __launch_bounds__(STEREO_WARPS * 32) k_stereo_match(const float* __restrict__ desc_l,
                                                      const asdm_kp* __restrict__ kps_l, int n_l,
                                                      const float* __restrict__ desc_r,
                                                      const asdm_kp* __restrict__ kps_r, int n_r,
                                                      const StereoLevels scale_factors, int n_levels,
                                                      int n_rows, float min_d, float max_d,
                                                      int32_t* __restrict__ best_r, float* __restrict__ best_dist,
                                                      uint32_t* __restrict__ ticket, int4* __restrict__ host_out) {
  // best_r | best_dist are one device array [2 * n_l (padded to 4)]; the CTA that takes the last ticket copies it to the
  // page-locked block in 16-byte pieces (1500 scattered 4-byte stores over PCIe cost more than the whole search)
  __shared__ __align__(16) float s_sq[STEREO_WARPS][STEREO_BATCH * STEREO_ROW];
  __shared__ int s_pend[STEREO_WARPS][64];
  // Per right keypoint {first row, last row of its band (Frame.cc:381-385), octave, x}: computed once per CTA into
  // shared memory (the level table is a kernel parameter: indexing it per candidate would cost a chain of constant
  // loads), so a step of the scan is one LDS.128 and four compares per lane.
  extern __shared__ __align__(16) int4 s_band[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int iR = threadIdx.x; iR < n_r; iR += blockDim.x) {
    const asdm_kp kpR = kps_r[iR];
    const int oc = min(max(kpR.octave, 0), n_levels - 1);
    float sfv = 0.f;
#pragma unroll
    for (int l = 0; l < ASDM_MAX_LEVELS; l++) sfv = l == oc ? scale_factors.v[l] : sfv;
    const float r = __fmul_rn(2.0f, sfv);
    s_band[iR] = make_int4((int)floorf(__fsub_rn(kpR.y, r)), (int)ceilf(__fadd_rn(kpR.y, r)), kpR.octave, __float_as_int(kpR.x));
  }
  __syncthreads();
  const int iL = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const bool live = iL < n_l;
  const asdm_kp kpL = live ? kps_l[iL] : asdm_kp{0.f, -1.f, 0.f, 0};
  const float vL = kpL.y, uL = kpL.x;
  int out_idx = -1;
  float out_d = -1.f;
  if (live && vL >= 0.f && vL < (float)n_rows) {
    const int row = (int)vL;  // vRowIndices[vL] (Frame.cc:406)
    const float minU = __fsub_rn(uL, max_d), maxU = __fsub_rn(uL, min_d);
    if (!(maxU < 0.f)) {
      const float4 l4 = __ldg(reinterpret_cast<const float4*>(desc_l + (size_t)iL * ASDM_DESC_DIM) + lane);
      float bd = ASDM_TH_HIGH;  // :417
      int bi = INT_MAX;
      int np = 0;  // pending survivors in s_pend[warp][0..np)
      float* sq = s_sq[warp];
      int* pend = s_pend[warp];
      // distances of the first `nb` pending survivors (ascending iR); lane c < nb ends up with candidate c's distance
      auto flush = [&](int nb) {
        for (int c = 0; c < nb; c++) {
          const int iR = pend[c];
          const float4 r4 = __ldg(reinterpret_cast<const float4*>(desc_r + (size_t)iR * ASDM_DESC_DIM) + lane);
          float4 s;
          float d;
          d = __fsub_rn(l4.x, r4.x); s.x = __fmul_rn(d, d);
          d = __fsub_rn(l4.y, r4.y); s.y = __fmul_rn(d, d);
          d = __fsub_rn(l4.z, r4.z); s.z = __fmul_rn(d, d);
          d = __fsub_rn(l4.w, r4.w); s.w = __fmul_rn(d, d);
          *reinterpret_cast<float4*>(sq + c * STEREO_ROW + 4 * lane) = s;
        }
        __syncwarp();
        if (lane < nb) {
          const float4* rowp = reinterpret_cast<const float4*>(sq + lane * STEREO_ROW);
          float acc = 0.f;
#pragma unroll 8
          for (int k = 0; k < ASDM_DESC_DIM / 4; k++) {
            const float4 x = rowp[k];
            acc = __fadd_rn(acc, x.x); acc = __fadd_rn(acc, x.y); acc = __fadd_rn(acc, x.z); acc = __fadd_rn(acc, x.w);
          }
          const int iR = pend[lane];
          if (acc < bd || (acc == bd && iR < bi)) {  // first minimum in ascending iR
            bd = acc;
            bi = iR;
          }
        }
        __syncwarp();
        // the survivors behind the batch move to the front
        const int rest = np - nb;
        int mv = 0;
        if (lane < rest) mv = pend[nb + lane];
        __syncwarp();
        if (lane < rest) pend[lane] = mv;
        np = rest;
        __syncwarp();
      };
      for (int base = 0; base < n_r; base += 32) {
        const int iR = base + lane;
        bool keep = false;
        if (iR < n_r) {
          const int4 B = s_band[iR];
          const float xR = __int_as_float(B.w);
          const bool in_band = row >= B.x && row <= B.y;
          const bool lvl_ok = !(B.z < kpL.octave - 1 || B.z > kpL.octave + 1);  // :428
          keep = in_band && lvl_ok && xR >= minU && xR <= maxU;
        }
        const unsigned m = __ballot_sync(0xffffffffu, keep);
        if (m) {
          if (keep) pend[np + __popc(m & ((1u << lane) - 1u))] = iR;  // np < STEREO_BATCH here, so at most 39 entries
          np += __popc(m);
          __syncwarp();
          while (np >= STEREO_BATCH) flush(STEREO_BATCH);
        }
      }
      if (np > 0) flush(np);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float od = __shfl_xor_sync(0xffffffffu, bd, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (od < bd || (od == bd && oi < bi)) {
          bd = od;
          bi = oi;
        }
      }
      const int thOrbDist = (int)((ASDM_TH_HIGH + ASDM_TH_LOW) / 2);  // Frame.cc:366 -> 1
      if (bi != INT_MAX && bd < (float)thOrbDist) {
        out_idx = bi;
        out_d = bd;
      }
    }
  }
  if (live && lane == 0) {
    best_r[iL] = out_idx;
    best_dist[iL] = out_d;
  }
  __shared__ int s_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(ticket, 1u) == gridDim.x - 1 ? 1 : 0;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const int n16 = (2 * ((n_l + 3) & ~3)) / 4;
  const int4* src = reinterpret_cast<const int4*>(best_r);
  for (int k = threadIdx.x; k < n16; k += blockDim.x) host_out[k] = __ldcg(src + k);
  if (threadIdx.x == 0) *ticket = 0u;
}

// Queries of SearchForInitialization: level-0 keypoints of F1 search F2 around their previous match
// (ORBmatcher.cc:431-437).
__global__ void __launch_bounds__(256) k_init_queries(const asdm_kp* __restrict__ kps1, int n1,
                                                      const float* __restrict__ prev_matched, float window,
                                                      asdm_proj_query* __restrict__ q) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n1) return;
  const asdm_kp kp = kps1[i];
  asdm_proj_query Q;
  Q.valid = kp.octave > 0 ? 0 : 1;
  Q.blocks = 0;
  Q.pad = 0;
  Q.u = prev_matched[2 * i];
  Q.v = prev_matched[2 * i + 1];
  Q.r = window;
  Q.min_level = kp.octave;
  Q.max_level = kp.octave;
  Q.angle = kp.angle;
  q[i] = Q;
}

// ---------------------------------------------------------------------------------------------------------
// Warp-per-query variant of k_proj_gather for LARGE windows (SearchForInitialization: 200 px windows, hundreds of
// candidates per query): the 32 lanes share each column's contiguous CSR range and compact the kept keypoints with
// ballots, which preserves the reference's scan order exactly as the thread-per-query kernel does.
__global__ void __launch_bounds__(256) k_proj_gather_warp(const asdm_proj_query* __restrict__ q, int nq,
                                                          const asdm_kp* __restrict__ kps, asdm_grid g, ProjGate G,
                                                          const int32_t* __restrict__ cell_start,
                                                          const int32_t* __restrict__ cell_items,
                                                          ProjSeg* __restrict__ seg, int32_t* __restrict__ ent_j,
                                                          int32_t* __restrict__ ent_q, int ent_cap,
                                                          int32_t* __restrict__ total, int32_t* __restrict__ overflow) {
  const int lane = threadIdx.x & 31;
  const int iq = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (iq >= nq) return;
  const asdm_proj_query Q = q[iq];
  ProjSeg s{0, 0};
  int x0, x1, y0, y1;
  if (Q.valid && area_cells(g, Q.u, Q.v, Q.r, x0, x1, y0, y1)) {
    const bool check_levels = (Q.min_level > 0) || (Q.max_level >= 0);  // Frame.cc:242
    int cnt = 0;
    for (int ix = x0; ix <= x1; ix++) {
      const int b = cell_start[ix * g.rows + y0], e = cell_start[ix * g.rows + y1 + 1];
      for (int p = b; p < e; p += 32) {
        const bool keep = p + lane < e &&
                          area_keep(kps[cell_items[p + lane]], Q.u, Q.v, Q.r, check_levels, Q.min_level, Q.max_level, G);
        cnt += __popc(__ballot_sync(0xffffffffu, keep));
      }
    }
    if (cnt > 0) {
      int off = 0;
      if (lane == 0) off = atomicAdd(total, cnt);
      off = __shfl_sync(0xffffffffu, off, 0);
      if (off + cnt > ent_cap) {
        if (lane == 0) *overflow = 1;
      } else {
        s.off = off;
        s.cnt = cnt;
        int w = off;
        for (int ix = x0; ix <= x1; ix++) {
          const int b = cell_start[ix * g.rows + y0], e = cell_start[ix * g.rows + y1 + 1];
          for (int p = b; p < e; p += 32) {
            int j = -1;
            bool keep = false;
            if (p + lane < e) {
              j = cell_items[p + lane];
              keep = area_keep(kps[j], Q.u, Q.v, Q.r, check_levels, Q.min_level, Q.max_level, G);
            }
            const unsigned m = __ballot_sync(0xffffffffu, keep);
            if (keep) {
              const int at = w + __popc(m & ((1u << lane) - 1u));
              ent_j[at] = j;
              ent_q[at] = iq;
            }
            w += __popc(m);
          }
        }
      }
    }
  }
  if (lane == 0) seg[iq] = s;
}

// ---------------------------------------------------------------------------------------------------------
// SearchForInitialization (ORBmatcher.cc:416-531) without the serial replay. The decision of query i1 (which
// keypoint it accepts, at what distance) depends on earlier queries only through vMatchedDistance (:456), and
// vMatchedDistance[j] at the time of i1 = the smallest accepted distance among earlier queries that took j (every
// accept lowers it, :456 + :480). So the decisions are the unique fixed point of  D(i) = f(D(i') : i' < i);  Jacobi
// sweeps over all queries in parallel reach it (after sweep t the first t queries are final; in practice a handful of
// sweeps), and a sweep that changes nothing proves it. Steals (:474-478) do not feed back into decisions: they only
// decide which accept survives, resolved at the end (the last accept of a keypoint wins).
#define INIT_K 8  // accepts remembered per keypoint and sweep; more -> overflow flag -> the serial kernel takes over
struct InitBuf {
  int32_t* dec_j;  // [n1] keypoint accepted by query i, -1 = none
  float* dec_d;    // [n1]
  int32_t* cnt;    // [n2] accepts of keypoint j in this sweep
  int32_t* lst_i;  // [n2][INIT_K] accepting queries (unordered)
  float* lst_d;    // [n2][INIT_K] their distances
};
// state: [0] converged, [1] buffer holding the fixed point, [2] overflow, [4 + t] sweep t changed a decision
#define INIT_STATE_CHG 4

__global__ void __launch_bounds__(256) k_init_sweep(int t, int n1, int n2, const ProjSeg* __restrict__ seg,
                                                    const int32_t* __restrict__ ent_j, const float* __restrict__ ent_d,
                                                    float nnratio, InitBuf prev, InitBuf cur, InitBuf next,
                                                    int32_t* __restrict__ state) {
  if (state[0]) return;
  if (t >= 2 && state[INIT_STATE_CHG + t - 1] == 0) {  // the previous sweep changed nothing: fixed point
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      state[1] = (t - 1) % 3;
      __threadfence();
      state[0] = 1;
    }
    return;
  }
  for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < n2; j += gridDim.x * blockDim.x) next.cnt[j] = 0;
  const int lane = threadIdx.x & 31;
  const int i1 = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (i1 >= n1) return;
  const ProjSeg s = seg[i1];
  const int BIG = INT_MAX;
  float b1 = 100.f, b2 = 100.f;
  int p1 = BIG, p2 = BIG, j1 = -1;
  for (int c = lane; c < s.cnt; c += 32) {
    const int j = ent_j[s.off + c];
    const float d = ent_d[s.off + c];
    if (!(d < 100.f)) continue;
    // vMatchedDistance[j] as query i1 would find it (:456)
    float md = 100.f;
    const int na = min(prev.cnt[j], INIT_K);
    for (int a = 0; a < na; a++)
      if (prev.lst_i[j * INIT_K + a] < i1) md = fminf(md, prev.lst_d[j * INIT_K + a]);
    if (md <= d) continue;
    if (two_min_before(d, c, b1, p1)) {
      b2 = b1; p2 = p1;
      b1 = d; p1 = c; j1 = j;
    } else if (two_min_before(d, c, b2, p2)) {
      b2 = d; p2 = c;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ob1 = __shfl_xor_sync(0xffffffffu, b1, o), ob2 = __shfl_xor_sync(0xffffffffu, b2, o);
    const int op1 = __shfl_xor_sync(0xffffffffu, p1, o), op2 = __shfl_xor_sync(0xffffffffu, p2, o);
    const int oj1 = __shfl_xor_sync(0xffffffffu, j1, o);
    if (two_min_before(ob1, op1, b1, p1)) {
      if (two_min_before(b1, p1, ob2, op2)) { b2 = b1; p2 = p1; } else { b2 = ob2; p2 = op2; }
      b1 = ob1; p1 = op1; j1 = oj1;
    } else {
      if (two_min_before(ob1, op1, b2, p2)) { b2 = ob1; p2 = op1; }
    }
  }
  if (lane == 0) {
    const bool accept = j1 >= 0 && b1 <= ASDM_TH_LOW && b1 < __fmul_rn(b2, nnratio);  // :470-472
    const int dj = accept ? j1 : -1;
    const float dd = accept ? b1 : -1.f;
    if (dj != prev.dec_j[i1] || __float_as_int(dd) != __float_as_int(prev.dec_d[i1])) state[INIT_STATE_CHG + t] = 1;
    cur.dec_j[i1] = dj;
    cur.dec_d[i1] = dd;
    if (accept) {
      const int slot = atomicAdd(&cur.cnt[j1], 1);
      if (slot < INIT_K) {
        cur.lst_i[j1 * INIT_K + slot] = i1;
        cur.lst_d[j1 * INIT_K + slot] = b1;
      } else {
        state[2] = 1;
      }
    }
  }
}

// Tail of SearchForInitialization on the fixed point: which accept of every keypoint survives (the last one, :474-482),
// the rotation histogram over ALL accepts including stolen ones (:486-496, :515), nmatches and vbPrevMatched (:526-528).
__global__ void __launch_bounds__(1024) k_init_finalize(const asdm_kp* __restrict__ kps1, int n1,
                                                        const asdm_kp* __restrict__ kps2, int check_ori, InitBuf b0,
                                                        InitBuf b1, InitBuf b2, const int32_t* __restrict__ state,
                                                        int32_t* __restrict__ matches12, float* __restrict__ dist12,
                                                        float* __restrict__ prev_matched,
                                                        int32_t* __restrict__ nmatches_out) {
  if (!state[0] || state[2]) return;  // not converged yet / overflow: the host continues or falls back
  const InitBuf B = state[1] == 0 ? b0 : (state[1] == 1 ? b1 : b2);
  __shared__ int s_hist[ASDM_HISTO_LENGTH];
  __shared__ int s_keep[3];
  __shared__ int s_live, s_removed;
  const int t = threadIdx.x, nt = blockDim.x;
  if (t < ASDM_HISTO_LENGTH) s_hist[t] = 0;
  if (t == 0) { s_live = 0; s_removed = 0; }
  __syncthreads();
  int live = 0;
  for (int i = t; i < n1; i += nt) {
    const int j = B.dec_j[i];
    int m = -1;
    float d = -1.f;
    if (j >= 0) {
      bool last = true;  // no later query accepted j
      const int na = min(B.cnt[j], INIT_K);
      for (int a = 0; a < na; a++) last &= B.lst_i[j * INIT_K + a] <= i;
      if (last) { m = j; d = B.dec_d[i]; live++; }
      if (check_ori) atomicAdd(&s_hist[asdm_rot_bin(kps1[i].angle, kps2[j].angle)], 1);
    }
    matches12[i] = m;
    dist12[i] = d;
  }
  atomicAdd(&s_live, live);
  __syncthreads();
  if (check_ori) {  // :500-523
    if (t == 0) asdm_three_maxima(s_hist, ASDM_HISTO_LENGTH, s_keep[0], s_keep[1], s_keep[2]);
    __syncthreads();
    int removed = 0;
    for (int i = t; i < n1; i += nt) {
      const int j = B.dec_j[i];
      if (j < 0) continue;
      const int b = asdm_rot_bin(kps1[i].angle, kps2[j].angle);
      if (b == s_keep[0] || b == s_keep[1] || b == s_keep[2]) continue;
      if (matches12[i] >= 0) {  // :515
        matches12[i] = -1;
        dist12[i] = -1.f;
        removed++;
      }
    }
    atomicAdd(&s_removed, removed);
  }
  __syncthreads();
  for (int i = t; i < n1; i += nt)  // :526-528
    if (matches12[i] >= 0) {
      prev_matched[2 * i] = kps2[matches12[i]].x;
      prev_matched[2 * i + 1] = kps2[matches12[i]].y;
    }
  if (t == 0) *nmatches_out = s_live - s_removed;
}
