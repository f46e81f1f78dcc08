/*
 * asd_oracle.cc -- CPU ORACLE (test infrastructure, NOT product code; see asd_oracle.h).
 *
 * Restatement over flat arrays of the ASD-SLAM reference matcher.  Each function names the
 * reference file:line it follows (paths relative to /root/reference/src/vslam/).
 * Build: g++ -O2 -ffp-contract=off -fno-fast-math (oracle/Makefile).
 */
#include "asd_oracle.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <map>
#include <utility>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

const float TH_HIGH = 1.5f; /* src/ORBmatcher.cc:37 */
const float TH_LOW = 0.5f;  /* src/ORBmatcher.cc:38 */
const int HISTO_LENGTH = ORA_HISTO_LENGTH;

/* src/ORBmatcher.cc:1637-1644: one fp32 accumulator, i ascending, (a-b)*(a-b) then add.
 * The CV_64F round trip at :1633-1636 is the identity on floats. */
inline float dd(const float* a, const float* b) {
  float sqd = 0.f;
  for (int i = 0; i < ORA_DESC_DIM; i += 4) {
    sqd += (a[i] - b[i]) * (a[i] - b[i]);
    sqd += (a[i + 1] - b[i + 1]) * (a[i + 1] - b[i + 1]);
    sqd += (a[i + 2] - b[i + 2]) * (a[i + 2] - b[i + 2]);
    sqd += (a[i + 3] - b[i + 3]) * (a[i + 3] - b[i + 3]);
  }
  return sqd;
}

/* e.g. src/ORBmatcher.cc:245-250 */
inline int rot_bin(float a1, float a2) {
  const float factor = 1.0f / HISTO_LENGTH;
  float rot = a1 - a2;
  if (rot < 0.0) rot += 360.0f;
  int bin = (int)std::round(rot * factor);
  if (bin == HISTO_LENGTH) bin = 0;
  return bin;
}

/* src/ORBmatcher.cc:1584-1625 */
void three_maxima(const std::vector<int>* histo, int L, int& ind1, int& ind2, int& ind3) {
  int max1 = 0, max2 = 0, max3 = 0;
  for (int i = 0; i < L; i++) {
    const int s = (int)histo[i].size();
    if (s > max1) {
      max3 = max2; max2 = max1; max1 = s;
      ind3 = ind2; ind2 = ind1; ind1 = i;
    } else if (s > max2) {
      max3 = max2; max2 = s;
      ind3 = ind2; ind2 = i;
    } else if (s > max3) {
      max3 = s; ind3 = i;
    }
  }
  if (max2 < 0.1f * (float)max1) {
    ind2 = -1; ind3 = -1;
  } else if (max3 < 0.1f * (float)max1) {
    ind3 = -1;
  }
}

/* Grid owned by a Frame/KeyFrame: mGrid[ix][iy] vectors in insertion order. */
struct Grid {
  ora_grid g;
  std::vector<std::vector<int32_t>> cells; /* ix*rows+iy */
};

/* src/Frame.cc:276-286 (KeyFrame.cc:132-143) */
inline bool pos_in_grid(const ora_kp& kp, const ora_grid& g, int& px, int& py) {
  px = (int)std::round((kp.x - g.min_x) * g.inv_w);
  py = (int)std::round((kp.y - g.min_y) * g.inv_h);
  if (px < 0 || px >= g.cols || py < 0 || py >= g.rows) return false;
  return true;
}

/* src/Frame.cc:123-138 */
void build_grid(const ora_kp* kps, int n, const ora_grid& g, Grid& out) {
  out.g = g;
  out.cells.assign((size_t)g.cols * g.rows, std::vector<int32_t>());
  for (int i = 0; i < n; i++) {
    int px, py;
    if (pos_in_grid(kps[i], g, px, py)) out.cells[(size_t)px * g.rows + py].push_back(i);
  }
}

/* src/Frame.cc:219-274; KeyFrame.cc:839-878 is the same with no level test */
void features_in_area(const ora_kp* kps, const Grid& G, float x, float y, float r, int minLevel, int maxLevel,
                      std::vector<int32_t>& vIndices) {
  vIndices.clear();
  const ora_grid& g = G.g;
  const int nMinCellX = std::max(0, (int)std::floor((x - g.min_x - r) * g.inv_w));
  if (nMinCellX >= g.cols) return;
  const int nMaxCellX = std::min((int)g.cols - 1, (int)std::ceil((x - g.min_x + r) * g.inv_w));
  if (nMaxCellX < 0) return;
  const int nMinCellY = std::max(0, (int)std::floor((y - g.min_y - r) * g.inv_h));
  if (nMinCellY >= g.rows) return;
  const int nMaxCellY = std::min((int)g.rows - 1, (int)std::ceil((y - g.min_y + r) * g.inv_h));
  if (nMaxCellY < 0) return;

  const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);

  for (int ix = nMinCellX; ix <= nMaxCellX; ix++) {
    for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
      const std::vector<int32_t>& vCell = G.cells[(size_t)ix * g.rows + iy];
      for (size_t j = 0; j < vCell.size(); j++) {
        const ora_kp& kp = kps[vCell[j]];
        if (bCheckLevels) {
          if (kp.octave < minLevel) continue;
          if (maxLevel >= 0)
            if (kp.octave > maxLevel) continue;
        }
        const float distx = kp.x - x;
        const float disty = kp.y - y;
        if (std::fabs(distx) < r && std::fabs(disty) < r) vIndices.push_back(vCell[j]);
      }
    }
  }
}

/* src/ORBmatcher.cc:136-153 */
inline bool check_dist_epipolar_line(const ora_kp& kp1, const ora_kp& kp2, const float* F12,
                                     const float* level_sigma2_2) {
  const float a = kp1.x * F12[0] + kp1.y * F12[3] + F12[6];
  const float b = kp1.x * F12[1] + kp1.y * F12[4] + F12[7];
  const float c = kp1.x * F12[2] + kp1.y * F12[5] + F12[8];
  const float num = a * kp2.x + b * kp2.y + c;
  const float den = a * a + b * b;
  if (den == 0) return false;
  const float dsqr = num * num / den;
  return dsqr < 3.84 * level_sigma2_2[kp2.octave];
}

/* Merge-join of two FeatureVectors by NodeId (src/ORBmatcher.cc:181-271 and twins). The
 * lower_bound jumps of the reference visit exactly the equal-id pairs in ascending order. */
template <class Fn>
void for_shared_nodes(const ora_featvec* a, const ora_featvec* b, Fn fn) {
  int ia = 0, ib = 0;
  while (ia < a->n_nodes && ib < b->n_nodes) {
    if (a->node_ids[ia] == b->node_ids[ib]) {
      fn(ia, ib);
      ia++; ib++;
    } else if (a->node_ids[ia] < b->node_ids[ib]) {
      while (ia < a->n_nodes && a->node_ids[ia] < b->node_ids[ib]) ia++;
    } else {
      while (ib < b->n_nodes && b->node_ids[ib] < a->node_ids[ia]) ib++;
    }
  }
}

}  // namespace

extern "C" {

float ora_descriptor_distance(const float* a, const float* b) { return dd(a, b); }

int ora_rot_bin(float angle1, float angle2) { return rot_bin(angle1, angle2); }

void ora_compute_three_maxima(const int32_t* sizes, int L, int32_t* ind1, int32_t* ind2, int32_t* ind3) {
  std::vector<std::vector<int>> h(L);
  for (int i = 0; i < L; i++) h[i].assign(sizes[i], 0);
  int a = -1, b = -1, c = -1;
  three_maxima(h.data(), L, a, b, c);
  *ind1 = a; *ind2 = b; *ind3 = c;
}

void ora_assign_features_to_grid(const ora_kp* kps, int n, const ora_grid* g, int32_t* cell_start,
                                 int32_t* cell_items, int32_t* cell_of) {
  Grid G;
  build_grid(kps, n, *g, G);
  for (int i = 0; i < n; i++) cell_of[i] = -1;
  int pos = 0;
  const int nc = g->cols * g->rows;
  for (int c = 0; c < nc; c++) {
    cell_start[c] = pos;
    for (int32_t i : G.cells[c]) {
      cell_items[pos++] = i;
      cell_of[i] = c;
    }
  }
  cell_start[nc] = pos;
}

int ora_get_features_in_area(const ora_kp* kps, int n, const ora_grid* g, const int32_t* cell_start,
                             const int32_t* cell_items, float x, float y, float r, int min_level,
                             int max_level, int32_t* out) {
  (void)n;
  Grid G;
  G.g = *g;
  const int nc = g->cols * g->rows;
  G.cells.resize(nc);
  for (int c = 0; c < nc; c++) G.cells[c].assign(cell_items + cell_start[c], cell_items + cell_start[c + 1]);
  std::vector<int32_t> v;
  features_in_area(kps, G, x, y, r, min_level, max_level, v);
  for (size_t k = 0; k < v.size(); k++) out[k] = v[k];
  return (int)v.size();
}

/* ---------------------------------------------------------------- a8: src/ORBmatcher.cc:156-297 */
int ora_search_by_bow_kf_f(const float* desc_kf, const ora_kp* kps_kf, int n_kf, const ora_featvec* fv_kf,
                           const uint8_t* valid_kf, const float* desc_f, const ora_kp* kps_f, int n_f,
                           const ora_featvec* fv_f, float nnratio, int check_ori, int32_t* match_f,
                           float* dist_f) {
  (void)n_kf;
  for (int j = 0; j < n_f; j++) match_f[j] = -1; /* :160 */
  if (dist_f) for (int j = 0; j < n_f; j++) dist_f[j] = -1.f;
  int nmatches = 0;
  std::vector<int> rotHist[HISTO_LENGTH];

  for_shared_nodes(fv_kf, fv_f, [&](int ka, int kb) {
    for (int iKF = fv_kf->off[ka]; iKF < fv_kf->off[ka + 1]; iKF++) { /* :190 */
      const int realIdxKF = fv_kf->idx[iKF];
      if (!valid_kf[realIdxKF]) continue; /* :194-200 */
      const float* dKF = desc_kf + (size_t)realIdxKF * ORA_DESC_DIM;
      float bestDist1 = 256;
      int bestIdxF = -1;
      float bestDist2 = 256;
      for (int iF = fv_f->off[kb]; iF < fv_f->off[kb + 1]; iF++) { /* :208 */
        const int realIdxF = fv_f->idx[iF];
        if (match_f[realIdxF] >= 0) continue; /* :212 */
        const float dist = dd(dKF, desc_f + (size_t)realIdxF * ORA_DESC_DIM);
        if (dist < bestDist1) {
          bestDist2 = bestDist1;
          bestDist1 = dist;
          bestIdxF = realIdxF;
        } else if (dist < bestDist2) {
          bestDist2 = dist;
        }
      }
      if (bestDist1 <= TH_LOW) {                 /* :231 */
        if (bestDist1 < nnratio * bestDist2) {   /* :234 */
          match_f[bestIdxF] = realIdxKF;         /* :237 */
          if (dist_f) dist_f[bestIdxF] = bestDist1;
          if (check_ori) rotHist[rot_bin(kps_kf[realIdxKF].angle, kps_f[bestIdxF].angle)].push_back(bestIdxF);
          nmatches++;
        }
      }
    }
  });

  if (check_ori) { /* :274-294 */
    int ind1 = -1, ind2 = -1, ind3 = -1;
    three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (size_t j = 0; j < rotHist[i].size(); j++) {
        match_f[rotHist[i][j]] = -1;
        if (dist_f) dist_f[rotHist[i][j]] = -1.f;
        nmatches--;
      }
    }
  }
  return nmatches;
}

/* ---------------------------------------------------------------- a9: src/ORBmatcher.cc:533-666 */
int ora_search_by_bow_kf_kf(const float* desc1, const ora_kp* kps1, int n1, const ora_featvec* fv1,
                            const uint8_t* valid1, const float* desc2, const ora_kp* kps2, int n2,
                            const ora_featvec* fv2, const uint8_t* valid2, float nnratio, int check_ori,
                            int32_t* match12, float* dist12) {
  for (int i = 0; i < n1; i++) match12[i] = -1; /* :545 */
  if (dist12) for (int i = 0; i < n1; i++) dist12[i] = -1.f;
  std::vector<uint8_t> vbMatched2(n2, 0);
  std::vector<int> rotHist[HISTO_LENGTH];
  int nmatches = 0;

  for_shared_nodes(fv1, fv2, [&](int ka, int kb) {
    for (int p1 = fv1->off[ka]; p1 < fv1->off[ka + 1]; p1++) { /* :565 */
      const int idx1 = fv1->idx[p1];
      if (!valid1[idx1]) continue; /* :569-573 */
      const float* d1 = desc1 + (size_t)idx1 * ORA_DESC_DIM;
      float bestDist1 = 256;
      int bestIdx2 = -1;
      float bestDist2 = 256;
      for (int p2 = fv2->off[kb]; p2 < fv2->off[kb + 1]; p2++) { /* :581 */
        const int idx2 = fv2->idx[p2];
        if (vbMatched2[idx2] || !valid2[idx2]) continue; /* :587-591 */
        const float dist = dd(d1, desc2 + (size_t)idx2 * ORA_DESC_DIM);
        if (dist < bestDist1) {
          bestDist2 = bestDist1;
          bestDist1 = dist;
          bestIdx2 = idx2;
        } else if (dist < bestDist2) {
          bestDist2 = dist;
        }
      }
      if (bestDist1 < TH_LOW) {                /* :609 strict */
        if (bestDist1 < nnratio * bestDist2) { /* :611 */
          match12[idx1] = bestIdx2;
          if (dist12) dist12[idx1] = bestDist1;
          vbMatched2[bestIdx2] = 1;
          if (check_ori) rotHist[rot_bin(kps1[idx1].angle, kps2[bestIdx2].angle)].push_back(idx1);
          nmatches++;
        }
      }
    }
  });

  if (check_ori) { /* :645-663 */
    int ind1 = -1, ind2 = -1, ind3 = -1;
    three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (size_t j = 0; j < rotHist[i].size(); j++) {
        match12[rotHist[i][j]] = -1;
        if (dist12) dist12[rotHist[i][j]] = -1.f;
        nmatches--;
      }
    }
  }
  return nmatches;
}

/* --------------------------------------------------------------- a11: src/ORBmatcher.cc:669-822 */
int ora_search_for_triangulation(const float* desc1, const ora_kp* kps1, int n1, const ora_featvec* fv1,
                                 const uint8_t* has_mp1, const float* desc2, const ora_kp* kps2, int n2,
                                 const ora_featvec* fv2, const uint8_t* has_mp2, const float* F12, float ex,
                                 float ey, const float* scale_factors2, const float* level_sigma2_2,
                                 int check_ori, int32_t* match12, float* dist12) {
  int nmatches = 0;
  std::vector<uint8_t> vbMatched2(n2, 0); /* :685 -- never set by the reference */
  for (int i = 0; i < n1; i++) match12[i] = -1; /* :686 */
  if (dist12) for (int i = 0; i < n1; i++) dist12[i] = -1.f;
  std::vector<int> rotHist[HISTO_LENGTH];

  for_shared_nodes(fv1, fv2, [&](int ka, int kb) {
    for (int p1 = fv1->off[ka]; p1 < fv1->off[ka + 1]; p1++) { /* :703 */
      const int idx1 = fv1->idx[p1];
      if (has_mp1[idx1]) continue; /* :708-712 */
      const ora_kp& kp1 = kps1[idx1];
      const float* d1 = desc1 + (size_t)idx1 * ORA_DESC_DIM;
      float bestDist = TH_LOW; /* :718 */
      int bestIdx2 = -1;
      for (int p2 = fv2->off[kb]; p2 < fv2->off[kb + 1]; p2++) { /* :721 */
        const int idx2 = fv2->idx[p2];
        if (vbMatched2[idx2] || has_mp2[idx2]) continue; /* :730 */
        const float dist = dd(d1, desc2 + (size_t)idx2 * ORA_DESC_DIM);
        if (dist > TH_LOW || dist > bestDist) continue; /* :737 */
        const ora_kp& kp2 = kps2[idx2];
        {
          const float distex = ex - kp2.x;
          const float distey = ey - kp2.y;
          if (distex * distex + distey * distey < 100 * scale_factors2[kp2.octave]) continue; /* :746 */
        }
        if (check_dist_epipolar_line(kp1, kp2, F12, level_sigma2_2)) { /* :750 */
          bestIdx2 = idx2;
          bestDist = dist;
        }
      }
      if (bestIdx2 >= 0) { /* :757 */
        match12[idx1] = bestIdx2;
        if (dist12) dist12[idx1] = bestDist;
        nmatches++;
        if (check_ori) rotHist[rot_bin(kp1.angle, kps2[bestIdx2].angle)].push_back(idx1);
      }
    }
  });

  if (check_ori) { /* :790-809 */
    int ind1 = -1, ind2 = -1, ind3 = -1;
    three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (size_t j = 0; j < rotHist[i].size(); j++) {
        match12[rotHist[i][j]] = -1;
        if (dist12) dist12[rotHist[i][j]] = -1.f;
        nmatches--;
      }
    }
  }
  return nmatches;
}

/* ------------------------------------- a4-a7: src/ORBmatcher.cc:44-122, 300-413, 1318-1452, 1455-1582 */
float ora_radius_by_viewing_cos(float view_cos) { /* :126-132 */
  if (view_cos > 0.998)
    return 2.5;
  else
    return 4.0;
}

float ora_local_map_window(float view_cos, float th, float scale_factor) { /* :48, :62-68 */
  const bool bFactor = th != 1.0;
  float r = ora_radius_by_viewing_cos(view_cos);
  if (bFactor) r *= th;
  return r * scale_factor;
}

int ora_predict_scale(float max_distance, float current_dist, float log_scale_factor, int n_levels) {
  /* MapPoint.cc:421-453 */
  float ratio = max_distance / current_dist;
  int nScale = (int)std::ceil(std::log(ratio) / log_scale_factor);
  if (nScale < 0)
    nScale = 0;
  else if (nScale >= n_levels)
    nScale = n_levels - 1;
  return nScale;
}

static int search_by_projection_impl(const float* desc_t, const ora_kp* kps_t, int n_t, const ora_grid* g,
                                     const ora_proj_query* q, const float* desc_q, int n_q, const uint8_t* blocked0,
                                     const ora_proj_params* p, int32_t* assign, float* dist_t, int32_t* query_best,
                                     float* query_dist);

int ora_search_by_projection(const float* desc_t, const ora_kp* kps_t, int n_t, const ora_grid* g,
                             const ora_proj_query* q, const float* desc_q, int n_q, const uint8_t* blocked0,
                             const ora_proj_params* p, int32_t* assign, float* dist_t) {
  return search_by_projection_impl(desc_t, kps_t, n_t, g, q, desc_q, n_q, blocked0, p, assign, dist_t, nullptr, nullptr);
}

/* per-query view: query_best[i] = keypoint chosen by query i or -1 (match phase of Fuse :825-1086, SearchBySim3 :1090-1314) */
int ora_search_by_projection_q(const float* desc_t, const ora_kp* kps_t, int n_t, const ora_grid* g,
                               const ora_proj_query* q, const float* desc_q, int n_q, const uint8_t* blocked0,
                               const ora_proj_params* p, int32_t* query_best, float* query_dist) {
  std::vector<int32_t> assign((size_t)(n_t > 0 ? n_t : 1));
  return search_by_projection_impl(desc_t, kps_t, n_t, g, q, desc_q, n_q, blocked0, p, assign.data(), nullptr, query_best,
                                   query_dist);
}

static int search_by_projection_impl(const float* desc_t, const ora_kp* kps_t, int n_t, const ora_grid* g,
                                     const ora_proj_query* q, const float* desc_q, int n_q, const uint8_t* blocked0,
                                     const ora_proj_params* p, int32_t* assign, float* dist_t, int32_t* query_best,
                                     float* query_dist) {
  if (query_best)
    for (int i = 0; i < n_q; i++) {
      query_best[i] = -1;
      if (query_dist) query_dist[i] = -1.f;
    }
  Grid G;
  build_grid(kps_t, n_t, *g, G);
  /* occupancy of F.mvpMapPoints[idx] as far as the descriptor loop can see it:
   * 0 free / non-blocking occupant, 1 blocking occupant */
  std::vector<uint8_t> blocking(n_t, 0);
  for (int j = 0; j < n_t; j++) {
    assign[j] = -1;
    if (dist_t) dist_t[j] = -1.f;
    blocking[j] = blocked0 ? blocked0[j] : 0;
  }
  int nmatches = 0;
  std::vector<int> rotHist[HISTO_LENGTH];
  std::vector<int32_t> vIndices;

  for (int iq = 0; iq < n_q; iq++) {
    const ora_proj_query& Q = q[iq];
    if (!Q.valid) continue;
    features_in_area(kps_t, G, Q.u, Q.v, Q.r, Q.min_level, Q.max_level, vIndices);
    if (vIndices.empty()) continue;
    const float* dq = desc_q + (size_t)iq * ORA_DESC_DIM;

    float bestDist = p->best_init;
    int bestLevel = -1;
    float bestDist2 = p->best_init;
    int bestLevel2 = -1;
    int bestIdx = -1;
    for (size_t k = 0; k < vIndices.size(); k++) {
      const int idx = vIndices[k];
      if (blocking[idx]) continue; /* :86-88, :385, :1393-1395, :1524 */
      if (p->chi2_max > 0) { /* Fuse: :919-926 */
        const float ex = Q.u - kps_t[idx].x;
        const float ey = Q.v - kps_t[idx].y;
        const float e2 = ex * ex + ey * ey;
        const int lvl = kps_t[idx].octave;
        if (lvl >= 0 && lvl < p->n_levels && e2 * p->inv_level_sigma2[lvl] > p->chi2_max) continue;
      }
      const float dist = dd(dq, desc_t + (size_t)idx * ORA_DESC_DIM);
      if (dist < bestDist) {
        bestDist2 = bestDist;
        bestDist = dist;
        bestLevel2 = bestLevel;
        bestLevel = kps_t[idx].octave;
        bestIdx = idx;
      } else if (p->mode == ORA_PROJ_RATIO_LEVEL && dist < bestDist2) { /* :102-106 */
        bestLevel2 = kps_t[idx].octave;
        bestDist2 = dist;
      }
    }
    if (bestDist <= p->th_accept) { /* :110, :404, :1408, :1538 */
      if (p->mode == ORA_PROJ_RATIO_LEVEL) {
        if (bestLevel == bestLevel2 && bestDist > p->nnratio * bestDist2) continue; /* :112 */
      }
      assign[bestIdx] = iq;
      if (query_best) {
        query_best[iq] = bestIdx;
        if (query_dist) query_dist[iq] = bestDist;
      }
      if (dist_t) dist_t[bestIdx] = bestDist;
      blocking[bestIdx] = Q.blocks;
      nmatches++;
      if (p->count_twice) nmatches++; /* :116-117 */
      if (p->check_ori) rotHist[rot_bin(Q.angle, kps_t[bestIdx].angle)].push_back(bestIdx); /* :1413-1423 */
    }
  }

  if (p->check_ori) { /* :1430-1449 */
    int ind1 = -1, ind2 = -1, ind3 = -1;
    three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++) {
      if (i != ind1 && i != ind2 && i != ind3) {
        for (size_t j = 0; j < rotHist[i].size(); j++) {
          assign[rotHist[i][j]] = -2; /* cleared by the rotation check (set to NULL at :1441), as opposed to -1 = untouched */
          if (dist_t) dist_t[rotHist[i][j]] = -1.f;
          nmatches--;
        }
      }
    }
  }
  return nmatches;
}

/* --------------------------------------------------------------- a10: src/ORBmatcher.cc:416-531 */
int ora_search_for_initialization(const float* desc1, const ora_kp* kps1, int n1, const float* desc2,
                                  const ora_kp* kps2, int n2, const ora_grid* g2, float* prev_matched,
                                  int window_size, float nnratio, int check_ori, int32_t* matches12,
                                  float* dist12) {
  Grid G;
  build_grid(kps2, n2, *g2, G);
  int nmatches = 0;
  for (int i = 0; i < n1; i++) matches12[i] = -1; /* :420 */
  if (dist12) for (int i = 0; i < n1; i++) dist12[i] = -1.f;
  std::vector<int> rotHist[HISTO_LENGTH];
  std::vector<float> vMatchedDistance(n2, 100); /* :427 */
  std::vector<int> vnMatches21(n2, -1);
  std::vector<int32_t> vIndices2;

  for (int i1 = 0; i1 < n1; i1++) {
    const ora_kp& kp1 = kps1[i1];
    const int level1 = kp1.octave;
    if (level1 > 0) continue; /* :433-435 */
    features_in_area(kps2, G, prev_matched[2 * i1], prev_matched[2 * i1 + 1], (float)window_size, level1, level1,
                     vIndices2); /* :437 */
    if (vIndices2.empty()) continue;
    const float* d1 = desc1 + (size_t)i1 * ORA_DESC_DIM;
    float bestDist = 100.0;
    float bestDist2 = 100.0;
    int bestIdx2 = -1;
    for (size_t k = 0; k < vIndices2.size(); k++) {
      const int i2 = vIndices2[k];
      const float dist = dd(d1, desc2 + (size_t)i2 * ORA_DESC_DIM);
      if (vMatchedDistance[i2] <= dist) continue; /* :456 */
      if (dist < bestDist) {
        bestDist2 = bestDist;
        bestDist = dist;
        bestIdx2 = i2;
      } else if (dist < bestDist2) {
        bestDist2 = dist;
      }
    }
    if (bestDist <= TH_LOW) {                /* :470 */
      if (bestDist < bestDist2 * nnratio) {  /* :472 */
        if (vnMatches21[bestIdx2] >= 0) {    /* :474-478 steal */
          matches12[vnMatches21[bestIdx2]] = -1;
          if (dist12) dist12[vnMatches21[bestIdx2]] = -1.f;
          nmatches--;
        }
        matches12[i1] = bestIdx2;
        if (dist12) dist12[i1] = bestDist;
        vnMatches21[bestIdx2] = i1;
        vMatchedDistance[bestIdx2] = bestDist;
        nmatches++;
        if (check_ori) rotHist[rot_bin(kps1[i1].angle, kps2[bestIdx2].angle)].push_back(i1);
      }
    }
  }

  if (check_ori) { /* :500-523 */
    int ind1 = -1, ind2 = -1, ind3 = -1;
    three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (size_t j = 0; j < rotHist[i].size(); j++) {
        const int idx1 = rotHist[i][j];
        if (matches12[idx1] >= 0) { /* :515 */
          matches12[idx1] = -1;
          if (dist12) dist12[idx1] = -1.f;
          nmatches--;
        }
      }
    }
  }

  for (int i1 = 0; i1 < n1; i1++) /* :526-528 */
    if (matches12[i1] >= 0) {
      prev_matched[2 * i1] = kps2[matches12[i1]].x;
      prev_matched[2 * i1 + 1] = kps2[matches12[i1]].y;
    }
  return nmatches;
}

/* ---------------------------------------------------------------------- Frame.cc:360-444 */
void ora_stereo_row_match(const float* desc_l, const ora_kp* kps_l, int n_l, const float* desc_r,
                          const ora_kp* kps_r, int n_r, const float* scale_factors, int n_rows, float min_d,
                          float max_d, int32_t* best_r, float* best_dist) {
  const int thOrbDist = (int)((TH_HIGH + TH_LOW) / 2); /* :366 */
  std::vector<std::vector<int32_t>> vRowIndices(n_rows);
  for (int iR = 0; iR < n_r; iR++) { /* :378-388 */
    const float kpY = kps_r[iR].y;
    const float r = 2.0f * scale_factors[kps_r[iR].octave];
    const int maxr = (int)std::ceil(kpY + r);
    const int minr = (int)std::floor(kpY - r);
    for (int yi = minr; yi <= maxr; yi++)
      if (yi >= 0 && yi < n_rows) vRowIndices[yi].push_back(iR);
  }
  for (int iL = 0; iL < n_l; iL++) { /* :399 */
    best_r[iL] = -1;
    if (best_dist) best_dist[iL] = -1.f;
    const ora_kp& kpL = kps_l[iL];
    const int levelL = kpL.octave;
    const float vL = kpL.y;
    const float uL = kpL.x;
    if (!(vL >= 0.f) || !(vL < (float)n_rows)) continue;
    const std::vector<int32_t>& vCandidates = vRowIndices[(size_t)vL]; /* :406 */
    if (vCandidates.empty()) continue;
    const float minU = uL - max_d;
    const float maxU = uL - min_d;
    if (maxU < 0) continue;
    float bestDist = TH_HIGH; /* :417 */
    int bestIdxR = 0;
    const float* dL = desc_l + (size_t)iL * ORA_DESC_DIM;
    for (size_t iC = 0; iC < vCandidates.size(); iC++) {
      const int iR = vCandidates[iC];
      const ora_kp& kpR = kps_r[iR];
      if (kpR.octave < levelL - 1 || kpR.octave > levelL + 1) continue; /* :428 */
      const float uR = kpR.x;
      if (uR >= minU && uR <= maxU) {
        const float dist = dd(dL, desc_r + (size_t)iR * ORA_DESC_DIM);
        if (dist < bestDist) {
          bestDist = dist;
          bestIdxR = iR;
        }
      }
    }
    if (bestDist < thOrbDist) { /* :447 */
      best_r[iL] = bestIdxR;
      if (best_dist) best_dist[iL] = bestDist;
    }
  }
}

/* ------------------------------------------------------------------------------- batched drivers */
void ora_batch_bow_kf_kf(const ora_keyframe* kfs, int n, const int32_t* ia, const int32_t* ib, int n_pairs,
                         float nnratio, int check_ori, int threads, int32_t* match12, float* dist12,
                         int32_t* nmatches) {
  (void)threads;
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads > 0 ? threads : 1)
#endif
  for (int p = 0; p < n_pairs; p++) {
    const ora_keyframe& A = kfs[ia[p]];
    const ora_keyframe& B = kfs[ib[p]];
    nmatches[p] = ora_search_by_bow_kf_kf(A.desc, A.kps, n, A.fv, A.mask, B.desc, B.kps, n, B.fv, B.mask, nnratio,
                                          check_ori, match12 + (size_t)p * n, dist12 ? dist12 + (size_t)p * n : 0);
  }
}

void ora_batch_triangulation(const ora_keyframe* kfs, int n, const int32_t* ia, const int32_t* ib, int n_pairs,
                             const float* F12, const float* epipoles, const float* scale_factors,
                             const float* level_sigma2, int check_ori, int threads, int32_t* match12,
                             float* dist12, int32_t* nmatches) {
  (void)threads;
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads > 0 ? threads : 1)
#endif
  for (int p = 0; p < n_pairs; p++) {
    const ora_keyframe& A = kfs[ia[p]];
    const ora_keyframe& B = kfs[ib[p]];
    nmatches[p] = ora_search_for_triangulation(A.desc, A.kps, n, A.fv, A.mask, B.desc, B.kps, n, B.fv, B.mask,
                                               F12 + 9 * (size_t)p, epipoles[2 * p], epipoles[2 * p + 1],
                                               scale_factors, level_sigma2, check_ori, match12 + (size_t)p * n,
                                               dist12 ? dist12 + (size_t)p * n : 0);
  }
}

// MapPoint.cc:271-338
void ora_distinctive(const float* const* kf_desc, int n_points, const int32_t* obs_off, const int32_t* obs_kf,
                     const int32_t* obs_row, int32_t* best, float* best_median) {
  for (int p = 0; p < n_points; p++) {
    const int N = obs_off[p + 1] - obs_off[p];
    best[p] = -1;
    if (best_median) best_median[p] = 100.f;
    if (N <= 0) continue;  // :285-286 / :300-301: mDescriptor is left alone
    std::vector<const float*> row((size_t)N);
    for (int i = 0; i < N; i++) {
      const int e = obs_off[p] + i;
      row[i] = kf_desc[obs_kf[e]] + (size_t)obs_row[e] * ORA_DESC_DIM;
    }
    std::vector<float> D((size_t)N * N);
    for (int i = 0; i < N; i++) {  // :306-318
      D[(size_t)i * N + i] = 0;
      for (int j = i + 1; j < N; j++) {
        const float d = ora_descriptor_distance(row[i], row[j]);
        D[(size_t)i * N + j] = d;
        D[(size_t)j * N + i] = d;
      }
    }
    float best_med = 100;  // :321
    int best_idx = 0;
    for (int i = 0; i < N; i++) {  // :323-334
      std::vector<float> v(D.begin() + (size_t)i * N, D.begin() + (size_t)(i + 1) * N);
      std::sort(v.begin(), v.end());
      const float median = v[(size_t)(0.5 * (N - 1))];
      if (median < best_med) {
        best_med = median;
        best_idx = i;
      }
    }
    best[p] = best_idx;
    if (best_median) best_median[p] = best_med;
  }
}

// FSift::distance (src/dbow2/DBoW2/FSift.cpp:86-101): float difference, float product, double running sum
static double fsift_distance(const float* a, const float* b) {
  double sqd = 0.;
  for (int i = 0; i < ORA_DESC_DIM; i++) {
    const float d = a[i] - b[i];
    const float p = d * d;
    sqd += p;
  }
  return sqd;
}

int ora_voc_transform(int L, int weighting, int scoring, int n_nodes, const int32_t* child_off, const int32_t* child_idx,
                      const float* node_desc, const double* weight, const int32_t* word_of_node, const float* desc, int n,
                      int levelsup, int32_t* word_id, int32_t* node_id, uint8_t* in_fv, uint32_t* bow_id, double* bow_val) {
  (void)n_nodes;
  std::map<uint32_t, double> v;  // DBoW2::BowVector
  const int nid_level = L - levelsup;  // TemplatedVocabulary.h:1224
  for (int f = 0; f < n; f++) {
    const float* feat = desc + (size_t)f * ORA_DESC_DIM;
    int nid = 0;
    bool nid_set = nid_level <= 0;  // :1225
    int final_id = 0, level = 0;
    do {  // :1230-1251
      ++level;
      const int off = child_off[final_id], kp = child_off[final_id + 1] - off;
      final_id = child_idx[off];
      double best_d = fsift_distance(feat, node_desc + (size_t)final_id * ORA_DESC_DIM);
      for (int c = 1; c < kp; c++) {
        const int id = child_idx[off + c];
        const double d = fsift_distance(feat, node_desc + (size_t)id * ORA_DESC_DIM);
        if (d < best_d) {
          best_d = d;
          final_id = id;
        }
      }
      if (level == nid_level) {
        nid = final_id;
        nid_set = true;
      }
    } while (child_off[final_id + 1] != child_off[final_id]);
    if (!nid_set) nid = final_id;
    const double w = weight[final_id];
    const uint32_t id = (uint32_t)word_of_node[final_id];
    if (word_id) word_id[f] = (int32_t)id;
    if (node_id) node_id[f] = nid;
    if (in_fv) in_fv[f] = w > 0;
    if (w > 0) {  // :1155 / :1183
      auto it = v.lower_bound(id);
      const bool exists = it != v.end() && !(id < it->first);
      if (weighting == 0 || weighting == 1) {  // TF_IDF, TF: addWeight
        if (exists) it->second += w;
        else v.insert(it, std::make_pair(id, w));
      } else if (!exists) {  // IDF, BINARY: addIfNotExist
        v.insert(it, std::make_pair(id, w));
      }
    }
  }
  const bool must = scoring != 5;  // every scoring object but DotProductScoring normalises (ScoringObject.h:72-89)
  if ((weighting == 0 || weighting == 1) && !v.empty() && !must) {  // :1163-1169
    const double nd = v.size();
    for (auto& e : v) e.second /= nd;
  }
  if (must) {  // BowVector::normalize (BowVector.cpp:61-84); L2Scoring alone asks for L2
    double norm = 0.0;
    if (scoring != 1) {
      for (auto& e : v) norm += fabs(e.second);
    } else {
      for (auto& e : v) norm += e.second * e.second;
      norm = sqrt(norm);
    }
    if (norm > 0.0)
      for (auto& e : v) e.second /= norm;
  }
  int k = 0;
  for (auto& e : v) {
    if (bow_id) bow_id[k] = e.first;
    if (bow_val) bow_val[k] = e.second;
    k++;
  }
  return k;
}

long ora_map_write_block(const float* desc, int n, int d, uint8_t* out) {
  // System.cc:337-343: descriptors(j, k) = mDescriptors.at<float>(k, j) into a column-major D x N matrix
  std::vector<float> m((size_t)d * n);
  for (int j = 0; j < d; j++)
    for (int k = 0; k < n; k++) m[(size_t)j + (size_t)k * d] = desc[(size_t)k * d + j];
  // visual_map_seri.cc:134-146
  long pos = 0;
  const int32_t width = d, count = n;
  memcpy(out + pos, &width, 4); pos += 4;
  memcpy(out + pos, &count, 4); pos += 4;
  for (int j = 0; j < count; j++)
    for (int k = 0; k < width; k++) {
      memcpy(out + pos, &m[(size_t)k + (size_t)j * d], 4);
      pos += 4;
    }
  return pos;
}

int ora_map_read_block(const uint8_t* block, long bytes, float* rows, int cap_rows, int* width) {
  if (bytes < 8) return -1;
  int32_t w, c;  // visual_map_seri.cc:272-273
  memcpy(&w, block, 4);
  memcpy(&c, block + 4, 4);
  if (w < 0 || c < 0 || 8 + (long)w * c * 4 > bytes) return -1;
  *width = w;
  long pos = 8;
  for (int j = 0; j < c; j++)  // :275-285
    for (int k = 0; k < w; k++) {
      if (j < cap_rows) memcpy(&rows[(size_t)j * w + k], block + pos, 4);
      pos += 4;
    }
  return c;
}

// ============================================================================ KeyFrameDatabase (BoW candidates)
// DBoW2::L1Scoring::score, ScoringObject.cpp:24-61. lower_bound on a sorted array == std::map::lower_bound.
double ora_bow_score_l1(const uint32_t* w1, const double* v1, int n1, const uint32_t* w2, const double* v2, int n2) {
  int i = 0, j = 0;
  double score = 0;
  while (i < n1 && j < n2) {
    const double vi = v1[i], wi = v2[j];
    if (w1[i] == w2[j]) {
      score += fabs(vi - wi) - fabs(vi) - fabs(wi);  // :41
      ++i;
      ++j;
    } else if (w1[i] < w2[j]) {
      i = (int)(std::lower_bound(w1 + i, w1 + n1, w2[j]) - w1);  // :48
    } else {
      j = (int)(std::lower_bound(w2 + j, w2 + n2, w1[i]) - w2);  // :52
    }
  }
  score = -score / 2.0;  // :56
  return score;
}

namespace {
// the inverted file as KeyFrameDatabase::add builds it (:41-47): word -> keyframes in add order, erased ones absent
std::map<uint32_t, std::vector<int> > inverted_file(int n_kf, const int32_t* bow_off, const uint32_t* bow_word,
                                                    const uint8_t* alive) {
  std::map<uint32_t, std::vector<int> > inv;
  for (int s = 0; s < n_kf; s++) {
    if (alive && !alive[s]) continue;
    for (int e = bow_off[s]; e < bow_off[s + 1]; e++) inv[bow_word[e]].push_back(s);
  }
  return inv;
}
}  // namespace

int ora_detect_loop_candidates(int n_kf, const int32_t* bow_off, const uint32_t* bow_word, const double* bow_val,
                               const uint8_t* alive, const uint32_t* q_word, const double* q_val, int nq,
                               const uint8_t* connected, const uint8_t* global_flag, const int32_t* neigh_off,
                               const int32_t* neigh_slot, float min_score, int only_global_map, int32_t* out,
                               int32_t* common, float* score) {
  const std::map<uint32_t, std::vector<int> > inv = inverted_file(n_kf, bow_off, bow_word, alive);
  std::vector<uint8_t> queried(n_kf, 0);  // mnLoopQuery == pKF->mnId
  std::vector<int> words(n_kf, 0);        // mnLoopWords
  std::vector<float> loop_score(n_kf, 0.f);
  std::vector<int> sharing;               // lKFsSharingWords
  for (int q = 0; q < nq; q++) {          // :93-113
    std::map<uint32_t, std::vector<int> >::const_iterator it = inv.find(q_word[q]);
    if (it == inv.end()) continue;
    for (size_t k = 0; k < it->second.size(); k++) {
      const int s = it->second[k];
      if (!queried[s]) {
        words[s] = 0;
        if (!(connected && connected[s])) {
          queried[s] = 1;
          sharing.push_back(s);
        }
      }
      words[s]++;
    }
  }
  if (common)
    for (int s = 0; s < n_kf; s++) common[s] = queried[s] ? words[s] : 0;
  if (score)
    for (int s = 0; s < n_kf; s++) score[s] = -1.f;
  if (sharing.empty()) return 0;  // :114-115
  int maxCommonWords = 0;         // :121-126
  for (size_t k = 0; k < sharing.size(); k++)
    if (words[sharing[k]] > maxCommonWords) maxCommonWords = words[sharing[k]];
  const int minCommonWords = maxCommonWords * 0.6f;  // :128
  std::vector<std::pair<float, int> > scoreAndMatch;
  for (size_t k = 0; k < sharing.size(); k++) {  // :133-150
    const int s = sharing[k];
    if (words[s] > minCommonWords) {
      const float si = (float)ora_bow_score_l1(q_word, q_val, nq, bow_word + bow_off[s], bow_val + bow_off[s],
                                               bow_off[s + 1] - bow_off[s]);
      loop_score[s] = si;
      if (score) score[s] = si;
      if (si >= min_score)
        if ((only_global_map && global_flag && global_flag[s]) || !only_global_map) scoreAndMatch.push_back(std::make_pair(si, s));
    }
  }
  if (scoreAndMatch.empty()) return 0;  // :151-152
  std::vector<std::pair<float, int> > accScoreAndMatch;
  float bestAccScore = min_score;  // :155
  for (size_t k = 0; k < scoreAndMatch.size(); k++) {  // :156-181
    const int s = scoreAndMatch[k].second;
    float bestScore = scoreAndMatch[k].first;
    float accScore = scoreAndMatch[k].first;
    int best = s;
    for (int e = neigh_off ? neigh_off[s] : 0; e < (neigh_off ? neigh_off[s + 1] : 0); e++) {
      const int s2 = neigh_slot[e];
      if (queried[s2] && words[s2] > minCommonWords) {  // :168
        accScore += loop_score[s2];
        if (loop_score[s2] > bestScore) {
          best = s2;
          bestScore = loop_score[s2];
        }
      }
    }
    accScoreAndMatch.push_back(std::make_pair(accScore, best));
    if (accScore > bestAccScore) bestAccScore = accScore;
  }
  const float minScoreToRetain = 0.55f * bestAccScore;  // :184
  std::vector<uint8_t> added(n_kf, 0);
  int n_out = 0;
  for (size_t k = 0; k < accScoreAndMatch.size(); k++)  // :189-200
    if (accScoreAndMatch[k].first > minScoreToRetain) {
      const int s = accScoreAndMatch[k].second;
      if (!added[s]) {
        out[n_out++] = s;
        added[s] = 1;
      }
    }
  return n_out;
}

int ora_detect_reloc_candidates(int n_kf, const int32_t* bow_off, const uint32_t* bow_word, const double* bow_val,
                                const uint8_t* alive, const uint32_t* q_word, const double* q_val, int nq,
                                const uint8_t* global_flag, const int32_t* neigh_off, const int32_t* neigh_slot,
                                int only_global_map, float* reloc_score, int32_t* out, int32_t* common) {
  const std::map<uint32_t, std::vector<int> > inv = inverted_file(n_kf, bow_off, bow_word, alive);
  std::vector<uint8_t> queried(n_kf, 0);  // mnRelocQuery == F->mnId
  std::vector<int> words(n_kf, 0);        // mnRelocWords
  std::vector<int> sharing;
  for (int q = 0; q < nq; q++) {  // :226-243
    std::map<uint32_t, std::vector<int> >::const_iterator it = inv.find(q_word[q]);
    if (it == inv.end()) continue;
    for (size_t k = 0; k < it->second.size(); k++) {
      const int s = it->second[k];
      if (!queried[s]) {
        words[s] = 0;
        queried[s] = 1;
        if ((only_global_map && global_flag && global_flag[s]) || !only_global_map) sharing.push_back(s);
      }
      words[s]++;
    }
  }
  if (common)
    for (int s = 0; s < n_kf; s++) common[s] = words[s];
  if (sharing.empty()) return 0;
  int maxCommonWords = 0;  // :249-254
  for (size_t k = 0; k < sharing.size(); k++)
    if (words[sharing[k]] > maxCommonWords) maxCommonWords = words[sharing[k]];
  const int minCommonWords = maxCommonWords * 0.8f;  // :256
  std::vector<std::pair<float, int> > scoreAndMatch;
  for (size_t k = 0; k < sharing.size(); k++) {  // :262-274
    const int s = sharing[k];
    if (words[s] > minCommonWords) {
      const float si = (float)ora_bow_score_l1(q_word, q_val, nq, bow_word + bow_off[s], bow_val + bow_off[s],
                                               bow_off[s + 1] - bow_off[s]);
      reloc_score[s] = si;
      scoreAndMatch.push_back(std::make_pair(si, s));
    }
  }
  if (scoreAndMatch.empty()) return 0;
  std::vector<std::pair<float, int> > accScoreAndMatch;
  float bestAccScore = 0;  // :280
  for (size_t k = 0; k < scoreAndMatch.size(); k++) {  // :282-303
    const int s = scoreAndMatch[k].second;
    float bestScore = scoreAndMatch[k].first;
    float accScore = bestScore;
    int best = s;
    for (int e = neigh_off ? neigh_off[s] : 0; e < (neigh_off ? neigh_off[s + 1] : 0); e++) {
      const int s2 = neigh_slot[e];
      if (!queried[s2]) continue;    // :291-292
      accScore += reloc_score[s2];   // :294 -- whatever score the neighbour last received
      if (reloc_score[s2] > bestScore) {
        best = s2;
        bestScore = reloc_score[s2];
      }
    }
    accScoreAndMatch.push_back(std::make_pair(accScore, best));
    if (accScore > bestAccScore) bestAccScore = accScore;
  }
  const float minScoreToRetain = 0.75f * bestAccScore;  // :305
  std::vector<uint8_t> added(n_kf, 0);
  int n_out = 0;
  for (size_t k = 0; k < accScoreAndMatch.size(); k++)
    if (accScoreAndMatch[k].first > minScoreToRetain) {
      const int s = accScoreAndMatch[k].second;
      if (!added[s]) {
        out[n_out++] = s;
        added[s] = 1;
      }
    }
  return n_out;
}

int ora_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

}  // extern "C"
