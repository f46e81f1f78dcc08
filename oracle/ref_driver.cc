// ref_driver.cc -- flat C entry points over the reference's OWN matcher, compiled from where it lies:
//   /root/reference/src/vslam/src/ORBmatcher.cc                  (unmodified, whole file)
//   Frame.cc / KeyFrame.cc / MapPoint.cc function bodies          (line ranges extracted by oracle/Makefile into
//                                                                  oracle/_ref/gen/*.inc at build time; not committed)
// against the from-scratch header shims in oracle/ref_shim (OpenCV 3.2, libtorch, ROS and g2o are not installed).
// The result, oracle/_ref/libref_orbmatcher.so, exists to PIN the CPU oracle (tests/test_oracle_vs_ref.py) and to
// generate the golden fixtures under tests/golden/.  TEST INFRASTRUCTURE ONLY: nothing in the product loads it, and
// it cannot be built on the GPU box (/root/reference does not exist there) -- the prebuilt .so travels instead.
//
// Each ref_* function takes the same flat arrays as its ora_* twin in asd_oracle.h, builds shim Frame / KeyFrame /
// MapPoint objects from them and calls the reference method.  Where an overload derives its search window from
// poses and 3-D points, the harness picks identity poses, fx=fy=1, cx=cy=0 and points (u, v, 1), so the reference's
// own projection code yields exactly the (u, v) passed in.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>

#include "ORBmatcher.h"  // ref_shim/ORBmatcher.h -> the reference's include/ORBmatcher.h
#include "asd_oracle.h"

namespace ORB_SLAM2 {
float Frame::fx = 1, Frame::fy = 1, Frame::cx = 0, Frame::cy = 0, Frame::invfx = 1, Frame::invfy = 1;
float Frame::mfGridElementWidthInv = 0, Frame::mfGridElementHeightInv = 0;
float Frame::mnMinX = 0, Frame::mnMaxX = 0, Frame::mnMinY = 0, Frame::mnMaxY = 0;
// the reference's own function bodies (see oracle/Makefile for the line ranges); MapPoint.cc relies on `using namespace std`
using std::sort;
#include "_ref/gen/frame_grid.inc"
#include "_ref/gen/keyframe_grid.inc"
#include "_ref/gen/mappoint_scale.inc"
#include "_ref/gen/mappoint_distinct.inc"

// protected members of the reference class, opened for direct checks
struct OpenMatcher : public ORBmatcher {
  OpenMatcher(float r, bool o) : ORBmatcher(r, o) {}
  using ORBmatcher::CheckDistEpipolarLine;
  using ORBmatcher::ComputeThreeMaxima;
  using ORBmatcher::RadiusByViewingCos;
};
}  // namespace ORB_SLAM2

using namespace ORB_SLAM2;

namespace {

cv::Mat desc_mat(const float* d, int n) {
  cv::Mat m(n, ORA_DESC_DIM, CV_32F);
  if (n > 0) memcpy(m.ptr<float>(0), d, (size_t)n * ORA_DESC_DIM * sizeof(float));
  return m;
}
cv::Mat desc_row(const float* d) { return desc_mat(d, 1); }

std::vector<cv::KeyPoint> keypoints(const ora_kp* k, int n) {
  std::vector<cv::KeyPoint> v((size_t)n);
  for (int i = 0; i < n; i++) {
    v[i].pt.x = k[i].x;
    v[i].pt.y = k[i].y;
    v[i].angle = k[i].angle;
    v[i].octave = k[i].octave;
  }
  return v;
}

void featvec(const ora_featvec* fv, DBoW2::FeatureVector& out) {
  out.clear();
  if (!fv) return;
  for (int k = 0; k < fv->n_nodes; k++)
    for (int e = fv->off[k]; e < fv->off[k + 1]; e++) out.addFeature(fv->node_ids[k], (unsigned)fv->idx[e]);
}

cv::Mat eye4() {
  cv::Mat T(4, 4, CV_32F);
  for (int i = 0; i < 4; i++) T.at<float>(i, i) = 1.f;
  return T;
}
cv::Mat vec3(float x, float y, float z) {
  cv::Mat v(3, 1, CV_32F);
  v.at<float>(0) = x; v.at<float>(1) = y; v.at<float>(2) = z;
  return v;
}

// ---- scene: where the harness puts cameras and map points -------------------------------------------------------
// mode 0 (default): identity poses, fx = fy = 1, cx = cy = 0, points (u, v, 1): every projection is exact in any
// arithmetic, so the flat oracle (which takes windows, not poses) can be compared with the reference bit for bit.
// mode 1 (ref_set_scene): EuRoC intrinsics, a random SE3 pose (a second one for the other frame / keyframe), a Sim3
// scale != 1, per-point depths 1.5 .. 11.5 m with ~6 % of the points BEHIND the camera, ~15 % outside the distance-
// invariance band (above and below), ~15 % of the normals beyond the 60 degree viewing gate. A point meant to project
// at (u, v) is placed at T^-1 * backproject(u, v, z); the reference's own arithmetic (shim cv::Mat products) then
// lands within rounding of (u, v). Used for the adaptor-level goldens: reference and adaptor see identical objects.
struct Scene {
  int mode = 0;
  uint32_t seed = 0;
  float fx = 1, fy = 1, cx = 0, cy = 0;
  double R[2][9] = {{1, 0, 0, 0, 1, 0, 0, 0, 1}, {1, 0, 0, 0, 1, 0, 0, 0, 1}}, t[2][3] = {{0, 0, 0}, {0, 0, 0}};
  double s = 1.0;  // scale of the Sim3 overloads (Scw = [sR | st])
} g_scene;

double hash01(uint32_t i, uint32_t salt) {
  uint32_t x = i * 2654435761u ^ (salt * 40503u + g_scene.seed * 69069u + 12345u);
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return (x >> 8) * (1.0 / 16777216.0);
}

void rodrigues(double ax, double ay, double az, double ang, double* R) {
  const double n = std::sqrt(ax * ax + ay * ay + az * az);
  ax /= n; ay /= n; az /= n;
  const double c = std::cos(ang), s = std::sin(ang), C = 1 - c;
  const double M[9] = {c + ax * ax * C, ax * ay * C - az * s, ax * az * C + ay * s, ay * ax * C + az * s, c + ay * ay * C,
                       ay * az * C - ax * s, az * ax * C - ay * s, az * ay * C + ax * s, c + az * az * C};
  for (int k = 0; k < 9; k++) R[k] = M[k];
}

cv::Mat scene_T(int cam, double scale = 1.0) {  // 4x4 [scale*R | scale*t]
  cv::Mat T(4, 4, CV_32F);
  for (int r = 0; r < 3; r++) {
    for (int c = 0; c < 3; c++) T.at<float>(r, c) = (float)(scale * g_scene.R[cam][3 * r + c]);
    T.at<float>(r, 3) = (float)(scale * g_scene.t[cam][r]);
  }
  T.at<float>(3, 3) = 1.f;
  return T;
}

void scene_center(int cam, double* O) {  // -R^T t
  for (int c = 0; c < 3; c++) {
    O[c] = 0;
    for (int r = 0; r < 3; r++) O[c] -= g_scene.R[cam][3 * r + c] * g_scene.t[cam][r];
  }
}

double scene_depth(int i) {
  if (g_scene.mode == 0) return 1.0;
  if (hash01((uint32_t)i, 1) < 0.06) return -(1.0 + 5.0 * hash01((uint32_t)i, 2));  // behind the camera
  return 1.5 + 10.0 * hash01((uint32_t)i, 3);
}

// camera-frame point that projects to (u, v) at the depth of point i
void scene_backproject(float u, float v, int i, double* Xc) {
  const double z = scene_depth(i);
  Xc[0] = ((double)u - g_scene.cx) * z / g_scene.fx;
  Xc[1] = ((double)v - g_scene.cy) * z / g_scene.fy;
  Xc[2] = z;
}

void scene_to_world(int cam, const double* Xc, double* Xw) {  // R^T (Xc - t)
  for (int c = 0; c < 3; c++) {
    Xw[c] = 0;
    for (int r = 0; r < 3; r++) Xw[c] += g_scene.R[cam][3 * r + c] * (Xc[r] - g_scene.t[cam][r]);
  }
}

cv::Mat scene_point(int cam, float u, float v, int i) {
  if (g_scene.mode == 0) return vec3(u, v, 1.f);
  double Xc[3], Xw[3];
  scene_backproject(u, v, i, Xc);
  scene_to_world(cam, Xc, Xw);
  return vec3((float)Xw[0], (float)Xw[1], (float)Xw[2]);
}

// distance-invariance band of point p seen from camera `cam` such that PredictScale returns `lvl`
// (MapPoint.cc:421-453: ceil(log(max/dist)/log(1.2)); the middle of the level's bin), or a band the point is outside of
void scene_band(MapPoint* p, int cam, int lvl, int i) {
  double O[3];
  scene_center(cam, O);
  const double dx = p->mWorldPos.at<float>(0) - O[0], dy = p->mWorldPos.at<float>(1) - O[1], dz = p->mWorldPos.at<float>(2) - O[2];
  const float dist = (float)std::sqrt(dx * dx + dy * dy + dz * dz);
  p->mfMaxDistance = dist * std::pow(1.2f, (float)lvl - 0.5f);
  p->mfMinDistance = 0.f;
  if (g_scene.mode == 0) return;
  const double h = hash01((uint32_t)i, 4);
  if (h < 0.08) p->mfMaxDistance = dist / 1.5f;       // farther than 1.2 * max
  else if (h < 0.15) p->mfMinDistance = dist * 1.6f;  // nearer than 0.8 * min
}

// viewing direction of point p from camera `cam`; in scene mode ~15 % are tilted 75 degrees away (cos < 0.5)
cv::Mat scene_normal(const MapPoint* p, int cam, int i) {
  double O[3];
  scene_center(cam, O);
  double n[3] = {p->mWorldPos.at<float>(0) - O[0], p->mWorldPos.at<float>(1) - O[1], p->mWorldPos.at<float>(2) - O[2]};
  const double d = std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
  for (double& x : n) x /= d;
  if (g_scene.mode && hash01((uint32_t)i, 5) < 0.15) {
    double q[3] = {n[1], -n[0], 0.0};  // perpendicular (n is never along z alone here)
    double qn = std::sqrt(q[0] * q[0] + q[1] * q[1]);
    if (qn < 1e-9) { q[0] = 1; q[1] = 0; qn = 1; }
    const double c = std::cos(75.0 * M_PI / 180.0), s = std::sin(75.0 * M_PI / 180.0);
    for (int k = 0; k < 3; k++) n[k] = c * n[k] + s * q[k] / qn;
  }
  return vec3((float)n[0], (float)n[1], (float)n[2]);
}

void levels(const float* sf, const float* sigma2, int n_levels, float log_scale, std::vector<float>& vsf,
            std::vector<float>& vs2, std::vector<float>& vis2, int& nl, float& lsf) {
  vsf.assign(sf, sf + n_levels);
  vs2.clear(); vis2.clear();
  for (int l = 0; l < n_levels; l++) {
    const float s2 = sigma2 ? sigma2[l] : sf[l] * sf[l];
    vs2.push_back(s2);
    vis2.push_back(1.0f / s2);
  }
  nl = n_levels;
  lsf = log_scale;
}

void fill_frame(Frame& F, const float* desc, const ora_kp* kps, int n, const ora_grid* g) {
  F.N = n;
  F.mvKeys = F.mvKeysUn = keypoints(kps, n);
  F.mDescriptors = desc_mat(desc, n);
  F.mvpMapPoints.assign((size_t)n, nullptr);
  F.mvbOutlier.assign((size_t)n, false);
  F.mTcw = scene_T(0);
  Frame::fx = g_scene.fx; Frame::fy = g_scene.fy; Frame::cx = g_scene.cx; Frame::cy = g_scene.cy;
  Frame::invfx = 1.f / Frame::fx; Frame::invfy = 1.f / Frame::fy;
  if (g) {
    Frame::mnMinX = g->min_x; Frame::mnMaxX = g->max_x; Frame::mnMinY = g->min_y; Frame::mnMaxY = g->max_y;
    Frame::mfGridElementWidthInv = g->inv_w; Frame::mfGridElementHeightInv = g->inv_h;
    F.AssignFeaturesToGrid();
  }
}

void set_keyframe_camera(KeyFrame& K, int cam) {
  K.Tcw = scene_T(cam);
  double O[3];
  scene_center(cam, O);
  K.Ow = vec3((float)O[0], (float)O[1], (float)O[2]);
  K.fx = g_scene.fx; K.fy = g_scene.fy; K.cx = g_scene.cx; K.cy = g_scene.cy;
  K.invfx = 1.f / K.fx; K.invfy = 1.f / K.fy;
}

void fill_keyframe(KeyFrame& K, const float* desc, const ora_kp* kps, int n, const ora_grid* g) {
  K.N = n;
  K.mvKeys = K.mvKeysUn = keypoints(kps, n);
  K.mDescriptors = desc_mat(desc, n);
  K.mvpMapPoints.assign((size_t)n, nullptr);
  set_keyframe_camera(K, 0);
  if (g) {
    K.mnMinX = (int)g->min_x; K.mnMaxX = (int)g->max_x; K.mnMinY = (int)g->min_y; K.mnMaxY = (int)g->max_y;
    K.mfGridElementWidthInv = g->inv_w; K.mfGridElementHeightInv = g->inv_h;
    K.mnGridCols = g->cols; K.mnGridRows = g->rows;
    K.mGrid.assign((size_t)g->cols, std::vector<std::vector<size_t> >((size_t)g->rows));
    K.AssignFeaturesToGrid();
  }
}

// one MapPoint per usable feature; unusable features alternate between "no map point" and "bad map point"
// (the reference treats both alike: `if(!pMP) continue; if(pMP->isBad()) continue;`)
struct Pool {
  std::vector<std::unique_ptr<MapPoint> > own;
  MapPoint* make(int id) {
    own.emplace_back(new MapPoint());
    own.back()->id = id;
    return own.back().get();
  }
};

void attach_valid(Pool& pool, KeyFrame& K, const uint8_t* valid, int n) {
  for (int i = 0; i < n; i++) {
    if (valid[i]) {
      K.mvpMapPoints[i] = pool.make(i);
      K.mvpMapPoints[i]->nObs = 1;
    } else if (i & 1) {
      MapPoint* p = pool.make(i);
      p->mbBad = true;
      K.mvpMapPoints[i] = p;
    }
  }
}

}  // namespace

extern "C" {

// ------------------------------------------------------------------------------------------- scene
void ref_set_scene(int mode, unsigned seed) {
  g_scene = Scene();
  if (!mode) return;
  g_scene.mode = 1;
  g_scene.seed = seed;
  g_scene.fx = 458.654f; g_scene.fy = 457.296f; g_scene.cx = 367.215f; g_scene.cy = 248.375f;  // EuRoC cam0
  for (int cam = 0; cam < 2; cam++) {
    rodrigues(hash01(cam, 10) - 0.5, hash01(cam, 11) - 0.5, hash01(cam, 12) + 0.2, 0.1 + 0.4 * hash01(cam, 13), g_scene.R[cam]);
    for (int k = 0; k < 3; k++) g_scene.t[cam][k] = 2.0 * hash01(cam, 14 + k) - 1.0;
  }
  g_scene.s = 0.9 + 0.4 * hash01(7, 17);
}

// ------------------------------------------------------------------------------------------- scalar pieces
float ref_descriptor_distance(const float* a, const float* b) {
  return ORBmatcher::DescriptorDistance(desc_row(a), desc_row(b));  // ORBmatcher.cc:1629-1650
}

void ref_compute_three_maxima(const int32_t* sizes, int L, int32_t* i1, int32_t* i2, int32_t* i3) {
  std::vector<std::vector<int> > h((size_t)L);
  for (int b = 0; b < L; b++) h[b].assign((size_t)sizes[b], 0);
  OpenMatcher m(0.6f, true);
  int a = -1, b = -1, c = -1;  // every call site initialises them to -1 (e.g. :276-278)
  m.ComputeThreeMaxima(h.data(), L, a, b, c);  // :1584-1625
  *i1 = a; *i2 = b; *i3 = c;
}

float ref_radius_by_viewing_cos(float view_cos) { return OpenMatcher(0.6f, true).RadiusByViewingCos(view_cos); }

int ref_check_dist_epipolar_line(float x1, float y1, float x2, float y2, int octave2, const float* F12,
                                 const float* level_sigma2, int n_levels) {
  KeyFrame K2;
  K2.mvLevelSigma2.assign(level_sigma2, level_sigma2 + n_levels);
  cv::Mat F(3, 3, CV_32F);
  memcpy(F.ptr<float>(0), F12, 9 * sizeof(float));
  cv::KeyPoint k1, k2;
  k1.pt.x = x1; k1.pt.y = y1; k2.pt.x = x2; k2.pt.y = y2; k2.octave = octave2;
  return OpenMatcher(0.6f, true).CheckDistEpipolarLine(k1, k2, F, &K2) ? 1 : 0;  // :136-153
}

int ref_predict_scale(float max_distance, float current_dist, float log_scale_factor, int n_levels) {
  MapPoint mp;
  mp.mfMaxDistance = max_distance;
  Frame F;
  F.mfLogScaleFactor = log_scale_factor;
  F.mnScaleLevels = n_levels;
  return mp.PredictScale(current_dist, &F);  // MapPoint.cc:438-453
}

// ------------------------------------------------------------------------------------------- grid
// Frame::AssignFeaturesToGrid + GetFeaturesInArea (Frame.cc:123-138, 219-286): returns the count written to out
int ref_frame_features_in_area(const ora_kp* kps, int n, const ora_grid* g, float x, float y, float r, int min_level,
                               int max_level, int32_t* out) {
  Frame F;
  fill_frame(F, nullptr, kps, 0, nullptr);
  F.N = n;
  F.mvKeys = F.mvKeysUn = keypoints(kps, n);
  Frame::mnMinX = g->min_x; Frame::mnMaxX = g->max_x; Frame::mnMinY = g->min_y; Frame::mnMaxY = g->max_y;
  Frame::mfGridElementWidthInv = g->inv_w; Frame::mfGridElementHeightInv = g->inv_h;
  F.AssignFeaturesToGrid();
  const std::vector<size_t> v = F.GetFeaturesInArea(x, y, r, min_level, max_level);
  for (size_t k = 0; k < v.size(); k++) out[k] = (int32_t)v[k];
  return (int)v.size();
}

// KeyFrame::AssignFeaturesToGrid + GetFeaturesInArea (KeyFrame.cc:116-143, 839-878)
int ref_keyframe_features_in_area(const ora_kp* kps, int n, const ora_grid* g, float x, float y, float r,
                                  int32_t* out) {
  KeyFrame K;
  fill_keyframe(K, nullptr, kps, 0, nullptr);
  K.N = n;
  K.mvKeys = K.mvKeysUn = keypoints(kps, n);
  K.mnMinX = (int)g->min_x; K.mnMaxX = (int)g->max_x; K.mnMinY = (int)g->min_y; K.mnMaxY = (int)g->max_y;
  K.mfGridElementWidthInv = g->inv_w; K.mfGridElementHeightInv = g->inv_h;
  K.mnGridCols = g->cols; K.mnGridRows = g->rows;
  K.mGrid.assign((size_t)g->cols, std::vector<std::vector<size_t> >((size_t)g->rows));
  K.AssignFeaturesToGrid();
  const std::vector<size_t> v = K.GetFeaturesInArea(x, y, r);
  for (size_t k = 0; k < v.size(); k++) out[k] = (int32_t)v[k];
  return (int)v.size();
}

// ------------------------------------------------------------------------------------------- BoW-gated searches
// ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches)  (ORBmatcher.cc:156-297)
int ref_search_by_bow_kf_f(const float* desc_kf, const ora_kp* kps_kf, int n_kf, const ora_featvec* fv_kf,
                           const uint8_t* valid_kf, const float* desc_f, const ora_kp* kps_f, int n_f,
                           const ora_featvec* fv_f, float nnratio, int check_ori, int32_t* match_f) {
  Pool pool;
  KeyFrame K;
  fill_keyframe(K, desc_kf, kps_kf, n_kf, nullptr);
  featvec(fv_kf, K.mFeatVec);
  attach_valid(pool, K, valid_kf, n_kf);
  Frame F;
  fill_frame(F, desc_f, kps_f, n_f, nullptr);
  featvec(fv_f, F.mFeatVec);
  std::vector<MapPoint*> out;
  ORBmatcher m(nnratio, check_ori != 0);
  const int nm = m.SearchByBoW(&K, F, out);
  for (int j = 0; j < n_f; j++) match_f[j] = out[j] ? out[j]->id : -1;
  return nm;
}

// ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12)  (ORBmatcher.cc:533-666)
int ref_search_by_bow_kf_kf(const float* desc1, const ora_kp* kps1, int n1, const ora_featvec* fv1,
                            const uint8_t* valid1, const float* desc2, const ora_kp* kps2, int n2,
                            const ora_featvec* fv2, const uint8_t* valid2, float nnratio, int check_ori,
                            int32_t* match12) {
  Pool pool;
  KeyFrame K1, K2;
  fill_keyframe(K1, desc1, kps1, n1, nullptr);
  fill_keyframe(K2, desc2, kps2, n2, nullptr);
  featvec(fv1, K1.mFeatVec);
  featvec(fv2, K2.mFeatVec);
  attach_valid(pool, K1, valid1, n1);
  attach_valid(pool, K2, valid2, n2);
  std::vector<MapPoint*> out;
  ORBmatcher m(nnratio, check_ori != 0);
  const int nm = m.SearchByBoW(&K1, &K2, out);
  for (int i = 0; i < n1; i++) match12[i] = out[i] ? out[i]->id : -1;  // map point of KF2 feature id
  return nm;
}

// ORBmatcher::SearchForTriangulation (ORBmatcher.cc:669-822). The epipole (:675-682) is produced by the reference's
// own code from Cw = (ex, ey, 1), R2w = I, t2w = 0, fx = fy = 1, cx = cy = 0.
int ref_search_for_triangulation(const float* desc1, const ora_kp* kps1, int n1, const ora_featvec* fv1,
                                 const uint8_t* has_mp1, const float* desc2, const ora_kp* kps2, int n2,
                                 const ora_featvec* fv2, const uint8_t* has_mp2, const float* F12, float ex,
                                 float ey, const float* scale_factors2, const float* level_sigma2_2, int n_levels,
                                 int check_ori, int32_t* match12) {
  Pool pool;
  KeyFrame K1, K2;
  fill_keyframe(K1, desc1, kps1, n1, nullptr);
  fill_keyframe(K2, desc2, kps2, n2, nullptr);
  featvec(fv1, K1.mFeatVec);
  featvec(fv2, K2.mFeatVec);
  for (int i = 0; i < n1; i++)
    if (has_mp1[i]) K1.mvpMapPoints[i] = pool.make(i);
  for (int i = 0; i < n2; i++)
    if (has_mp2[i]) K2.mvpMapPoints[i] = pool.make(i);
  K1.Ow = vec3(ex, ey, 1.f);
  float lsf;
  levels(scale_factors2, level_sigma2_2, n_levels, std::log(1.2f), K2.mvScaleFactors, K2.mvLevelSigma2,
         K2.mvInvLevelSigma2, K2.mnScaleLevels, lsf);
  K1.mvScaleFactors = K2.mvScaleFactors; K1.mvLevelSigma2 = K2.mvLevelSigma2;
  cv::Mat F(3, 3, CV_32F);
  memcpy(F.ptr<float>(0), F12, 9 * sizeof(float));
  std::vector<std::pair<size_t, size_t> > pairs;
  ORBmatcher m(0.6f, check_ori != 0);
  const int nm = m.SearchForTriangulation(&K1, &K2, F, pairs, false);
  for (int i = 0; i < n1; i++) match12[i] = -1;
  for (size_t k = 0; k < pairs.size(); k++) match12[pairs[k].first] = (int32_t)pairs[k].second;
  return nm;
}

// ORBmatcher::SearchForInitialization (ORBmatcher.cc:416-531)
int ref_search_for_initialization(const float* desc1, const ora_kp* kps1, int n1, const float* desc2,
                                  const ora_kp* kps2, int n2, const ora_grid* g2, float* prev_matched,
                                  int window_size, float nnratio, int check_ori, int32_t* matches12) {
  Frame F1, F2;
  fill_frame(F1, desc1, kps1, n1, nullptr);
  fill_frame(F2, desc2, kps2, n2, g2);
  std::vector<cv::Point2f> prev((size_t)n1);
  for (int i = 0; i < n1; i++) prev[i] = cv::Point2f(prev_matched[2 * i], prev_matched[2 * i + 1]);
  std::vector<int> out;
  ORBmatcher m(nnratio, check_ori != 0);
  const int nm = m.SearchForInitialization(F1, F2, prev, out, window_size);
  for (int i = 0; i < n1; i++) {
    matches12[i] = out[i];
    prev_matched[2 * i] = prev[i].x;
    prev_matched[2 * i + 1] = prev[i].y;
  }
  return nm;
}

// ------------------------------------------------------------------------------------------- projection searches
// Shared inputs: q[i].valid / .blocks / .u / .v / .angle as in ora_proj_query (r, min_level, max_level are NOT read:
// the reference derives them itself from level[i], aux[i] and th), desc_q = one descriptor per query,
// blocked0[j] = target keypoint j already holds a map point with observations.
// assign[j] = query whose map point ends up on keypoint j, or -1.

// a4: SearchByProjection(Frame&, vector<MapPoint*>&, th)  (ORBmatcher.cc:44-122); aux = mTrackViewCos
int ref_sbp_local_map(const float* desc_t, const ora_kp* kps_t, int n_t, const ora_grid* g, const ora_proj_query* q,
                      const int32_t* level, const float* view_cos, const float* desc_q, int n_q,
                      const uint8_t* blocked0, const float* scale_factors, int n_levels, float th, float nnratio,
                      int32_t* assign) {
  Pool pool;
  Frame F;
  fill_frame(F, desc_t, kps_t, n_t, g);
  F.mvScaleFactors.assign(scale_factors, scale_factors + n_levels);
  for (int j = 0; j < n_t; j++)
    if (blocked0 && blocked0[j]) {
      F.mvpMapPoints[j] = pool.make(-2);
      F.mvpMapPoints[j]->nObs = 1;
    }
  std::vector<MapPoint*> mps((size_t)n_q);
  for (int i = 0; i < n_q; i++) {
    MapPoint* p = pool.make(i);
    p->mbTrackInView = q[i].valid != 0;
    p->mTrackProjX = q[i].u;
    p->mTrackProjY = q[i].v;
    p->mnTrackScaleLevel = level[i];
    p->mTrackViewCos = view_cos[i];
    p->nObs = q[i].blocks ? 1 : 0;
    p->mDescriptor = desc_row(desc_q + (size_t)i * ORA_DESC_DIM);
    mps[i] = p;
  }
  ORBmatcher m(nnratio, true);
  const int nm = m.SearchByProjection(F, mps, th);
  for (int j = 0; j < n_t; j++) assign[j] = (F.mvpMapPoints[j] && F.mvpMapPoints[j]->id >= 0) ? F.mvpMapPoints[j]->id : -1;
  return nm;
}

// a5: SearchByProjection(Frame& Cur, const Frame& Last, th, bMono=true)  (ORBmatcher.cc:1318-1452);
// level[i] = LastFrame.mvKeys[i].octave; invalid queries alternate between "no map point" and "outlier"
int ref_sbp_frame_frame(const float* desc_t, const ora_kp* kps_t, int n_t, const ora_grid* g, const ora_proj_query* q,
                        const int32_t* level, const float* desc_q, int n_q, const uint8_t* blocked0,
                        const float* scale_factors, int n_levels, float th, float nnratio, int check_ori,
                        int32_t* assign) {
  Pool pool;
  Frame Cur, Last;
  fill_frame(Cur, desc_t, kps_t, n_t, g);
  Cur.mvScaleFactors.assign(scale_factors, scale_factors + n_levels);
  for (int j = 0; j < n_t; j++)
    if (blocked0 && blocked0[j]) {
      Cur.mvpMapPoints[j] = pool.make(-2);
      Cur.mvpMapPoints[j]->nObs = 1;
    }
  Last.N = n_q;
  Last.mTcw = scene_T(1);
  Last.mvKeys.resize((size_t)n_q);
  Last.mvKeysUn.resize((size_t)n_q);
  Last.mvpMapPoints.assign((size_t)n_q, nullptr);
  Last.mvbOutlier.assign((size_t)n_q, false);
  for (int i = 0; i < n_q; i++) {
    Last.mvKeys[i].octave = Last.mvKeysUn[i].octave = level[i];
    Last.mvKeys[i].angle = Last.mvKeysUn[i].angle = q[i].angle;
    if (!q[i].valid && !(i & 1)) continue;
    MapPoint* p = pool.make(i);
    p->mWorldPos = scene_point(0, q[i].u, q[i].v, i);
    p->nObs = q[i].blocks ? 1 : 0;
    p->mDescriptor = desc_row(desc_q + (size_t)i * ORA_DESC_DIM);
    Last.mvpMapPoints[i] = p;
    if (!q[i].valid) Last.mvbOutlier[i] = true;
  }
  ORBmatcher m(nnratio, check_ori != 0);
  const int nm = m.SearchByProjection(Cur, Last, th, true);
  for (int j = 0; j < n_t; j++) assign[j] = (Cur.mvpMapPoints[j] && Cur.mvpMapPoints[j]->id >= 0) ? Cur.mvpMapPoints[j]->id : -1;
  return nm;
}

static void set_distance_for_level(MapPoint* p, float u, float v, int lvl) {
  // PredictScale (MapPoint.cc:421-453) returns ceil(log(max/dist)/log(1.2)); aim at the middle of the level's bin
  const float dist = std::sqrt(u * u + v * v + 1.f);
  p->mfMaxDistance = dist * std::pow(1.2f, (float)lvl - 0.5f);
  p->mfMinDistance = 0.f;
}

// band for a point whose distance to the observing camera is `dist` (scene mode; see scene_band)
static void band_for_dist(MapPoint* p, float dist, int lvl, int i) {
  p->mfMaxDistance = dist * std::pow(1.2f, (float)lvl - 0.5f);
  p->mfMinDistance = 0.f;
  const double h = hash01((uint32_t)i, 4);
  if (h < 0.08) p->mfMaxDistance = dist / 1.5f;
  else if (h < 0.15) p->mfMinDistance = dist * 1.6f;
}

// world position, distance band and (optionally) normal of candidate point i that should project to (u, v) in camera 0
static void place_point(MapPoint* p, float u, float v, int lvl, int i, bool with_normal) {
  if (g_scene.mode == 0) {  // exactly the objects the identity-pose goldens were generated from
    p->mWorldPos = vec3(u, v, 1.f);
    set_distance_for_level(p, u, v, lvl);
    if (with_normal) {
      const float d = std::sqrt(u * u + v * v + 1.f);
      p->mNormalVector = vec3(u / d, v / d, 1.f / d);
    }
    return;
  }
  p->mWorldPos = scene_point(0, u, v, i);
  scene_band(p, 0, lvl, i);
  if (with_normal) p->mNormalVector = scene_normal(p, 0, i);
}

// a6: SearchByProjection(Frame& Cur, KeyFrame*, sAlreadyFound, th, ORBdist)  (ORBmatcher.cc:1455-1582);
// level[i] = level PredictScale must return; invalid queries cycle through null / bad / already found
int ref_sbp_frame_keyframe(const float* desc_t, const ora_kp* kps_t, int n_t, const ora_grid* g,
                           const ora_proj_query* q, const int32_t* level, const float* desc_q, int n_q,
                           const uint8_t* blocked0, const float* scale_factors, int n_levels, float th, float orb_dist,
                           float nnratio, int check_ori, int32_t* assign) {
  Pool pool;
  Frame Cur;
  fill_frame(Cur, desc_t, kps_t, n_t, g);
  Cur.mvScaleFactors.assign(scale_factors, scale_factors + n_levels);
  Cur.mnScaleLevels = n_levels;
  Cur.mfLogScaleFactor = std::log(1.2f);
  for (int j = 0; j < n_t; j++)
    if (blocked0 && blocked0[j]) Cur.mvpMapPoints[j] = pool.make(-2);
  KeyFrame K;
  K.N = n_q;
  K.mvKeysUn.resize((size_t)n_q);
  K.mvpMapPoints.assign((size_t)n_q, nullptr);
  std::set<MapPoint*> found;
  for (int i = 0; i < n_q; i++) {
    K.mvKeysUn[i].angle = q[i].angle;
    if (!q[i].valid && i % 3 == 0) continue;
    MapPoint* p = pool.make(i);
    place_point(p, q[i].u, q[i].v, level[i], i, false);
    p->mDescriptor = desc_row(desc_q + (size_t)i * ORA_DESC_DIM);
    K.mvpMapPoints[i] = p;
    if (!q[i].valid) {
      if (i % 3 == 1) p->mbBad = true;
      else found.insert(p);
    }
  }
  ORBmatcher m(nnratio, check_ori != 0);
  const int nm = m.SearchByProjection(Cur, &K, found, th, orb_dist);
  for (int j = 0; j < n_t; j++) assign[j] = (Cur.mvpMapPoints[j] && Cur.mvpMapPoints[j]->id >= 0) ? Cur.mvpMapPoints[j]->id : -1;
  return nm;
}

// a7: SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th)  (ORBmatcher.cc:300-413), Scw = identity
int ref_sbp_keyframe_sim3(const float* desc_t, const ora_kp* kps_t, int n_t, const ora_grid* g,
                          const ora_proj_query* q, const int32_t* level, const float* desc_q, int n_q,
                          const uint8_t* blocked0, const float* scale_factors, int n_levels, int th, float nnratio,
                          int32_t* assign) {
  Pool pool;
  KeyFrame K;
  fill_keyframe(K, desc_t, kps_t, n_t, g);
  K.mvScaleFactors.assign(scale_factors, scale_factors + n_levels);
  K.mnScaleLevels = n_levels;
  K.mfLogScaleFactor = std::log(1.2f);
  std::vector<MapPoint*> matched((size_t)n_t, nullptr);
  for (int j = 0; j < n_t; j++)
    if (blocked0 && blocked0[j]) matched[j] = pool.make(-2);
  std::vector<MapPoint*> pts((size_t)n_q);
  for (int i = 0; i < n_q; i++) {
    MapPoint* p = pool.make(i);
    place_point(p, q[i].u, q[i].v, level[i], i, true);
    p->mDescriptor = desc_row(desc_q + (size_t)i * ORA_DESC_DIM);
    if (!q[i].valid) p->mbBad = true;
    pts[i] = p;
  }
  ORBmatcher m(nnratio, true);
  const int nm = m.SearchByProjection(&K, scene_T(0, g_scene.s), pts, matched, th);
  for (int j = 0; j < n_t; j++) assign[j] = (matched[j] && matched[j]->id >= 0) ? matched[j]->id : -1;
  return nm;
}

// ------------------------------------------------------------------------------------------- Fuse / SearchBySim3
// Shared set-up: target keyframe with occupant map points (occ_obs[j] = 0: none, k > 0: a map point with k
// observations), candidate map points at (u, v, 1) with nObs = mp_obs[i], normal towards the camera.
namespace {
struct FuseRig {
  Pool pool;
  KeyFrame K;
  std::vector<MapPoint*> pts;
  FuseRig(const float* desc_t, const ora_kp* kps_t, int n_t, const ora_grid* g, const ora_proj_query* q,
          const int32_t* level, const float* desc_q, int n_q, const int32_t* occ_obs, const int32_t* mp_obs,
          const float* scale_factors, int n_levels) {
    fill_keyframe(K, desc_t, kps_t, n_t, g);
    float lsf;
    levels(scale_factors, nullptr, n_levels, std::log(1.2f), K.mvScaleFactors, K.mvLevelSigma2, K.mvInvLevelSigma2,
           K.mnScaleLevels, lsf);
    K.mfLogScaleFactor = lsf;
    for (int j = 0; j < n_t; j++)
      if (occ_obs[j] > 0) {
        K.mvpMapPoints[j] = pool.make(1000000 + j);
        K.mvpMapPoints[j]->nObs = occ_obs[j];
        K.mvpMapPoints[j]->mObservations[&K] = (size_t)j;
      }
    pts.assign((size_t)n_q, nullptr);
    for (int i = 0; i < n_q; i++) {
      if (!q[i].valid && (i % 3) == 0) continue;  // null entry (Fuse(KF, vector) tolerates them)
      MapPoint* p = pool.make(i);
      place_point(p, q[i].u, q[i].v, level[i], i, true);
      p->mDescriptor = desc_row(desc_q + (size_t)i * ORA_DESC_DIM);
      p->nObs = mp_obs[i];
      if (!q[i].valid) p->mbBad = true;
      pts[i] = p;
    }
  }
};
}  // namespace

// Fuse(KeyFrame*, vpMapPoints, th)  (ORBmatcher.cc:825-960).  action[i]: -1 nothing; 4*idx+0 observation added at
// keypoint idx; 4*idx+1 candidate replaced by the keypoint's occupant; 4*idx+2 occupant replaced by the candidate.
int ref_fuse(const float* desc_t, const ora_kp* kps_t, int n_t, const ora_grid* g, const ora_proj_query* q,
             const int32_t* level, const float* desc_q, int n_q, const int32_t* occ_obs, const int32_t* mp_obs,
             const float* scale_factors, int n_levels, float th, int32_t* action) {
  FuseRig R(desc_t, kps_t, n_t, g, q, level, desc_q, n_q, occ_obs, mp_obs, scale_factors, n_levels);
  ORBmatcher m(0.6f, true);
  const int nf = m.Fuse(&R.K, R.pts, th);
  // final state of the graph as seen from every candidate: where it sits in the keyframe (4*idx), whether it was
  // replaced (4*idx_of_replacer+1, or 4*replacer_candidate+3 when the replacer is another candidate), or nothing
  for (int i = 0; i < n_q; i++) {
    action[i] = -1;
    MapPoint* p = R.pts[i];
    if (!p) continue;
    if (p->mObservations.count(&R.K)) action[i] = 4 * (int)p->mObservations[&R.K] + 0;
    else if (p->mpReplaced) {
      MapPoint* r = p->mpReplaced;
      action[i] = r->id >= 1000000 ? 4 * (r->id - 1000000) + 1 : 4 * r->id + 3;
    }
  }
  return nf;
}

// Fuse(KeyFrame*, Scw, vpPoints, th, vpReplacePoint)  (ORBmatcher.cc:963-1086), Scw = identity.
// replace[i] = keypoint whose occupant lands in vpReplacePoint[i], added[i] = keypoint that received the candidate.
int ref_fuse_sim3(const float* desc_t, const ora_kp* kps_t, int n_t, const ora_grid* g, const ora_proj_query* q,
                  const int32_t* level, const float* desc_q, int n_q, const int32_t* occ_obs, const int32_t* mp_obs,
                  const float* scale_factors, int n_levels, float th, int32_t* replace, int32_t* added) {
  FuseRig R(desc_t, kps_t, n_t, g, q, level, desc_q, n_q, occ_obs, mp_obs, scale_factors, n_levels);
  for (int i = 0; i < n_q; i++)
    if (!R.pts[i]) {  // this overload dereferences every entry: give the null slots a bad point
      R.pts[i] = R.pool.make(i);
      R.pts[i]->mbBad = true;
    }
  std::vector<MapPoint*> repl((size_t)n_q, nullptr);
  ORBmatcher m(0.8f, true);
  const int nf = m.Fuse(&R.K, scene_T(0, g_scene.s), R.pts, th, repl);
  for (int i = 0; i < n_q; i++) {
    replace[i] = repl[i] ? repl[i]->GetIndexInKeyFrame(&R.K) : -1;
    added[i] = R.pts[i]->mObservations.count(&R.K) ? (int)R.pts[i]->mObservations[&R.K] : -1;
  }
  return nf;
}

// SearchBySim3 (ORBmatcher.cc:1090-1314) with identity poses and s12 = 1, R12 = I, t12 = 0: map point i of KF1 sits
// at (u1[i], v1[i], 1) (its projection into KF2) and map point j of KF2 at (u2[j], v2[j], 1).
int ref_search_by_sim3(const float* desc1, const ora_kp* kps1, int n1, const float* desc2, const ora_kp* kps2, int n2,
                       const ora_grid* g, const uint8_t* valid1, const uint8_t* valid2, const float* uv1,
                       const float* uv2, const int32_t* level1, const int32_t* level2, const float* mpdesc1,
                       const float* mpdesc2, const uint8_t* already1, const float* scale_factors, int n_levels, float th,
                       int32_t* match12) {
  Pool pool;
  KeyFrame K1, K2;
  fill_keyframe(K1, desc1, kps1, n1, g);
  fill_keyframe(K2, desc2, kps2, n2, g);
  float lsf;
  levels(scale_factors, nullptr, n_levels, std::log(1.2f), K1.mvScaleFactors, K1.mvLevelSigma2, K1.mvInvLevelSigma2,
         K1.mnScaleLevels, lsf);
  K1.mfLogScaleFactor = lsf;
  K2.mvScaleFactors = K1.mvScaleFactors; K2.mnScaleLevels = n_levels; K2.mfLogScaleFactor = lsf;
  // relative Sim3 (camera 2 -> camera 1): identity in mode 0; in scene mode a genuine one, KF2 on the second pose
  double R12d[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, t12d[3] = {0, 0, 0}, s12d = 1.0;
  if (g_scene.mode) {
    set_keyframe_camera(K2, 1);
    rodrigues(0.3, -0.5, 0.8, 0.21, R12d);
    t12d[0] = 0.3; t12d[1] = -0.2; t12d[2] = 0.12;
    s12d = 1.1;
  }
  for (int i = 0; i < n1; i++)
    if (valid1[i] || (i & 1)) {
      MapPoint* p = pool.make(i);
      p->mbBad = !valid1[i];
      if (g_scene.mode == 0) {
        p->mWorldPos = vec3(uv1[2 * i], uv1[2 * i + 1], 1.f);
        set_distance_for_level(p, uv1[2 * i], uv1[2 * i + 1], level1[i]);
      } else {  // seen in camera 2 at uv1: Xc2 -> Xc1 = s12 R12 Xc2 + t12 -> world through camera 1 (pose 0)
        double X2[3], X1[3], Xw[3];
        scene_backproject(uv1[2 * i], uv1[2 * i + 1], i, X2);
        for (int r = 0; r < 3; r++) X1[r] = s12d * (R12d[3 * r] * X2[0] + R12d[3 * r + 1] * X2[1] + R12d[3 * r + 2] * X2[2]) + t12d[r];
        scene_to_world(0, X1, Xw);
        p->mWorldPos = vec3((float)Xw[0], (float)Xw[1], (float)Xw[2]);
        band_for_dist(p, (float)std::sqrt(X2[0] * X2[0] + X2[1] * X2[1] + X2[2] * X2[2]), level1[i], i);
      }
      p->mDescriptor = desc_row(mpdesc1 + (size_t)i * ORA_DESC_DIM);
      K1.mvpMapPoints[i] = p;
    }
  for (int j = 0; j < n2; j++)
    if (valid2[j] || (j & 1)) {
      MapPoint* p = pool.make(j);
      p->mbBad = !valid2[j];
      if (g_scene.mode == 0) {
        p->mWorldPos = vec3(uv2[2 * j], uv2[2 * j + 1], 1.f);
        set_distance_for_level(p, uv2[2 * j], uv2[2 * j + 1], level2[j]);
      } else {  // seen in camera 1 at uv2: Xc1 -> Xc2 = (1/s12) R12^T (Xc1 - t12) -> world through camera 2 (pose 1)
        double X1[3], X2[3], Xw[3];
        scene_backproject(uv2[2 * j], uv2[2 * j + 1], 100000 + j, X1);
        for (int c = 0; c < 3; c++)
          X2[c] = (R12d[c] * (X1[0] - t12d[0]) + R12d[3 + c] * (X1[1] - t12d[1]) + R12d[6 + c] * (X1[2] - t12d[2])) / s12d;
        scene_to_world(1, X2, Xw);
        p->mWorldPos = vec3((float)Xw[0], (float)Xw[1], (float)Xw[2]);
        band_for_dist(p, (float)std::sqrt(X1[0] * X1[0] + X1[1] * X1[1] + X1[2] * X1[2]), level2[j], 100000 + j);
      }
      p->mDescriptor = desc_row(mpdesc2 + (size_t)j * ORA_DESC_DIM);
      p->mObservations[&K2] = (size_t)j;
      K2.mvpMapPoints[j] = p;
    }
  std::vector<MapPoint*> m12((size_t)n1, nullptr);
  std::vector<MapPoint*> pre;
  for (int i = 0; i < n1; i++)
    if (already1[i] && K2.mvpMapPoints[i % n2] && !K2.mvpMapPoints[i % n2]->mbBad) {
      m12[i] = K2.mvpMapPoints[i % n2];  // an earlier SearchByBoW match (LoopClosing.cc:304)
      pre.push_back(m12[i]);
    }
  cv::Mat R12(3, 3, CV_32F);
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) R12.at<float>(r, c) = (float)R12d[3 * r + c];
  ORBmatcher m(0.85f, true);
  const float s12 = (float)s12d;
  const int nf = m.SearchBySim3(&K1, &K2, m12, s12, R12, vec3((float)t12d[0], (float)t12d[1], (float)t12d[2]), th);
  for (int i = 0; i < n1; i++) match12[i] = m12[i] ? m12[i]->id : -1;
  return nf;
}

// Descriptor stage of Frame::ComputeStereoMatches (Frame.cc:360-447; dead code upstream, BASELINE config 3). The
// reference's own text runs here (oracle/Makefile extracts :366-397, :399-444 and :447): this function only supplies the
// members it names as locals. minD = 0 and maxD = mbf / mb in the reference (:391-393): mb = 1, mbf = max_d gives
// maxD = max_d exactly. Right keypoints must keep their row bands inside the image (the reference indexes vRowIndices
// without a bounds check, :384-387; its extractor keeps keypoints 19 px off the border).
void ref_stereo_row_match(const float* desc_l, const ora_kp* kps_l, int n_l, const float* desc_r, const ora_kp* kps_r,
                          int n_r, const float* scale_factors, int n_levels, int n_rows, float max_d, int32_t* best_r,
                          float* best_dist) {
  using namespace std;
  struct Level0 { int rows; };
  struct Extractor { vector<Level0> mvImagePyramid; } extractor;
  extractor.mvImagePyramid.push_back(Level0{n_rows});
  Extractor* mpORBextractorLeft = &extractor;
  const int N = n_l;
  vector<cv::KeyPoint> mvKeys = keypoints(kps_l, n_l), mvKeysRight = keypoints(kps_r, n_r);
  vector<float> mvScaleFactors(scale_factors, scale_factors + n_levels);
  cv::Mat mDescriptors = desc_mat(desc_l, n_l), mDescriptorsRight = desc_mat(desc_r, n_r);
  const float mb = 1.0f, mbf = max_d;
  for (int i = 0; i < n_l; i++) {
    best_r[i] = -1;
    best_dist[i] = -1.f;
  }
#include "_ref/gen/frame_stereo_a.inc"
#include "_ref/gen/frame_stereo_b.inc"
#include "_ref/gen/frame_stereo_c.inc"
        {
            best_r[iL] = (int32_t)bestIdxR;
            best_dist[iL] = bestDist;
        }
    }
}

// MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:271-338) over a batch of map points. Observations of point p
// are (obs_kf[e], obs_row[e]) for e in [obs_off[p], obs_off[p+1]), ascending and unique in obs_kf -- the reference
// walks a std::map<KeyFrame*, size_t>, and the keyframes live in one array here so pointer order is index order.
// out_desc: the 128 floats the reference leaves in mDescriptor (untouched rows when a point has no observation).
void ref_distinctive(const float* const* kf_desc, const int32_t* kf_n, int n_kf, int n_points, const int32_t* obs_off,
                     const int32_t* obs_kf, const int32_t* obs_row, float* out_desc) {
  std::vector<KeyFrame> kfs((size_t)n_kf);
  for (int k = 0; k < n_kf; k++) kfs[k].mDescriptors = desc_mat(kf_desc[k], kf_n[k]);
  for (int p = 0; p < n_points; p++) {
    MapPoint mp;
    for (int e = obs_off[p]; e < obs_off[p + 1]; e++) mp.mObservations[&kfs[obs_kf[e]]] = (size_t)obs_row[e];
    mp.ComputeDistinctiveDescriptors();
    if (!mp.mDescriptor.empty()) memcpy(out_desc + (size_t)p * ORA_DESC_DIM, mp.mDescriptor.ptr<float>(0), ORA_DESC_DIM * sizeof(float));
  }
}

}  // extern "C"
