// ORBmatcher_cuda.cc -- drop-in replacement for the reference's src/vslam/src/ORBmatcher.cc.
//
// It defines the SAME class, ORB_SLAM2::ORBmatcher, declared by the reference's own include/ORBmatcher.h
// (ORBmatcher.h:43-110: constructor, 11 search methods, DescriptorDistance, the three protected helpers and the
// TH_LOW / TH_HIGH / HISTO_LENGTH constants), so Tracking / LocalMapping / LoopClosing / MapPoint compile and call
// it unchanged.  Every descriptor loop, gate, ratio test, greedy assignment and rotation histogram runs in
// libasdmatch_cuda.so (include/asdmatch.h); this file only
//   1. evaluates, per query, the gates that precede the descriptor loop in the reference (pose projection, image
//      bounds, distance invariance, viewing angle, PredictScale) with the same cv::Mat expressions the reference
//      uses, so the window (u, v, r, level range) is bit-identical,
//   2. flattens MapPoint* state into masks / query arrays, and
//   3. maps the returned index lists back onto MapPoint* containers.
// Frame / KeyFrame descriptors live on the device (asdm::DeviceSet, uploaded once, see asdm_device.h).
// There is no CPU matching path here: a failing GPU call aborts.
#include <cmath>
#include <cstring>
#include <set>
#include <vector>

#include "ORBmatcher.h"  // the reference's header (through oracle/ref_shim/ORBmatcher.h in the test build)
#include "asdm_device.h"

using namespace std;

namespace ORB_SLAM2 {

const float ORBmatcher::TH_HIGH = ASDM_TH_HIGH;
const float ORBmatcher::TH_LOW = ASDM_TH_LOW;
const int ORBmatcher::HISTO_LENGTH = ASDM_HISTO_LENGTH;

ORBmatcher::ORBmatcher(float nnratio, bool checkOri) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}

namespace {

asdm_grid frame_grid() {  // Frame keeps its bounds in static floats (Frame.cc:34, Frame.h:156-157, 181-184)
  asdm_grid g;
  g.min_x = Frame::mnMinX; g.min_y = Frame::mnMinY; g.max_x = Frame::mnMaxX; g.max_y = Frame::mnMaxY;
  g.inv_w = Frame::mfGridElementWidthInv; g.inv_h = Frame::mfGridElementHeightInv;
  g.cols = FRAME_GRID_COLS; g.rows = FRAME_GRID_ROWS;
  return g;
}

asdm_grid keyframe_grid(const KeyFrame* pKF) {  // KeyFrame keeps integer bounds (KeyFrame.h:217-220); exact in float
  asdm_grid g;
  g.min_x = (float)pKF->mnMinX; g.min_y = (float)pKF->mnMinY; g.max_x = (float)pKF->mnMaxX; g.max_y = (float)pKF->mnMaxY;
  g.inv_w = pKF->mfGridElementWidthInv; g.inv_h = pKF->mfGridElementHeightInv;
  g.cols = pKF->mnGridCols; g.rows = pKF->mnGridRows;
  return g;
}

asdm::DeviceSet& dev(const Frame& F) {
  if (!F.mDev) {
    const asdm_grid g = frame_grid();
    F.mDev = asdm::Upload(F.mDescriptors, F.mvKeysUn, &g);
  }
  return *F.mDev;
}

asdm::DeviceSet& dev(const KeyFrame* pKF) {
  if (!pKF->mDev) {
    const asdm_grid g = keyframe_grid(pKF);
    pKF->mDev = asdm::Upload(pKF->mDescriptors, pKF->mvKeysUn, &g);
  }
  return *pKF->mDev;
}

// Page-locked rows, one buffer per calling thread, kept across calls: the projection kernel reads rows that lie in
// page-locked memory in place over PCIe, so gathering the map-point descriptors straight into such a buffer saves the
// library's own staging copy of the same megabyte (measured on the C2 call: 0.16 ms -> 0.07 ms).
struct PinnedRows {
  float* p = nullptr;
  size_t cap = 0;  // rows
  ~PinnedRows() { asdm_host_free(p); }
  float* row(size_t i) {
    if (i >= cap) {
      const size_t want = std::max<size_t>(2 * cap, 4096);
      void* np = nullptr;
      asdm::Check(asdm_host_alloc(&np, want * ASDM_DESC_DIM * sizeof(float)), "asdm_host_alloc");
      if (p) memcpy(np, p, cap * ASDM_DESC_DIM * sizeof(float));
      asdm_host_free(p);
      p = (float*)np;
      cap = want;
    }
    return p + i * ASDM_DESC_DIM;
  }
};
PinnedRows& thread_rows() {
  static thread_local PinnedRows rows;
  return rows;
}

// MapPoint::GetDescriptor() rows gathered for one call (map-point descriptors are 1 x 128 Mats scattered on the heap).
// One QueryPack is alive per thread at a time: the searches below are synchronous calls.
struct QueryPack {
  vector<asdm_proj_query> q;
  PinnedRows& desc = thread_rows();
  vector<int> src;  // index of the query in the caller's container
  void add(int source, float u, float v, float r, int min_level, int max_level, bool blocks, float angle,
           const cv::Mat& d) {
    asdm_proj_query Q;
    memset(&Q, 0, sizeof Q);
    Q.valid = 1;
    Q.blocks = blocks ? 1 : 0;
    Q.u = u; Q.v = v; Q.r = r;
    Q.min_level = min_level; Q.max_level = max_level;
    Q.angle = angle;
    memcpy(desc.row(q.size()), d.ptr<float>(0), ASDM_DESC_DIM * sizeof(float));
    q.push_back(Q);
    src.push_back(source);
  }
  int n() const { return (int)q.size(); }
};

asdm_proj_params proj_params(int mode, float th_accept, float best_init, float nnratio, bool check_ori, bool twice) {
  asdm_proj_params p;
  memset(&p, 0, sizeof p);
  p.mode = mode;
  p.th_accept = th_accept;
  p.best_init = best_init;
  p.nnratio = nnratio;
  p.check_ori = check_ori ? 1 : 0;
  p.count_twice = twice ? 1 : 0;
  return p;
}

}  // namespace

// ------------------------------------------------------------------------------------------------------------------
// Kept for the callers outside the matcher (MapPoint::ComputeDistinctiveDescriptors, MapPoint.cc:305-316): the
// reference's arithmetic (ORBmatcher.cc:1637-1644) on the two rows.  No Search* / Fuse method below calls it.
float ORBmatcher::DescriptorDistance(const cv::Mat& a, const cv::Mat& b) {
  const float* pa = a.ptr<float>(0);
  const float* pb = b.ptr<float>(0);
  float sqd = 0.f;
  for (int i = 0; i < ASDM_DESC_DIM; i++) {
    const float d = pa[i] - pb[i];
    sqd += d * d;  // compile with -ffp-contract=off
  }
  return sqd;
}

float ORBmatcher::RadiusByViewingCos(const float& viewCos) { return viewCos > 0.998 ? 2.5f : 4.0f; }  // :126-132

// The two protected helpers below are part of the class declaration; the GPU kernels carry their own copies
// (asdm_three_maxima, tri_geo_ok).  They are defined so that the reference header links unchanged.
bool ORBmatcher::CheckDistEpipolarLine(const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Mat& F12,
                                       const KeyFrame* pKF2) {
  const float a = kp1.pt.x * F12.at<float>(0, 0) + kp1.pt.y * F12.at<float>(1, 0) + F12.at<float>(2, 0);
  const float b = kp1.pt.x * F12.at<float>(0, 1) + kp1.pt.y * F12.at<float>(1, 1) + F12.at<float>(2, 1);
  const float c = kp1.pt.x * F12.at<float>(0, 2) + kp1.pt.y * F12.at<float>(1, 2) + F12.at<float>(2, 2);
  const float num = a * kp2.pt.x + b * kp2.pt.y + c;
  const float den = a * a + b * b;
  if (den == 0) return false;
  return num * num / den < 3.84 * pKF2->mvLevelSigma2[kp2.octave];
}

void ORBmatcher::ComputeThreeMaxima(vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3) {
  int m1 = 0, m2 = 0, m3 = 0;
  for (int i = 0; i < L; i++) {
    const int s = (int)histo[i].size();
    if (s > m1) { m3 = m2; m2 = m1; m1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
    else if (s > m2) { m3 = m2; m2 = s; ind3 = ind2; ind2 = i; }
    else if (s > m3) { m3 = s; ind3 = i; }
  }
  if (m2 < 0.1f * (float)m1) ind2 = ind3 = -1;
  else if (m3 < 0.1f * (float)m1) ind3 = -1;
}

// ------------------------------------------------------------------------------------------------------------------
// SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches)   reference :156-297
int ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame& F, vector<MapPoint*>& vpMapPointMatches) {
  const vector<MapPoint*> vpMapPointsKF = pKF->GetMapPointMatches();
  vpMapPointMatches = vector<MapPoint*>(F.N, static_cast<MapPoint*>(NULL));
  asdm::DeviceSet& dk = dev(pKF);
  asdm::DeviceSet& df = dev(F);
  asdm::AttachFeatureVector(dk, pKF->mFeatVec);
  asdm::AttachFeatureVector(df, F.mFeatVec);
  vector<uint8_t> valid(vpMapPointsKF.size() + 1, 0);
  for (size_t i = 0; i < vpMapPointsKF.size(); i++) {
    MapPoint* pMP = vpMapPointsKF[i];
    valid[i] = (pMP && !pMP->isBad()) ? 1 : 0;  // :194-200
  }
  vector<int32_t> match((size_t)F.N + 1, -1);
  int32_t nm = 0;
  asdm::Check(asdm_search_by_bow_kf_f(asdm::Context(), dk.set, valid.data(), df.set, mfNNratio,
                                      mbCheckOrientation ? 1 : 0, match.data(), NULL, &nm),
              "asdm_search_by_bow_kf_f");
  for (int j = 0; j < F.N; j++)
    if (match[j] >= 0) vpMapPointMatches[j] = vpMapPointsKF[match[j]];
  return nm;
}

// SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12)   reference :533-666
int ORBmatcher::SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, vector<MapPoint*>& vpMatches12) {
  const vector<MapPoint*> vpMapPoints1 = pKF1->GetMapPointMatches();
  const vector<MapPoint*> vpMapPoints2 = pKF2->GetMapPointMatches();
  vpMatches12 = vector<MapPoint*>(vpMapPoints1.size(), static_cast<MapPoint*>(NULL));
  asdm::DeviceSet& d1 = dev(pKF1);
  asdm::DeviceSet& d2 = dev(pKF2);
  asdm::AttachFeatureVector(d1, pKF1->mFeatVec);
  asdm::AttachFeatureVector(d2, pKF2->mFeatVec);
  vector<uint8_t> v1(vpMapPoints1.size() + 1, 0), v2(vpMapPoints2.size() + 1, 0);
  for (size_t i = 0; i < vpMapPoints1.size(); i++) v1[i] = (vpMapPoints1[i] && !vpMapPoints1[i]->isBad()) ? 1 : 0;
  for (size_t i = 0; i < vpMapPoints2.size(); i++) v2[i] = (vpMapPoints2[i] && !vpMapPoints2[i]->isBad()) ? 1 : 0;
  vector<int32_t> match(vpMapPoints1.size() + 1, -1);
  int32_t nm = 0;
  asdm::Check(asdm_search_by_bow_kf_kf(asdm::Context(), d1.set, v1.data(), d2.set, v2.data(), mfNNratio,
                                       mbCheckOrientation ? 1 : 0, match.data(), NULL, &nm),
              "asdm_search_by_bow_kf_kf");
  for (size_t i = 0; i < vpMapPoints1.size(); i++)
    if (match[i] >= 0) vpMatches12[i] = vpMapPoints2[match[i]];
  return nm;
}

// SearchForTriangulation   reference :669-822
int ORBmatcher::SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12,
                                       vector<pair<size_t, size_t> >& vMatchedPairs, const bool bOnlyStereo) {
  (void)bOnlyStereo;  // the fork is monocular: the reference body never reads it either
  // epipole of camera 1 in image 2 (:675-682), same expressions
  cv::Mat Cw = pKF1->GetCameraCenter();
  cv::Mat R2w = pKF2->GetRotation();
  cv::Mat t2w = pKF2->GetTranslation();
  cv::Mat C2 = R2w * Cw + t2w;
  const float invz = 1.0f / C2.at<float>(2);
  const float ex = pKF2->fx * C2.at<float>(0) * invz + pKF2->cx;
  const float ey = pKF2->fy * C2.at<float>(1) * invz + pKF2->cy;

  asdm::DeviceSet& d1 = dev(pKF1);
  asdm::DeviceSet& d2 = dev(pKF2);
  asdm::AttachFeatureVector(d1, pKF1->mFeatVec);
  asdm::AttachFeatureVector(d2, pKF2->mFeatVec);
  vector<uint8_t> has1((size_t)pKF1->N + 1, 0), has2((size_t)pKF2->N + 1, 0);
  for (int i = 0; i < pKF1->N; i++) has1[i] = pKF1->GetMapPoint(i) ? 1 : 0;  // :708-712
  for (int i = 0; i < pKF2->N; i++) has2[i] = pKF2->GetMapPoint(i) ? 1 : 0;  // :730
  float F[9];
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) F[3 * r + c] = F12.at<float>(r, c);
  vector<int32_t> match((size_t)pKF1->N + 1, -1);
  int32_t nm = 0;
  asdm::Check(asdm_search_for_triangulation(asdm::Context(), d1.set, has1.data(), d2.set, has2.data(), F, ex, ey,
                                            pKF2->mvScaleFactors.data(), pKF2->mvLevelSigma2.data(),
                                            (int)pKF2->mvScaleFactors.size(), mbCheckOrientation ? 1 : 0, match.data(),
                                            NULL, &nm),
              "asdm_search_for_triangulation");
  vMatchedPairs.clear();
  vMatchedPairs.reserve(nm);
  for (int i = 0; i < pKF1->N; i++)  // :814-819: pairs in ascending idx1
    if (match[i] >= 0) vMatchedPairs.push_back(make_pair((size_t)i, (size_t)match[i]));
  return nm;
}

// SearchForInitialization   reference :416-531
int ORBmatcher::SearchForInitialization(Frame& F1, Frame& F2, vector<cv::Point2f>& vbPrevMatched,
                                        vector<int>& vnMatches12, int windowSize) {
  asdm::DeviceSet& d1 = dev(F1);
  asdm::DeviceSet& d2 = dev(F2);
  const int n1 = (int)F1.mvKeysUn.size();
  vector<float> prev(2 * (size_t)n1 + 2, 0.f);
  for (int i = 0; i < n1; i++) {
    prev[2 * i] = vbPrevMatched[i].x;
    prev[2 * i + 1] = vbPrevMatched[i].y;
  }
  vector<int32_t> match((size_t)n1 + 1, -1);
  int32_t nm = 0;
  asdm::Check(asdm_search_for_initialization(asdm::Context(), d1.set, d2.set, prev.data(), windowSize, mfNNratio,
                                             mbCheckOrientation ? 1 : 0, match.data(), NULL, &nm),
              "asdm_search_for_initialization");
  vnMatches12 = vector<int>(n1, -1);
  for (int i = 0; i < n1; i++) {
    vnMatches12[i] = match[i];
    vbPrevMatched[i].x = prev[2 * i];  // :526-528 (only matched entries change on the device)
    vbPrevMatched[i].y = prev[2 * i + 1];
  }
  return nm;
}

// ------------------------------------------------------------------------------------------------------------------
// SearchByProjection(Frame&, vector<MapPoint*>&, th)   reference :44-122   (tracking the local map)
int ORBmatcher::SearchByProjection(Frame& F, const vector<MapPoint*>& vpMapPoints, const float th) {
  const bool bFactor = th != 1.0;
  QueryPack P;
  for (size_t iMP = 0; iMP < vpMapPoints.size(); iMP++) {
    MapPoint* pMP = vpMapPoints[iMP];
    if (!pMP->mbTrackInView) continue;
    if (pMP->isBad()) continue;
    const int& nPredictedLevel = pMP->mnTrackScaleLevel;
    float r = RadiusByViewingCos(pMP->mTrackViewCos);  // :62-66
    if (bFactor) r *= th;
    P.add((int)iMP, pMP->mTrackProjX, pMP->mTrackProjY, r * F.mvScaleFactors[nPredictedLevel], nPredictedLevel - 1,
          nPredictedLevel, pMP->Observations() > 0, 0.f, pMP->GetDescriptor());
  }
  vector<uint8_t> blocked((size_t)F.N + 1, 0);
  for (int j = 0; j < F.N; j++) blocked[j] = (F.mvpMapPoints[j] && F.mvpMapPoints[j]->Observations() > 0) ? 1 : 0;  // :86-88
  const asdm_proj_params pp = proj_params(ASDM_PROJ_RATIO_LEVEL, TH_HIGH, 256.f, mfNNratio, false, true);
  vector<int32_t> assign((size_t)F.N + 1, -1);
  int32_t nm = 0;
  asdm::Check(asdm_search_by_projection(asdm::Context(), dev(F).set, P.q.data(), P.n(), P.desc.p, NULL, NULL,
                                        blocked.data(), &pp, assign.data(), NULL, &nm),
              "asdm_search_by_projection");
  for (int j = 0; j < F.N; j++)
    if (assign[j] >= 0) F.mvpMapPoints[j] = vpMapPoints[P.src[assign[j]]];  // :115
  return nm;  // doubled, as the reference's two increments (:116-117)
}

// SearchByProjection(Frame& Cur, const Frame& Last, th, bMono)   reference :1318-1452   (motion model)
int ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono) {
  const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3);
  const cv::Mat tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
  const cv::Mat twc = -Rcw.t() * tcw;
  const cv::Mat Rlw = LastFrame.mTcw.rowRange(0, 3).colRange(0, 3);
  const cv::Mat tlw = LastFrame.mTcw.rowRange(0, 3).col(3);
  const cv::Mat tlc = Rlw * twc + tlw;
  const bool bForward = tlc.at<float>(2) > CurrentFrame.mb && !bMono;
  const bool bBackward = -tlc.at<float>(2) > CurrentFrame.mb && !bMono;

  QueryPack P;
  for (int i = 0; i < LastFrame.N; i++) {
    MapPoint* pMP = LastFrame.mvpMapPoints[i];
    if (!pMP || LastFrame.mvbOutlier[i]) continue;
    cv::Mat x3Dw = pMP->GetWorldPos();
    cv::Mat x3Dc = Rcw * x3Dw + tcw;
    const float xc = x3Dc.at<float>(0);
    const float yc = x3Dc.at<float>(1);
    const float invzc = 1.0 / x3Dc.at<float>(2);
    if (invzc < 0) continue;
    float u = CurrentFrame.fx * xc * invzc + CurrentFrame.cx;
    float v = CurrentFrame.fy * yc * invzc + CurrentFrame.cy;
    if (u < CurrentFrame.mnMinX || u > CurrentFrame.mnMaxX) continue;
    if (v < CurrentFrame.mnMinY || v > CurrentFrame.mnMaxY) continue;
    const int nLastOctave = LastFrame.mvKeys[i].octave;
    const float radius = th * CurrentFrame.mvScaleFactors[nLastOctave];
    int lo, hi;  // the three GetFeaturesInArea variants of :1375-1380
    if (bForward) { lo = nLastOctave; hi = -1; }
    else if (bBackward) { lo = 0; hi = nLastOctave; }
    else { lo = nLastOctave - 1; hi = nLastOctave + 1; }
    P.add(i, u, v, radius, lo, hi, pMP->Observations() > 0, LastFrame.mvKeysUn[i].angle, pMP->GetDescriptor());
  }
  const int N = CurrentFrame.N;
  vector<uint8_t> blocked((size_t)N + 1, 0);
  for (int j = 0; j < N; j++)
    blocked[j] = (CurrentFrame.mvpMapPoints[j] && CurrentFrame.mvpMapPoints[j]->Observations() > 0) ? 1 : 0;  // :1393-1395
  const asdm_proj_params pp = proj_params(ASDM_PROJ_BEST_ONLY, TH_HIGH, 100.f, mfNNratio, mbCheckOrientation, false);
  vector<int32_t> assign((size_t)N + 1, -1);
  int32_t nm = 0;
  asdm::Check(asdm_search_by_projection(asdm::Context(), dev(CurrentFrame).set, P.q.data(), P.n(), P.desc.p, NULL,
                                        NULL, blocked.data(), &pp, assign.data(), NULL, &nm),
              "asdm_search_by_projection");
  for (int j = 0; j < N; j++) {
    if (assign[j] >= 0) CurrentFrame.mvpMapPoints[j] = LastFrame.mvpMapPoints[P.src[assign[j]]];  // :1410
    else if (assign[j] == -2) CurrentFrame.mvpMapPoints[j] = static_cast<MapPoint*>(NULL);       // :1441
  }
  return nm;
}

// SearchByProjection(Frame& Cur, KeyFrame*, sAlreadyFound, th, ORBdist)   reference :1455-1582   (relocalisation)
int ORBmatcher::SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const set<MapPoint*>& sAlreadyFound,
                                   const float th, const float ORBdist) {
  const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3);
  const cv::Mat tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
  const cv::Mat Ow = -Rcw.t() * tcw;
  const vector<MapPoint*> vpMPs = pKF->GetMapPointMatches();
  QueryPack P;
  for (size_t i = 0, iend = vpMPs.size(); i < iend; i++) {
    MapPoint* pMP = vpMPs[i];
    if (!pMP || pMP->isBad() || sAlreadyFound.count(pMP)) continue;
    cv::Mat x3Dw = pMP->GetWorldPos();
    cv::Mat x3Dc = Rcw * x3Dw + tcw;
    const float xc = x3Dc.at<float>(0);
    const float yc = x3Dc.at<float>(1);
    const float invzc = 1.0 / x3Dc.at<float>(2);
    const float u = CurrentFrame.fx * xc * invzc + CurrentFrame.cx;
    const float v = CurrentFrame.fy * yc * invzc + CurrentFrame.cy;
    if (u < CurrentFrame.mnMinX || u > CurrentFrame.mnMaxX) continue;
    if (v < CurrentFrame.mnMinY || v > CurrentFrame.mnMaxY) continue;
    cv::Mat PO = x3Dw - Ow;
    float dist3D = cv::norm(PO);
    const float maxDistance = pMP->GetMaxDistanceInvariance();
    const float minDistance = pMP->GetMinDistanceInvariance();
    if (dist3D < minDistance || dist3D > maxDistance) continue;
    int nPredictedLevel = pMP->PredictScale(dist3D, &CurrentFrame);
    const float radius = th * CurrentFrame.mvScaleFactors[nPredictedLevel];
    P.add((int)i, u, v, radius, nPredictedLevel - 1, nPredictedLevel + 1, true, pKF->mvKeysUn[i].angle,
          pMP->GetDescriptor());
  }
  const int N = CurrentFrame.N;
  vector<uint8_t> blocked((size_t)N + 1, 0);
  for (int j = 0; j < N; j++) blocked[j] = CurrentFrame.mvpMapPoints[j] ? 1 : 0;  // :1524: any occupant blocks
  const asdm_proj_params pp = proj_params(ASDM_PROJ_BEST_ONLY, ORBdist, 256.f, mfNNratio, mbCheckOrientation, false);
  vector<int32_t> assign((size_t)N + 1, -1);
  int32_t nm = 0;
  asdm::Check(asdm_search_by_projection(asdm::Context(), dev(CurrentFrame).set, P.q.data(), P.n(), P.desc.p, NULL,
                                        NULL, blocked.data(), &pp, assign.data(), NULL, &nm),
              "asdm_search_by_projection");
  for (int j = 0; j < N; j++) {
    if (assign[j] >= 0) CurrentFrame.mvpMapPoints[j] = vpMPs[P.src[assign[j]]];
    else if (assign[j] == -2) CurrentFrame.mvpMapPoints[j] = NULL;  // :1571
  }
  return nm;
}

namespace {
// Sim3 decomposition shared by :300-413 and :963-1086
void decompose_sim3(const cv::Mat& Scw, cv::Mat& Rcw, cv::Mat& tcw, cv::Mat& Ow) {
  cv::Mat sRcw = Scw.rowRange(0, 3).colRange(0, 3);
  const float scw = sqrt(sRcw.row(0).dot(sRcw.row(0)));
  Rcw = sRcw / scw;
  tcw = Scw.rowRange(0, 3).col(3) / scw;
  Ow = -Rcw.t() * tcw;
}

// The gates every KeyFrame-side projection shares (:331-368, :852-888, :995-1033): returns false when the point is
// rejected before the descriptor loop; otherwise the window centre and the predicted level.
bool project_into_keyframe(KeyFrame* pKF, MapPoint* pMP, const cv::Mat& Rcw, const cv::Mat& tcw, const cv::Mat& Ow,
                           bool invz_double, float& u, float& v, int& level) {
  cv::Mat p3Dw = pMP->GetWorldPos();
  cv::Mat p3Dc = Rcw * p3Dw + tcw;
  if (p3Dc.at<float>(2) < 0.0f) return false;
  const float invz = invz_double ? (float)(1.0 / p3Dc.at<float>(2)) : 1 / p3Dc.at<float>(2);  // :340 / :1000
  const float x = p3Dc.at<float>(0) * invz;
  const float y = p3Dc.at<float>(1) * invz;
  u = pKF->fx * x + pKF->cx;
  v = pKF->fy * y + pKF->cy;
  if (!pKF->IsInImage(u, v)) return false;
  const float maxDistance = pMP->GetMaxDistanceInvariance();
  const float minDistance = pMP->GetMinDistanceInvariance();
  cv::Mat PO = p3Dw - Ow;
  const float dist3D = cv::norm(PO);
  if (dist3D < minDistance || dist3D > maxDistance) return false;
  cv::Mat Pn = pMP->GetNormal();
  if (PO.dot(Pn) < 0.5 * dist3D) return false;  // viewing angle below 60 degrees
  level = pMP->PredictScale(dist3D, pKF);
  return true;
}
}  // namespace

// SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th)   reference :300-413   (loop closing)
int ORBmatcher::SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const vector<MapPoint*>& vpPoints,
                                   vector<MapPoint*>& vpMatched, int th) {
  cv::Mat Rcw, tcw, Ow;
  decompose_sim3(Scw, Rcw, tcw, Ow);
  set<MapPoint*> spAlreadyFound(vpMatched.begin(), vpMatched.end());
  spAlreadyFound.erase(static_cast<MapPoint*>(NULL));
  QueryPack P;
  for (int iMP = 0, iend = (int)vpPoints.size(); iMP < iend; iMP++) {
    MapPoint* pMP = vpPoints[iMP];
    if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
    float u, v;
    int lvl;
    if (!project_into_keyframe(pKF, pMP, Rcw, tcw, Ow, false, u, v, lvl)) continue;
    P.add(iMP, u, v, th * pKF->mvScaleFactors[lvl], lvl - 1, lvl, true, 0.f, pMP->GetDescriptor());
  }
  const int N = pKF->N;
  vector<uint8_t> blocked((size_t)N + 1, 0);
  for (int j = 0; j < N && j < (int)vpMatched.size(); j++) blocked[j] = vpMatched[j] ? 1 : 0;  // :385
  const asdm_proj_params pp = proj_params(ASDM_PROJ_BEST_ONLY, TH_LOW, 256.f, mfNNratio, false, false);
  vector<int32_t> assign((size_t)N + 1, -1);
  int32_t nm = 0;
  asdm::Check(asdm_search_by_projection(asdm::Context(), dev(pKF).set, P.q.data(), P.n(), P.desc.p, NULL, NULL,
                                        blocked.data(), &pp, assign.data(), NULL, &nm),
              "asdm_search_by_projection");
  for (int j = 0; j < N; j++)
    if (assign[j] >= 0) vpMatched[j] = vpPoints[P.src[assign[j]]];  // :406
  return nm;
}

// ------------------------------------------------------------------------------------------------------------------
// Fuse(KeyFrame*, vpMapPoints, th)   reference :825-960.  Two phases: the match phase (window, octave and chi-square
// gates, best descriptor) has no cross-query state and runs on the GPU for every candidate; the graph mutations are
// then applied on the host in the reference's order, re-evaluating at apply time the two gates a mutation can change
// (isBad after Replace, IsInKeyFrame after AddObservation).
int ORBmatcher::Fuse(KeyFrame* pKF, const vector<MapPoint*>& vpMapPoints, const float th) {
  cv::Mat Rcw = pKF->GetRotation();
  cv::Mat tcw = pKF->GetTranslation();
  cv::Mat Ow = pKF->GetCameraCenter();
  QueryPack P;
  const int nMPs = (int)vpMapPoints.size();
  for (int i = 0; i < nMPs; i++) {
    MapPoint* pMP = vpMapPoints[i];
    if (!pMP) continue;
    if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
    float u, v;
    int lvl;
    if (!project_into_keyframe(pKF, pMP, Rcw, tcw, Ow, false, u, v, lvl)) continue;
    P.add(i, u, v, th * pKF->mvScaleFactors[lvl], lvl - 1, lvl, false, 0.f, pMP->GetDescriptor());
  }
  asdm_proj_params pp = proj_params(ASDM_PROJ_BEST_ONLY, TH_LOW, 256.f, mfNNratio, false, false);
  pp.chi2_max = 5.99;  // :919-926
  pp.n_levels = (int)std::min<size_t>(pKF->mvInvLevelSigma2.size(), 32);
  for (int l = 0; l < pp.n_levels; l++) pp.inv_level_sigma2[l] = pKF->mvInvLevelSigma2[l];
  vector<int32_t> best((size_t)P.n() + 1, -1);
  int32_t nm = 0;
  asdm::Check(asdm_search_by_projection_q(asdm::Context(), dev(pKF).set, P.q.data(), P.n(), P.desc.p, NULL, NULL,
                                          NULL, &pp, best.data(), NULL, &nm),
              "asdm_search_by_projection_q");
  int nFused = 0;
  for (int k = 0; k < P.n(); k++) {
    if (best[k] < 0) continue;
    MapPoint* pMP = vpMapPoints[P.src[k]];
    if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;  // changed by an earlier mutation of this call
    const int bestIdx = best[k];
    MapPoint* pMPinKF = pKF->GetMapPoint(bestIdx);
    if (pMPinKF) {
      if (!pMPinKF->isBad()) {
        if (pMPinKF->Observations() > pMP->Observations()) pMP->Replace(pMPinKF);
        else pMPinKF->Replace(pMP);
      }
    } else {
      pMP->AddObservation(pKF, bestIdx);
      pKF->AddMapPoint(pMP, bestIdx);
    }
    nFused++;
  }
  return nFused;
}

// Fuse(KeyFrame*, Scw, vpPoints, th, vpReplacePoint)   reference :963-1086
int ORBmatcher::Fuse(KeyFrame* pKF, cv::Mat Scw, const vector<MapPoint*>& vpPoints, float th,
                     vector<MapPoint*>& vpReplacePoint) {
  cv::Mat Rcw, tcw, Ow;
  decompose_sim3(Scw, Rcw, tcw, Ow);
  const set<MapPoint*> spAlreadyFound = pKF->GetMapPoints();
  QueryPack P;
  const int nPoints = (int)vpPoints.size();
  for (int iMP = 0; iMP < nPoints; iMP++) {
    MapPoint* pMP = vpPoints[iMP];
    if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
    float u, v;
    int lvl;
    if (!project_into_keyframe(pKF, pMP, Rcw, tcw, Ow, true, u, v, lvl)) continue;
    P.add(iMP, u, v, th * pKF->mvScaleFactors[lvl], lvl - 1, lvl, false, 0.f, pMP->GetDescriptor());
  }
  const asdm_proj_params pp = proj_params(ASDM_PROJ_BEST_ONLY, TH_LOW, 100.f, mfNNratio, false, false);
  vector<int32_t> best((size_t)P.n() + 1, -1);
  int32_t nm = 0;
  asdm::Check(asdm_search_by_projection_q(asdm::Context(), dev(pKF).set, P.q.data(), P.n(), P.desc.p, NULL, NULL,
                                          NULL, &pp, best.data(), NULL, &nm),
              "asdm_search_by_projection_q");
  int nFused = 0;
  for (int k = 0; k < P.n(); k++) {
    if (best[k] < 0) continue;
    const int iMP = P.src[k];
    MapPoint* pMP = vpPoints[iMP];
    // spAlreadyFound is a snapshot taken before the loop (:979) and isBad() is not changed by this overload
    MapPoint* pMPinKF = pKF->GetMapPoint(best[k]);
    if (pMPinKF) {
      if (!pMPinKF->isBad()) vpReplacePoint[iMP] = pMPinKF;
    } else {
      pMP->AddObservation(pKF, best[k]);
      pKF->AddMapPoint(pMP, best[k]);
    }
    nFused++;
  }
  return nFused;
}

// SearchBySim3   reference :1090-1314: two independent projection searches (1 -> 2 and 2 -> 1) and a mutual check
int ORBmatcher::SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, vector<MapPoint*>& vpMatches12, const float& s12,
                             const cv::Mat& R12, const cv::Mat& t12, const float th) {
  cv::Mat R1w = pKF1->GetRotation();
  cv::Mat t1w = pKF1->GetTranslation();
  cv::Mat R2w = pKF2->GetRotation();
  cv::Mat t2w = pKF2->GetTranslation();
  cv::Mat sR12 = s12 * R12;
  cv::Mat sR21 = (1.0 / s12) * R12.t();
  cv::Mat t21 = -sR21 * t12;
  const vector<MapPoint*> vpMapPoints1 = pKF1->GetMapPointMatches();
  const int N1 = (int)vpMapPoints1.size();
  const vector<MapPoint*> vpMapPoints2 = pKF2->GetMapPointMatches();
  const int N2 = (int)vpMapPoints2.size();
  vector<bool> vbAlreadyMatched1(N1, false), vbAlreadyMatched2(N2, false);
  for (int i = 0; i < N1; i++) {
    MapPoint* pMP = vpMatches12[i];
    if (pMP) {
      vbAlreadyMatched1[i] = true;
      int idx2 = pMP->GetIndexInKeyFrame(pKF2);
      if (idx2 >= 0 && idx2 < N2) vbAlreadyMatched2[idx2] = true;
    }
  }
  // one direction: points of `from` (camera frame Rw, tw) mapped by (sR, t) into `to`
  struct Dir {
    static void run(ORBmatcher* self, KeyFrame* to, const vector<MapPoint*>& pts, const vector<bool>& already,
                    const cv::Mat& Rw, const cv::Mat& tw, const cv::Mat& sR, const cv::Mat& t, float th,
                    vector<int>& vnMatch) {
      QueryPack P;
      for (int i = 0; i < (int)pts.size(); i++) {
        MapPoint* pMP = pts[i];
        if (!pMP || already[i]) continue;
        if (pMP->isBad()) continue;
        cv::Mat p3Dw = pMP->GetWorldPos();
        cv::Mat p3Dc_from = Rw * p3Dw + tw;
        cv::Mat p3Dc = sR * p3Dc_from + t;
        if (p3Dc.at<float>(2) < 0.0) continue;
        const float invz = 1.0 / p3Dc.at<float>(2);
        const float x = p3Dc.at<float>(0) * invz;
        const float y = p3Dc.at<float>(1) * invz;
        const float u = to->fx * x + to->cx;  // both keyframes share the calibration (:1094-1097 reads pKF1's)
        const float v = to->fy * y + to->cy;
        if (!to->IsInImage(u, v)) continue;
        const float maxDistance = pMP->GetMaxDistanceInvariance();
        const float minDistance = pMP->GetMinDistanceInvariance();
        const float dist3D = cv::norm(p3Dc);
        if (dist3D < minDistance || dist3D > maxDistance) continue;
        const int lvl = pMP->PredictScale(dist3D, to);
        P.add(i, u, v, th * to->mvScaleFactors[lvl], lvl - 1, lvl, false, 0.f, pMP->GetDescriptor());
      }
      const asdm_proj_params pp = proj_params(ASDM_PROJ_BEST_ONLY, TH_HIGH, 100.f, self->mfNNratio, false, false);
      vector<int32_t> best((size_t)P.n() + 1, -1);
      int32_t nm = 0;
      asdm::Check(asdm_search_by_projection_q(asdm::Context(), dev(to).set, P.q.data(), P.n(), P.desc.p, NULL, NULL,
                                              NULL, &pp, best.data(), NULL, &nm),
                  "asdm_search_by_projection_q");
      for (int k = 0; k < P.n(); k++)
        if (best[k] >= 0) vnMatch[P.src[k]] = best[k];
    }
  };
  vector<int> vnMatch1(N1, -1), vnMatch2(N2, -1);
  Dir::run(this, pKF2, vpMapPoints1, vbAlreadyMatched1, R1w, t1w, sR21, t21, th, vnMatch1);  // :1136-1212
  Dir::run(this, pKF1, vpMapPoints2, vbAlreadyMatched2, R2w, t2w, sR12, t12, th, vnMatch2);  // :1216-1292
  int nFound = 0;
  for (int i1 = 0; i1 < N1; i1++) {  // :1296-1311
    const int idx2 = vnMatch1[i1];
    if (idx2 >= 0 && vnMatch2[idx2] == i1) {
      vpMatches12[i1] = vpMapPoints2[idx2];
      nFound++;
    }
  }
  return nFound;
}

}  // namespace ORB_SLAM2
