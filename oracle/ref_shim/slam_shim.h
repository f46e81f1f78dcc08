// slam_shim.h -- minimal Frame / KeyFrame / MapPoint carrying exactly the state the reference's ORBmatcher.cc reads
// (SURVEY.md appendix B), so that the UNMODIFIED reference ORBmatcher.cc links against it.  Member names and types
// mirror the reference declarations (Frame.h:109-186, KeyFrame.h:150-255, MapPoint.h:41-160) because the reference
// source refers to them by name.  The bodies of the grid / scale functions are NOT written here: oracle/Makefile
// extracts the reference's own definitions (Frame.cc, KeyFrame.cc, MapPoint.cc line ranges) into oracle/_ref/gen/
// and ref_driver.cc compiles them.  TEST INFRASTRUCTURE ONLY.
#pragma once
#include <cmath>
#include <map>
#include <memory>
#include <set>
#include <string>
#include <vector>

#include "cv_shim.h"
#include "FeatureVector.h"  // the reference's own src/dbow2/include/FeatureVector.h (std::map based, header only)

#define FRAME_GRID_ROWS 48
#define FRAME_GRID_COLS 64

namespace asdm { struct DeviceSet; }  // asd-slam_b200/host/asdm_device.h; the member the integration adds

using namespace std;  // the reference's headers rely on a using-directive leaked by its own Frame.h / KeyFrame.h

namespace ORB_SLAM2 {

class KeyFrame;
class Frame;

class MapPoint {
 public:
  // tracking scratch fields (MapPoint.h:95-100)
  float mTrackProjX = 0, mTrackProjY = 0, mTrackProjXR = 0;
  bool mbTrackInView = false;
  int mnTrackScaleLevel = 0;
  float mTrackViewCos = 0;

  bool isBad() { return mbBad; }
  int Observations() { return nObs; }
  cv::Mat GetDescriptor() { return mDescriptor.clone(); }
  cv::Mat GetWorldPos() { return mWorldPos.clone(); }
  cv::Mat GetNormal() { return mNormalVector.clone(); }
  void ComputeDistinctiveDescriptors();  // body: the reference's MapPoint.cc:271-338 (oracle/_ref/gen)
  float GetMinDistanceInvariance();
  float GetMaxDistanceInvariance();
  int PredictScale(const float& currentDist, KeyFrame* pKF);
  int PredictScale(const float& currentDist, Frame* pF);
  bool IsInKeyFrame(KeyFrame* pKF) { return mObservations.count(pKF); }
  int GetIndexInKeyFrame(KeyFrame* pKF) { return mObservations.count(pKF) ? (int)mObservations[pKF] : -1; }
  // Graph mutations of Fuse.  Replace follows what the reference's MapPoint::Replace does to the state the matcher
  // can observe afterwards (MapPoint.cc:205-245): this point turns bad, its observations move to pMP (or are erased
  // from keyframes that already see pMP).  Defined after KeyFrame below.
  void Replace(MapPoint* pMP);
  void AddObservation(KeyFrame* pKF, size_t idx) {
    if (mObservations.count(pKF)) return;
    mObservations[pKF] = idx;
    nObs++;
  }

  bool mbBad = false;
  int nObs = 0;
  cv::Mat mDescriptor, mWorldPos, mNormalVector;
  float mfMinDistance = 0, mfMaxDistance = 0;
  std::map<KeyFrame*, size_t> mObservations;
  std::vector<cv::Mat> mappointAllDescriptors;  // MapPoint.h: side buffer the fork fills in ComputeDistinctiveDescriptors
  MapPoint* mpReplaced = nullptr;
  int id = -1;
};

class Frame {
 public:
  bool PosInGrid(const cv::KeyPoint& kp, int& posX, int& posY);
  vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r, const int minLevel = -1,
                                   const int maxLevel = -1) const;
  void AssignFeaturesToGrid();

  static float fx, fy, cx, cy, invfx, invfy;
  float mbf = 0, mb = 0, mThDepth = 0;
  int N = 0;
  std::vector<cv::KeyPoint> mvKeys, mvKeysRight, mvKeysUn;
  DBoW2::FeatureVector mFeatVec;
  cv::Mat mDescriptors, mDescriptorsRight;
  std::vector<MapPoint*> mvpMapPoints;
  std::vector<bool> mvbOutlier;
  static float mfGridElementWidthInv, mfGridElementHeightInv;
  std::vector<std::size_t> mGrid[FRAME_GRID_COLS][FRAME_GRID_ROWS];
  cv::Mat mTcw;
  int mnScaleLevels = 0;
  float mfScaleFactor = 0, mfLogScaleFactor = 0;
  vector<float> mvScaleFactors, mvInvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
  static float mnMinX, mnMaxX, mnMinY, mnMaxY;
  mutable std::shared_ptr<asdm::DeviceSet> mDev;  // device-resident descriptors, shared by Frame copies
};

class KeyFrame {
 public:
  bool PosInGrid(const cv::KeyPoint& kp, int& posX, int& posY);
  void AssignFeaturesToGrid();
  std::vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r) const;
  bool IsInImage(const float& x, const float& y) const;
  bool isBad() { return mbBad; }
  bool mbBad = false;
  std::vector<MapPoint*> GetMapPointMatches() { return mvpMapPoints; }
  MapPoint* GetMapPoint(const size_t& idx) { return mvpMapPoints[idx]; }
  std::set<MapPoint*> GetMapPoints() {
    std::set<MapPoint*> s;
    for (MapPoint* p : mvpMapPoints)
      if (p && !p->isBad()) s.insert(p);
    return s;
  }
  void AddMapPoint(MapPoint* pMP, const size_t& idx) { mvpMapPoints[idx] = pMP; }
  cv::Mat GetRotation() { return Tcw.rowRange(0, 3).colRange(0, 3).clone(); }
  cv::Mat GetTranslation() { return Tcw.rowRange(0, 3).col(3).clone(); }
  cv::Mat GetCameraCenter() { return Ow.clone(); }

  int mnGridCols = FRAME_GRID_COLS, mnGridRows = FRAME_GRID_ROWS;
  float mfGridElementWidthInv = 0, mfGridElementHeightInv = 0;
  float fx = 0, fy = 0, cx = 0, cy = 0, invfx = 0, invfy = 0, mbf = 0, mb = 0, mThDepth = 0;
  int N = 0;
  std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
  cv::Mat mDescriptors;
  DBoW2::FeatureVector mFeatVec;
  int mnScaleLevels = 0;
  float mfScaleFactor = 0, mfLogScaleFactor = 0;
  std::vector<float> mvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
  int mnMinX = 0, mnMinY = 0, mnMaxX = 0, mnMaxY = 0;  // integer bounds (KeyFrame.h:217-220)
  std::vector<MapPoint*> mvpMapPoints;
  std::vector<std::vector<std::vector<size_t> > > mGrid;
  cv::Mat Tcw, Ow;
  mutable std::shared_ptr<asdm::DeviceSet> mDev;  // device-resident descriptors (handed over by the Frame)
};

inline void MapPoint::Replace(MapPoint* pMP) {
  if (pMP == this) return;
  std::map<KeyFrame*, size_t> obs = mObservations;
  mObservations.clear();
  mbBad = true;
  mpReplaced = pMP;
  for (std::map<KeyFrame*, size_t>::iterator it = obs.begin(); it != obs.end(); ++it) {
    KeyFrame* pKF = it->first;
    if (!pMP->IsInKeyFrame(pKF)) {
      pKF->mvpMapPoints[it->second] = pMP;
      pMP->AddObservation(pKF, it->second);
    } else {
      pKF->mvpMapPoints[it->second] = static_cast<MapPoint*>(NULL);
    }
  }
}

}  // namespace ORB_SLAM2
