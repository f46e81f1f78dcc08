// Include shim for the reference's KeyFrameDatabase.cc (compiled unmodified from where it lies): that file does
// `#include "KeyFrameDatabase.h"`; with -I oracle/ref_shim/kfdb first it lands here.  The include guards of the
// reference's KeyFrame.h / Frame.h / ORBVocabulary.h (libtorch, ROS, g2o, OpenCV behind them) are pre-defined so they
// expand to nothing; below are a KeyFrame / Frame / ORBVocabulary carrying exactly the members KeyFrameDatabase.cc reads
// (names as in KeyFrame.h:150-200, Frame.h), the vocabulary's score() forwarding to the reference's own
// DBoW2::L1Scoring (ScoringObject.cpp, compiled from the reference).  TEST INFRASTRUCTURE ONLY.
#pragma once
#define KEYFRAME_H
#define FRAME_H
#define ORBVOCABULARY_H
#include <list>
#include <mutex>
#include <set>
#include <vector>

#include "BowVector.h"      // the reference's src/dbow2/include/BowVector.h
#include "ScoringObject.h"  // the reference's src/dbow2/include/ScoringObject.h

using namespace std;  // the reference's headers rely on a using-directive leaked by its own Frame.h / KeyFrame.h

namespace ORB_SLAM2 {

class KeyFrame {
 public:
  long unsigned int mnId = 0;
  DBoW2::BowVector mBowVec;
  long unsigned int mnLoopQuery = 0;  // KeyFrame.cc:150
  int mnLoopWords = 0;
  float mLoopScore = 0;               // uninitialised in the reference
  long unsigned int mnRelocQuery = (long unsigned int)-1;
  int mnRelocWords = 0;
  float mRelocScore = 0;              // uninitialised in the reference
  std::set<KeyFrame*> GetConnectedKeyFrames() { return connected; }
  std::vector<KeyFrame*> GetBestCovisibilityKeyFrames(const int& N) {
    if ((int)ordered.size() < N) return ordered;
    return std::vector<KeyFrame*>(ordered.begin(), ordered.begin() + N);  // KeyFrame.cc GetBestCovisibilityKeyFrames
  }
  bool GetGlobalMapFlag() { return global; }
  std::set<KeyFrame*> connected;
  std::vector<KeyFrame*> ordered;
  bool global = false;
  int slot = -1;
};

class Frame {
 public:
  long unsigned int mnId = 0;
  DBoW2::BowVector mBowVec;
};

class ORBVocabulary {
 public:
  unsigned int size() const { return n_words; }
  double score(const DBoW2::BowVector& a, const DBoW2::BowVector& b) const { return l1.score(a, b); }
  DBoW2::L1Scoring l1;
  unsigned int n_words = 0;
};

}  // namespace ORB_SLAM2

#ifndef ASD_REF_KFDB_H
#error "pass -DASD_REF_KFDB_H='\"/root/reference/src/vslam/include/KeyFrameDatabase.h\"'"
#endif
#include ASD_REF_KFDB_H
