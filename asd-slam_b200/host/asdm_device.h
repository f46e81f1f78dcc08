// asdm_device.h -- host-side glue between the ORB-SLAM2-derived classes of ASD-SLAM and libasdmatch_cuda.so.
//
// The reference keeps descriptors in cv::Mat members (Frame.h:147, KeyFrame.h:199).  The integration adds ONE member
// to Frame and to KeyFrame:
//     mutable std::shared_ptr<asdm::DeviceSet> mDev;
// filled on first use (or eagerly at construction) and shared by Frame copies / the KeyFrame built from a Frame
// (Frame.cc:41-61, KeyFrame.cc:146-174 copy the shared_ptr instead of re-uploading).
//
// Header only; needs only <opencv2/core> types (cv::Mat, cv::KeyPoint) and include/asdmatch.h.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <map>
#include <memory>
#include <vector>

#include "../../include/asdmatch.h"

namespace asdm {

struct DeviceSet {
  asdm_set* set = nullptr;
  bool has_fv = false;
  ~DeviceSet() {
    if (set) asdm_set_release(set);
  }
};

[[noreturn]] inline void Die(const char* what, int rc, asdm_ctx* ctx) {
  // the reference matcher has no error path (int return = match count); a failing GPU call is fatal, never a
  // silent CPU fallback
  std::fprintf(stderr, "asdmatch: %s failed (%d): %s\n", what, rc, ctx ? asdm_last_error(ctx) : "no context");
  std::abort();
}

// One context per process (the fork runs tracking, mapping and loop closing on one thread; the library serialises
// calls on a context anyway).  ASDM_DEVICE selects the GPU.
inline asdm_ctx* Context() {
  static asdm_ctx* ctx = [] {
    const char* e = std::getenv("ASDM_DEVICE");
    asdm_ctx* c = nullptr;
    const int rc = asdm_create(e ? std::atoi(e) : 0, &c);
    if (rc != ASDM_OK) Die("asdm_create (a B200 is required; there is no CPU fallback)", rc, nullptr);
    return c;
  }();
  return ctx;
}

inline void Check(int rc, const char* what) {
  if (rc != ASDM_OK) Die(what, rc, Context());
}

// cv::KeyPoint -> the four fields the matcher reads
template <class KP>
inline void PackKeypoints(const std::vector<KP>& in, std::vector<asdm_kp>& out) {
  out.resize(in.size());
  for (size_t i = 0; i < in.size(); i++) {
    out[i].x = in[i].pt.x;
    out[i].y = in[i].pt.y;
    out[i].angle = in[i].angle;
    out[i].octave = in[i].octave;
  }
}

// Uploads descriptors (N x 128 CV_32F, any row stride), undistorted keypoints and the grid geometry once.
template <class MAT, class KP>
inline std::shared_ptr<DeviceSet> Upload(const MAT& desc, const std::vector<KP>& keys_un, const asdm_grid* grid) {
  std::vector<asdm_kp> kps;
  PackKeypoints(keys_un, kps);
  const int n = (int)keys_un.size();
  std::shared_ptr<DeviceSet> d(new DeviceSet());
  const float* base = n > 0 ? desc.template ptr<float>(0) : nullptr;
  const int stride = n > 1 ? (int)((const char*)desc.template ptr<float>(1) - (const char*)base) : 0;
  Check(asdm_set_create(Context(), base, n, stride, kps.data(), grid, &d->set), "asdm_set_create");
  return d;
}

// DBoW2::FeatureVector (std::map<NodeId, std::vector<unsigned>>) -> CSR, attached once per set
template <class FV>
inline void AttachFeatureVector(DeviceSet& d, const FV& fv) {
  if (d.has_fv) return;
  std::vector<uint32_t> ids;
  std::vector<int32_t> off(1, 0), idx;
  for (typename FV::const_iterator it = fv.begin(); it != fv.end(); ++it) {
    ids.push_back((uint32_t)it->first);
    for (size_t k = 0; k < it->second.size(); k++) idx.push_back((int32_t)it->second[k]);
    off.push_back((int32_t)idx.size());
  }
  asdm_featvec c;
  c.n_nodes = (int32_t)ids.size();
  c.node_ids = ids.data();
  c.off = off.data();
  c.idx = idx.data();
  Check(asdm_set_featvec(d.set, &c), "asdm_set_featvec");
  d.has_fv = true;
}

}  // namespace asdm
