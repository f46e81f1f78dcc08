// ref_map_driver.cc -- the reference's OWN descriptor (de)serialisation loops for `chamo.map`, extracted by line range
// at build time (oracle/Makefile -> _ref/gen/*.inc; no reference source is committed) and run over a minimal
// column-major matrix (Eigen::Matrix<float, Dynamic, Dynamic> is column-major by default; Eigen itself is not in this
// image, and visual_map.h needs OpenCV + glog besides).  TEST INFRASTRUCTURE ONLY.
//   vm_io.inc      visual_map_seri.cc:5-37    putToFile / getFromFile helpers
//   vm_write.inc   visual_map_seri.cc:134-146 descriptor block writer of vm::save_visual_map
//   vm_read.inc    visual_map_seri.cc:272-285 descriptor block reader of vm::loader_visual_map
//   sys_fill.inc   System.cc:333-343          System::saveToVisualMap: mDescriptors (N x 128) -> frame_p->descriptors (D x N)
#include <cstdint>
#include <cstring>
#include <fstream>
#include <string>
#include <vector>

#include "cv_shim.h"

namespace {

struct ColMajorF {  // the slice of Eigen::MatrixXf these loops use
  int r = 0, c = 0;
  std::vector<float> d;
  int rows() const { return r; }
  int cols() const { return c; }
  void resize(int rr, int cc) { r = rr; c = cc; d.assign((size_t)rr * cc, 0.f); }
  float& operator()(int i, int j) { return d[(size_t)i + (size_t)j * r]; }
};
struct FrameLike { ColMajorF descriptors; };
struct KeyFrameLike { cv::Mat mDescriptors; };

}  // namespace

namespace vm {
#include "_ref/gen/vm_io.inc"
}
using namespace vm;

extern "C" {

// System::saveToVisualMap + vm::save_visual_map for the descriptors of one keyframe: returns the bytes written
long ref_map_write_block(const float* desc, int n, const char* path) {
  KeyFrameLike kf;
  kf.mDescriptors = cv::Mat(n, 128, CV_32F);
  if (n) memcpy(kf.mDescriptors.ptr<float>(0), desc, (size_t)n * 128 * sizeof(float));
  KeyFrameLike* pKF = &kf;
  FrameLike fr;
  FrameLike* frame_p = &fr;
  {
#include "_ref/gen/sys_fill.inc"
  }
  std::fstream output(path, std::ios::out | std::ios::trunc | std::ios::binary);
  {
#include "_ref/gen/vm_write.inc"
  }
  const long pos = (long)output.tellp();
  output.close();
  return pos;
}

// vm::loader_visual_map for one block; out_rows receives descriptor j at out_rows[j*128..] (= descriptors(:, j), what
// KeyFrame::setData should be handed as a CV_32F N x 128 Mat); returns N, dims[0] = desc_width
int ref_map_read_block(const char* path, float* out_rows, int cap_rows, int32_t* dims) {
  std::fstream input(path, std::ios::in | std::ios::binary);
  FrameLike fr;
  FrameLike* frame_p = &fr;
  {
#include "_ref/gen/vm_read.inc"
  }
  dims[0] = fr.descriptors.rows();
  dims[1] = fr.descriptors.cols();
  for (int j = 0; j < fr.descriptors.cols() && j < cap_rows; j++)
    for (int k = 0; k < fr.descriptors.rows(); k++) out_rows[(size_t)j * fr.descriptors.rows() + k] = fr.descriptors(k, j);
  return fr.descriptors.cols();
}

}  // extern "C"
