// cv_shim.h -- the sliver of OpenCV that the reference's ORBmatcher.cc touches, written from scratch so that the
// UNMODIFIED reference source (/root/reference/src/vslam/src/ORBmatcher.cc) compiles here without OpenCV 3.2
// (absent from this image; see oracle/Makefile `ref`).  TEST INFRASTRUCTURE ONLY -- never part of the product.
//
// Only dense 2-D CV_32F / CV_64F matrices exist.  Sub-matrix accessors (row, col, rowRange, colRange) return
// copies: the matcher only ever reads through them.  Matrix products accumulate in double and round once to float,
// which is what OpenCV 3.2's GEMMSingleMul<float,double> does for the tiny (3x3, 3x1) CV_32F operands on this path.
#pragma once
#include <cassert>
#include <cmath>
#include <limits>
#include <memory>
#include <cstring>
#include <vector>

#define CV_32F 5
#define CV_64F 6
#define CV_32FC1 CV_32F
#define CV_64FC1 CV_64F

namespace cv {

template <typename T>
struct Point_ {
  T x, y;
  Point_() : x(0), y(0) {}
  Point_(T x_, T y_) : x(x_), y(y_) {}
};
typedef Point_<float> Point2f;
typedef Point_<int> Point2i;
typedef Point2i Point;

struct KeyPoint {
  Point2f pt;
  float size = 0, angle = -1, response = 0;
  int octave = 0, class_id = -1;
};

#ifdef ASD_CV_VIEWS
// ---- header semantics, as cv::Mat has them (OpenCV 3.2 core/mat.hpp): copies, row(), col(), rowRange(), colRange() and
// reshape() are HEADERS over shared, reference-counted data; only clone(), create() and convertTo() allocate, and
// convertTo allocates its destination once. oracle/Makefile builds the reference a second time with -DASD_CV_VIEWS
// (oracle/_ref/libref_orbmatcher_views.so): the goldens must come out bit-identical, and bench.py reports the reference's
// speed in both modes so the share of this stand-in in the reference's cost is known.
class Mat {
 public:
  int rows = 0, cols = 0;
  Mat() {}
  Mat(int r, int c, int type) { create(r, c, type); }
  void create(int r, int c, int type) {
    rows = r; cols = c; type_ = type; step_ = c; off_ = 0;
    buf_ = std::make_shared<Buf>();
    buf_->f.assign(type == CV_32F ? (size_t)r * c : 0, 0.f);
    buf_->d.assign(type == CV_64F ? (size_t)r * c : 0, 0.0);
  }
  int type() const { return type_; }
  bool empty() const { return rows == 0 || cols == 0; }
  void release() { create(0, 0, type_); }
  static Mat zeros(int r, int c, int type) { return Mat(r, c, type); }
  Mat clone() const {
    Mat o(rows, cols, type_);
    for (int i = 0; i < rows; i++)
      for (int j = 0; j < cols; j++) o.set(i, j, get(i, j));
    return o;
  }
  template <typename T> T& at(int i, int j) { return data<T>()[(size_t)i * step_ + j]; }
  template <typename T> const T& at(int i, int j) const { return data<T>()[(size_t)i * step_ + j]; }
  // single-index access of a vector (n x 1 or 1 x n)
  template <typename T> T& at(int i) { return cols == 1 ? data<T>()[(size_t)i * step_] : data<T>()[i]; }
  template <typename T> const T& at(int i) const { return cols == 1 ? data<T>()[(size_t)i * step_] : data<T>()[i]; }
  template <typename T> T* ptr(int i = 0) { return data<T>() + (size_t)i * step_; }
  template <typename T> const T* ptr(int i = 0) const { return data<T>() + (size_t)i * step_; }

  Mat rowRange(int a, int b) const { return block(a, b, 0, cols); }
  Mat colRange(int a, int b) const { return block(0, rows, a, b); }
  Mat row(int i) const { return block(i, i + 1, 0, cols); }
  Mat col(int j) const { return block(0, rows, j, j + 1); }
  Mat t() const {
    Mat o(cols, rows, type_);
    for (int i = 0; i < rows; i++)
      for (int j = 0; j < cols; j++) o.set(j, i, get(i, j));
    return o;
  }
  Mat reshape(int cn, int r) const {  // a header when the data is continuous (all the matcher ever reshapes is a row)
    assert(cn == 1 && r > 0);
    (void)cn;
    if (!(rows == 1 || step_ == cols)) return clone().reshape(cn, r);
    Mat o = *this;
    o.rows = r;
    o.cols = rows * cols / r;
    o.step_ = o.cols;
    return o;
  }
  void convertTo(Mat& dst, int rtype) const {
    const Mat src = *this;  // dst may alias this
    dst.create(src.rows, src.cols, rtype);
    for (int i = 0; i < src.rows; i++)
      for (int j = 0; j < src.cols; j++) dst.set(i, j, src.get(i, j));
  }
  template <typename T> operator std::vector<T>() const {
    std::vector<T> v((size_t)rows * cols);
    for (int i = 0; i < rows; i++)
      for (int j = 0; j < cols; j++) v[(size_t)i * cols + j] = (T)get(i, j);
    return v;
  }
  double dot(const Mat& o) const {
    double s = 0;
    for (int i = 0; i < rows; i++)
      for (int j = 0; j < cols; j++) s += get(i, j) * o.get(i, j);
    return s;
  }
  double get(int i, int j) const {
    const size_t at = off_ + (size_t)i * step_ + j;
    return type_ == CV_32F ? (double)buf_->f[at] : buf_->d[at];
  }
  void set(int i, int j, double v) {
    const size_t at = off_ + (size_t)i * step_ + j;
    if (type_ == CV_32F) buf_->f[at] = (float)v;
    else buf_->d[at] = v;
  }

 private:
  struct Buf {
    std::vector<float> f;
    std::vector<double> d;
  };
  template <typename T> T* data();
  template <typename T> const T* data() const;
  Mat block(int r0, int r1, int c0, int c1) const {
    Mat o = *this;
    o.rows = r1 - r0;
    o.cols = c1 - c0;
    o.off_ = off_ + (size_t)r0 * step_ + c0;
    return o;
  }
  int type_ = CV_32F;
  int step_ = 0;     // elements between consecutive rows
  size_t off_ = 0;   // first element inside the shared buffer
  std::shared_ptr<Buf> buf_;
};
template <> inline float* Mat::data<float>() { assert(type_ == CV_32F); return buf_->f.data() + off_; }
template <> inline const float* Mat::data<float>() const { assert(type_ == CV_32F); return buf_->f.data() + off_; }
template <> inline double* Mat::data<double>() { assert(type_ == CV_64F); return buf_->d.data() + off_; }
template <> inline const double* Mat::data<double>() const { assert(type_ == CV_64F); return buf_->d.data() + off_; }
#else
class Mat {
 public:
  int rows = 0, cols = 0;
  Mat() {}
  Mat(int r, int c, int type) { create(r, c, type); }
  void create(int r, int c, int type) {
    rows = r; cols = c; type_ = type;
    f_.assign(type == CV_32F ? (size_t)r * c : 0, 0.f);
    d_.assign(type == CV_64F ? (size_t)r * c : 0, 0.0);
  }
  int type() const { return type_; }
  bool empty() const { return rows == 0 || cols == 0; }
  void release() { create(0, 0, type_); }
  static Mat zeros(int r, int c, int type) { return Mat(r, c, type); }  // create() zero-fills
  Mat clone() const { return *this; }
  template <typename T> T& at(int i, int j) { return data<T>()[(size_t)i * cols + j]; }
  template <typename T> const T& at(int i, int j) const { return data<T>()[(size_t)i * cols + j]; }
  // single-index access of a vector (n x 1 or 1 x n)
  template <typename T> T& at(int i) { return data<T>()[i]; }
  template <typename T> const T& at(int i) const { return data<T>()[i]; }
  template <typename T> T* ptr(int i = 0) { return data<T>() + (size_t)i * cols; }
  template <typename T> const T* ptr(int i = 0) const { return data<T>() + (size_t)i * cols; }

  Mat rowRange(int a, int b) const { return block(a, b, 0, cols); }
  Mat colRange(int a, int b) const { return block(0, rows, a, b); }
  Mat row(int i) const { return block(i, i + 1, 0, cols); }
  Mat col(int j) const { return block(0, rows, j, j + 1); }
  Mat t() const {
    Mat o(cols, rows, type_);
    for (int i = 0; i < rows; i++)
      for (int j = 0; j < cols; j++) o.set(j, i, get(i, j));
    return o;
  }
  Mat reshape(int cn, int r) const {
    assert(cn == 1 && r > 0);
    (void)cn;
    Mat o = *this;
    o.rows = r;
    o.cols = rows * cols / r;
    return o;
  }
  // cv::Mat::convertTo without scaling: saturate_cast between float and double is a plain C++ conversion
  void convertTo(Mat& dst, int rtype) const {
    Mat o(rows, cols, rtype);
    for (int i = 0; i < rows; i++)
      for (int j = 0; j < cols; j++) o.set(i, j, get(i, j));
    dst = o;
  }
  // Mat -> std::vector<T> (cv::Mat::operator std::vector<_Tp>: copyTo with conversion to _Tp)
  template <typename T> operator std::vector<T>() const {
    std::vector<T> v((size_t)rows * cols);
    for (int i = 0; i < rows; i++)
      for (int j = 0; j < cols; j++) v[(size_t)i * cols + j] = (T)get(i, j);
    return v;
  }
  double dot(const Mat& o) const {
    double s = 0;
    for (int i = 0; i < rows; i++)
      for (int j = 0; j < cols; j++) s += get(i, j) * o.get(i, j);
    return s;
  }
  double get(int i, int j) const {
    return type_ == CV_32F ? (double)f_[(size_t)i * cols + j] : d_[(size_t)i * cols + j];
  }
  void set(int i, int j, double v) {
    if (type_ == CV_32F) f_[(size_t)i * cols + j] = (float)v;
    else d_[(size_t)i * cols + j] = v;
  }

 private:
  template <typename T> T* data();
  template <typename T> const T* data() const;
  Mat block(int r0, int r1, int c0, int c1) const {
    Mat o(r1 - r0, c1 - c0, type_);
    for (int i = r0; i < r1; i++)
      for (int j = c0; j < c1; j++) o.set(i - r0, j - c0, get(i, j));
    return o;
  }
  int type_ = CV_32F;
  std::vector<float> f_;
  std::vector<double> d_;
};
template <> inline float* Mat::data<float>() { assert(type_ == CV_32F); return f_.data(); }
template <> inline const float* Mat::data<float>() const { assert(type_ == CV_32F); return f_.data(); }
template <> inline double* Mat::data<double>() { assert(type_ == CV_64F); return d_.data(); }
template <> inline const double* Mat::data<double>() const { assert(type_ == CV_64F); return d_.data(); }

#endif  // ASD_CV_VIEWS

// element-wise results are rounded to the matrix type once per element, as OpenCV's float kernels do
inline Mat operator*(const Mat& a, const Mat& b) {
  assert(a.cols == b.rows);
  Mat o(a.rows, b.cols, a.type());
  for (int i = 0; i < a.rows; i++)
    for (int j = 0; j < b.cols; j++) {
      double s = 0;
      for (int k = 0; k < a.cols; k++) s += a.get(i, k) * b.get(k, j);
      o.set(i, j, s);
    }
  return o;
}
inline Mat ew(const Mat& a, const Mat& b, int op) {
  assert(a.rows == b.rows && a.cols == b.cols);
  Mat o(a.rows, a.cols, a.type());
  for (int i = 0; i < a.rows; i++)
    for (int j = 0; j < a.cols; j++) {
      if (a.type() == CV_32F) {
        const float x = (float)a.get(i, j), y = (float)b.get(i, j);
        o.set(i, j, op ? x - y : x + y);
      } else {
        o.set(i, j, op ? a.get(i, j) - b.get(i, j) : a.get(i, j) + b.get(i, j));
      }
    }
  return o;
}
inline Mat operator+(const Mat& a, const Mat& b) { return ew(a, b, 0); }
inline Mat operator-(const Mat& a, const Mat& b) { return ew(a, b, 1); }
inline Mat scale(const Mat& a, double s) {
  Mat o(a.rows, a.cols, a.type());
  for (int i = 0; i < a.rows; i++)
    for (int j = 0; j < a.cols; j++) o.set(i, j, a.get(i, j) * s);
  return o;
}
inline Mat operator-(const Mat& a) { return scale(a, -1.0); }
inline Mat operator*(double s, const Mat& a) { return scale(a, s); }
inline Mat operator*(const Mat& a, double s) { return scale(a, s); }
inline Mat operator/(const Mat& a, double s) { return scale(a, 1.0 / s); }
inline double norm(const Mat& a) { return std::sqrt(a.dot(a)); }

}  // namespace cv

#include "cv_fs_shim.h"
