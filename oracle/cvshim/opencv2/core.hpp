// oracle/cvshim/opencv2/core.hpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// A minimal stand-in for the slice of the OpenCV C++ API that the reference's
// src/ORBextractor.cc touches, so that file can be compiled UNMODIFIED from
// /root/reference into oracle/_ref/ (OpenCV C++ is not installed in this image).
// The five image primitives forward to oracle/cvprims.hpp, whose arithmetic is pinned
// against cv2 4.13.0.  Nothing here is shipped or linked into the product library.
#pragma once
#include <algorithm>
#include <cassert>
#include <climits>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <sstream>
#include <string>
#include <vector>

#include "../../cvprims.hpp"

typedef unsigned char uchar;
#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5
#define CV_PI 3.1415926535897932384626433832795

static inline int cvRound(double v) { return cvprims::cv_round(v); }
static inline int cvRound(float v) { return cvprims::cv_round(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { return cvprims::cv_floor(v); }
static inline int cvCeil(double v) { return cvprims::cv_ceil(v); }

namespace cv {

template <typename T>
struct Point_ {
  T x, y;
  Point_() : x(0), y(0) {}
  Point_(T _x, T _y) : x(_x), y(_y) {}
  template <typename U>
  Point_& operator*=(U s) {
    x = (T)(x * s);
    y = (T)(y * s);
    return *this;
  }
};
typedef Point_<int> Point2i;
typedef Point_<int> Point;
typedef Point_<float> Point2f;

struct Size {
  int width, height;
  Size() : width(0), height(0) {}
  Size(int w, int h) : width(w), height(h) {}
};
struct Rect {
  int x, y, width, height;
  Rect(int _x, int _y, int w, int h) : x(_x), y(_y), width(w), height(h) {}
};

struct KeyPoint {
  Point2f pt;
  float size, angle, response;
  int octave, class_id;
  KeyPoint() : size(0), angle(-1), response(0), octave(0), class_id(-1) {}
  KeyPoint(float x, float y, float _size, float _angle = -1, float _response = 0, int _octave = 0,
           int _class_id = -1)
      : pt(x, y), size(_size), angle(_angle), response(_response), octave(_octave), class_id(_class_id) {}
};

struct MatZeros {
  int rows, cols, type;
};

// Single-channel matrix, u8 (images, descriptors) or f32 (poses, points), with a malloc'd,
// reference-counted buffer (malloc on purpose: the extractor parity harness replaces operator new
// with a monotonic arena that is reset per call).
class Mat {
  struct Block {
    int refs;
  };
  Block* blk_;
  int type_;

  void retain() {
    if (blk_) ++blk_->refs;
  }
  void drop() {
    if (blk_ && --blk_->refs == 0) free(blk_);
    blk_ = nullptr;
  }

 public:
  uchar* data;
  int rows, cols;
  size_t step;

  static size_t esz(int type) { return type == CV_32F ? 4 : 1; }
  size_t elemSize() const { return esz(type_); }

  Mat() : blk_(nullptr), type_(CV_8U), data(nullptr), rows(0), cols(0), step(0) {}
  Mat(int r, int c, int type) : blk_(nullptr), type_(CV_8U), data(nullptr), rows(0), cols(0), step(0) { create(r, c, type); }
  Mat(Size s, int type) : blk_(nullptr), type_(CV_8U), data(nullptr), rows(0), cols(0), step(0) { create(s.height, s.width, type); }
  Mat(const MatZeros& z) : blk_(nullptr), type_(CV_8U), data(nullptr), rows(0), cols(0), step(0) { *this = z; }
  // wrap caller-owned memory
  Mat(int r, int c, int type, void* ext, size_t st = 0)
      : blk_(nullptr), type_(type), data((uchar*)ext), rows(r), cols(c), step(st ? st : (size_t)c * esz(type)) {}
  Mat(const Mat& o) : blk_(o.blk_), type_(o.type_), data(o.data), rows(o.rows), cols(o.cols), step(o.step) { retain(); }
  Mat& operator=(const Mat& o) {
    if (this != &o) {
      Block* nb = o.blk_;
      if (nb) ++nb->refs;
      drop();
      blk_ = nb;
      type_ = o.type_;
      data = o.data; rows = o.rows; cols = o.cols; step = o.step;
    }
    return *this;
  }
  Mat& operator=(const MatZeros& z) {
    create(z.rows, z.cols, z.type);
    for (int y = 0; y < rows; ++y) memset(data + (size_t)y * step, 0, (size_t)cols * elemSize());
    return *this;
  }
  ~Mat() { drop(); }

  void create(int r, int c, int type) {
    assert(type == CV_8U || type == CV_32F);
    if (data && r == rows && c == cols && type == type_) return;
    drop();
    type_ = type;
    rows = r; cols = c; step = (size_t)c * esz(type);
    size_t bytes = sizeof(Block) + 64 + (size_t)r * step;
    blk_ = (Block*)malloc(bytes);
    blk_->refs = 1;
    data = (uchar*)blk_ + 64;
  }
  void create(Size s, int type) { create(s.height, s.width, type); }
  void release() {
    drop();
    data = nullptr; rows = cols = 0; step = 0;
  }
  static MatZeros zeros(int r, int c, int type) { return MatZeros{r, c, type}; }
  static Mat ones(int r, int c, int type) {
    Mat m(r, c, type);
    for (int y = 0; y < r; ++y)
      for (int x = 0; x < c; ++x) {
        if (type == CV_32F) m.at<float>(y, x) = 1.f;
        else m.at<uchar>(y, x) = 1;
      }
    return m;
  }
  // u8 -> f32 (exact) or same-type copy; always a fresh buffer, so `view.convertTo(view, CV_32F)`
  // detaches the view from the image it pointed into, as OpenCV does when the type changes
  void convertTo(Mat& dst, int type) const {
    Mat out(rows, cols, type);
    for (int y = 0; y < rows; ++y)
      for (int x = 0; x < cols; ++x) {
        const float v = type_ == CV_32F ? at<float>(y, x) : (float)at<uchar>(y, x);
        if (type == CV_32F) out.at<float>(y, x) = v;
        else out.at<uchar>(y, x) = (uchar)v;
      }
    dst = out;
  }
  Mat reshape(int) const { return *this; }  // N x 2 x 1ch <-> N x 1 x 2ch share one layout (see undistortPoints)

  Mat operator()(const Rect& r) const {
    Mat m(*this);
    m.data = data + (size_t)r.y * step + (size_t)r.x * elemSize();
    m.rows = r.height; m.cols = r.width;
    return m;
  }
  Mat rowRange(int a, int b) const { return (*this)(Rect(0, a, cols, b - a)); }
  Mat colRange(int a, int b) const { return (*this)(Rect(a, 0, b - a, rows)); }
  Mat row(int y) const { return rowRange(y, y + 1); }
  Mat col(int x) const { return colRange(x, x + 1); }
  Mat clone() const {
    Mat m(rows, cols, type_);
    for (int y = 0; y < rows; ++y) memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, (size_t)cols * elemSize());
    return m;
  }
  void copyTo(Mat& dst) const { dst = clone(); }
  template <typename T> T& at(int y, int x) { return *(T*)(data + (size_t)y * step + x * sizeof(T)); }
  template <typename T> const T& at(int y, int x) const { return *(const T*)(data + (size_t)y * step + x * sizeof(T)); }
  // one index: element i of a row or column vector
  template <typename T> T& at(int i) { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
  template <typename T> const T& at(int i) const { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
  uchar* ptr(int y = 0) { return data + (size_t)y * step; }
  const uchar* ptr(int y = 0) const { return data + (size_t)y * step; }
  template <typename T> T* ptr(int y = 0) { return (T*)(data + (size_t)y * step); }
  template <typename T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step); }
  int type() const { return type_; }
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  size_t step1() const { return step / elemSize(); }
  Size size() const { return Size(cols, rows); }

  // ---- f32 algebra (poses and points in the matcher).  Plain float arithmetic evaluated eagerly;
  // NOT bit-pinned to OpenCV's gemm: both sides of every matcher comparison project with this code.
  Mat t() const {
    assert(type_ == CV_32F);
    Mat m(cols, rows, CV_32F);
    for (int y = 0; y < rows; ++y)
      for (int x = 0; x < cols; ++x) m.at<float>(x, y) = at<float>(y, x);
    return m;
  }
  double dot(const Mat& o) const {
    assert(type_ == CV_32F && o.type_ == CV_32F && rows * cols == o.rows * o.cols);
    double s = 0;
    const int n = rows * cols;
    for (int i = 0; i < n; ++i) s += (double)lin(i) * (double)o.lin(i);
    return s;
  }
  float lin(int i) const { return at<float>(i / cols, i % cols); }
};

inline Mat operator*(const Mat& a, const Mat& b) {
  assert(a.type() == CV_32F && b.type() == CV_32F && a.cols == b.rows);
  Mat m(a.rows, b.cols, CV_32F);
  for (int y = 0; y < a.rows; ++y)
    for (int x = 0; x < b.cols; ++x) {
      float s = 0.f;
      for (int k = 0; k < a.cols; ++k) s += a.at<float>(y, k) * b.at<float>(k, x);
      m.at<float>(y, x) = s;
    }
  return m;
}
template <typename F>
inline Mat mat_zip(const Mat& a, const Mat& b, F f) {
  assert(a.type() == CV_32F && b.type() == CV_32F && a.rows == b.rows && a.cols == b.cols);
  Mat m(a.rows, a.cols, CV_32F);
  for (int y = 0; y < a.rows; ++y)
    for (int x = 0; x < a.cols; ++x) m.at<float>(y, x) = f(a.at<float>(y, x), b.at<float>(y, x));
  return m;
}
template <typename F>
inline Mat mat_map(const Mat& a, F f) {
  assert(a.type() == CV_32F);
  Mat m(a.rows, a.cols, CV_32F);
  for (int y = 0; y < a.rows; ++y)
    for (int x = 0; x < a.cols; ++x) m.at<float>(y, x) = f(a.at<float>(y, x));
  return m;
}
inline Mat operator+(const Mat& a, const Mat& b) { return mat_zip(a, b, [](float p, float q) { return p + q; }); }
inline Mat operator-(const Mat& a, const Mat& b) { return mat_zip(a, b, [](float p, float q) { return p - q; }); }
inline Mat operator-(const Mat& a) { return mat_map(a, [](float p) { return -p; }); }
inline Mat operator*(double s, const Mat& a) { return mat_map(a, [s](float p) { return (float)(p * s); }); }
inline Mat operator*(const Mat& a, double s) { return s * a; }
inline Mat operator/(const Mat& a, double s) { return mat_map(a, [s](float p) { return (float)(p / s); }); }
inline double norm(const Mat& a) { return std::sqrt(a.dot(a)); }
enum { NORM_L1 = 2 };
// cv::norm(a, b, NORM_L1) on f32: OpenCV accumulates |a - b| in double; the stereo refinement feeds it
// integer-valued patches (sum < 2^24), for which every summation order gives the same value
inline double norm(const Mat& a, const Mat& b, int type) {
  assert(type == NORM_L1 && a.type() == CV_32F && b.type() == CV_32F && a.rows == b.rows && a.cols == b.cols);
  (void)type;
  double s = 0;
  for (int y = 0; y < a.rows; ++y)
    for (int x = 0; x < a.cols; ++x) s += std::fabs((double)(a.at<float>(y, x) - b.at<float>(y, x)));
  return s;
}

// cv::Mat_<T>(r, c) << a, b, c
template <typename T>
class Mat_ : public Mat {
  int fill_;
 public:
  Mat_(int r, int c) : Mat(r, c, sizeof(T) == 4 ? CV_32F : CV_8U), fill_(0) {}
  Mat_& operator<<(T v) { return (*this), v; }
  Mat_& operator,(T v) {
    this->template at<T>(fill_ / cols, fill_ % cols) = v;
    ++fill_;
    return *this;
  }
};

// cv::undistortPoints as Frame.cc:438 / :466 call it: src = dst = an N x 2 CV_32F matrix (the shim's
// reshape() is the identity, the memory layout of N x 2 x 1 channel and N x 1 x 2 channels being the
// same), cameraMatrix = P = mK (3 x 3 CV_32F), distCoeffs 4 x 1 or 5 x 1 CV_32F, R empty.  Forwards to
// oracle/cvprims.hpp (pinned against cv2 4.13.0).
inline void undistortPoints(const Mat& src, Mat& dst, const Mat& K, const Mat& D, const Mat& R, const Mat& P) {
  assert(src.cols == 2 && R.data == nullptr && K.rows == 3 && K.cols == 3 && P.rows == 3 && P.cols == 3);
  float k9[9], p9[9], d[14] = {0};
  for (int i = 0; i < 9; ++i) { k9[i] = K.at<float>(i / 3, i % 3); p9[i] = P.at<float>(i / 3, i % 3); }
  const int nd = D.rows * D.cols;
  for (int i = 0; i < nd && i < 14; ++i) d[i] = D.rows == 1 ? D.at<float>(0, i) : D.at<float>(i, 0);
  std::vector<float> in((size_t)src.rows * 2), out((size_t)src.rows * 2);
  for (int i = 0; i < src.rows; ++i) { in[2 * i] = src.at<float>(i, 0); in[2 * i + 1] = src.at<float>(i, 1); }
  cvprims::undistort_points(in.data(), src.rows, k9, d, nd, p9, out.data());
  if (dst.data != src.data || dst.rows != src.rows) dst.create(src.rows, 2, CV_32F);
  for (int i = 0; i < src.rows; ++i) { dst.at<float>(i, 0) = out[2 * i]; dst.at<float>(i, 1) = out[2 * i + 1]; }
}

// cv::FileStorage / cv::FileNode: compile-only stand-ins (DBoW2's TemplatedVocabulary has virtual
// save/load members over them; the oracle only uses its text loader).  Every read yields "absent".
class FileNode {
 public:
  FileNode operator[](const char*) const { return FileNode(); }
  FileNode operator[](const std::string&) const { return FileNode(); }
  FileNode operator[](int) const { return FileNode(); }
  size_t size() const { return 0; }
  operator int() const { return 0; }
  operator float() const { return 0.f; }
  operator double() const { return 0.0; }
  operator std::string() const { return std::string(); }
};
class FileStorage {
 public:
  enum { READ = 0, WRITE = 1 };
  FileStorage() {}
  FileStorage(const std::string&, int) {}
  bool isOpened() const { return false; }
  void release() {}
  FileNode operator[](const char*) const { return FileNode(); }
  FileNode operator[](const std::string&) const { return FileNode(); }
};
template <typename T>
inline FileStorage& operator<<(FileStorage& fs, const T&) { return fs; }

class _InputArray {
  const Mat* m_;
 public:
  _InputArray(const Mat& m) : m_(&m) {}
  bool empty() const { return m_->empty(); }
  Mat getMat() const { return *m_; }
};
class _OutputArray {
  Mat* m_;
 public:
  _OutputArray(Mat& m) : m_(&m) {}
  void create(int r, int c, int type) const { m_->create(r, c, type); }
  void release() const { m_->release(); }
  Mat getMat() const { return *m_; }
  bool empty() const { return m_->empty(); }
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;

enum { INTER_LINEAR = 1 };
enum { BORDER_REFLECT_101 = 4, BORDER_ISOLATED = 16 };

inline float fastAtan2(float y, float x) { return cvprims::fast_atan2_deg(y, x); }

inline void resize(const Mat& src, Mat& dst, Size sz, double, double, int interp) {
  assert(interp == INTER_LINEAR);
  (void)interp;
  dst.create(sz, CV_8U);
  cvprims::resize_linear_u8(src.data, src.cols, src.rows, src.step, dst.data, dst.cols, dst.rows, dst.step);
}

inline void copyMakeBorder(const Mat& src, Mat& dst, int top, int bottom, int left, int right, int type) {
  assert((type & ~BORDER_ISOLATED) == BORDER_REFLECT_101 && top == bottom && top == left && top == right);
  (void)bottom; (void)left; (void)right; (void)type;
  dst.create(src.rows + 2 * top, src.cols + 2 * top, CV_8U);
  cvprims::make_border_reflect101(src.data, src.cols, src.rows, src.step, dst.data, dst.step, top);
}

inline void GaussianBlur(const Mat& src, Mat& dst, Size k, double sx, double sy, int border) {
  assert(k.width == 7 && k.height == 7 && sx == 2 && sy == 2 && border == BORDER_REFLECT_101);
  (void)k; (void)sx; (void)sy; (void)border;
  Mat s = src;  // keeps the buffer alive when dst aliases src
  dst.create(s.rows, s.cols, CV_8U);
  cvprims::gaussian_blur7_u8(s.data, s.cols, s.rows, s.step, dst.data, dst.step);
}

inline void FAST(const Mat& img, std::vector<KeyPoint>& kps, int threshold, bool nonmax) {
  assert(nonmax);
  (void)nonmax;
  std::vector<cvprims::FastKp> f;
  cvprims::fast9_16(img.data, img.cols, img.rows, img.step, threshold, f);
  kps.clear();
  for (const auto& k : f) kps.push_back(KeyPoint((float)k.x, (float)k.y, 7.f, -1.f, (float)k.score));
}

struct KeyPointsFilter {  // only reached from the reference's unused ComputeKeyPointsOld
  static void retainBest(std::vector<KeyPoint>& k, int n) {
    if (n >= 0 && (int)k.size() > n) {
      std::stable_sort(k.begin(), k.end(), [](const KeyPoint& a, const KeyPoint& b) { return a.response > b.response; });
      k.resize(n);
    }
  }
};

}  // namespace cv
