/* ORBextractor.h -- source-compatible ORB_SLAM2::ORBextractor over the B200 C ABI (orbfront.h).
 *
 * Drop-in for the reference's include/ORBextractor.h:46-112 + src/ORBextractor.cc: same
 * constructor, operator(), inline getters and public mvImagePyramid, so Tracking.cc:132-138 and
 * Frame.cc:69-75,265-267,534,624-641 compile against it unchanged.  Header-only on purpose:
 * liborbfront.so has no OpenCV dependency; this file uses whatever cv:: the includer provides
 * (OpenCV in the reference build, oracle/cvshim in tests/cpp).
 *
 * Member layout: every data member of the reference class is kept, in declaration order
 * (getters are inline and Frame.cc reads mvImagePyramid directly), and one opaque handle is
 * appended last.  All extraction work runs on the GPU; there is no CPU fallback -- construction
 * throws std::runtime_error when no CUDA device is usable.
 */
#ifndef ORBEXTRACTOR_H
#define ORBEXTRACTOR_H

#include <cassert>
#include <cstdlib>
#include <list>
#include <stdexcept>
#include <string>
#include <vector>

#include <opencv2/core/core.hpp>

#include "orbfront.h"

namespace ORB_SLAM2 {

class ORBextractor {
 public:
  enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

  /* src/ORBextractor.cc:410-470; the scale/quota tables are computed once in the library with the
   * reference's expression types and read back, so getters return bit-identical floats. */
  ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);
  ~ORBextractor() { orbx_destroy(impl_); }
  ORBextractor(const ORBextractor&) = delete;
  ORBextractor& operator=(const ORBextractor&) = delete;

  /* Compute the ORB features and descriptors on an image; mask is ignored as in the reference
   * (include/ORBextractor.h:59).  src/ORBextractor.cc:1051-1113. */
  void operator()(cv::InputArray image, cv::InputArray mask, std::vector<cv::KeyPoint>& keypoints,
                  cv::OutputArray descriptors);

  int inline GetLevels() { return nlevels; }
  float inline GetScaleFactor() { return scaleFactor; }
  std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
  std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
  std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
  std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

  /* Level l is a h_l x w_l ROI of a (h_l+38) x (w_l+38) buffer whose 19-px frame is
   * BORDER_REFLECT_101 (src/ORBextractor.cc:1115-1140), copied back after every call because
   * Frame::ComputeStereoMatches reads it (Frame.cc:534,624-641). */
  std::vector<cv::Mat> mvImagePyramid;

  /* Extension (not in the reference): monocular / RGB-D callers never read mvImagePyramid, so
   * they can skip the per-frame device-to-host copy of the pyramid. */
  void SetPyramidDownload(bool enabled) { download_pyramid_ = enabled; }
  /* Extension: the C-ABI handle, for batch / device-pointer use (orbx_extract_batch). */
  orbx_handle* Handle() { return impl_; }
  /* Extension: the whole of Frame::ComputeStereoMatches (src/Frame.cc:527-700) on the pyramids the two
   * extractors still hold on the device after operator() -- row-band descriptor search, 11x11
   * sliding-window refinement, parabola fit, median rejection.  With it a stereo Frame never reads
   * mvImagePyramid, so both extractors can SetPyramidDownload(false). */
  static void ComputeStereoMatches(ORBextractor& left, ORBextractor& right, const std::vector<cv::KeyPoint>& keysLeft,
                                   const cv::Mat& descLeft, const std::vector<cv::KeyPoint>& keysRight, const cv::Mat& descRight,
                                   float mb, float mbf, std::vector<float>& mvuRight, std::vector<float>& mvDepth) {
    assert((keysLeft.empty() || (size_t)descLeft.step == 32) && (keysRight.empty() || (size_t)descRight.step == 32));
    mvuRight.assign(keysLeft.size(), -1.0f);
    mvDepth.assign(keysLeft.size(), -1.0f);
    Check(orbx_stereo_matches(left.impl_, 0, right.impl_, 0, reinterpret_cast<const orbx_keypoint*>(keysLeft.data()), descLeft.data,
                              (int)keysLeft.size(), reinterpret_cast<const orbx_keypoint*>(keysRight.data()), descRight.data,
                              (int)keysRight.size(), mb, mbf, mvuRight.data(), mvDepth.data()),
          "ORBextractor::ComputeStereoMatches");
  }

 protected:
  std::vector<cv::Point> pattern;  // kept for layout; the pattern lives in the library

  int nfeatures;
  double scaleFactor;
  int nlevels;
  int iniThFAST;
  int minThFAST;

  std::vector<int> mnFeaturesPerLevel;

  std::vector<int> umax;

  std::vector<float> mvScaleFactor;
  std::vector<float> mvInvScaleFactor;
  std::vector<float> mvLevelSigma2;
  std::vector<float> mvInvLevelSigma2;

  // appended (not in the reference)
  orbx_handle* impl_ = nullptr;
  bool download_pyramid_ = true;

  static void Check(int rc, const char* what) {
    if (rc != ORB_OK) throw std::runtime_error(std::string(what) + ": " + orb_last_error());
  }
};

inline ORBextractor::ORBextractor(int _nfeatures, float _scaleFactor, int _nlevels, int _iniThFAST, int _minThFAST)
    : nfeatures(_nfeatures), scaleFactor(_scaleFactor), nlevels(_nlevels), iniThFAST(_iniThFAST), minThFAST(_minThFAST) {
  static_assert(sizeof(cv::KeyPoint) == sizeof(orbx_keypoint), "cv::KeyPoint must be 7 x 4 bytes");
  orbx_config cfg;
  cfg.nfeatures = _nfeatures;
  cfg.scale_factor = _scaleFactor;
  cfg.nlevels = _nlevels;
  cfg.ini_th_fast = _iniThFAST;
  cfg.min_th_fast = _minThFAST;
  cfg.device = 0;
  if (const char* e = getenv("ORBFRONT_DEVICE")) cfg.device = atoi(e);
  cfg.max_width = cfg.max_height = cfg.max_batch = 0;  // sized on the first frame
  Check(orbx_create(&cfg, &impl_), "ORBextractor");
  mvScaleFactor.resize(nlevels);
  mvInvScaleFactor.resize(nlevels);
  mvLevelSigma2.resize(nlevels);
  mvInvLevelSigma2.resize(nlevels);
  mnFeaturesPerLevel.resize(nlevels);
  orbx_get_scale_table(impl_, 0, mvScaleFactor.data());
  orbx_get_scale_table(impl_, 1, mvInvScaleFactor.data());
  orbx_get_scale_table(impl_, 2, mvLevelSigma2.data());
  orbx_get_scale_table(impl_, 3, mvInvLevelSigma2.data());
  orbx_get_level_quota(impl_, mnFeaturesPerLevel.data());
  mvImagePyramid.resize(nlevels);
}

inline void ORBextractor::operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint>& _keypoints,
                                     cv::OutputArray _descriptors) {
  if (_image.empty()) return;  // :1054-1055, outputs untouched
  cv::Mat image = _image.getMat();
  assert(image.type() == CV_8UC1);  // :1058
  const int EDGE = 19;
  int n = 0;
  Check(orbx_extract_batch(impl_, image.data, 1, image.cols, image.rows, (size_t)image.step,
                           (size_t)image.step * (size_t)image.rows, 0),
        "ORBextractor::operator()");
  Check(orbx_batch_counts(impl_, &n), "ORBextractor::operator()");
  _keypoints.clear();
  if (n == 0) {
    _descriptors.release();  // :1072
  } else {
    _keypoints.resize((size_t)n);
    _descriptors.create(n, 32, CV_8U);  // :1076
    cv::Mat d = _descriptors.getMat();
    assert((size_t)d.step == 32);
    int got = 0;
    Check(orbx_batch_read(impl_, 0, reinterpret_cast<orbx_keypoint*>(_keypoints.data()), d.data, n, &got),
          "ORBextractor::operator()");
  }
  if (download_pyramid_) {
    for (int l = 0; l < nlevels; ++l) {
      int w = 0, h = 0;
      Check(orbx_level_size(impl_, l, &w, &h), "ORBextractor pyramid");
      cv::Mat temp(h + 2 * EDGE, w + 2 * EDGE, CV_8U);
      Check(orbx_get_level(impl_, 0, l, temp.data, (size_t)temp.step, 1), "ORBextractor pyramid");
      mvImagePyramid[l] = temp(cv::Rect(EDGE, EDGE, w, h));
    }
  }
}

}  // namespace ORB_SLAM2

#endif  // ORBEXTRACTOR_H
