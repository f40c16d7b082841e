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
#include <cstring>
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

  /* Extension: the rest of the Frame constructors after extraction in one device call --
   * Frame::UndistortKeyPoints (src/Frame.cc:419-449), ComputeStereoFromRGBD (:703-727, when imDepth is
   * not empty; CV_32F as Tracking::GrabImageRGBD leaves it) and AssignFeaturesToGrid (:245-260).  K and
   * distCoef are Frame::mK / mDistCoef (CV_32F), the four grid numbers Frame's statics.  mvuRight /
   * mvDepth are only written with a depth image.  Grid is Frame::mGrid (any [cols][rows] array of
   * index vectors). */
  template <class Grid>
  static void FinishFrame(const std::vector<cv::KeyPoint>& mvKeys, const cv::Mat& K, const cv::Mat& distCoef, float mbf,
                          float mnMinX, float mnMinY, float gridElementWidthInv, float gridElementHeightInv,
                          const cv::Mat& imDepth, std::vector<cv::KeyPoint>& mvKeysUn, std::vector<float>& mvuRight,
                          std::vector<float>& mvDepth, Grid& mGrid, int gridCols = 64, int gridRows = 48, int device = 0) {
    const orbx_camera cam = Camera(K, distCoef, mbf);
    const orbx_frame_bounds fb = {mnMinX, mnMinY, gridElementWidthInv, gridElementHeightInv, gridCols, gridRows};
    const int n = (int)mvKeys.size();
    std::vector<float> xy(2 * (size_t)n);
    std::vector<int32_t> off((size_t)gridCols * gridRows + 1);
    std::vector<uint32_t> idx((size_t)n);
    const bool rgbd = !imDepth.empty();
    if (rgbd) { mvuRight.assign((size_t)n, -1.0f); mvDepth.assign((size_t)n, -1.0f); }
    Check(orbx_frame_finish(device, reinterpret_cast<const orbx_keypoint*>(mvKeys.data()), n, &cam, &fb,
                            rgbd ? reinterpret_cast<const float*>(imDepth.data) : nullptr, imDepth.cols, imDepth.rows,
                            (size_t)imDepth.step, 0, xy.data(), rgbd ? mvuRight.data() : nullptr, rgbd ? mvDepth.data() : nullptr,
                            off.data(), idx.data(), nullptr),
          "ORBextractor::FinishFrame");
    mvKeysUn = mvKeys;
    for (int i = 0; i < n; ++i) { mvKeysUn[i].pt.x = xy[2 * (size_t)i]; mvKeysUn[i].pt.y = xy[2 * (size_t)i + 1]; }
    for (int ix = 0; ix < gridCols; ++ix)
      for (int iy = 0; iy < gridRows; ++iy) {
        const int c = ix * gridRows + iy;
        mGrid[ix][iy].assign(idx.begin() + off[c], idx.begin() + off[c + 1]);
      }
  }
  /* Extension: Frame::ComputeImageBounds (src/Frame.cc:451-485) for a cols x rows image. */
  static void ComputeImageBounds(const cv::Mat& K, const cv::Mat& distCoef, int cols, int rows, float& mnMinX, float& mnMaxX,
                                 float& mnMinY, float& mnMaxY, int device = 0) {
    const orbx_camera cam = Camera(K, distCoef, 0.f);
    if (cam.k1 != 0.0f) {
      float m[8] = {0.f, 0.f, (float)cols, 0.f, 0.f, (float)rows, (float)cols, (float)rows};
      Check(orbx_undistort_points(device, m, 4, &cam, m), "ORBextractor::ComputeImageBounds");
      mnMinX = std::min(m[0], m[4]);
      mnMaxX = std::max(m[2], m[6]);
      mnMinY = std::min(m[1], m[3]);
      mnMaxY = std::max(m[5], m[7]);
    } else {
      mnMinX = 0.0f; mnMaxX = (float)cols; mnMinY = 0.0f; mnMaxY = (float)rows;
    }
  }

 protected:
  static orbx_camera Camera(const cv::Mat& K, const cv::Mat& distCoef, float bf) {
    orbx_camera c;
    c.fx = K.at<float>(0, 0); c.fy = K.at<float>(1, 1); c.cx = K.at<float>(0, 2); c.cy = K.at<float>(1, 2);
    const int nd = distCoef.rows * distCoef.cols;  // 4 x 1 or 5 x 1 (src/Tracking.cc:66-77)
    const float* d = reinterpret_cast<const float*>(distCoef.data);
    c.k1 = nd > 0 ? d[0] : 0.f; c.k2 = nd > 1 ? d[1] : 0.f; c.p1 = nd > 2 ? d[2] : 0.f; c.p2 = nd > 3 ? d[3] : 0.f;
    c.k3 = nd > 4 ? d[4] : 0.f;
    c.bf = bf;
    return c;
  }
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
  int out_cap_ = 0, cap_w_ = 0, cap_h_ = 0;
  std::vector<uint8_t> desc_stage_;

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
  Check(orbx_set_pyramid_staging(impl_, download_pyramid_ ? 1 : 0), "ORBextractor::operator()");
  if (out_cap_ == 0 || image.cols != cap_w_ || image.rows != cap_h_) {
    Check(orbx_frame_capacity(impl_, image.cols, image.rows, &out_cap_), "ORBextractor::operator()");
    cap_w_ = image.cols; cap_h_ = image.rows;
  }
  Check(orbx_extract_batch(impl_, image.data, 1, image.cols, image.rows, (size_t)image.step,
                           (size_t)image.step * (size_t)image.rows, 0),
        "ORBextractor::operator()");
  // one synchronisation for count + keypoints + descriptors (page-locked staging inside the handle)
  if (desc_stage_.size() < (size_t)out_cap_ * 32) desc_stage_.resize((size_t)out_cap_ * 32);
  _keypoints.resize((size_t)out_cap_);
  Check(orbx_read_frame0(impl_, reinterpret_cast<orbx_keypoint*>(_keypoints.data()), desc_stage_.data(), out_cap_, &n),
        "ORBextractor::operator()");
  _keypoints.resize((size_t)n);
  if (n == 0) {
    _descriptors.release();  // :1072
  } else {
    _descriptors.create(n, 32, CV_8U);  // :1076
    cv::Mat d = _descriptors.getMat();
    assert((size_t)d.step == 32);
    memcpy(d.data, desc_stage_.data(), (size_t)n * 32);
  }
  if (download_pyramid_) {
    // every bordered level arrived in ONE transfer that ran beside the later stages; the Mats are views into the
    // handle's page-locked staging, valid until this extractor is called again (Frame.cc reads them only in its constructor)
    for (int l = 0; l < nlevels; ++l) {
      const uint8_t* p00 = nullptr;
      int w = 0, h = 0;
      size_t step = 0;
      Check(orbx_get_pyramid_host(impl_, l, &p00, &w, &h, &step), "ORBextractor pyramid");
      cv::Mat temp(h + 2 * EDGE, w + 2 * EDGE, CV_8U, (void*)(p00 - (size_t)EDGE * step - EDGE), step);
      mvImagePyramid[l] = temp(cv::Rect(EDGE, EDGE, w, h));
    }
  }
}

}  // namespace ORB_SLAM2

#endif  // ORBEXTRACTOR_H
