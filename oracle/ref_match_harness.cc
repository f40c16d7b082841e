// oracle/ref_match_harness.cc -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// C entry points (refm_*) over the reference's OWN matcher: /root/reference/src/ORBmatcher.cc is
// compiled UNMODIFIED next to this file, against the reference's own include/ORBmatcher.h (reached
// through a build-time symlink in oracle/_ref/env so that its `#include "MapPoint.h"` etc. resolve
// to tests/cpp/mock instead of the reference's Eigen/g2o-dependent headers) and oracle/cvshim.
// The scenes are tests/cpp/matcher_scenes.inc, the same file the product's drop-in harness builds,
// so both sides see identical objects.  Output: oracle/_ref/libmatchref.so (git-ignored).
#include "ORBmatcher.h"

float ORB_SLAM2::Frame::mnMinX, ORB_SLAM2::Frame::mnMinY, ORB_SLAM2::Frame::mnMaxX, ORB_SLAM2::Frame::mnMaxY;
float ORB_SLAM2::Frame::mfGridElementWidthInv, ORB_SLAM2::Frame::mfGridElementHeightInv;
#define SCN(name) refm_##name
#include "../tests/cpp/matcher_scenes.inc"
