// slam_ref.h -- the reference's REAL Frame / KeyFrame / MapPoint / ORBmatcher for test programs (test infrastructure, not product).
//
// Brings the reference's own, unmodified headers (include/Frame.h, KeyFrame.h, MapPoint.h, ORBmatcher.h, ORBextractor.h,
// ORBVocabulary.h + vendored DBoW2) into a translation unit under the namespace ORB_SLAM2_REFS, the way oracle/_ref/libslam_ref.so is
// compiled (see prelude.h for what is stood in), and then clears the three include guards that the product's drop-in headers share with
// the reference's, so that include/ORBextractor.h, ORBmatcher.h and ORBVocabulary.h of this repository can follow in the same file as
// namespace ORB_SLAM2.  Needs -I oracle/ref_shim -I <reference>/include -I <reference>; include it before the product headers.
#ifndef ORBX_SLAM_REF_H
#define ORBX_SLAM_REF_H
#define ORB_SLAM2 ORB_SLAM2_REFS
#include "prelude.h"
#include <Frame.h>
#include <KeyFrame.h>
#include <MapPoint.h>
#include <ORBmatcher.h>
#undef ORB_SLAM2
#undef ORBMATCHER_H
#undef ORBEXTRACTOR_H
#undef ORBVOCABULARY_H
#endif
