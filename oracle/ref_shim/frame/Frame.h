// Wrapper found first on the include path when the reference's sources say #include "Frame.h" in the build of oracle/_ref/libframe_ref.so
// (test infrastructure, not product).  MapPoint and KeyFrame are the stand-ins of ref_shim/slam/slam_objects.h (their include guards blank
// out the real headers, which drag in Map, ServerMap, boost and ROS messages); Frame itself, ORBextractor, ORBmatcher, ORBVocabulary and
// the vendored DBoW2 are the reference's own, unmodified files, reached with #include_next.
#ifndef ORBX_FRAME_WRAPPER_H
#define ORBX_FRAME_WRAPPER_H
#define ORBX_REAL_FRAME
#include "../slam/slam_objects.h"
using namespace std;      // the reference's headers rely on a using-directive that one of the real object headers leaks
#include_next "Frame.h"
namespace ORB_SLAM2
{
ORBX_PREDICT_SCALE_FRAME
}
#endif
