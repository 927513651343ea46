// Wrapper for #include "ORBmatcher.h" in the libframe_ref.so build (test infrastructure): the stand-in MapPoint / KeyFrame and the
// reference's own Frame first, then the reference's unmodified include/ORBmatcher.h.
#include <Frame.h>      // angle brackets: found through the include path, so that its own #include_next continues from there
#include_next "ORBmatcher.h"
