// Wrapper for #include "ORBmatcher.h" in the libmappoint_ref.so build (test infrastructure): the reference's own MapPoint and the
// stand-in KeyFrame / Frame first, then the reference's unmodified include/ORBmatcher.h.
#include <MapPoint.h>      // angle brackets: found through the include path, so that its own #include_next continues from there
#include_next "ORBmatcher.h"
