// Wrapper found first on the include path when the reference's src/ORBmatcher.cc says #include "ORBmatcher.h" (test infrastructure):
// it brings in the stand-in objects -- whose include guards blank out the real MapPoint.h / KeyFrame.h / Frame.h -- and then hands
// over to the reference's own, unmodified include/ORBmatcher.h.
#include "slam_objects.h"
using namespace std;      // the reference's headers rely on a using-directive that one of the real object headers leaks
#include_next "ORBmatcher.h"
