// prelude_dropin.h -- force-included (-include) in the build of oracle/_ref/libframe_dropin.so (test infrastructure, not product).
//
// The swap a maintainer makes, on the files of the path that compile here: the reference's UNMODIFIED src/Frame.cc, src/KeyFrame.cc and
// src/MapPoint.cc, built against THIS REPOSITORY'S include/ORBextractor.h and include/ORBmatcher.h in place of the reference's headers of the
// same names.  The product headers are read first; they define the include guards ORBEXTRACTOR_H and ORBMATCHER_H, so the reference's
// `#include "ORBextractor.h"` / `"ORBmatcher.h"` (Frame.h:31, Frame.cc:23, KeyFrame.cc:23, MapPoint.cc:22) become empty and every use --
// (*mpORBextractorLeft)(im, cv::Mat(), mvKeys, mDescriptors), GetLevels() .. GetInverseScaleSigmaSquares(), mvImagePyramid in
// ComputeStereoMatches, ORBmatcher::TH_HIGH / TH_LOW / DescriptorDistance -- binds to the product's classes.  No ORBextractor.cc and no
// ORBmatcher.cc of the reference is linked.  The third drop-in header, include/ORBVocabulary.h, replaces the reference's ORBVocabulary.h
// (= DBoW2's TemplatedVocabulary) in the same way: Frame::ComputeBoW and KeyFrame::ComputeBoW (Frame.cc:410-417, KeyFrame.cc:59-68) then call the
// product's transform with the reference's own DBoW2::BowVector / FeatureVector members (the vendored BowVector.h / FeatureVector.h stay).
#pragma once
#include "../slam_full/prelude.h"
#include "../../../include/ORBextractor.h"
#include "../../../include/ORBmatcher.h"
#include "../../../include/ORBVocabulary.h"
