// orb_internal.h -- CPU ORACLE internals (test infrastructure, not product).
#pragma once
#include <stdint.h>
#include <stddef.h>
#include <cmath>
#include <vector>

namespace orbo {

struct XYR { int x, y, r; };

// cvRound: round half to even (lrint under the default rounding mode).
static inline int cv_round(double v) { return (int)std::lrint(v); }
static inline int cv_round(float v) { return (int)std::lrintf(v); }

void resize_linear_u8(const uint8_t* src, int sw, int sh, size_t sstride, uint8_t* dst, int dw, int dh, size_t dstride);
void border_reflect101(const uint8_t* src, int w, int h, size_t sstride, uint8_t* dst, int b, size_t dstride);
void gauss7(const uint8_t* src, int w, int h, size_t sstride, uint8_t* dst, size_t dstride);
int fast_raw(const uint8_t* p, size_t stride);
void fast_detect(const uint8_t* roi, int w, int h, size_t stride, int threshold, bool nms, std::vector<XYR>& out);
float fast_atan2(float y, float x);

// DistributeOctTree (ORBextractor.cc:539-763) on integer-valued keys; returns indices into `keys` in output order.
std::vector<int> distribute_octtree(const std::vector<XYR>& keys, int minX, int maxX, int minY, int maxY, int N, int tie_mode);

int descriptor_distance(const uint8_t* a, const uint8_t* b);

} // namespace orbo
