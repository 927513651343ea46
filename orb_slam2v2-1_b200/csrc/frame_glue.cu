// frame_glue.cu -- the per-keypoint maps of the Frame constructors that sit between the extractor and the matcher (SURVEY.md
// section 8(f) rank 1): Frame::UndistortKeyPoints (reference src/Frame.cc:419-449), ComputeImageBounds (:451-479) and
// ComputeStereoFromRGBD (:658-679).  cv::undistortPoints is OpenCV's (un-vendored); the kernel follows its published
// algorithm in double precision, operation for operation.  THIS FILE IS COMPILED WITH --fmad=false (see the Makefile): a
// contracted multiply-add would change the last bit of the iteration and with it the float the reference stores.
#include "matcher_internal.cuh"
#include <string.h>
#include <algorithm>
#include <vector>

namespace orbx {

struct Camera { double fx, fy, cx, cy, ifx, ify, k[12]; };

// cvUndistortPointsInternal, the path taken by undistortPoints(src, dst, K, dist, noArray(), K): 5 iterations, no EPS test
__device__ __forceinline__ float2 undistort_one(float pu, float pv, const Camera& C)
{
    const double u = pu, v = pv;
    double x = (u - C.cx) * C.ifx, y = (v - C.cy) * C.ify;
    const double x0 = x, y0 = y;
    const double* k = C.k;
    for (int j = 0; j < 5; j++) {
        const double r2 = x * x + y * y;
        const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
        if (icdist < 0) { x = (u - C.cx) * C.ifx; y = (v - C.cy) * C.ify; break; }
        const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
        const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
        x = (x0 - deltaX) * icdist;
        y = (y0 - deltaY) * icdist;
    }
    return make_float2((float)(C.fx * x + C.cx), (float)(C.fy * y + C.cy));
}

// stride = floats between consecutive points in `in` / `out` (2 for packed points, 7 for orbx_kp records)
__global__ void __launch_bounds__(256) k_undistort(const float* in, int n, int stride, Camera C, float* out)      // in may be out
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float2 r = undistort_one(in[(size_t)i * stride], in[(size_t)i * stride + 1], C);
    out[(size_t)i * stride] = r.x;
    out[(size_t)i * stride + 1] = r.y;
}

__global__ void __launch_bounds__(256) k_stereo_from_rgbd(const orbx_kp* __restrict__ kps, const orbx_kp* __restrict__ kpsUn, int n,
                                                          const uint8_t* __restrict__ imDepth, int w, int h, size_t strideBytes, float bf,
                                                          float* __restrict__ uRight, float* __restrict__ depth)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int v = (int)kps[i].y, u = (int)kps[i].x;          // imDepth.at<float>(v, u) with float arguments (:672)
    float ur = -1.0f, dp = -1.0f;
    if (u >= 0 && v >= 0 && u < w && v < h) {
        const float d = *reinterpret_cast<const float*>(imDepth + (size_t)v * strideBytes + sizeof(float) * (size_t)u);
        if (d > 0) { dp = d; ur = kpsUn[i].x - bf / d; }
    }
    uRight[i] = ur;
    depth[i] = dp;
}

static bool make_camera(const char* who, float fx, float fy, float cx, float cy, const float* dist, int ndist, Camera& C)
{
    if (!dist || ndist < 4 || ndist > 12 || fx == 0 || fy == 0) { set_error("%s: bad camera (need 4..12 distortion coefficients, fx, fy != 0)", who); return false; }
    C.fx = fx; C.fy = fy; C.cx = cx; C.cy = cy;
    C.ifx = 1.0 / C.fx;
    C.ify = 1.0 / C.fy;
    for (int i = 0; i < 12; i++) C.k[i] = i < ndist ? (double)dist[i] : 0.0;
    return true;
}

static int undistort(orbx_matcher* m, const float* in, int n, int stride, const Camera& C, float* out)
{
    ORBX_CUDA(cudaSetDevice(m->device));
    const size_t bytes = sizeof(float) * (size_t)n * stride;
    Carver cv(m->arena);
    const size_t oI = cv.take(bytes);
    int rc = m->arena.reserve(cv.total());
    if (!rc) rc = m->arena.mirror();
    if (rc) return rc;
    float* d = (float*)(m->arena.base + oI);
    Xfer x(m->arena);
    x.up(oI, in, bytes);
    ORBX_CUDA(x.send());
    k_undistort<<<(n + 255) / 256, 256, 0, m->stream>>>(d, n, stride, C, d);          // in place, like the reference (:438)
    ORBX_CUDA(cudaGetLastError());
    m->lastLaunches = 1;
    x.down(out, oI, bytes);
    ORBX_CUDA(x.fetch());
    return ORBX_OK;
}

// Frame::UndistortKeyPoints (src/Frame.cc:419-449) on keypoint records that already live on the device (orbm_frame_from_extractor):
// in place; a camera without distortion (mDistCoef[0] == 0, :421-425) leaves them untouched.
int undistort_keypoints_device(cudaStream_t s, orbx_kp* d_kps, int n, float fx, float fy, float cx, float cy, const float* dist, int ndist)
{
    if (n <= 0 || !dist || ndist < 1 || dist[0] == 0.0f) return ORBX_OK;
    Camera C;
    if (!make_camera("orbm_frame_from_extractor", fx, fy, cx, cy, dist, ndist, C)) return ORBX_ERR_ARG;
    k_undistort<<<(n + 255) / 256, 256, 0, s>>>(reinterpret_cast<const float*>(d_kps), n, 7, C, reinterpret_cast<float*>(d_kps));
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

}  // namespace orbx

using namespace orbx;

extern "C" {

int orbm_undistort_points(orbx_matcher* m, const float* pts, int n, float fx, float fy, float cx, float cy, const float* dist, int ndist,
                          float* out)
{
    Camera C;
    if (!m || n < 0 || (n > 0 && (!pts || !out))) { set_error("orbm_undistort_points: bad argument"); return ORBX_ERR_ARG; }
    if (!make_camera("orbm_undistort_points", fx, fy, cx, cy, dist, ndist, C)) return ORBX_ERR_ARG;
    if (n == 0) return ORBX_OK;
    return undistort(m, pts, n, 2, C, out);
}

int orbm_undistort_keypoints(orbx_matcher* m, const orbx_kp* kps, int n, float fx, float fy, float cx, float cy, const float* dist, int ndist,
                             orbx_kp* out)
{
    Camera C;
    if (!m || n < 0 || (n > 0 && (!kps || !out))) { set_error("orbm_undistort_keypoints: bad argument"); return ORBX_ERR_ARG; }
    if (!make_camera("orbm_undistort_keypoints", fx, fy, cx, cy, dist, ndist, C)) return ORBX_ERR_ARG;
    if (n == 0) return ORBX_OK;
    if (dist[0] == 0.0) {                                      // :421-425: mvKeysUn = mvKeys
        if (out != kps) memcpy(out, kps, sizeof(orbx_kp) * (size_t)n);
        m->lastLaunches = 0;
        return ORBX_OK;
    }
    static_assert(sizeof(orbx_kp) == 7 * sizeof(float), "orbx_kp is 7 words");
    return undistort(m, reinterpret_cast<const float*>(kps), n, 7, C, reinterpret_cast<float*>(out));
}

int orbm_image_bounds(orbx_matcher* m, int cols, int rows, float fx, float fy, float cx, float cy, const float* dist, int ndist, float* bounds4)
{
    Camera C;
    if (!m || !bounds4 || cols <= 0 || rows <= 0) { set_error("orbm_image_bounds: bad argument"); return ORBX_ERR_ARG; }
    if (!make_camera("orbm_image_bounds", fx, fy, cx, cy, dist, ndist, C)) return ORBX_ERR_ARG;
    if (dist[0] != 0.0) {
        const float mat[8] = {0.0f, 0.0f, (float)cols, 0.0f, 0.0f, (float)rows, (float)cols, (float)rows};      // :455-459
        float u[8];
        int rc = undistort(m, mat, 4, 2, C, u);
        if (rc) return rc;
        bounds4[0] = std::min(u[0], u[4]);                     // :466-469
        bounds4[1] = std::max(u[2], u[6]);
        bounds4[2] = std::min(u[1], u[3]);
        bounds4[3] = std::max(u[5], u[7]);
    } else {
        bounds4[0] = 0.0f; bounds4[1] = (float)cols; bounds4[2] = 0.0f; bounds4[3] = (float)rows;
        m->lastLaunches = 0;
    }
    return ORBX_OK;
}

int orbm_stereo_from_rgbd(orbx_matcher* m, const orbx_kp* kps, const orbx_kp* kps_un, int n, const float* im_depth, int w, int h,
                          size_t stride_bytes, float bf, float* u_right, float* depth)
{
    if (!m || n < 0 || (n > 0 && (!kps || !kps_un || !u_right || !depth)) || !im_depth || w <= 0 || h <= 0 ||
        stride_bytes < sizeof(float) * (size_t)w || (stride_bytes & 3)) {
        set_error("orbm_stereo_from_rgbd: bad argument");
        return ORBX_ERR_ARG;
    }
    if (n == 0) return ORBX_OK;
    ORBX_CUDA(cudaSetDevice(m->device));
    Carver cv(m->arena);
    const size_t oK = cv.take(sizeof(orbx_kp) * (size_t)n), oU = cv.take(sizeof(orbx_kp) * (size_t)n), oD = cv.take(stride_bytes * (size_t)h),
                 oR = cv.take(sizeof(float) * (size_t)n), oZ = cv.take(sizeof(float) * (size_t)n);
    int rc = m->arena.reserve(cv.total());
    if (rc) return rc;
    uint8_t* B = m->arena.base;
    cudaStream_t s = m->stream;
    ORBX_CUDA(cudaMemcpyAsync(B + oK, kps, sizeof(orbx_kp) * (size_t)n, cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(B + oU, kps_un, sizeof(orbx_kp) * (size_t)n, cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(B + oD, im_depth, stride_bytes * (size_t)(h - 1) + sizeof(float) * (size_t)w, cudaMemcpyHostToDevice, s));
    k_stereo_from_rgbd<<<(n + 255) / 256, 256, 0, s>>>((const orbx_kp*)(B + oK), (const orbx_kp*)(B + oU), n, B + oD, w, h, stride_bytes, bf,
                                                       (float*)(B + oR), (float*)(B + oZ));
    ORBX_CUDA(cudaGetLastError());
    m->lastLaunches = 1;
    ORBX_CUDA(cudaMemcpyAsync(u_right, B + oR, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaMemcpyAsync(depth, B + oZ, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    return ORBX_OK;
}

}  // extern "C"
