/* orbx.h -- C ABI of liborbx.so: the sm_100a CUDA implementation of ORB-SLAM2's data-parallel front end.
 *
 * This is the drop-in boundary.  Everything above it (include/ORBextractor.h,
 * include/ORBmatcher.h, the Python ctypes mirror in orb_slam2v2-1_b200/orbx) only forwards to these entry
 * points; everything below it is hand-written CUDA.  There is NO CPU fallback: every compute entry point
 * returns ORBX_ERR_CUDA when no device is usable.
 *
 * Conventions: plain C types only; every function returns an int status (0 = ORBX_OK) unless stated;
 * no exception crosses the ABI; the text of the last failure of the calling thread is orbx_last_error().
 * One extractor handle owns one CUDA stream and one device workspace: calls on DIFFERENT handles may run
 * concurrently from different host threads (the reference extracts left/right images on two std::threads,
 * reference src/Frame.cc:78-81); one handle must not be used from two threads at once (the reference's
 * ORBextractor is not re-entrant either: it mutates mvImagePyramid, src/ORBextractor.cc:1107-1132).
 *
 * Reference citations are relative to kimwin2/ORB_SLAM2v2-1.
 */
#ifndef ORBX_H_
#define ORBX_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
    ORBX_OK = 0,
    ORBX_ERR_ARG = 1,      /* bad argument (null pointer, non-positive size, unsupported shape) */
    ORBX_ERR_CUDA = 2,     /* CUDA runtime failure, including "no device" */
    ORBX_ERR_CAPACITY = 3, /* caller's output buffer too small; *n still receives the true count */
    ORBX_ERR_STATE = 4     /* call order violated (e.g. fetch before extract) */
};

/* Same field order and size (28 bytes) as cv::KeyPoint {pt.x, pt.y, size, angle, response, octave, class_id}. */
typedef struct orbx_kp {
    float x, y, size, angle, response;
    int32_t octave, class_id;
} orbx_kp;

typedef struct orbx_extractor orbx_extractor;

/* ---------------------------------------------------------------------------------------------------------
 * ORBextractor  (reference include/ORBextractor.h:45-111, src/ORBextractor.cc)
 * ------------------------------------------------------------------------------------------------------- */

/* Replaces ORBextractor::ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)
 * (src/ORBextractor.cc:410-470).  device = CUDA ordinal.  max_batch = largest number of frames one call will
 * carry (workspace is sized for it at the first call of each image shape). */
int orbx_create(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST,
                int device, int max_batch, orbx_extractor** out);
void orbx_destroy(orbx_extractor* ex);

/* descriptor rotation arithmetic of computeOrbDescriptor (src/ORBextractor.cc:119-120):
 * 0 = each product and sum rounded separately (portable definition, default),
 * 1 = FMA-contracted the way g++ -O3 -march=native compiles the reference on an FMA host. */
int orbx_set_fma_mode(orbx_extractor* ex, int fma_mode);

/* Tuning / test knob for DistributeOctTree (src/ORBextractor.cc:539-763): the quadtree kernel exists in three forms that
 * give identical results and differ in CTA size; 0 = chosen per launch from the quotas and the batch size (default),
 * 1 = 128 threads, 2 = 256 threads, 3 = 512 threads. */
int orbx_set_octree_form(orbx_extractor* ex, int form);

/* Replaces the by-value accessors GetLevels / GetScaleFactors / GetInverseScaleFactors / GetScaleSigmaSquares /
 * GetInverseScaleSigmaSquares (include/ORBextractor.h:63-83).  Each non-null pointer receives nlevels floats. */
int orbx_levels(const orbx_extractor* ex);
int orbx_scale_tables(const orbx_extractor* ex, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2);
/* mnFeaturesPerLevel (src/ORBextractor.cc:435-446) and umax[0..15] (:454-469); protected members of the reference
 * class, exported for the parity tests. */
int orbx_features_per_level(const orbx_extractor* ex, int* out_nlevels);
int orbx_umax(const orbx_extractor* ex, int* out16);
/* Upper bound on the keypoints one frame can produce (a quadtree pass may overshoot the per-level quota,
 * src/ORBextractor.cc:669-737): callers size kps/desc with it. */
int orbx_max_keypoints(const orbx_extractor* ex);

/* Replaces ORBextractor::operator()(image, mask, keypoints, descriptors) (src/ORBextractor.cc:1043-1105) for one
 * 8-bit single-channel HOST image (row stride in bytes).  Synchronous.  kps receives *n keypoints (level-0 pixel
 * units, level-ascending, quadtree list order inside a level), desc receives *n rows of 32 bytes.  cap = capacity of
 * kps/desc in keypoints.  w<=0 or h<=0 or img==NULL -> *n = 0, ORBX_OK (the reference returns silently on an empty
 * image, :1046-1047).  The mask argument of the reference is ignored there (:1043 "Mask is ignored") and absent here. */
int orbx_extract(orbx_extractor* ex, const uint8_t* img, int w, int h, size_t stride,
                 orbx_kp* kps, uint8_t* desc, int cap, int* n);

/* Throughput path: B frames of one shape per call (independent frames).
 * imgs[b] = HOST pointer of frame b.  kps/desc are HOST arrays of B*cap entries (frame b at offset b*cap);
 * n receives B counts.  Synchronous: returns when the results are in the caller's buffers. */
int orbx_extract_batch(orbx_extractor* ex, const uint8_t* const* imgs, int B, int w, int h, size_t stride,
                       orbx_kp* kps, uint8_t* desc, int cap, int* n);

/* Device-resident variant: d_imgs = DEVICE pointer to B frames (frame b at d_imgs + b*frame_stride, rows pitch bytes
 * apart; pitch must be a multiple of 4).  Enqueues the whole extraction on the handle's stream and returns without
 * synchronising; results stay in the handle's device workspace until orbx_fetch / the next extract call. */
int orbx_extract_device(orbx_extractor* ex, const uint8_t* d_imgs, int B, int w, int h, size_t pitch, size_t frame_stride);
/* Colour input: cvtColor(im, im, CV_RGB2GRAY | CV_BGR2GRAY | CV_RGBA2GRAY | CV_BGRA2GRAY) of Tracking::GrabImageStereo / RGBD /
 * Monocular (src/Tracking.cc:280-305, :320-334, :350-364) fused into the ingestion: channels = 3 or 4 interleaved bytes per
 * pixel, rgb = 1 when byte 0 is red (mbRGB), 0 when it is blue.  OpenCV's 8-bit arithmetic, (R*9798 + G*19235 + B*3735 + 2^14) >> 15,
 * bit-exact; then exactly orbx_extract_batch / orbx_extract_device on the grey planes (which the handle owns: orbx_pyramid_level 0
 * returns the grey image).  Device form: pointer, pitch and frame stride multiples of 4 (3 channels) or 16 (4 channels). */
int orbx_extract_color_batch(orbx_extractor* ex, const uint8_t* const* imgs, int B, int w, int h, size_t stride, int channels, int rgb,
                             orbx_kp* kps, uint8_t* desc, int cap, int* n);
int orbx_extract_color_device(orbx_extractor* ex, const uint8_t* d_imgs, int B, int w, int h, size_t pitch, size_t frame_stride,
                              int channels, int rgb);

/* ---- Multi-GPU driver.  The reference extracts one image per thread (src/Frame.cc:78-81); frames are
 * independent, so a pool shards a batch over GPUs with no collective: the batch is cut into chunks of `chunk` consecutive frames,
 * chunk c goes to devices[c mod ndevices] (chunk = 1: frame i -> GPU i mod G), and every device runs `lanes` host threads, each with
 * its own extractor handle (stream + workspace + page-locked staging buffer), so that copies and kernels of different chunks overlap.
 * Results are written to the caller's arrays at the frame's own index: the output bytes are the same for every device count, chunk
 * size and lane count.  devices = NULL or ndevices = 0: every visible GPU.  The pool entry points may be called from one host thread at a
 * time; several batches may be in flight through the submit / wait pair below. */
typedef struct orbx_pool orbx_pool;
int orbx_pool_create(const int* devices, int ndevices, int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST,
                     int chunk, int lanes, orbx_pool** out);
void orbx_pool_destroy(orbx_pool* pool);      /* lets the batches in flight run to their end first */
int orbx_pool_devices(const orbx_pool* pool);
int orbx_pool_set_fma_mode(orbx_pool* pool, int fma_mode);
/* Plans an image shape on every handle of the pool (allocates the workspaces once, outside any timed region) and returns the
 * per-frame keypoint capacity for that shape: with cap equal to it and page-locked result arrays, orbx_pool_extract_batch copies
 * straight into the caller's memory. */
int orbx_pool_plan(orbx_pool* pool, int w, int h, int* max_keypoints);
/* orbx_extract_batch over the pool: same arguments, same results, B unbounded.  Synchronous (= submit + wait). */
int orbx_pool_extract_batch(orbx_pool* pool, const uint8_t* const* imgs, int B, int w, int h, size_t stride,
                            orbx_kp* kps, uint8_t* desc, int cap, int* n);
/* The same, asynchronous: the batch is queued and the call returns.  Lanes start on a queued batch as soon as they find no chunk left in
 * the batches before it, so consecutive batches follow each other without the fill / drain gap of a synchronous call (0.4 ms per call at
 * 32-frame chunks: 150 k -> 165 k frames/s at 512 VGA frames per batch with two batches in flight).  The frames and the result arrays
 * must stay valid and untouched until orbx_pool_wait(pool, ticket) has returned (the imgs pointer array itself is copied); wait returns the
 * batch's status and forgets the ticket.  Batches complete in any order; waiting for them in any order is fine. */
int orbx_pool_submit_batch(orbx_pool* pool, const uint8_t* const* imgs, int B, int w, int h, size_t stride,
                           orbx_kp* kps, uint8_t* desc, int cap, int* n, unsigned long long* ticket);
int orbx_pool_wait(orbx_pool* pool, unsigned long long ticket);

/* Waits for the handle's stream and copies the results of the last orbx_extract_device call to HOST buffers laid out
 * as in orbx_extract_batch. */
int orbx_fetch(orbx_extractor* ex, orbx_kp* kps, uint8_t* desc, int cap, int* n);
/* Device pointers of the last results: kps = B*stride_kp orbx_kp, desc = B*stride_kp*32 bytes, counts = B ints
 * (valid until the next extract call on this handle; read them on orbx_stream()). */
int orbx_results_device(orbx_extractor* ex, const orbx_kp** d_kps, const uint8_t** d_desc, const int** d_counts, int* stride_kp);
int orbx_sync(orbx_extractor* ex);
/* The handle's cudaStream_t (as void*), so that callers can record CUDA events around the launches. */
void* orbx_stream(orbx_extractor* ex);
/* Per-stage device timing: with profiling on, every extract call records a CUDA event on the handle's stream after each
 * stage; orbx_stage_ms waits for the last call and returns the 5 stage durations in milliseconds:
 * pyramid (nlevels-1 launches), cell FAST, quadtree, Gaussian blur, orientation+descriptors (1 launch each). */
int orbx_set_profiling(orbx_extractor* ex, int on);
int orbx_stage_ms(orbx_extractor* ex, float* out5);
/* Number of kernels the last extract call launched (for bench.py's gpu_launches). */
int orbx_last_launches(const orbx_extractor* ex);

/* Replaces reads of the public member ORBextractor::mvImagePyramid (include/ORBextractor.h:85; read by
 * Frame::ComputeStereoMatches, src/Frame.cc:488,578,590,595).  Level size of the last call: */
int orbx_level_size(const orbx_extractor* ex, int level, int* w, int* h);
/* Copies level `level` of frame `frame` of the last call to HOST memory.  border = 19 reproduces the reference's padded
 * (w+38)x(h+38) buffer (BORDER_REFLECT_101 of the level itself, src/ORBextractor.cc:1113-1128); border = 0 gives the
 * w x h interior that mvImagePyramid[level] views.  dst_stride in bytes. */
int orbx_pyramid_level(orbx_extractor* ex, int frame, int level, int border, uint8_t* dst, size_t dst_stride);
/* Host mirror of frame 0's pyramid, for callers that read mvImagePyramid after every call (Frame::ComputeStereoMatches): with
 * the mirror on, each extract call also copies the 8 interiors into ONE page-locked buffer of the handle (two copies on the copy
 * engine, beside the FAST and quadtree kernels; no extra synchronisation), and orbx_pyramid_mirror returns the host address and
 * row pitch of a level inside it.  The pointers stay valid until the next extract call or shape change on the handle; read them
 * after a synchronous extract call or orbx_fetch / orbx_sync. */
int orbx_set_pyramid_mirror(orbx_extractor* ex, int on);
int orbx_pyramid_mirror(orbx_extractor* ex, int level, const uint8_t** ptr, size_t* pitch);
/* Zero-copy device view of the same interior (for the device stereo matcher). */
int orbx_pyramid_device(orbx_extractor* ex, int frame, int level, const uint8_t** d_ptr, size_t* pitch);

/* Layout of that device pyramid for the whole batch: plane of (frame, level) = *d_base + frame * *frame_stride. */
int orbx_pyramid_layout(orbx_extractor* ex, int level, const uint8_t** d_base, size_t* pitch, size_t* frame_stride);
/* Frames processed by the last extract call on this handle (0 = none yet). */
int orbx_last_batch(const orbx_extractor* ex);

/* Stage dumps of the last call (parity tests only): blurred level (w x h) and the FAST candidates handed to
 * DistributeOctTree in reference order as (x, y, response) int triples relative to minBorder = 16. */
int orbx_debug_blurred_level(orbx_extractor* ex, int frame, int level, uint8_t* dst, size_t dst_stride);
int orbx_debug_candidates(orbx_extractor* ex, int frame, int level, int* xyr, int cap, int* n);
int orbx_debug_level_count(orbx_extractor* ex, int frame, int level, int* n);

/* ---------------------------------------------------------------------------------------------------------
 * ORBmatcher  (reference include/ORBmatcher.h:37-102, src/ORBmatcher.cc; Frame::ComputeStereoMatches)
 * Descriptors are rows of 32 bytes (CV_8U N x 32, continuous), as ORBmatcher::DescriptorDistance requires
 * (src/ORBmatcher.cc:1649-1665).
 * ------------------------------------------------------------------------------------------------------- */

typedef struct orbx_matcher orbx_matcher;
/* A matcher context = one CUDA stream + growable device scratch.  ORBmatcher itself is stateless apart from
 * (nnratio, checkOri) (src/ORBmatcher.cc:41-43); those stay arguments. */
/* device = CUDA ordinal, or -1 for the calling thread's current device.  Contexts are recycled: orbm_destroy parks the
 * context (stream and arenas alive) in a per-thread free list and the next orbm_create for that device takes it back, so the
 * reference's habit of building an ORBmatcher on the stack per call (src/Tracking.cc:742,907,1012,1329,1502,1541) costs
 * nothing.  orbm_trim really frees the calling thread's parked contexts; they are also freed when the thread exits. */
int orbm_create(int device, orbx_matcher** out);
void orbm_destroy(orbx_matcher* m);
void orbm_trim(void);
int orbm_device(const orbx_matcher* m);
void* orbm_stream(orbx_matcher* m);
/* Number of kernels the last matcher call launched. */
int orbm_last_launches(const orbx_matcher* m);

/* ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1649-1665) for n independent row pairs (a[i], b[i]) -> dist[i]. */
int orbm_descriptor_distance(orbx_matcher* m, const uint8_t* a, const uint8_t* b, int n, int* dist);

/* Best / second-best scan shared by the SearchBy* routines (inner loop src/ORBmatcher.cc:205-226): for every query
 * row scan ALL train rows in index order; strict '<' for best, else-if '<' for second, both start at 256.
 * best_idx = -1 when nt == 0.  HOST buffers. */
int orbm_match_bruteforce(orbx_matcher* m, const uint8_t* q, int nq, const uint8_t* t, int nt,
                          int* best_idx, int* best_dist, int* second_dist);
/* The same scan followed by the acceptance test of src/ORBmatcher.cc:228-230: accepted[i] = best <= th_low &&
 * (float)best < nnratio * (float)second. */
int orbm_match_ratio(orbx_matcher* m, const uint8_t* q, int nq, const uint8_t* t, int nt, int th_low, float nnratio,
                     int* best_idx, int* best_dist, int* second_dist, uint8_t* accepted);
/* Device-resident variant (all pointers DEVICE memory; enqueued on orbm_stream(), not synchronised). */
int orbm_match_bruteforce_device(orbx_matcher* m, const uint8_t* d_q, int nq, const uint8_t* d_t, int nt,
                                 int* d_best_idx, int* d_best_dist, int* d_second_dist);

/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  The scan of orbm_match_bruteforce_device -- same
 * kernel, same results -- with the kernel's scratch in a fixed file-scope device array instead of a context's grown buffers:
 * no handle, and no cudaMalloc / cudaFree / allocating library call on any path.  Works on the calling thread's CURRENT device
 * and enqueues on `stream` (a cudaStream_t, NULL = the default stream) without synchronising; inputs and outputs are device
 * pointers as in orbm_match_bruteforce_device.
 * Limits: nq <= 16384 queries and nt < 2^22 train rows; anything larger returns ORBX_ERR_ARG (use the handle-based entry).
 * The array is 1 MiB + 4 KiB per GPU, held by every GPU that loads the library until the process exits.  All calls on a device
 * SHARE it and it keeps the last call's contents (what must start at zero is cleared on `stream` by each call): two calls that
 * overlap on the GPU -- different streams, or different threads without ordering -- race; order them on one stream or with
 * events.  The first call on each device takes the array's address (cudaGetSymbolAddress): make it outside stream capture. */
int orbm_match_bruteforce_device_noalloc(const uint8_t* d_q, int nq, const uint8_t* d_t, int nt,
                                         int* d_best_idx, int* d_best_dist, int* d_second_dist, void* stream);

/* Candidate-list scan (the windowed / BoW-bucketed searches): query i scans train rows
 * cand_idx[cand_off[i] .. cand_off[i+1]) in that order.  second_idx = -1 when there is none.  HOST buffers. */
int orbm_match_candidates(orbx_matcher* m, const uint8_t* q, int nq, const uint8_t* t, int nt,
                          const int* cand_off, const int* cand_idx,
                          int* best_idx, int* best_dist, int* second_idx, int* second_dist);
/* All candidate distances, for the order-dependent replays (SearchForInitialization's vMatchedDistance skip,
 * src/ORBmatcher.cc:444): dist[c] = distance(query owning slot c, train cand_idx[c]). */
int orbm_candidate_distances(orbx_matcher* m, const uint8_t* q, int nq, const uint8_t* t, int nt,
                             const int* cand_off, const int* cand_idx, int* dist);

/* Frame grid (src/Frame.cc:230-245 AssignFeaturesToGrid, :397-407 PosInGrid) and Frame::GetFeaturesInArea
 * (:342-395) for a batch of queries, on the device.  Produces the CSR candidate lists in the reference's order
 * (cell column outer, cell row inner, insertion order inside a cell).  cand_idx receives at most cap entries;
 * cand_off has nq+1 entries and always holds the true totals (ORBX_ERR_CAPACITY if cand_off[nq] > cap). */
int orbm_grid_candidates(orbx_matcher* m, const orbx_kp* kps, int n, float minX, float maxX, float minY, float maxY,
                         const float* qxy, int nq, const float* qr, int minLevel, int maxLevel, const int* q_level,
                         int* cand_off, int* cand_idx, int cap);

/* ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:405-520).  prev_matched: in/out n1 x 2 floats
 * (vbPrevMatched); matches12: n1 ints (-1 = none).  *nmatches receives the return value of the reference routine. */
int orbm_search_for_initialization(orbx_matcher* m, const orbx_kp* k1, const uint8_t* d1, int n1,
                                   const orbx_kp* k2, const uint8_t* d2, int n2,
                                   float minX, float maxX, float minY, float maxY,
                                   float* prev_matched, int* matches12, int windowSize, float nnratio, int checkOri,
                                   int* nmatches);

/* ORBmatcher::SearchByProjection(Frame &F, const vector<MapPoint*> &vpMapPoints, const float th) (src/ORBmatcher.cc:45-129),
 * the TrackLocalMap search.  SLAM objects are passed as the plain arrays the routine reads from them:
 *   frame F: kps = mvKeysUn, desc = mDescriptors, u_right = mvuRight, occupied[j] = (F.mvpMapPoints[j] && Observations() > 0)
 *            on entry, bounds mnMinX..mnMaxY, sf = mvScaleFactors (nlevels floats);
 *   map points that passed mbTrackInView && !isBad() (:52-56), in vpMapPoints order: proj_xy = (mTrackProjX, mTrackProjY),
 *            proj_xr = mTrackProjXR, level = mnTrackScaleLevel, view_cos = mTrackViewCos, mp_desc = GetDescriptor() rows,
 *            mp_has_obs[i] = Observations() > 0 (a feature taken by such a point is skipped by later points, :88-90).
 * feature_match[j] (n ints) = index of the point this call stored into F.mvpMapPoints[j], -1 = untouched.
 * Candidate windows and Hamming distances run on the device; the order-dependent bookkeeping is replayed on the host. */
int orbm_search_by_projection(orbx_matcher* m, const orbx_kp* kps, const uint8_t* desc, const float* u_right, const uint8_t* occupied,
                              int n, float minX, float maxX, float minY, float maxY, const float* sf, int nlevels,
                              const float* proj_xy, const float* proj_xr, const int* level, const float* view_cos,
                              const uint8_t* mp_desc, const uint8_t* mp_has_obs, int nmp, float th, float nnratio,
                              int* feature_match, int* nmatches);

/* ORBmatcher::SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, const float th, const bool bMono)
 * (src/ORBmatcher.cc:1330-1472), the TrackWithMotionModel search.  Points of the last frame that passed the pMP /
 * mvbOutlier / invzc / image-bounds filters (:1354-1378), in LastFrame order: proj_xy = (u, v), proj_ur = u - mbf*invzc (:1411),
 * last_octave / last_angle = LastFrame.mvKeys[i].octave / mvKeysUn[i].angle, descriptor, has_obs.  mode: 0 = neither,
 * 1 = bForward, 2 = bBackward (:1351-1352).  Outputs as above (after the rotation-consistency pass :1448-1468). */
int orbm_search_by_projection_frame(orbx_matcher* m, const orbx_kp* kps, const uint8_t* desc, const float* u_right,
                                    const uint8_t* occupied, int n, float minX, float maxX, float minY, float maxY,
                                    const float* sf, int nlevels, const float* proj_xy, const float* proj_ur,
                                    const int* last_octave, const float* last_angle, const uint8_t* mp_desc,
                                    const uint8_t* mp_has_obs, int nmp, float th, int mode, int checkOri,
                                    int* feature_match, int* nmatches);

/* ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches) (src/ORBmatcher.cc:159-288; variant 0) and
 * SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12) (:522-655; variant 1).  A DBoW2::FeatureVector is passed as CSR
 * (node = ascending node ids, off = nn+1 offsets, idx = feature indices in stored order).  valid1[i] = feature i of the
 * first key frame has a MapPoint that is not bad (:190-196); valid2 likewise for variant 1 (:571-577; may be NULL for
 * variant 0).  variant 0: matches has n2 entries, matches[j] = feature of the key frame whose MapPoint went to feature j
 * of the frame (accept best <= TH_LOW); variant 1: n1 entries, matches[i] = matched feature of the second key frame
 * (accept best < TH_LOW).  -1 = none. */
int orbm_search_by_bow(orbx_matcher* m, int variant, const orbx_kp* k1, const uint8_t* d1, int n1, const uint8_t* valid1,
                       const int* node1, const int* off1, const int* idx1, int nn1,
                       const orbx_kp* k2, const uint8_t* d2, int n2, const uint8_t* valid2,
                       const int* node2, const int* off2, const int* idx2, int nn2,
                       float nnratio, int checkOri, int* matches, int* nmatches);

/* ---- A Frame that lives on the device.  The reference's Frame keeps mvKeysUn, mDescriptors and the
 * 64 x 48 grid of AssignFeaturesToGrid (src/Frame.cc:230-245, 397-407) on the host and every windowed search walks them; here they are
 * built once per frame in HBM -- the grid by a counting sort on the device that keeps insertion order -- and the searches below take
 * the handle, so a search moves only the MapPoint side and is ONE launch (window enumeration in GetFeaturesInArea's order fused with
 * the Hamming distances), one copy back and the greedy host replay.  At most 8192 keypoints per frame.  A frame belongs to the
 * matcher's device. ---- */
typedef struct orbm_frame orbm_frame;
/* From host arrays: kps = mvKeysUn, desc = mDescriptors (n x 32), bounds = mnMinX, mnMaxX, mnMinY, mnMaxY. */
int orbm_frame_create(orbx_matcher* m, const orbx_kp* kps, const uint8_t* desc, int n, float minX, float maxX, float minY, float maxY,
                      orbm_frame** out);
/* Straight from the device results of frame `frame` of the extractor's last call, without a host round trip of the descriptors:
 * Frame::UndistortKeyPoints (src/Frame.cc:419-449) runs on the device records (dist_coef = mDistCoef as floats; NULL or k1 == 0 leaves
 * the keypoints as extracted, :421-425), then the grid is built.  orbm_frame_keypoints returns the (undistorted) keypoints. */
int orbm_frame_from_extractor(orbx_matcher* m, orbx_extractor* ex, int frame, float fx, float fy, float cx, float cy, const float* dist_coef,
                              int ndist, float minX, float maxX, float minY, float maxY, orbm_frame** out);
void orbm_frame_destroy(orbm_frame* f);
int orbm_frame_size(const orbm_frame* f);
int orbm_frame_keypoints(const orbm_frame* f, orbx_kp* out);
/* The device grid, for tests: cell_start = 64 * 48 + 1 ints over column-major cells (cell = ix * 48 + iy), entries = n feature indices. */
int orbm_frame_grid(orbx_matcher* m, const orbm_frame* f, int* cell_start, int* entries);
/* orbm_search_by_projection / orbm_search_by_projection_frame / orbm_search_for_initialization with the frame side resident on the
 * device: same arguments without the frame's keypoints, descriptors and bounds; same results, index for index. */
int orbm_search_by_projection_dev(orbx_matcher* m, const orbm_frame* f, const float* u_right, const uint8_t* occupied, const float* sf, int nlevels,
                                  const float* proj_xy, const float* proj_xr, const int* level, const float* view_cos, const uint8_t* mp_desc,
                                  const uint8_t* mp_has_obs, int nmp, float th, float nnratio, int* feature_match, int* nmatches);
int orbm_search_by_projection_frame_dev(orbx_matcher* m, const orbm_frame* f, const float* u_right, const uint8_t* occupied, const float* sf,
                                        int nlevels, const float* proj_xy, const float* proj_ur, const int* last_octave, const float* last_angle,
                                        const uint8_t* mp_desc, const uint8_t* mp_has_obs, int nmp, float th, int mode, int checkOri,
                                        int* feature_match, int* nmatches);
int orbm_search_for_initialization_dev(orbx_matcher* m, const orbm_frame* f1, const orbm_frame* f2, float* prev_matched, int* matches12,
                                       int windowSize, float nnratio, int checkOri, int* nmatches);

/* ---- the LocalMapping / LoopClosing / relocalisation searches.  As above, the caller runs the reference's own skip tests
 * and projection (isBad, spAlreadyFound, depth sign, IsInImage, distance range, viewing angle, PredictScale) and passes, per
 * surviving MapPoint, the projection, the predicted level and GetDescriptor(). ---- */

/* The inner loop shared by Fuse (src/ORBmatcher.cc:886-935, :1059-1077) and SearchBySim3 (:1183-1203, :1263-1283), for nq
 * queries in ONE kernel: KeyFrame::GetFeaturesInArea(u, v, r) (src/KeyFrame.cc:572-611) -> skip[idx] (optional, fixed for the
 * call) -> octave in [qlevel-1, qlevel] -> optional chi-square gate of Fuse (inv_level_sigma2 != NULL: needs u_right and qur,
 * :898-922) -> FIRST minimum of the Hamming distance in visiting order.  qlevel[i] < 0 skips query i.
 * best_idx[i] = -1 / best_dist[i] = 256 when no candidate survives. */
int orbm_search_in_area(orbx_matcher* m, const orbx_kp* kps, const uint8_t* desc, const float* u_right, const uint8_t* skip, int n,
                        float minX, float maxX, float minY, float maxY, const float* qxy, const float* qur, const float* qr,
                        const int* qlevel, const uint8_t* qdesc, int nq, const float* inv_level_sigma2, int nlevels,
                        int* best_idx, int* best_dist);

/* ORBmatcher::Fuse(KeyFrame*, vpMapPoints, th) (src/ORBmatcher.cc:827-975; stereo_chi2 = 1, proj_ur = u - bf*invz) and
 * Fuse(KeyFrame*, Scw, vpPoints, th, vpReplacePoint) (:977-1102; stereo_chi2 = 0, u_right / proj_ur / inv_level_sigma2 unused).
 * best_idx[i] = key-frame feature MapPoint i fuses into (bestDist <= TH_LOW), -1 = none; *nfused = how many.  Replace /
 * AddObservation on that feature (:940-957, :1083-1095) stay with the caller: the search reads no state they change. */
int orbm_fuse(orbx_matcher* m, const orbx_kp* kps, const uint8_t* desc, const float* u_right, int n,
              float minX, float maxX, float minY, float maxY, const float* sf, const float* inv_level_sigma2, int nlevels,
              const float* proj_xy, const float* proj_ur, const int* level, const uint8_t* mp_desc, int nmp, float th, int stereo_chi2,
              int* best_idx, int* nfused);

/* ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:1104-1328).  proj12_xy / level12 / mp1_desc: n1 entries, MapPoint of feature i1
 * of key frame 1 projected into key frame 2 (level12[i1] < 0: no point, already matched, bad, or failed :1160-1173);
 * proj21_* likewise for key frame 2 into 1.  match12[i1] = i2 where both directions agree (:1305-1322), else -1. */
int orbm_search_by_sim3(orbx_matcher* m, const orbx_kp* k1, const uint8_t* d1, int n1, const orbx_kp* k2, const uint8_t* d2, int n2,
                        float minX, float maxX, float minY, float maxY, const float* sf, int nlevels,
                        const float* proj12_xy, const int* level12, const uint8_t* mp1_desc,
                        const float* proj21_xy, const int* level21, const uint8_t* mp2_desc, float th, int* match12, int* nfound);

/* ORBmatcher::SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th) (src/ORBmatcher.cc:290-403), loop detection.
 * matched[j] = vpMatched[j] != NULL on entry.  feature_match[j] = point stored into vpMatched[j] by this call, -1 = untouched. */
int orbm_search_by_projection_sim3(orbx_matcher* m, const orbx_kp* kps, const uint8_t* desc, const uint8_t* matched, int n,
                                   float minX, float maxX, float minY, float maxY, const float* sf, int nlevels,
                                   const float* proj_xy, const int* level, const uint8_t* mp_desc, int nmp, float th,
                                   int* feature_match, int* nmatches);

/* ORBmatcher::SearchByProjection(Frame &CurrentFrame, KeyFrame*, sAlreadyFound, th, ORBdist) (src/ORBmatcher.cc:1474-1601),
 * relocalisation.  occupied[j] = CurrentFrame.mvpMapPoints[j] != NULL; level = nPredictedLevel; kf_angle = pKF->mvKeysUn[i].angle. */
int orbm_search_by_projection_kf(orbx_matcher* m, const orbx_kp* kps, const uint8_t* desc, const uint8_t* occupied, int n,
                                 float minX, float maxX, float minY, float maxY, const float* sf, int nlevels,
                                 const float* proj_xy, const int* level, const float* kf_angle, const uint8_t* mp_desc, int nmp,
                                 float th, int ORBdist, int checkOri, int* feature_match, int* nmatches);

/* ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:657-825) with CheckDistEpipolarLine (:140-157).  has_mp[i] = the key
 * frame's feature i already has a MapPoint; u_right = mvuRight; feature vectors as CSR (see orbm_search_by_bow); F12 row-major
 * 3x3 floats; (ex, ey) = epipole in image 2 (:667-672); sf2 / sigma2_2 = pKF2->mvScaleFactors / mvLevelSigma2.
 * matches12[i1] = i2 or -1 (vMatchedPairs = the pairs with matches12[i1] >= 0, in i1 order). */
int orbm_search_for_triangulation(orbx_matcher* m, const orbx_kp* k1, const uint8_t* d1, int n1, const uint8_t* has_mp1, const float* u_right1,
                                  const int* node1, const int* off1, const int* idx1, int nn1,
                                  const orbx_kp* k2, const uint8_t* d2, int n2, const uint8_t* has_mp2, const float* u_right2,
                                  const int* node2, const int* off2, const int* idx2, int nn2,
                                  const float* F12, float ex, float ey, const float* sf2, const float* sigma2_2, int nlevels,
                                  int onlyStereo, int checkOri, int* matches12, int* nmatches);

/* MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:252-312) for npoints MapPoints in one launch.  The descriptors observing
 * point p (pKF->mDescriptors.row(idx) over mObservations in map order, bad key frames left out, :270-276) are rows off[p] ..
 * off[p+1] of desc (32 bytes each); best[p] = index WITHIN those rows of the descriptor with the least median Hamming distance
 * to the others (:297-311; the first such row), -1 for a point without observations (the reference returns early, :266-278). */
int orbm_distinctive_descriptors(orbx_matcher* m, const uint8_t* desc, const int* off, int npoints, int* best);

/* Frame::ComputeStereoMatches (src/Frame.cc:481-655).  left/right = extractor handles whose LAST call processed the
 * two images (their device pyramids are read in place; frame = index inside that batch).  Keypoints/descriptors are
 * HOST arrays as returned by orbx_extract.  maxD is explicit (the reference reads an uninitialised mb, src/Frame.cc:511
 * vs :114; the intended value is fx).  uRight/depth receive nl floats (-1 = no match); *nmatched the survivors. */
int orbm_stereo_matches(orbx_matcher* m, orbx_extractor* left, orbx_extractor* right, int frame,
                        const orbx_kp* kl, const uint8_t* dl, int nl,
                        const orbx_kp* kr, const uint8_t* dr, int nr,
                        float bf, float minD, float maxD, float* uRight, float* depth, int* nmatched);

/* The same association for every stereo pair of a batch, device-resident end to end: pair b = frame b of the LAST extract
 * call on `left` and on `right` (equal batch sizes and shapes).  Keypoints and descriptors are read from the handles'
 * device results, the pyramids in place; candidate search, SAD refinement and the median cut all run on the device.
 * uRight/depth: HOST arrays of B * cap floats (pair b at offset b * cap, -1 = no match), cap >= the largest left keypoint
 * count; nmatched: B ints. */
int orbm_stereo_matches_batch(orbx_matcher* m, orbx_extractor* left, orbx_extractor* right, float bf, float minD, float maxD,
                              float* uRight, float* depth, int cap, int* nmatched);

/* ---- Frame glue between the extractor and the matcher (src/Frame.cc). ----
 * cv::undistortPoints(pts, pts, mK, mDistCoef, cv::Mat(), mK) as called at src/Frame.cc:438 and :464: dist = (k1, k2, p1, p2[, k3, ...]),
 * 4 to 12 coefficients, OpenCV's 5-iteration fixed point in double precision; bit-exact against cv2 4.13 on the golden vectors. */
int orbm_undistort_points(orbx_matcher* m, const float* pts, int n, float fx, float fy, float cx, float cy, const float* dist, int ndist,
                          float* out);
/* Frame::UndistortKeyPoints (src/Frame.cc:419-449): out = mvKeysUn (a copy of kps when k1 == 0; out may alias kps). */
int orbm_undistort_keypoints(orbx_matcher* m, const orbx_kp* kps, int n, float fx, float fy, float cx, float cy, const float* dist, int ndist,
                             orbx_kp* out);
/* Frame::ComputeImageBounds (src/Frame.cc:451-479): bounds4 = (mnMinX, mnMaxX, mnMinY, mnMaxY). */
int orbm_image_bounds(orbx_matcher* m, int cols, int rows, float fx, float fy, float cx, float cy, const float* dist, int ndist, float* bounds4);
/* Frame::ComputeStereoFromRGBD (src/Frame.cc:658-679): im_depth = the CV_32F depth image (h rows of w floats, stride_bytes apart);
 * u_right / depth receive n floats (-1 where the depth is not positive). */
int orbm_stereo_from_rgbd(orbx_matcher* m, const orbx_kp* kps, const orbx_kp* kps_un, int n, const float* im_depth, int w, int h,
                          size_t stride_bytes, float bf, float* u_right, float* depth);

/* ---- ORBVocabulary (include/ORBVocabulary.h:31-32 = DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB>) ----
 * The tree is handed over as arrays, numbered as loadFromTextFile numbers it (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1351-1436):
 * node 0 = root, node i (file line i) has parent[i] < i, a 32-byte descriptor and a weight; children keep file order; the nodes
 * without children are the words, numbered in file order.  scoring / weighting are DBoW2's enums (L1_NORM = 0 ... DOT_PRODUCT = 5;
 * TF_IDF = 0, TF, IDF, BINARY).  Device memory: 40 bytes per node. */
typedef struct orbx_vocabulary orbx_vocabulary;
int orbv_create(int device, int k, int L, int scoring, int weighting, const int* parent, const uint8_t* desc, const double* weight, int nnodes,
                orbx_vocabulary** out);
void orbv_destroy(orbx_vocabulary* v);
int orbv_size(const orbx_vocabulary* v);                  /* number of words, TemplatedVocabulary::size() */
/* transform(feature, word_id, weight, nid, levelsup) (:1231-1270) for n descriptors in one launch.  node_id = the ancestor at level
 * L - levelsup; for a leaf that sits above that level the reference leaves it uninitialised - here it is the leaf itself. */
int orbv_transform_raw(orbx_vocabulary* v, const uint8_t* desc, int n, int levelsup, int* word_id, double* weight, int* node_id);
/* transform(features, BowVector, FeatureVector, levelsup) (:1140-1207) as Frame::ComputeBoW calls it (src/Frame.cc:410-417, levelsup 4).
 * BowVector: *nbow (id, value) pairs in map order into bow_id / bow_val (capacity n).  FeatureVector as CSR in map order: *nfv nodes
 * into fv_node (capacity n), fv_off (capacity n + 1), feature indices into fv_idx (capacity n) - the layout orbm_search_by_bow takes. */
int orbv_transform(orbx_vocabulary* v, const uint8_t* desc, int n, int levelsup, int* bow_id, double* bow_val, int* nbow,
                   int* fv_node, int* fv_off, int* fv_idx, int* nfv);

/* Page-locked host memory for the image / result arrays of the batch calls (orbx_extract_batch writes results straight into
 * page-locked caller arrays, and contiguous page-locked frames travel as one copy).  write_combined = 1 asks for
 * write-combined pages: not cached or snooped by the CPU, meant for buffers the host only WRITES (camera frames on their way
 * to the device); never for result arrays.  Returns ORBX_ERR_CUDA when the allocation fails. */
int orbx_alloc_pinned(size_t bytes, int write_combined, void** out);
int orbx_free_pinned(void* p);

/* Text of the last error raised on the calling thread ("" if none). */
const char* orbx_last_error(void);
/* Library/build identification, e.g. "orbx 0.1 sm_100a". */
const char* orbx_version(void);

#ifdef __cplusplus
}
#endif
#endif /* ORBX_H_ */
