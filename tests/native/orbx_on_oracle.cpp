// orbx_on_oracle.cpp -- TEST INFRASTRUCTURE, never shipped: the entry points of include/orbx.h that the drop-in headers call in
// tests/native/reference_matcher_test.cpp, implemented by the CPU oracle (oracle/orb_oracle.h).  Linked INSTEAD of liborbx.so it lets
// that test run without a GPU: the headers' flattening code (include/ORBmatcher.h, ORBmatcher_reference_api.inl) and the oracle's
// searches against the UNMODIFIED reference src/ORBmatcher.cc.  The product never sees this file (no CPU fallback exists in liborbx.so).
#include <string.h>
#include <vector>
#include "orbx.h"
#include "orb_oracle.h"

struct orbx_extractor {
    orbo_extractor* o;
    int nlevels, nfeatures;
    std::vector<std::vector<uint8_t> > levels;      // host "mirror" of the last call's pyramid interiors
    std::vector<int> lw, lh;
    bool mirror;
};
struct orbx_matcher { int unused; };

static const orbo_kp* K(const orbx_kp* k) { return reinterpret_cast<const orbo_kp*>(k); }     // same 28-byte layout

extern "C" {

const char* orbx_last_error(void) { return "orbx_on_oracle: no error text"; }

int orbx_create(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST, int, int, orbx_extractor** out)
{
    orbx_extractor* ex = new orbx_extractor;
    ex->o = orbo_create(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST);
    ex->nlevels = nlevels;
    ex->nfeatures = nfeatures;
    ex->mirror = true;
    *out = ex;
    return ORBX_OK;
}
void orbx_destroy(orbx_extractor* ex)
{
    if (!ex) return;
    orbo_destroy(ex->o);
    delete ex;
}
int orbx_scale_tables(const orbx_extractor* ex, float* s, float* is, float* g, float* ig) { orbo_scale_tables(ex->o, s, is, g, ig); return ORBX_OK; }
int orbx_max_keypoints(const orbx_extractor* ex) { return ex->nfeatures + 64 * ex->nlevels; }
int orbx_extract(orbx_extractor* ex, const uint8_t* img, int w, int h, size_t stride, orbx_kp* kps, uint8_t* desc, int cap, int* n)
{
    if (!img || w <= 0 || h <= 0) { *n = 0; return ORBX_OK; }
    const int got = orbo_extract(ex->o, img, w, h, stride, reinterpret_cast<orbo_kp*>(kps), desc, cap);
    if (got > cap) return ORBX_ERR_CAPACITY;
    *n = got;
    ex->levels.assign(ex->nlevels, std::vector<uint8_t>());
    ex->lw.assign(ex->nlevels, 0);
    ex->lh.assign(ex->nlevels, 0);
    for (int l = 0; l < ex->nlevels; l++) {
        int lw, lh;
        orbo_level_size(ex->o, l, &lw, &lh);
        std::vector<uint8_t> pad((size_t)(lw + 38) * (lh + 38));
        orbo_level_padded(ex->o, l, pad.data());
        ex->levels[l].resize((size_t)lw * lh);
        for (int y = 0; y < lh; y++) memcpy(&ex->levels[l][(size_t)y * lw], &pad[(size_t)(y + 19) * (lw + 38) + 19], lw);
        ex->lw[l] = lw; ex->lh[l] = lh;
    }
    return ORBX_OK;
}
int orbx_extract_color_batch(orbx_extractor*, const uint8_t* const*, int, int, int, size_t, int, int, orbx_kp*, uint8_t*, int, int*) { return ORBX_ERR_STATE; }
int orbx_level_size(const orbx_extractor* ex, int level, int* w, int* h)
{
    if (level < 0 || level >= (int)ex->lw.size()) return ORBX_ERR_ARG;
    *w = ex->lw[level]; *h = ex->lh[level];
    return ORBX_OK;
}
int orbx_set_pyramid_mirror(orbx_extractor* ex, int on) { ex->mirror = on != 0; return ORBX_OK; }
int orbx_pyramid_mirror(orbx_extractor* ex, int level, const uint8_t** ptr, size_t* pitch)
{
    if (level < 0 || level >= (int)ex->levels.size()) return ORBX_ERR_ARG;
    *ptr = ex->levels[level].data(); *pitch = (size_t)ex->lw[level];
    return ORBX_OK;
}

int orbm_create(int, orbx_matcher** out) { *out = new orbx_matcher; return ORBX_OK; }
void orbm_destroy(orbx_matcher* m) { delete m; }

int orbm_search_for_initialization(orbx_matcher*, const orbx_kp* k1, const uint8_t* d1, int n1, const orbx_kp* k2, const uint8_t* d2, int n2, float minX,
                                   float maxX, float minY, float maxY, float* prev, int* m12, int windowSize, float nnratio, int checkOri, int* nmatches)
{
    *nmatches = orbo_search_for_initialization(K(k1), d1, n1, K(k2), d2, n2, minX, maxX, minY, maxY, prev, m12, windowSize, nnratio, checkOri);
    return ORBX_OK;
}
int orbm_search_by_projection(orbx_matcher*, const orbx_kp* kps, const uint8_t* desc, const float* u_right, const uint8_t* occupied, int n, float minX,
                              float maxX, float minY, float maxY, const float* sf, int nlevels, const float* xy, const float* xr, const int* level,
                              const float* view_cos, const uint8_t* mp_desc, const uint8_t* has_obs, int nmp, float th, float nnratio, int* fm,
                              int* nmatches)
{
    *nmatches = orbo_search_by_projection(K(kps), desc, u_right, occupied, n, minX, maxX, minY, maxY, sf, nlevels, xy, xr, level, view_cos, mp_desc, has_obs,
                                          nmp, th, nnratio, fm);
    return ORBX_OK;
}
int orbm_search_by_projection_frame(orbx_matcher*, const orbx_kp* kps, const uint8_t* desc, const float* u_right, const uint8_t* occupied, int n,
                                    float minX, float maxX, float minY, float maxY, const float* sf, int nlevels, const float* xy, const float* ur,
                                    const int* oct, const float* ang, const uint8_t* mp_desc, const uint8_t* has_obs, int nmp, float th, int mode,
                                    int checkOri, int* fm, int* nmatches)
{
    *nmatches = orbo_search_by_projection_frame(K(kps), desc, u_right, occupied, n, minX, maxX, minY, maxY, sf, nlevels, xy, ur, oct, ang, mp_desc, has_obs,
                                                nmp, th, mode, checkOri, fm);
    return ORBX_OK;
}
int orbm_search_by_bow(orbx_matcher*, int variant, const orbx_kp* k1, const uint8_t* d1, int n1, const uint8_t* valid1, const int* node1, const int* off1,
                       const int* idx1, int nn1, const orbx_kp* k2, const uint8_t* d2, int n2, const uint8_t* valid2, const int* node2, const int* off2,
                       const int* idx2, int nn2, float nnratio, int checkOri, int* matches, int* nmatches)
{
    *nmatches = orbo_search_by_bow(variant, K(k1), d1, n1, valid1, node1, off1, idx1, nn1, K(k2), d2, n2, valid2, node2, off2, idx2, nn2, nnratio, checkOri,
                                   matches);
    return ORBX_OK;
}
int orbm_fuse(orbx_matcher*, const orbx_kp* kps, const uint8_t* desc, const float* u_right, int n, float minX, float maxX, float minY, float maxY,
              const float* sf, const float* ils2, int nlevels, const float* xy, const float* ur, const int* level, const uint8_t* mp_desc, int nmp, float th,
              int stereo_chi2, int* best_idx, int* nfused)
{
    *nfused = orbo_fuse(K(kps), desc, u_right, n, minX, maxX, minY, maxY, sf, ils2, nlevels, xy, ur, level, mp_desc, nmp, th, stereo_chi2, best_idx);
    return ORBX_OK;
}
int orbm_search_by_sim3(orbx_matcher*, const orbx_kp* k1, const uint8_t* d1, int n1, const orbx_kp* k2, const uint8_t* d2, int n2, float minX, float maxX,
                        float minY, float maxY, const float* sf, int nlevels, const float* xy12, const int* l12, const uint8_t* dd1, const float* xy21,
                        const int* l21, const uint8_t* dd2, float th, int* match12, int* nfound)
{
    *nfound = orbo_search_by_sim3(K(k1), d1, n1, K(k2), d2, n2, minX, maxX, minY, maxY, sf, nlevels, xy12, l12, dd1, xy21, l21, dd2, th, match12);
    return ORBX_OK;
}
int orbm_search_by_projection_sim3(orbx_matcher*, const orbx_kp* kps, const uint8_t* desc, const uint8_t* matched, int n, float minX, float maxX, float minY,
                                   float maxY, const float* sf, int nlevels, const float* xy, const int* level, const uint8_t* mp_desc, int nmp, float th,
                                   int* fm, int* nmatches)
{
    *nmatches = orbo_search_by_projection_sim3(K(kps), desc, matched, n, minX, maxX, minY, maxY, sf, nlevels, xy, level, mp_desc, nmp, th, fm);
    return ORBX_OK;
}
int orbm_search_by_projection_kf(orbx_matcher*, const orbx_kp* kps, const uint8_t* desc, const uint8_t* occupied, int n, float minX, float maxX, float minY,
                                 float maxY, const float* sf, int nlevels, const float* xy, const int* level, const float* ang, const uint8_t* mp_desc,
                                 int nmp, float th, int ORBdist, int checkOri, int* fm, int* nmatches)
{
    *nmatches = orbo_search_by_projection_kf(K(kps), desc, occupied, n, minX, maxX, minY, maxY, sf, nlevels, xy, level, ang, mp_desc, nmp, th, ORBdist,
                                             checkOri, fm);
    return ORBX_OK;
}
int orbm_search_for_triangulation(orbx_matcher*, const orbx_kp* k1, const uint8_t* d1, int n1, const uint8_t* has1, const float* ur1, const int* node1,
                                  const int* off1, const int* idx1, int nn1, const orbx_kp* k2, const uint8_t* d2, int n2, const uint8_t* has2,
                                  const float* ur2, const int* node2, const int* off2, const int* idx2, int nn2, const float* F12, float ex, float ey,
                                  const float* sf2, const float* sig2, int nlevels, int onlyStereo, int checkOri, int* m12, int* nmatches)
{
    *nmatches = orbo_search_for_triangulation(K(k1), d1, n1, has1, ur1, node1, off1, idx1, nn1, K(k2), d2, n2, has2, ur2, node2, off2, idx2, nn2, F12, ex, ey,
                                              sf2, sig2, nlevels, onlyStereo, checkOri, m12);
    return ORBX_OK;
}

// ---- ORBVocabulary (include/ORBVocabulary.h)
int orbv_create(int, int k, int L, int scoring, int weighting, const int* parent, const uint8_t* desc, const double* weight, int nnodes, orbx_vocabulary** out)
{
    *out = reinterpret_cast<orbx_vocabulary*>(orbo_voc_create(k, L, scoring, weighting, parent, desc, weight, nnodes));
    return *out ? ORBX_OK : ORBX_ERR_ARG;
}
void orbv_destroy(orbx_vocabulary* v) { if (v) orbo_voc_destroy(reinterpret_cast<orbo_vocabulary*>(v)); }
int orbv_size(const orbx_vocabulary* v) { return v ? orbo_voc_size(reinterpret_cast<const orbo_vocabulary*>(v)) : 0; }
int orbv_transform(orbx_vocabulary* v, const uint8_t* desc, int n, int levelsup, int* bow_id, double* bow_val, int* nbow, int* fv_node, int* fv_off,
                   int* fv_idx, int* nfv)
{
    *nbow = orbo_voc_transform(reinterpret_cast<const orbo_vocabulary*>(v), desc, n, levelsup, bow_id, bow_val, fv_node, fv_off, fv_idx, nfv);
    return ORBX_OK;
}

}  // extern "C"
