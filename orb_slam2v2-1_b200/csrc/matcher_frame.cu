// matcher_frame.cu -- frame-level searches of the C ABI: Frame grid + GetFeaturesInArea, SearchForInitialization,
// ComputeStereoMatches.  (being written: entry points present so that the ABI is complete)
#include "matcher_internal.cuh"

using namespace orbx;

extern "C" {

int orbm_grid_candidates(orbx_matcher*, const orbx_kp*, int, float, float, float, float, const float*, int, const float*, int, int,
                         const int*, int*, int*, int)
{ set_error("orbm_grid_candidates: not implemented yet"); return ORBX_ERR_STATE; }

int orbm_search_for_initialization(orbx_matcher*, const orbx_kp*, const uint8_t*, int, const orbx_kp*, const uint8_t*, int,
                                   float, float, float, float, float*, int*, int, float, int, int*)
{ set_error("orbm_search_for_initialization: not implemented yet"); return ORBX_ERR_STATE; }

int orbm_stereo_matches(orbx_matcher*, orbx_extractor*, orbx_extractor*, int, const orbx_kp*, const uint8_t*, int,
                        const orbx_kp*, const uint8_t*, int, float, float, float, float*, float*, int*)
{ set_error("orbm_stereo_matches: not implemented yet"); return ORBX_ERR_STATE; }

}
