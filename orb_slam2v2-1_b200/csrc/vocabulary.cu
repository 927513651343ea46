// vocabulary.cu -- ORBVocabulary::transform (reference include/ORBVocabulary.h:31-32 = DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB>;
// Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1140-1270) as called by Frame::ComputeBoW / KeyFrame::ComputeBoW (src/Frame.cc:410-417):
// every descriptor of a frame descends the k-ary tree by Hamming distance (FORB::distance, FORB.cpp:81-100) in ONE launch, a warp per
// descriptor, the children of a node compared lane-parallel.  The BowVector / FeatureVector maps are assembled on the host in feature
// order (their floating-point sums depend on it).
#include "matcher_internal.cuh"
#include <math.h>
#include <string.h>
#include <map>
#include <vector>

struct orbx_vocabulary {
    int device = 0, k = 0, L = 0, scoring = 0, weighting = 0, nnodes = 0, nwords = 0;
    cudaStream_t stream = nullptr;
    int* d_childStart = nullptr;       // nnodes + 1
    int* d_childId = nullptr;          // node ids in child-list order (children of node i at childStart[i] .. childStart[i+1])
    uint4* d_childDesc = nullptr;      // their descriptors in the same order: the children of a node are contiguous
    int* d_wordId = nullptr;           // nnodes (-1 for inner nodes)
    double* d_weight = nullptr;        // nnodes
    orbx::Arena arena;
};

namespace orbx {

__global__ void __launch_bounds__(256) k_voc_descend(const uint4* __restrict__ feat, int n, const int* __restrict__ childStart,
                                                     const int* __restrict__ childId, const uint4* __restrict__ childDesc,
                                                     const int* __restrict__ wordId, const double* __restrict__ weight, int nidLevel,
                                                     int* __restrict__ outWord, double* __restrict__ outWeight, int* __restrict__ outNode)
{
    const int f = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (f >= n) return;
    const uint4 qa = __ldg(&feat[2 * (size_t)f]), qb = __ldg(&feat[2 * (size_t)f + 1]);
    int node = 0, level = 0, nid = nidLevel <= 0 ? 0 : -1;                         // :1239-1243
    for (;;) {
        const int s = childStart[node], e = childStart[node + 1];
        if (s == e) break;                                                          // isLeaf()
        ++level;
        uint32_t key = 0xFFFFFFFFu;
        for (int c = s + lane; c < e; c += 32) {
            const int d = hamming256(qa, qb, __ldg(&childDesc[2 * (size_t)c]), __ldg(&childDesc[2 * (size_t)c + 1]));
            key = min(key, ((uint32_t)d << 16) | (uint32_t)(c - s));                // first minimum in child order (`d < best_d`)
        }
        key = __reduce_min_sync(0xFFFFFFFFu, key);
        node = childId[s + (int)(key & 0xFFFFu)];
        if (level == nidLevel) nid = node;                                          // :1264-1265
    }
    if (lane == 0) {
        outWord[f] = wordId[node];
        outWeight[f] = weight[node];
        outNode[f] = nid < 0 ? node : nid;      // a leaf above nidLevel: the reference leaves *nid unset; report the leaf
    }
}

static int descend(orbx_vocabulary* v, const uint8_t* desc, int n, int levelsup, int* word, double* weight, int* node)
{
    ORBX_CUDA(cudaSetDevice(v->device));
    Carver cv(v->arena);
    const size_t oF = cv.take((size_t)n * 32), oW = cv.take(sizeof(int) * (size_t)n), oV = cv.take(sizeof(double) * (size_t)n),
                 oN = cv.take(sizeof(int) * (size_t)n);
    int rc = v->arena.reserve(cv.total());
    if (!rc) rc = v->arena.mirror();
    if (rc) return rc;
    uint8_t* B = v->arena.base;
    cudaStream_t s = v->stream;
    Xfer x(v->arena);
    x.up(oF, desc, (size_t)n * 32);
    ORBX_CUDA(x.send());
    k_voc_descend<<<(n + 7) / 8, 256, 0, s>>>((const uint4*)(B + oF), n, v->d_childStart, v->d_childId, v->d_childDesc, v->d_wordId, v->d_weight,
                                              v->L - levelsup, (int*)(B + oW), (double*)(B + oV), (int*)(B + oN));
    ORBX_CUDA(cudaGetLastError());
    x.down(word, oW, sizeof(int) * (size_t)n);
    x.down(weight, oV, sizeof(double) * (size_t)n);
    x.down(node, oN, sizeof(int) * (size_t)n);
    ORBX_CUDA(x.fetch());
    return ORBX_OK;
}

}  // namespace orbx

using namespace orbx;

extern "C" {

int orbv_create(int device, int k, int L, int scoring, int weighting, const int* parent, const uint8_t* desc, const double* weight, int nnodes,
                orbx_vocabulary** out)
{
    if (!out) { set_error("orbv_create: bad argument"); return ORBX_ERR_ARG; }
    *out = nullptr;
    if (!parent || !desc || !weight || nnodes < 2 || L < 1 || scoring < 0 || scoring > 5 || weighting < 0 || weighting > 3) {
        set_error("orbv_create: bad argument (a vocabulary has a root and at least one word)");
        return ORBX_ERR_ARG;
    }
    std::vector<int> start(nnodes + 1, 0);
    for (int i = 1; i < nnodes; i++) {
        if (parent[i] < 0 || parent[i] >= i) { set_error("orbv_create: parent[%d] = %d (parents must precede their children)", i, parent[i]); return ORBX_ERR_ARG; }
        start[parent[i] + 1]++;
    }
    for (int i = 0; i < nnodes; i++) {
        if (start[i + 1] > 65535) { set_error("orbv_create: node %d has more than 65535 children", i); return ORBX_ERR_ARG; }
        start[i + 1] += start[i];
    }
    const int nchild = nnodes - 1;
    std::vector<int> childId(nchild), fill(start.begin(), start.end() - 1), wordId(nnodes, -1);
    for (int i = 1; i < nnodes; i++) childId[fill[parent[i]]++] = i;                 // children in node-id (= file) order, :1404-1405
    std::vector<uint8_t> childDesc((size_t)nchild * 32);
    for (int c = 0; c < nchild; c++) memcpy(&childDesc[(size_t)c * 32], desc + (size_t)childId[c] * 32, 32);
    int nwords = 0;
    for (int i = 1; i < nnodes; i++)
        if (start[i + 1] == start[i]) wordId[i] = nwords++;                          // :1424-1431
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) { set_error("orbv_create: cudaSetDevice(%d): %s", device, cudaGetErrorString(e)); return ORBX_ERR_CUDA; }
    orbx_vocabulary* v = new orbx_vocabulary;
    v->device = device; v->k = k; v->L = L; v->scoring = scoring; v->weighting = weighting; v->nnodes = nnodes; v->nwords = nwords;
    auto bail = [&](cudaError_t err, const char* what) { set_error("orbv_create: %s: %s", what, cudaGetErrorString(err)); };
#define ORBV_TRY(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { bail(e_, #call); orbv_destroy(v); return ORBX_ERR_CUDA; } } while (0)
    ORBV_TRY(cudaStreamCreateWithFlags(&v->stream, cudaStreamNonBlocking));
    v->arena.stream = v->stream;
    ORBV_TRY(cudaMalloc(&v->d_childStart, sizeof(int) * (size_t)(nnodes + 1)));
    ORBV_TRY(cudaMalloc(&v->d_childId, sizeof(int) * (size_t)nchild));
    ORBV_TRY(cudaMalloc(&v->d_childDesc, (size_t)nchild * 32));
    ORBV_TRY(cudaMalloc(&v->d_wordId, sizeof(int) * (size_t)nnodes));
    ORBV_TRY(cudaMalloc(&v->d_weight, sizeof(double) * (size_t)nnodes));
    ORBV_TRY(cudaMemcpyAsync(v->d_childStart, start.data(), sizeof(int) * (size_t)(nnodes + 1), cudaMemcpyHostToDevice, v->stream));
    ORBV_TRY(cudaMemcpyAsync(v->d_childId, childId.data(), sizeof(int) * (size_t)nchild, cudaMemcpyHostToDevice, v->stream));
    ORBV_TRY(cudaMemcpyAsync(v->d_childDesc, childDesc.data(), (size_t)nchild * 32, cudaMemcpyHostToDevice, v->stream));
    ORBV_TRY(cudaMemcpyAsync(v->d_wordId, wordId.data(), sizeof(int) * (size_t)nnodes, cudaMemcpyHostToDevice, v->stream));
    ORBV_TRY(cudaMemcpyAsync(v->d_weight, weight, sizeof(double) * (size_t)nnodes, cudaMemcpyHostToDevice, v->stream));
    ORBV_TRY(cudaStreamSynchronize(v->stream));
#undef ORBV_TRY
    *out = v;
    return ORBX_OK;
}

void orbv_destroy(orbx_vocabulary* v)
{
    if (!v) return;
    cudaSetDevice(v->device);
    if (v->stream) cudaStreamSynchronize(v->stream);
    cudaFree(v->d_childStart); cudaFree(v->d_childId); cudaFree(v->d_childDesc); cudaFree(v->d_wordId); cudaFree(v->d_weight);
    v->arena.release();
    if (v->stream) cudaStreamDestroy(v->stream);
    delete v;
}

int orbv_size(const orbx_vocabulary* v) { return v ? v->nwords : 0; }

int orbv_transform_raw(orbx_vocabulary* v, const uint8_t* desc, int n, int levelsup, int* word_id, double* weight, int* node_id)
{
    if (!v || n < 0 || (n > 0 && (!desc || !word_id || !weight || !node_id))) { set_error("orbv_transform_raw: bad argument"); return ORBX_ERR_ARG; }
    if (n == 0) return ORBX_OK;
    return descend(v, desc, n, levelsup, word_id, weight, node_id);
}

int orbv_transform(orbx_vocabulary* voc, const uint8_t* desc, int n, int levelsup, int* bow_id, double* bow_val, int* nbow,
                   int* fv_node, int* fv_off, int* fv_idx, int* nfv)
{
    if (!voc || n < 0 || !nbow || !nfv || !fv_off || (n > 0 && (!desc || !bow_id || !bow_val || !fv_node || !fv_idx))) {
        set_error("orbv_transform: bad argument");
        return ORBX_ERR_ARG;
    }
    *nbow = 0;
    *nfv = 0;
    fv_off[0] = 0;
    if (n == 0) return ORBX_OK;
    std::vector<int> wid(n), nid(n);
    std::vector<double> w(n);
    int rc = descend(voc, desc, n, levelsup, wid.data(), w.data(), nid.data());
    if (rc) return rc;
    // TemplatedVocabulary.h:1152-1206 on the per-feature results, in feature order
    std::map<unsigned, double> v;
    std::map<unsigned, std::vector<unsigned> > fv;
    const bool must = voc->scoring != 5, l2 = voc->scoring == 1;      // ScoringObject.h:73-90: only DotProduct skips the norm, only L2 uses L2
    const bool tf = voc->weighting == 0 || voc->weighting == 1;
    for (int i = 0; i < n; i++) {
        if (!(w[i] > 0)) continue;                                    // stopped word (:1170)
        if (tf) {                                                     // BowVector::addWeight
            std::map<unsigned, double>::iterator it = v.lower_bound((unsigned)wid[i]);
            if (it != v.end() && it->first == (unsigned)wid[i]) it->second += w[i];
            else v.insert(it, std::make_pair((unsigned)wid[i], w[i]));
        } else {                                                      // BowVector::addIfNotExist
            v.insert(std::make_pair((unsigned)wid[i], w[i]));
        }
        fv[(unsigned)nid[i]].push_back((unsigned)i);                  // FeatureVector::addFeature
    }
    if (tf && !v.empty() && !must) {
        const double nd = (double)v.size();
        for (std::map<unsigned, double>::iterator it = v.begin(); it != v.end(); ++it) it->second /= nd;
    }
    if (must) {                                                       // BowVector::normalize (BowVector.cpp:62-84)
        double norm = 0.0;
        if (!l2) for (std::map<unsigned, double>::iterator it = v.begin(); it != v.end(); ++it) norm += fabs(it->second);
        else {
            for (std::map<unsigned, double>::iterator it = v.begin(); it != v.end(); ++it) norm += it->second * it->second;
            norm = sqrt(norm);
        }
        if (norm > 0.0) for (std::map<unsigned, double>::iterator it = v.begin(); it != v.end(); ++it) it->second /= norm;
    }
    int nb = 0;
    for (std::map<unsigned, double>::iterator it = v.begin(); it != v.end(); ++it, ++nb) { bow_id[nb] = (int)it->first; bow_val[nb] = it->second; }
    int nn = 0, pos = 0;
    for (std::map<unsigned, std::vector<unsigned> >::iterator it = fv.begin(); it != fv.end(); ++it) {
        fv_node[nn] = (int)it->first;
        for (size_t j = 0; j < it->second.size(); j++) fv_idx[pos++] = (int)it->second[j];
        fv_off[++nn] = pos;
    }
    *nbow = nb;
    *nfv = nn;
    return ORBX_OK;
}

}  // extern "C"
