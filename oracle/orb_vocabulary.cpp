// orb_vocabulary.cpp -- CPU ORACLE (test infrastructure, not product).
// Restatement of ORBVocabulary::transform(features, BowVector, FeatureVector, levelsup) as Frame::ComputeBoW / KeyFrame::ComputeBoW call
// it (Frame.cc:410-417): DBoW2::TemplatedVocabulary::transform (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1140-1207), the tree descent
// (:1231-1270), FORB::distance (FORB.cpp:81-100, the same bit count as ORBmatcher::DescriptorDistance), BowVector::addWeight /
// addIfNotExist / normalize (BowVector.cpp:34-84), FeatureVector::addFeature (FeatureVector.cpp:31-45), and the node / word numbering of
// loadFromTextFile (:1351-1436: node ids in file order, children in file order, word ids = leaves in file order).
// The reference ships no vocabulary file (Vocabulary/ORBvoc.txt is downloaded separately), so tests build synthetic trees.
//
// One indeterminacy is resolved explicitly: when a leaf sits above level L - levelsup the reference leaves *nid uninitialised
// (:1239-1265 never assign it); here the node id reported for such a feature is the leaf itself.
#include "orb_oracle.h"
#include "orb_internal.h"
#include <cmath>
#include <map>
#include <vector>

struct orbo_vocabulary {
    int k, L, scoring, weighting;
    std::vector<std::vector<int>> children;
    std::vector<uint8_t> desc;
    std::vector<double> weight;
    std::vector<int> word_id;
};

extern "C" {

orbo_vocabulary* orbo_voc_create(int k, int L, int scoring, int weighting, const int* parent, const uint8_t* desc, const double* weight, int nnodes)
{
    orbo_vocabulary* v = new orbo_vocabulary;
    v->k = k; v->L = L; v->scoring = scoring; v->weighting = weighting;
    v->children.resize(nnodes);
    v->desc.assign(desc, desc + 32 * (size_t)nnodes);
    v->weight.assign(weight, weight + nnodes);
    v->word_id.assign(nnodes, -1);
    for (int i = 1; i < nnodes; i++) v->children[parent[i]].push_back(i);
    int nwords = 0;
    for (int i = 1; i < nnodes; i++)
        if (v->children[i].empty()) v->word_id[i] = nwords++;
    return v;
}
void orbo_voc_destroy(orbo_vocabulary* v) { delete v; }
// TemplatedVocabulary::size() (:838-842): the number of words
int orbo_voc_size(const orbo_vocabulary* v)
{
    int n = 0;
    for (size_t i = 1; i < v->word_id.size(); i++) n += v->word_id[i] >= 0;
    return n;
}

// transform(feature, word_id, weight, nid, levelsup) (:1231-1270)
void orbo_voc_transform_raw(const orbo_vocabulary* v, const uint8_t* desc, int n, int levelsup, int* word_id, double* weight, int* node_id)
{
    for (int f = 0; f < n; f++) {
        const uint8_t* feature = desc + 32 * (size_t)f;
        const int nid_level = v->L - levelsup;
        int nid = -1;
        if (nid_level <= 0) nid = 0;
        int final_id = 0, current_level = 0;
        do {
            ++current_level;
            const std::vector<int>& nodes = v->children[final_id];
            final_id = nodes[0];
            double best_d = orbo::descriptor_distance(feature, &v->desc[32 * (size_t)final_id]);
            for (size_t c = 1; c < nodes.size(); c++) {
                const int id = nodes[c];
                const double d = orbo::descriptor_distance(feature, &v->desc[32 * (size_t)id]);
                if (d < best_d) { best_d = d; final_id = id; }
            }
            if (current_level == nid_level) nid = final_id;
        } while (!v->children[final_id].empty());
        if (nid < 0) nid = final_id;                      // see the header: undefined in the reference
        word_id[f] = v->word_id[final_id];
        weight[f] = v->weight[final_id];
        node_id[f] = nid;
    }
}

// transform(features, v, fv, levelsup) (:1140-1207).  Outputs: the BowVector as (id, value) in map order, the FeatureVector as CSR
// (node ids ascending, offsets, feature indices in insertion order).  Returns the number of BowVector entries.
int orbo_voc_transform(const orbo_vocabulary* voc, const uint8_t* desc, int n, int levelsup, int* bow_id, double* bow_val,
                       int* fv_node, int* fv_off, int* fv_idx, int* nfv)
{
    std::vector<int> wid(n), nid(n);
    std::vector<double> w(n);
    orbo_voc_transform_raw(voc, desc, n, levelsup, wid.data(), w.data(), nid.data());
    std::map<unsigned, double> v;
    std::map<unsigned, std::vector<unsigned>> fv;
    const bool must = voc->scoring != 5;                   // DotProductScoring does not normalise (ScoringObject.h:73-90)
    const bool l2 = voc->scoring == 1;
    const bool tf = voc->weighting == 0 || voc->weighting == 1;      // TF_IDF, TF
    for (int i = 0; i < n; i++) {
        if (!(w[i] > 0)) continue;                        // stopped word
        if (tf) v[wid[i]] += w[i];                       // addWeight
        else v.insert(std::make_pair((unsigned)wid[i], w[i]));      // addIfNotExist
        fv[nid[i]].push_back(i);
    }
    if (tf && !v.empty() && !must) {
        const double nd = (double)v.size();
        for (auto& e : v) e.second /= nd;
    }
    if (must) {                                           // BowVector::normalize
        double norm = 0.0;
        if (!l2) { for (auto& e : v) norm += std::fabs(e.second); }
        else { for (auto& e : v) norm += e.second * e.second; norm = std::sqrt(norm); }
        if (norm > 0.0) for (auto& e : v) e.second /= norm;
    }
    int nb = 0;
    for (auto& e : v) { bow_id[nb] = (int)e.first; bow_val[nb] = e.second; nb++; }
    int nn = 0, pos = 0;
    fv_off[0] = 0;
    for (auto& e : fv) {
        fv_node[nn] = (int)e.first;
        for (unsigned i : e.second) fv_idx[pos++] = (int)i;
        fv_off[++nn] = pos;
    }
    *nfv = nn;
    return nb;
}

}  // extern "C"
