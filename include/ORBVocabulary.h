// ORBVocabulary.h -- drop-in for the part of the reference's ORBVocabulary the front end uses (kimwin2/ORB_SLAM2v2-1
// include/ORBVocabulary.h:31-32: DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB>): loadFromTextFile
// (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1351-1436) and transform(features, BowVector, FeatureVector, levelsup) (:1140-1207) as
// Frame::ComputeBoW / KeyFrame::ComputeBoW call it (src/Frame.cc:410-417), loadFromBinaryFile (:1467-1511; System.cc:84) and
// score(v1, v2) (:1213-1218 -> ScoringObject.cpp; KeyFrameDatabase.cc:133,249, LoopClosing.cc:144).  The tree lives on the device (orbv_* in
// include/orbx.h); scoring two BowVectors is a merge of two sorted maps and stays on the host.  BowVector / FeatureVector below are the
// std::map types DBoW2's classes derive from, so the reference's DBoW2::BowVector / FeatureVector members bind to every signature here.
#ifndef ORBVOCABULARY_H
#define ORBVOCABULARY_H

#include <cfloat>
#include <cmath>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>
#include "orbx.h"
#include "orbx_cv_shim.h"

namespace ORB_SLAM2
{

typedef std::map<unsigned int, double> BowVector;                         // DBoW2::BowVector: WordId -> WordValue
typedef std::map<unsigned int, std::vector<unsigned int> > FeatureVector; // DBoW2::FeatureVector: NodeId -> feature indices

class ORBVocabulary
{
public:
    explicit ORBVocabulary(int device = 0) : v_(nullptr), device_(device), scoring_(0) {}
    ~ORBVocabulary() { orbv_destroy(v_); }
    ORBVocabulary(const ORBVocabulary&) = delete;
    ORBVocabulary& operator=(const ORBVocabulary&) = delete;

    // Same text format and return convention as the reference (System.cc:79-92 checks the bool).
    bool loadFromTextFile(const std::string& filename)
    {
        std::ifstream f(filename.c_str());
        if (!f.is_open()) return false;
        std::string s;
        std::getline(f, s);
        std::stringstream ss(s);
        int k = -1, L = -1, n1 = -1, n2 = -1;
        ss >> k >> L >> n1 >> n2;
        if (k < 0 || k > 20 || L < 1 || L > 10 || n1 < 0 || n1 > 5 || n2 < 0 || n2 > 3) return false;      // :1372-1376
        std::vector<int> parent(1, 0);
        std::vector<uint8_t> desc(32, 0);
        std::vector<double> weight(1, 0.0);
        while (std::getline(f, s)) {
            std::stringstream sn(s);
            int pid, leaf;
            // an empty (last) line is skipped.  The reference turns it into one more child of the root whose leaf flag, weight and
            // descriptor are never written (TemplatedVocabulary.h:1394-1431 after a failed extraction; tests/test_reference_frame.py)
            if (!(sn >> pid >> leaf)) continue;
            parent.push_back(pid);
            for (int i = 0; i < 32; i++) { int b = 0; sn >> b; desc.push_back((uint8_t)b); }               // FORB::fromString
            double w = 0;
            sn >> w;
            weight.push_back(w);
        }
        return build(k, L, n1, n2, parent, desc, weight);
    }

    // The binary format this fork adds (saveToBinaryFile :1516-1536, loadFromBinaryFile :1467-1511): nb_nodes (root included), size_node, k, L,
    // scoring, weighting, then one record per node 1 .. nb_nodes - 1 = parent (int32), 32 descriptor bytes, weight (FLOAT, so weights come
    // back rounded to 32 bits), is-leaf byte.  Exactly nb_nodes - 1 records are read.  (The reference loops `while(!f.eof())` and so files the
    // last record once more as an extra node -- a duplicate last sibling that no descent can select, `d < best_d` being strict -- which makes
    // its size() one larger when that record is a word; tests/test_reference_frame.py.)
    bool loadFromBinaryFile(const std::string& filename)
    {
        std::ifstream f(filename.c_str(), std::ios::in | std::ios::binary);
        if (!f.is_open()) return false;
        unsigned int nb_nodes = 0, size_node = 0;
        int k = -1, L = -1, n1 = -1, n2 = -1;
        f.read((char*)&nb_nodes, sizeof nb_nodes);
        f.read((char*)&size_node, sizeof size_node);
        f.read((char*)&k, sizeof k);
        f.read((char*)&L, sizeof L);
        f.read((char*)&n1, sizeof n1);
        f.read((char*)&n2, sizeof n2);
        if (!f || nb_nodes == 0 || size_node < 41 || k < 0 || L < 1 || n1 < 0 || n1 > 5 || n2 < 0 || n2 > 3) return false;
        std::vector<int> parent(1, 0);
        std::vector<uint8_t> desc(32, 0);
        std::vector<double> weight(1, 0.0);
        std::vector<char> buf(size_node);
        for (unsigned int nid = 1; nid < nb_nodes; nid++) {
            f.read(buf.data(), size_node);
            if (!f) return false;                                                                          // truncated file
            int pid;
            float w;
            std::memcpy(&pid, buf.data(), 4);
            std::memcpy(&w, buf.data() + 36, 4);
            parent.push_back(pid);
            desc.insert(desc.end(), (const uint8_t*)buf.data() + 4, (const uint8_t*)buf.data() + 36);
            weight.push_back((double)w);
        }
        return build(k, L, n1, n2, parent, desc, weight);
    }

    // score between two BowVectors with the vocabulary's scoring type (TemplatedVocabulary::score :1213-1218 = ScoringObject.cpp:20-311).
    // A merge over the two sorted maps in word order, accumulated in the same order and with the same expressions as DBoW2, so the
    // doubles are DBoW2's -- bit for bit when this header is compiled with the floating-point contraction DBoW2 was compiled with (the
    // reference's -O3 -march=native turns `s += vi * wi` into one fma; tests/test_reference_frame.py builds both sides that way).
    // L1_NORM = 0, L2_NORM, CHI_SQUARE, KL, BHATTACHARYYA, DOT_PRODUCT.
    double score(const BowVector& v1, const BowVector& v2) const
    {
        BowVector::const_iterator a = v1.begin(), b = v2.begin();
        const BowVector::const_iterator ae = v1.end(), be = v2.end();
        const double LOG_EPS = std::log(DBL_EPSILON);                                                      // ScoringObject.cpp:18
        double s = 0;
        while (a != ae && b != be) {
            const double vi = a->second, wi = b->second;
            if (a->first == b->first) {
                switch (scoring_) {
                case 0: s += std::fabs(vi - wi) - std::fabs(vi) - std::fabs(wi); break;                    // :40
                case 2: if (vi + wi != 0.0) s += vi * wi / (vi + wi); break;                               // :142
                case 3: if (vi != 0 && wi != 0) s += vi * std::log(vi / wi); break;                        // :196
                case 4: s += std::sqrt(vi * wi); break;                                                    // :246
                default: s += vi * wi; break;                                                              // L2 :91, dot product :291
                }
                ++a;
                ++b;
            } else if (a->first < b->first) {
                if (scoring_ == 3) { s += vi * (std::log(vi) - LOG_EPS); ++a; }                            // :202-206: KL walks v1 entry by entry
                else a = v1.lower_bound(b->first);
            } else
                b = v2.lower_bound(a->first);
        }
        switch (scoring_) {
        case 0: return -s / 2.0;                                                                           // :66
        case 1: return s >= 1 ? 1.0 : 1.0 - std::sqrt(1.0 - s);                                           // :115-118
        case 2: return 2. * s;                                                                             // :170
        case 3:
            for (; a != ae; ++a)
                if (a->second != 0) s += a->second * (std::log(a->second) - LOG_EPS);                      // :216-218
            return s;
        default: return s;                                                                                 // Bhattacharyya, dot product
        }
    }

    bool empty() const { return orbv_size(v_) == 0; }
    unsigned int size() const { return (unsigned int)orbv_size(v_); }

    // vCurrentDesc = Converter::toDescriptorVector(mDescriptors) in the reference; the rows are gathered into one upload here.
    void transform(const std::vector<cv::Mat>& features, BowVector& v, FeatureVector& fv, int levelsup) const
    {
        std::vector<uint8_t> rows(features.size() * 32);
        for (size_t i = 0; i < features.size(); i++)
            for (int b = 0; b < 32; b++) rows[i * 32 + b] = features[i].ptr<uint8_t>(0)[b];
        transform(rows.data(), (int)features.size(), v, fv, levelsup);
    }
    // The same on a continuous N x 32 descriptor matrix (Frame::mDescriptors), without the per-row cv::Mat headers.
    void transform(const uint8_t* descriptors, int n, BowVector& v, FeatureVector& fv, int levelsup) const
    {
        v.clear();
        fv.clear();
        if (!v_ || n == 0) return;                                                                         // :1147-1150
        std::vector<int> bid(n), fnode(n), foff(n + 1), fidx(n);
        std::vector<double> bval(n);
        int nb = 0, nf = 0;
        if (orbv_transform(v_, descriptors, n, levelsup, bid.data(), bval.data(), &nb, fnode.data(), foff.data(), fidx.data(), &nf) != ORBX_OK)
            throw std::runtime_error(std::string("orbv_transform: ") + orbx_last_error());
        for (int i = 0; i < nb; i++) v.insert(v.end(), std::make_pair((unsigned int)bid[i], bval[i]));
        for (int a = 0; a < nf; a++) {
            std::vector<unsigned int>& dst = fv.insert(fv.end(), std::make_pair((unsigned int)fnode[a], std::vector<unsigned int>()))->second;
            for (int p = foff[a]; p < foff[a + 1]; p++) dst.push_back((unsigned int)fidx[p]);
        }
    }

    orbx_vocabulary* handle() { return v_; }

private:
    bool build(int k, int L, int scoring, int weighting, const std::vector<int>& parent, const std::vector<uint8_t>& desc, const std::vector<double>& weight)
    {
        orbv_destroy(v_);
        v_ = nullptr;
        scoring_ = scoring;
        if (orbv_create(device_, k, L, scoring, weighting, parent.data(), desc.data(), weight.data(), (int)parent.size(), &v_) != ORBX_OK)
            throw std::runtime_error(std::string("orbv_create: ") + orbx_last_error());                    // no CPU fallback
        return true;
    }
    orbx_vocabulary* v_;
    int device_, scoring_;
};

}  // namespace ORB_SLAM2

#endif
