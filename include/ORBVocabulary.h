// ORBVocabulary.h -- drop-in for the part of the reference's ORBVocabulary the front end uses (kimwin2/ORB_SLAM2v2-1
// include/ORBVocabulary.h:31-32: DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB>): loadFromTextFile
// (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1351-1436) and transform(features, BowVector, FeatureVector, levelsup) (:1140-1207) as
// Frame::ComputeBoW / KeyFrame::ComputeBoW call it (src/Frame.cc:410-417).  The tree lives on the device (orbv_* in include/orbx.h);
// scoring two BowVectors (KeyFrameDatabase, LoopClosing) stays with DBoW2 on the host - both types below convert from / to DBoW2's maps.
#ifndef ORBVOCABULARY_H
#define ORBVOCABULARY_H

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
    explicit ORBVocabulary(int device = 0) : v_(nullptr), device_(device) {}
    ~ORBVocabulary() { orbv_destroy(v_); }
    ORBVocabulary(const ORBVocabulary&) = delete;
    ORBVocabulary& operator=(const ORBVocabulary&) = delete;

    // Same text format and return convention as the reference (System.cc:68-76 checks the bool).
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
        orbv_destroy(v_);
        v_ = nullptr;
        if (orbv_create(device_, k, L, n1, n2, parent.data(), desc.data(), weight.data(), (int)parent.size(), &v_) != ORBX_OK)
            throw std::runtime_error(std::string("orbv_create: ") + orbx_last_error());                    // no CPU fallback
        return true;
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
    orbx_vocabulary* v_;
    int device_;
};

}  // namespace ORB_SLAM2

#endif
