// Exercises the drop-in C++ headers (include/ORBextractor.h, include/ORBmatcher.h) the way Frame.cc uses the reference
// classes.  usage: dropin_test in.raw w h nfeatures out.bin [vocabulary.txt]
// out.bin: int32 n | n x 28-byte keypoints | n x 32 descriptor bytes | int32 levels | per level int32 w,h + bytes |
//          int32 nmatch_self | n int32 matches (SearchBruteForce of the descriptors against themselves) | int32 d01 |
//          int32 nFused | n int32 (Fuse, chi-square gate) | int32 nFused | n int32 (Fuse, Sim3) | int32 nFound | n int32 (SearchBySim3)
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "ORBextractor.h"
#include "ORBmatcher.h"
#include "ORBVocabulary.h"

int main(int argc, char** argv)
{
    if (argc < 6) return 2;
    const int w = atoi(argv[2]), h = atoi(argv[3]), nf = atoi(argv[4]);
    std::vector<uint8_t> img((size_t)w * h);
    FILE* f = fopen(argv[1], "rb");
    if (!f || fread(img.data(), 1, img.size(), f) != img.size()) return 3;
    fclose(f);
    try {
        ORB_SLAM2::ORBextractor* ex = new ORB_SLAM2::ORBextractor(nf, 1.2f, 8, 20, 7);     // Tracking.cc:113-119
        cv::Mat im(h, w, CV_8UC1, img.data());
        std::vector<cv::KeyPoint> keys;
        cv::Mat desc;
        (*ex)(im, cv::Mat(), keys, desc);                                                    // Frame.cc:247-253
        ORB_SLAM2::ORBmatcher matcher(0.9, true);
        std::vector<int> m;
        const int nm = keys.empty() ? 0 : matcher.SearchBruteForce(desc, desc, m);
        const int d01 = keys.size() > 1 ? ORB_SLAM2::ORBmatcher::DescriptorDistance(desc.row(0), desc.row(1)) : -1;
        FILE* o = fopen(argv[5], "wb");
        const int n = (int)keys.size();
        fwrite(&n, 4, 1, o);
        static_assert(sizeof(cv::KeyPoint) == 28, "cv::KeyPoint layout");
        fwrite(keys.data(), 28, n, o);
        for (int i = 0; i < n; i++) fwrite(desc.ptr<uint8_t>(i), 1, 32, o);
        const int L = ex->GetLevels();
        fwrite(&L, 4, 1, o);
        for (int l = 0; l < L; l++) {
            const cv::Mat& p = ex->mvImagePyramid[l];
            fwrite(&p.cols, 4, 1, o);
            fwrite(&p.rows, 4, 1, o);
            for (int y = 0; y < p.rows; y++) fwrite(p.ptr<uint8_t>(y), 1, p.cols, o);
        }
        fwrite(&nm, 4, 1, o);
        m.resize(n, -1);
        fwrite(m.data(), 4, n, o);
        fwrite(&d01, 4, 1, o);
        // the map-side searches through the C++ wrappers, every feature re-projected onto itself one pixel to the right:
        // Fuse (mono chi-square gate), Fuse with a Sim3 (no gate), SearchBySim3 both ways on the same key frame
        ORB_SLAM2::FrameView view = { &keys, &desc, 0.0f, (float)w, 0.0f, (float)h };
        std::vector<float> sf = ex->GetScaleFactors(), isg = ex->GetInverseScaleSigmaSquares(), uR(n, -1.0f);
        ORB_SLAM2::ORBmatcher::ProjectedPoints P;
        for (int i = 0; i < n; i++) {
            P.xy.push_back(keys[i].pt.x + 1.0f);
            P.xy.push_back(keys[i].pt.y);
            P.xr.push_back(-1.0f);
            P.level.push_back(keys[i].octave);
            P.descriptors.insert(P.descriptors.end(), desc.ptr<uint8_t>(i), desc.ptr<uint8_t>(i) + 32);
        }
        std::vector<int> f1, f2, s12;
        const int nf1 = matcher.Fuse(view, uR, sf, isg, P, f1, 3.0f);
        const int nf2 = matcher.Fuse(view, sf, P, f2, 4.0f);
        const int ns = matcher.SearchBySim3(view, view, sf, P, P, s12, 7.5f);
        fwrite(&nf1, 4, 1, o);
        fwrite(f1.data(), 4, n, o);
        fwrite(&nf2, 4, 1, o);
        fwrite(f2.data(), 4, n, o);
        fwrite(&ns, 4, 1, o);
        fwrite(s12.data(), 4, n, o);
        // Frame::ComputeBoW (Frame.cc:410-417) through the ORBVocabulary drop-in, when a vocabulary file is given
        int nb = -1;
        if (argc > 6) {
            ORB_SLAM2::ORBVocabulary voc;
            if (!voc.loadFromTextFile(argv[6])) return 4;
            std::vector<cv::Mat> vCurrentDesc;                                              // Converter::toDescriptorVector
            for (int i = 0; i < n; i++) vCurrentDesc.push_back(desc.row(i));
            ORB_SLAM2::BowVector bow;
            ORB_SLAM2::FeatureVector fv;
            voc.transform(vCurrentDesc, bow, fv, 4);
            nb = (int)bow.size();
            fwrite(&nb, 4, 1, o);
            for (ORB_SLAM2::BowVector::const_iterator it = bow.begin(); it != bow.end(); ++it) { fwrite(&it->first, 4, 1, o); fwrite(&it->second, 8, 1, o); }
            const int nf = (int)fv.size();
            fwrite(&nf, 4, 1, o);
            for (ORB_SLAM2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it) {
                const int cnt = (int)it->second.size();
                fwrite(&it->first, 4, 1, o);
                fwrite(&cnt, 4, 1, o);
                fwrite(it->second.data(), 4, cnt, o);
            }
        }
        fclose(o);
        printf("ok %d keypoints, %d self matches, scale[1]=%g\n", n, nm, ex->GetScaleFactors()[1]);
        delete ex;
    } catch (const std::exception& e) {
        fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
    return 0;
}
