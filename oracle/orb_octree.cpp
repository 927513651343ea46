// orb_octree.cpp -- CPU ORACLE (test infrastructure, not product).
// Restatement of ExtractorNode::DivideNode (ORBextractor.cc:481-537) and
// ORBextractor::DistributeOctTree (ORBextractor.cc:539-763).
//
// The reference keeps nodes in a std::list built with push_front and, in its "largest first"
// phase, sorts pair<int size, ExtractorNode*> (ORBextractor.cc:684): ties on size are broken by
// HEAP ADDRESS, which depends on the allocator.  tie_mode 0 (the portable definition the CUDA
// path is held to) breaks ties by node creation order, later-created node first -- what a
// monotone allocator would give.  tie_mode 1 keeps the address comparison on real list nodes
// so the number of (frame, level) cases where this libc disagrees can be counted.
#include "orb_internal.h"
#include "orb_oracle.h"
#include <list>
#include <algorithm>
#include <utility>

namespace orbo {
namespace {

// 28-byte key like cv::KeyPoint so that the allocation pattern in tie_mode 1 resembles the reference's.
struct Key { float x, y, size, angle, response; int octave; int idx; };

struct Node {
    std::vector<Key> keys;
    int ULx = 0, ULy = 0, URx = 0, URy = 0, BLx = 0, BLy = 0, BRx = 0, BRy = 0;
    std::list<Node>::iterator self;
    bool noMore = false;
    long seq = 0;

    // DivideNode: children in the order TL, TR, BL, BR.
    void divide(Node c[4]) const
    {
        const int halfX = (int)std::ceil(static_cast<float>(URx - ULx) / 2);
        const int halfY = (int)std::ceil(static_cast<float>(BRy - ULy) / 2);
        const int midX = ULx + halfX, midY = ULy + halfY;
        c[0].ULx = ULx;  c[0].ULy = ULy;  c[0].URx = midX; c[0].URy = ULy;
        c[0].BLx = ULx;  c[0].BLy = midY; c[0].BRx = midX; c[0].BRy = midY;
        c[1].ULx = midX; c[1].ULy = ULy;  c[1].URx = URx;  c[1].URy = URy;
        c[1].BLx = midX; c[1].BLy = midY; c[1].BRx = URx;  c[1].BRy = midY;
        c[2].ULx = ULx;  c[2].ULy = midY; c[2].URx = midX; c[2].URy = midY;
        c[2].BLx = BLx;  c[2].BLy = BLy;  c[2].BRx = midX; c[2].BRy = BLy;
        c[3].ULx = midX; c[3].ULy = midY; c[3].URx = URx;  c[3].URy = midY;
        c[3].BLx = midX; c[3].BLy = BLy;  c[3].BRx = BRx;  c[3].BRy = BRy;
        for (int q = 0; q < 4; q++) c[q].keys.reserve(keys.size());
        for (const Key& k : keys) {
            const bool left = k.x < (float)midX, top = k.y < (float)midY;
            c[left ? (top ? 0 : 2) : (top ? 1 : 3)].keys.push_back(k);
        }
        for (int q = 0; q < 4; q++) if (c[q].keys.size() == 1) c[q].noMore = true;
    }
};

typedef std::pair<int, Node*> SizeNode;

struct BySizeThenSeq {
    bool operator()(const SizeNode& a, const SizeNode& b) const
    { return a.first != b.first ? a.first < b.first : a.second->seq < b.second->seq; }
};

} // namespace

std::vector<int> distribute_octtree(const std::vector<XYR>& in, int minX, int maxX, int minY, int maxY, int N, int tie_mode)
{
    std::vector<int> result;
    const int nIni = (int)std::round(static_cast<float>(maxX - minX) / (maxY - minY));
    if (nIni <= 0) return result;               // the reference divides by zero here; callers reject such shapes
    const float hX = static_cast<float>(maxX - minX) / nIni;

    std::list<Node> nodes;
    long seq = 0;
    std::vector<Node*> roots(nIni);
    for (int i = 0; i < nIni; i++) {
        Node r;
        r.ULx = (int)(hX * static_cast<float>(i));     r.ULy = 0;
        r.URx = (int)(hX * static_cast<float>(i + 1)); r.URy = 0;
        r.BLx = r.ULx; r.BLy = maxY - minY;
        r.BRx = r.URx; r.BRy = maxY - minY;
        r.keys.reserve(in.size());
        r.seq = seq++;
        nodes.push_back(r);
        roots[i] = &nodes.back();
    }
    for (size_t i = 0; i < in.size(); i++) {
        Key k{(float)in[i].x, (float)in[i].y, 7.f, -1.f, (float)in[i].r, 0, (int)i};
        size_t root = (size_t)(k.x / hX);
        if (root >= roots.size()) root = roots.size() - 1;   // unreachable for x < maxX-minX; guards the oracle only
        roots[root]->keys.push_back(k);
    }
    for (auto it = nodes.begin(); it != nodes.end();) {
        if (it->keys.size() == 1) { it->noMore = true; ++it; }
        else if (it->keys.empty()) it = nodes.erase(it);
        else ++it;
    }

    std::vector<SizeNode> expandable;
    expandable.reserve(nodes.size() * 4);
    // push the non-empty children of a split node to the front of the list (TL,TR,BL,BR order)
    auto adopt = [&](Node c[4], int* nToExpand) {
        for (int q = 0; q < 4; q++) {
            if (c[q].keys.empty()) continue;
            c[q].seq = seq++;
            nodes.push_front(c[q]);
            if (c[q].keys.size() > 1) {
                if (nToExpand) ++*nToExpand;
                expandable.push_back(std::make_pair((int)c[q].keys.size(), &nodes.front()));
                nodes.front().self = nodes.begin();
            }
        }
    };

    bool finished = false;
    while (!finished) {
        int prevSize = (int)nodes.size();
        int nToExpand = 0;
        expandable.clear();
        for (auto it = nodes.begin(); it != nodes.end();) {
            if (it->noMore) { ++it; continue; }
            Node c[4];
            it->divide(c);
            adopt(c, &nToExpand);
            it = nodes.erase(it);
        }
        if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) {
            finished = true;
        } else if ((int)nodes.size() + nToExpand * 3 > N) {
            while (!finished) {
                prevSize = (int)nodes.size();
                std::vector<SizeNode> prev = expandable;
                expandable.clear();
                if (tie_mode == 1) std::sort(prev.begin(), prev.end());            // (size, address), as the reference
                else std::sort(prev.begin(), prev.end(), BySizeThenSeq());         // (size, creation order)
                for (int j = (int)prev.size() - 1; j >= 0; j--) {
                    Node c[4];
                    prev[j].second->divide(c);
                    adopt(c, nullptr);
                    nodes.erase(prev[j].second->self);
                    if ((int)nodes.size() >= N) break;
                }
                if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) finished = true;
            }
        }
    }

    result.reserve(nodes.size());
    for (const Node& nd : nodes) {
        const Key* best = &nd.keys[0];
        float maxResponse = best->response;
        for (size_t k = 1; k < nd.keys.size(); k++)
            if (nd.keys[k].response > maxResponse) { best = &nd.keys[k]; maxResponse = nd.keys[k].response; }
        result.push_back(best->idx);
    }
    return result;
}

} // namespace orbo

extern "C" int orbo_distribute(const int* xyr, int n, int minX, int maxX, int minY, int maxY, int N, int tie_mode,
                               int* out_idx, int cap)
{
    std::vector<orbo::XYR> keys(n);
    for (int i = 0; i < n; i++) keys[i] = orbo::XYR{xyr[3 * i], xyr[3 * i + 1], xyr[3 * i + 2]};
    std::vector<int> r = orbo::distribute_octtree(keys, minX, maxX, minY, maxY, N, tie_mode);
    for (size_t i = 0; i < r.size() && (int)i < cap; i++) out_idx[i] = r[i];
    return (int)r.size();
}
