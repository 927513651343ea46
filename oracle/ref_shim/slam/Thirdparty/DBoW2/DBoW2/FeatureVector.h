// Stand-in for DBoW2's FeatureVector (test infrastructure, not product): the reference's ORBmatcher.cc only walks it as the
// std::map<node id, feature indices> it derives from.  Written from the DBoW2 public interface; nothing is taken from the reference.
#pragma once
#include <map>
#include <vector>
namespace DBoW2 {
typedef unsigned int NodeId;
class FeatureVector : public std::map<NodeId, std::vector<unsigned int> > {
public:
    void addFeature(NodeId id, unsigned int i_feature) { (*this)[id].push_back(i_feature); }
};
}  // namespace DBoW2
