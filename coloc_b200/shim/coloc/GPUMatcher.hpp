// coloc/GPUMatcher.hpp -- drop-in for the reference's include/coloc/GPUMatcher.hpp:45-272.
// Same class name, template shape and method signatures, so FeatureMatcher<bool, GPUMatcher>
// (FeatureMatcher.hpp:23-34, coloc.hpp:65) compiles unchanged; everything forwards to the C ABI.
// Kept conventions: computeMatches(FeatureMap&, PairWiseMatches&) returns EXIT_SUCCESS (=0) (:154);
// pair matches are IndMatch(query idx, train idx) ascending in the query index (:215-220); map matches are
// IndMatch(MAP idx, frame idx) (:263-266); thresholds 40 (pairs, :180) / 60 (map-map :162) / opts.thresh (map, :256).
#pragma once
#include <cstdint>
#include <cstdlib>
#include <memory>
#include <vector>
#include "coloc_b200.h"
#include "GPUDetector.hpp"   // clbCheck + the boundary types

namespace coloc {

#ifdef CLB_SHIM_STANDALONE
// what matchSceneWithMap needs from colocData (include/coloc/colocData.hpp:45-122)
struct colocData { FeatureMap regions; };
using openMVG::Pair;
using openMVG::matching::IndMatch;
using openMVG::matching::IndMatches;
using openMVG::matching::PairWiseMatches;
#endif

template <typename T>
class GPUMatcher {
public:
    unsigned int maxkpNum;

    GPUMatcher(MatcherOptions opts) : maxkpNum(opts.maxkp), matchThreshold((uint8_t)opts.thresh) {
        clb_matcher_opts o; o.distRatio = opts.distRatio; o.thresh = opts.thresh; o.maxkp = (int32_t)opts.maxkp;
        clbCheck(clb_matcher_create(&o, 0, &m), "clb_matcher_create");
    }
    ~GPUMatcher() { freeGPUMemory(); }
    void freeGPUMemory() { if (m) clb_matcher_destroy(m); m = nullptr; }          // :102-108

    void setMapData(int kpNum, void* desc) {                                        // :110-117
        clbCheck(clb_set_map(m, static_cast<const uint64_t*>(desc), kpNum), "clb_set_map");
        kpMap = kpNum;
    }
    void setTrainingImage(int kpNum, std::vector<uint64_t> const& desc) {           // :120-127
        train.assign(desc.begin(), desc.begin() + (size_t)8 * kpNum); kpTrain = kpNum;
    }
    void setQueryImage(int kpNum, void* desc) {                                     // :130-141
        const uint64_t* d = static_cast<const uint64_t*>(desc);
        query.assign(d, d + (size_t)8 * kpNum); kpQuery = kpNum;
    }

    // all unordered pairs I<J (Utils::handlePairs -> exhaustivePairs, Pair_Builder.hpp:25-33)   :143-155
    T computeMatches(FeatureMap& regions, PairWiseMatches& putativeMatches) {
        const int numImages = static_cast<int>(regions.size());
        for (int i = 0; i < numImages; ++i)
            for (int j = i + 1; j < numImages; ++j) {
                const Pair pairIdx(i, j);
                IndMatches pairMatches;
                computeMatchesPair(pairIdx, regions, pairMatches);
                if (!pairMatches.empty()) putativeMatches.insert({pairIdx, std::move(pairMatches)});
            }
        return EXIT_SUCCESS;
    }

    void matchMapFeatures(std::unique_ptr<openMVG::features::AKAZE_Binary_Regions>& map1,
                          std::unique_ptr<openMVG::features::AKAZE_Binary_Regions>& map2, IndMatches& commonFeatures) {   // :157-163
        commonFeatures = computeMatches(const_cast<void*>(map1->DescriptorRawData()), const_cast<void*>(map2->DescriptorRawData()),
                                        (int)map1->RegionCount(), (int)map2->RegionCount(), 60);
    }

    void computeMatchesPair(const Pair& pairIdx, FeatureMap& regions, IndMatches& putativeMatches) {                    // :165-172
        putativeMatches = computeMatches(const_cast<void*>(regions[pairIdx.first]->DescriptorRawData()),
                                         const_cast<void*>(regions[pairIdx.second]->DescriptorRawData()),
                                         (int)regions[pairIdx.first]->RegionCount(), (int)regions[pairIdx.second]->RegionCount());
    }

    void matchSceneWithMap(int& droneId, colocData& data, IndMatches& mapMatches) {                                     // :174-178
        setQueryImage((int)data.regions[droneId]->RegionCount(), const_cast<void*>(data.regions[droneId]->DescriptorRawData()));
        mapMatches = matchFeaturesWithMap();
    }

    // :180-226.  The reference copies (nT+8) rows, reading 512 B past the host vector (:188-191); the ABI needs no padding.
    IndMatches computeMatches(void* h_descriptorsQuery, void* h_descriptorsTraining, int numKPQuery, int numKPTraining,
                              uint8_t threshold = 40) {
        h_matches.resize((size_t)(numKPQuery > 0 ? numKPQuery : 0));
        clbCheck(clb_match_2nn(m, static_cast<const uint64_t*>(h_descriptorsQuery), numKPQuery,
                               static_cast<const uint64_t*>(h_descriptorsTraining), numKPTraining, threshold,
                               h_matches.data(), nullptr, nullptr), "clb_match_2nn");
        IndMatches matches;
        for (int i = 0; i < numKPQuery; ++i)
            if (h_matches[i] != -1) matches.emplace_back(i, h_matches[i]);
        return matches;
    }

    IndMatches computeMatchesPreset() {                                             // :229-250 (training vs query set earlier)
        h_matches.resize(kpQuery);
        clbCheck(clb_match_2nn(m, query.data(), (int)kpQuery, train.data(), (int)kpTrain, matchThreshold, h_matches.data(), nullptr, nullptr),
                 "clb_match_2nn");
        IndMatches matches;
        for (size_t i = 0; i < kpQuery; ++i)
            if (h_matches[i] != -1) matches.emplace_back(h_matches[i], (uint32_t)i);
        return matches;
    }

    IndMatches matchFeaturesWithMap() {                                             // :252-271
        h_matches.resize(kpQuery);
        clbCheck(clb_match_map(m, query.data(), (int)kpQuery, matchThreshold, h_matches.data(), nullptr, nullptr), "clb_match_map");
        IndMatches matches;
        for (size_t i = 0; i < kpQuery; ++i)
            if (h_matches[i] != -1) matches.emplace_back(h_matches[i], (uint32_t)i);   // i_ = MAP idx, j_ = FRAME idx
        return matches;
    }

private:
    clb_matcher* m = nullptr;
    unsigned int kpTrain = 0, kpQuery = 0, kpMap = 0;
    uint8_t matchThreshold;
    std::vector<uint64_t> train, query;
    std::vector<int32_t> h_matches;
};

}  // namespace coloc
