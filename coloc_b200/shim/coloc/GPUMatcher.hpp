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

#if defined(CLB_SHIM_STANDALONE) || defined(CLB_SHIM_OPENMVG)
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

    // COLOC_B200_DEVICES=0,1,2,... in the environment shards the map database over those GPUs (clb_sharded_*)
    // without a change to the host code: setMapData cuts the rows over the devices, matchSceneWithMap / matchFeaturesWithMap
    // sweep every shard and merge.  useDevices() does the same programmatically.  Pair matching stays on the first device.
    GPUMatcher(MatcherOptions opts) : maxkpNum(opts.maxkp), matchThreshold((uint8_t)opts.thresh) {
        o.distRatio = opts.distRatio; o.thresh = opts.thresh; o.maxkp = (int32_t)opts.maxkp;
        clbCheck(clb_matcher_create(&o, 0, &m), "clb_matcher_create");
        if (const char* env = std::getenv("COLOC_B200_DEVICES")) {
            std::vector<int> devs;
            for (const char* p = env; *p;) { char* e = nullptr; const long v = std::strtol(p, &e, 10); if (e == p) break; devs.push_back((int)v); p = (*e == ',') ? e + 1 : e; }
            if (devs.size() > 1) useDevices(devs);
        }
    }
    ~GPUMatcher() { freeGPUMemory(); }
    void freeGPUMemory() {                                                          // :102-108
        if (sharded) clb_sharded_destroy(sharded);
        sharded = nullptr;
        if (fmap) clb_fmap_destroy(fmap);
        fmap = nullptr;
        if (m) clb_matcher_destroy(m);
        m = nullptr;
    }
    void useDevices(const std::vector<int>& devices) {
        if (sharded) { clb_sharded_destroy(sharded); sharded = nullptr; }
        if (devices.size() > 1) clbCheck(clb_sharded_create(&o, devices.data(), (int)devices.size(), &sharded), "clb_sharded_create");
    }

    void setMapData(int kpNum, void* desc) {                                        // :110-117
        if (sharded) clbCheck(clb_sharded_set_map(sharded, static_cast<const uint64_t*>(desc), kpNum), "clb_sharded_set_map");
        else         clbCheck(clb_set_map(m, static_cast<const uint64_t*>(desc), kpNum), "clb_set_map");
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
    // The reference uploads both descriptor sets again for every pair (:188-196); here every Regions goes to the device
    // FeatureMap once and the whole pair list is one clb_match_pairs call (one sweep launch + one merge launch).
    T computeMatches(FeatureMap& regions, PairWiseMatches& putativeMatches) {
        const int numImages = static_cast<int>(regions.size());
        if (!fmap) clbCheck(clb_fmap_create(0, &fmap), "clb_fmap_create");
        std::vector<int> count((size_t)numImages);
        for (int i = 0; i < numImages; ++i) {
            auto& r = regions.at((openMVG::IndexT)i);
            count[i] = (int)r->RegionCount();
            clbCheck(clb_fmap_put(fmap, (uint32_t)i, nullptr, static_cast<const uint64_t*>(r->DescriptorRawData()), count[i]), "clb_fmap_put");
        }
        std::vector<clb_pair> pairs;
        size_t rows = 0;
        for (int i = 0; i < numImages; ++i)
            for (int j = i + 1; j < numImages; ++j) {
                clb_pair p; p.query_idx = (uint32_t)i; p.train_idx = (uint32_t)j; p.q_lo = 0; p.q_hi = -1;
                pairs.push_back(p); rows += (size_t)count[i];
            }
        h_matches.resize(rows);
        clbCheck(clb_match_pairs(m, fmap, pairs.data(), (int)pairs.size(), 40, 0.f, h_matches.data(), nullptr, nullptr), "clb_match_pairs");
        size_t off = 0;
        for (const clb_pair& p : pairs) {
            IndMatches pairMatches;
            for (int q = 0; q < count[p.query_idx]; ++q)
                if (h_matches[off + q] != -1) pairMatches.emplace_back(q, h_matches[off + q]);
            off += (size_t)count[p.query_idx];
            if (!pairMatches.empty()) putativeMatches.insert({Pair(p.query_idx, p.train_idx), std::move(pairMatches)});
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

    // The CPU build's pair matcher (CPUMatcher::computeMatchesPair, CPUMatcher.hpp:67-76 -> matching::DistanceRatioMatch):
    // ratio test, IndMatch dedup and (x,y) dedup, with the 2-NN search on the GPU.  IndMatch(i_ = idx in pair.first, j_ = idx in pair.second).
    T computeMatchesPairRatio(const Pair& pairIdx, FeatureMap& regions, IndMatches& putativeMatches, float distRatio = 0.8f) {
        const auto& rI = regions.at(pairIdx.first);
        const auto& rJ = regions.at(pairIdx.second);
        const int nI = (int)rI->RegionCount(), nJ = (int)rJ->RegionCount();
        std::vector<float> fI((size_t)4 * nI), fJ((size_t)4 * nJ);
        for (int k = 0; k < nI; ++k) { fI[4 * k] = rI->Features()[k].x(); fI[4 * k + 1] = rI->Features()[k].y(); }
        for (int k = 0; k < nJ; ++k) { fJ[4 * k] = rJ->Features()[k].x(); fJ[4 * k + 1] = rJ->Features()[k].y(); }
        std::vector<uint32_t> out((size_t)2 * (nJ > 0 ? nJ : 1));
        int n = 0;
        clbCheck(clb_distance_ratio_match(m, static_cast<const uint64_t*>(rI->DescriptorRawData()), fI.data(), nI,
                                          static_cast<const uint64_t*>(rJ->DescriptorRawData()), fJ.data(), nJ, distRatio,
                                          out.data(), nJ > 0 ? nJ : 1, &n), "clb_distance_ratio_match");
        putativeMatches.clear();
        for (int k = 0; k < n; ++k) putativeMatches.emplace_back(out[2 * k], out[2 * k + 1]);
        return EXIT_SUCCESS;
    }

    // matching::Filter over 2-NN lists in both directions (matching_filters.hpp:118-213): filter 1 = symmetric cross-check,
    // 2 = distance ratio, 3 = both.  IndMatch(i_ = idx in pair.first, j_ = idx in pair.second).
    T computeMatchesPairFiltered(const Pair& pairIdx, FeatureMap& regions, IndMatches& putativeMatches, int filter = 1, float distRatio = 0.8f) {
        const auto& rI = regions.at(pairIdx.first);
        const auto& rJ = regions.at(pairIdx.second);
        const int nI = (int)rI->RegionCount(), nJ = (int)rJ->RegionCount();
        std::vector<uint32_t> out((size_t)2 * (nI > 0 ? nI : 1));
        int n = 0;
        clbCheck(clb_match_filtered(m, static_cast<const uint64_t*>(rI->DescriptorRawData()), nI, static_cast<const uint64_t*>(rJ->DescriptorRawData()), nJ,
                                    filter, distRatio, out.data(), nI > 0 ? nI : 1, &n), "clb_match_filtered");
        putativeMatches.clear();
        for (int k = 0; k < n; ++k) putativeMatches.emplace_back(out[2 * k], out[2 * k + 1]);
        return EXIT_SUCCESS;
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
        if (sharded) clbCheck(clb_sharded_match_map(sharded, query.data(), (int)kpQuery, matchThreshold, h_matches.data(), nullptr, nullptr), "clb_sharded_match_map");
        else         clbCheck(clb_match_map(m, query.data(), (int)kpQuery, matchThreshold, h_matches.data(), nullptr, nullptr), "clb_match_map");
        IndMatches matches;
        for (size_t i = 0; i < kpQuery; ++i)
            if (h_matches[i] != -1) matches.emplace_back(h_matches[i], (uint32_t)i);   // i_ = MAP idx, j_ = FRAME idx
        return matches;
    }

private:
    clb_matcher_opts o;
    clb_matcher* m = nullptr;
    clb_sharded* sharded = nullptr;   // map database cut over several GPUs (COLOC_B200_DEVICES / useDevices)
    clb_fmap* fmap = nullptr;      // device-side copy of the FeatureMap used by computeMatches(FeatureMap&, ...)
    unsigned int kpTrain = 0, kpQuery = 0, kpMap = 0;
    uint8_t matchThreshold;
    std::vector<uint64_t> train, query;
    std::vector<int32_t> h_matches;
};

}  // namespace coloc
