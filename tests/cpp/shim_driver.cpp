// shim_driver.cpp -- exercises the drop-in C++ shim exactly the way coloc.hpp drives the reference classes:
// FeatureDetector<bool,GPUDetector>::detectFeaturesFile for two vehicles, FeatureMatcher<bool,GPUMatcher>::
// computeMatches over the FeatureMap, setMapData + matchSceneWithMap.  Dumps raw results for the python test.
// Build: g++ -std=c++14 -DCLB_SHIM_STANDALONE -Iinclude -Icoloc_b200/shim tests/cpp/shim_driver.cpp -Lcoloc_b200 -lcoloc_b200
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include "coloc/GPUDetector.hpp"
#include "coloc/GPUMatcher.hpp"

// the policy-template shells of FeatureDetector.hpp:21-32 / FeatureMatcher.hpp:23-34
namespace coloc {
template <typename T, template <typename> class ProcessorType>
class FeatureDetector : public ProcessorType<T> {
public:
    FeatureDetector(DetectorOptions opts) : ProcessorType<T>(opts) {}
};
template <typename T, template <typename> class ProcessorType>
class FeatureMatcher : public ProcessorType<T> {
public:
    FeatureMatcher(MatcherOptions opts) : ProcessorType<T>(opts) {}
};
}  // namespace coloc

template <class V> static void dump(FILE* f, const V& v) {
    const uint64_t n = v.size();
    fwrite(&n, 8, 1, f);
    if (n) fwrite(v.data(), sizeof(v[0]), n, f);
}

int main(int argc, char** argv) {
    if (argc < 6) { fprintf(stderr, "usage: shim_driver a.pgm b.pgm w h out.bin\n"); return 2; }
    using namespace coloc;
    DetectorOptions dopts; dopts.scale_factor = 1.2f; dopts.scale_levels = 8; dopts.width = atoi(argv[3]); dopts.height = atoi(argv[4]);
    dopts.maxkp = 5000; dopts.thresh = 40;                       // coloc_node.cpp:73-85
    MatcherOptions mopts; mopts.distRatio = 0.8f; mopts.thresh = 60; mopts.maxkp = 5000;
    FeatureDetector<bool, GPUDetector> detector(dopts);
    FeatureMatcher<bool, GPUMatcher> matcher(mopts);
    FeatureMap regions;
    std::string a = argv[1], b = argv[2];
    FILE* f = fopen(argv[5], "wb");
    if (detector.detectFeaturesFile(0, regions, a) != EXIT_SUCCESS) return 3;
    dump(f, detector.kps); dump(f, detector.desc);
    if (detector.detectFeaturesFile(1, regions, b) != EXIT_SUCCESS) return 3;
    dump(f, detector.kps); dump(f, detector.desc);
    dump(f, regions[0]->Features());
    {   // a ROS-topic style frame (GPUDetector.hpp:188-212): same bytes as file b through detectFeaturesTopic
        std::vector<unsigned char> px; int pw = 0, ph = 0;
        if (FILE* g = fopen(b.c_str(), "rb")) {
            int maxv = 0;
            if (fscanf(g, "P5 %d %d %d", &pw, &ph, &maxv) == 3 && fgetc(g) != EOF) { px.resize((size_t)pw * ph); if (fread(px.data(), 1, px.size(), g) != px.size()) px.clear(); }
            fclose(g);
        }
        if (px.empty()) return 6;
        struct Img { unsigned char* data; int cols, rows; };
        struct Msg { Img image; } msg{{px.data(), pw, ph}};
        detector.detectFeaturesTopic(2, regions, &msg);
        if (regions[2]->RegionCount() != regions[1]->RegionCount() ||
            memcmp(regions[2]->DescriptorRawData(), regions[1]->DescriptorRawData(), regions[1]->RegionCount() * 64) != 0) return 6;
        for (size_t i = 0; i < regions[1]->RegionCount(); ++i)
            if (regions[2]->Features()[i].x() != regions[1]->Features()[i].x() || regions[2]->Features()[i].y() != regions[1]->Features()[i].y()) return 6;
        regions.erase(2);
    }
    openMVG::matching::PairWiseMatches putative;
    if (matcher.computeMatches(regions, putative) != EXIT_SUCCESS) return 4;
    dump(f, putative[{0, 1}]);
    // the CPU build's semantics (ratio + dedups) on the same pair
    openMVG::matching::IndMatches ratioMatches;
    if (matcher.computeMatchesPairRatio({0, 1}, regions, ratioMatches) != EXIT_SUCCESS) return 5;
    dump(f, ratioMatches);
    // map tracking: vehicle 0's descriptors as the "map", vehicle 1 as the frame
    matcher.setMapData((int)regions[0]->RegionCount(), const_cast<void*>(regions[0]->DescriptorRawData()));
    colocData data; data.regions = std::move(regions);
    int drone = 1;
    openMVG::matching::IndMatches mapMatches;
    matcher.matchSceneWithMap(drone, data, mapMatches);
    dump(f, mapMatches);
    fclose(f);
    printf("shim ok: %zu/%zu keypoints, %zu pair matches, %zu map matches\n", data.regions[0]->RegionCount(),
           data.regions[1]->RegionCount(), putative[{0, 1}].size(), mapMatches.size());
    detector.freeGPUMemory();
    return 0;
}
