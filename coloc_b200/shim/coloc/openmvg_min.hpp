// openmvg_min.hpp -- stand-ins for the few openMVG / coloc types that cross the detector / matcher
// boundary, used ONLY when the shim is built without openMVG (-DCLB_SHIM_STANDALONE: tests, demos).
// Names, members and semantics follow the originals so the shim code is identical in both builds:
//   openMVG::features::SIOPointFeature        lib/openMVG/src/openMVG/features/feature.hpp:68-93
//   features::AKAZE_Binary_Regions            features/regions_factory.hpp:26, binary_regions.hpp:24-134
//   matching::IndMatch / IndMatches / PairWiseMatches   matching/indMatch.hpp:25-95
//   Pair                                      types.hpp
//   coloc::FeatureMap / DetectorOptions / MatcherOptions   include/coloc/colocData.hpp:19-42
#pragma once
#include <array>
#include <cstdint>
#include <map>
#include <memory>
#include <utility>
#include <vector>

namespace openMVG {
using IndexT = uint32_t;
using Pair = std::pair<IndexT, IndexT>;
namespace features {
struct SIOPointFeature {
    float x_, y_, scale_, orientation_;
    SIOPointFeature(float x = 0, float y = 0, float s = 0, float o = 0) : x_(x), y_(y), scale_(s), orientation_(o) {}
    float x() const { return x_; }
    float y() const { return y_; }
    float scale() const { return scale_; }
    float orientation() const { return orientation_; }
};
template <typename T, std::size_t N> using Descriptor = std::array<T, N>;
struct AKAZE_Binary_Regions {
    using FeatureT = SIOPointFeature;
    using DescriptorT = Descriptor<unsigned char, 64>;
    std::vector<FeatureT>& Features() { return feats_; }
    const std::vector<FeatureT>& Features() const { return feats_; }
    std::vector<DescriptorT>& Descriptors() { return descs_; }
    const std::vector<DescriptorT>& Descriptors() const { return descs_; }
    const void* DescriptorRawData() const { return descs_.data(); }
    std::size_t RegionCount() const { return feats_.size(); }
private:
    std::vector<FeatureT> feats_;
    std::vector<DescriptorT> descs_;
};
}  // namespace features
namespace matching {
struct IndMatch {
    IndexT i_, j_;
    IndMatch(IndexT i = 0, IndexT j = 0) : i_(i), j_(j) {}
    friend bool operator==(const IndMatch& a, const IndMatch& b) { return a.i_ == b.i_ && a.j_ == b.j_; }
};
using IndMatches = std::vector<IndMatch>;
using PairWiseMatches = std::map<Pair, IndMatches>;
}  // namespace matching
}  // namespace openMVG

namespace coloc {
using FeatureMap = std::map<openMVG::IndexT, std::unique_ptr<openMVG::features::AKAZE_Binary_Regions>>;
struct DetectorOptions { float scale_factor; uint8_t scale_levels; unsigned int width, height, maxkp; uint8_t thresh; };
struct MatcherOptions { float distRatio; int thresh; unsigned int maxkp; };
}  // namespace coloc
