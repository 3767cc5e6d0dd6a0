// ref_cpu.cpp -- harness that compiles the REFERENCE's own CPU code, unmodified, from where it
// lies under /root/reference (include paths only; no reference source is copied here).
// TEST INFRASTRUCTURE: built by oracle/Makefile into oracle/_ref/libcoloc_ref_cpu.so.
//   * KFAST<mt,true>            include/coloc/KFAST.h:503-540
//   * featureAngle/fastAtan2    include/coloc/FeatureAngle.h:160-247
//   * ArrayMatcherBruteForce<uchar,Hamming<uchar>> + NNdistanceRatio
//       lib/openMVG/src/openMVG/matching/matcher_brute_force.hpp:100-200, metric_hamming.hpp:80-99,
//       matching_filters.hpp:39-60  (what CPUMatcher.hpp:67-76 -> regions_matcher.hpp:137-184 runs)
//   * matching::DistanceRatioMatch on two AKAZE_Binary_Regions (regions_matcher.cpp:19-30: ratio filter,
//       IndMatch dedup, (x,y) dedup) and Binary_Regions::Save / Load (.feat / .desc, binary_regions.hpp:51-66)
//   * the two-view error functors AC-RANSAC evaluates per hypothesis (solver_fundamental_kernel.cpp:156-192,
//       solver_homography_kernel.hpp:59-63)
#include <cstring>
#include <thread>
#include <future>
#include <vector>
#include <cstdint>
#include <cfloat>
#include <cmath>
#include <algorithm>
#include "coloc/Keypoint.h"
#include "coloc/KFAST.h"
#include "coloc/FeatureAngle.h"
#include "openMVG/matching/matcher_brute_force.hpp"
#include "openMVG/matching/metric_hamming.hpp"
#include "openMVG/matching/matching_filters.hpp"
#include "openMVG/features/feature.hpp"
#include "openMVG/features/descriptor.hpp"
#include "openMVG/features/binary_regions.hpp"
#include "openMVG/features/regions_factory.hpp"
#include "openMVG/matching/regions_matcher.hpp"
#include "openMVG/multiview/solver_fundamental_kernel.hpp"
#include "openMVG/multiview/solver_homography_kernel.hpp"

static_assert(sizeof(Keypoint) == 20, "Keypoint.h layout");

extern "C" {

// returns the keypoint count; writes at most cap entries (x, y, score; angle/scale untouched = 0)
int ref_kfast(const uint8_t* img, int w, int h, int stride, int thresh, int multithread, Keypoint* out, int cap) {
    std::vector<Keypoint> kps;
    if (multithread) KFAST<true, true>(img, w, h, stride, kps, (uint8_t)thresh);
    else             KFAST<false, true>(img, w, h, stride, kps, (uint8_t)thresh);
    const int n = (int)kps.size();
    for (int i = 0; i < n && i < cap; ++i) { std::memset(&out[i], 0, sizeof(Keypoint)); out[i].x = kps[i].x; out[i].y = kps[i].y; out[i].score = kps[i].score; }
    return n;
}

float ref_feature_angle(const uint8_t* img, int x, int y, int stride) { return featureAngle(img, x, y, stride); }
float ref_fast_atan2(float y, float x) { return fastAtan2(y, x); }
// bulk version for exhaustive sweeps
void ref_fast_atan2_many(const float* y, const float* x, float* out, long n) { for (long i = 0; i < n; ++i) out[i] = fastAtan2(y[i], x[i]); }

int ref_hamming(const uint8_t* a, const uint8_t* b, int nbytes) {
    openMVG::matching::Hamming<unsigned char> m;
    return (int)m(a, b, (size_t)nbytes);
}

// openMVG brute-force 2-NN (database = T rows, queries = Q rows, 64 B each).
// idx/dist: nQ*2 entries (neighbour order as returned); ratio_ok: nQ flags from NNdistanceRatio(ratio).
// Returns 0 on success, 1 if SearchNeighbours refused (empty input), as matching_test.cpp:123-131.
int ref_bruteforce_2nn(const uint8_t* T, int nT, const uint8_t* Q, int nQ, float ratio,
                       int32_t* idx, int32_t* dist, uint8_t* ratio_ok) {
    using namespace openMVG::matching;
    using Matcher = ArrayMatcherBruteForce<unsigned char, Hamming<unsigned char>>;
    Matcher m;
    if (!m.Build(T, nT, 64)) return 1;
    IndMatches ind;
    std::vector<Matcher::DistanceType> d;
    if (!m.SearchNeighbours(Q, nQ, &ind, &d, 2)) return 1;
    for (int i = 0; i < nQ * 2; ++i) { if (idx) idx[i] = (int32_t)ind[i].j_; if (dist) dist[i] = (int32_t)d[i]; }
    if (ratio_ok) {
        std::vector<int> ok;
        NNdistanceRatio(d.begin(), d.end(), 2, ok, ratio);
        std::memset(ratio_ok, 0, nQ);
        for (int i : ok) ratio_ok[i] = 1;
    }
    return 0;
}

// matching::Filter (matching_filters.hpp:118-213) over openMVG's own brute-force 2-NN lists in both directions.
// filter: 1 symmetric, 2 distance ratio, 3 both.  pairs: (i_, j_) per kept match; returns the count (or -1).
int ref_filter_matches(const uint8_t* A, int nA, const uint8_t* B, int nB, int filter, float ratio, uint32_t* pairs, int cap) {
    using namespace openMVG::matching;
    using Matcher = ArrayMatcherBruteForce<unsigned char, Hamming<unsigned char>>;
    Matcher mB, mA;
    if (!mB.Build(B, nB, 64) || !mA.Build(A, nA, 64)) return -1;
    IndMatches i01, i10;
    std::vector<Matcher::DistanceType> d01, d10;
    const int NN = 2;
    if (!mB.SearchNeighbours(A, nA, &i01, &d01, NN) || !mA.SearchNeighbours(B, nB, &i10, &d10, NN)) return -1;
    std::vector<int> v01(i01.size()), v10(i10.size());
    std::vector<float> f01(d01.begin(), d01.end()), f10(d10.begin(), d10.end());
    for (size_t k = 0; k < i01.size(); ++k) v01[k] = (int)i01[k].j_;
    for (size_t k = 0; k < i10.size(); ++k) v10[k] = (int)i10[k].j_;
    std::vector<IndMatch> out;
    Filter(NN, v01, f01, v10, f10, out, (eMatchFilter)filter, ratio);
    int n = 0;
    for (const IndMatch& m : out) { if (n < cap) { pairs[2 * n] = m.i_; pairs[2 * n + 1] = m.j_; } ++n; }
    return n;
}

static void fill_regions(openMVG::features::AKAZE_Binary_Regions& r, const float* feats, const uint8_t* desc, int n) {
    r.Features().resize(n); r.Descriptors().resize(n);
    for (int i = 0; i < n; ++i) {
        r.Features()[i] = openMVG::features::SIOPointFeature(feats[4 * i], feats[4 * i + 1], feats[4 * i + 2], feats[4 * i + 3]);
        std::memcpy(&r.Descriptors()[i], desc + 64 * (size_t)i, 64);
    }
}

// Binary_Regions<SIOPointFeature,64>::Save: feats = n x (x, y, scale, orientation), desc = n x 64 B.  0 = ok.
int ref_regions_save(const float* feats, const uint8_t* desc, int n, const char* feat_path, const char* desc_path) {
    openMVG::features::AKAZE_Binary_Regions r;
    fill_regions(r, feats, desc, n);
    return r.Save(feat_path, desc_path) ? 0 : 1;
}

// Binary_Regions::Load; returns the region count (or -1), copies at most cap rows out
int ref_regions_load(const char* feat_path, const char* desc_path, float* feats, uint8_t* desc, int cap, int* n_desc) {
    openMVG::features::AKAZE_Binary_Regions r;
    if (!r.Load(feat_path, desc_path)) return -1;
    const int n = (int)r.Features().size();
    if (n_desc) *n_desc = (int)r.Descriptors().size();
    for (int i = 0; i < n && i < cap; ++i) {
        const auto& f = r.Features()[i];
        feats[4 * i] = f.x(); feats[4 * i + 1] = f.y(); feats[4 * i + 2] = f.scale(); feats[4 * i + 3] = f.orientation();
    }
    for (int i = 0; i < (int)r.Descriptors().size() && i < cap; ++i) std::memcpy(desc + 64 * (size_t)i, &r.Descriptors()[i], 64);
    return n;
}

// matching::DistanceRatioMatch(ratio, BRUTE_FORCE_HAMMING, regions_I (database), regions_J (queries), out) exactly as
// CPUMatcher.hpp:67-76 calls it.  out = (i_, j_) pairs; returns the number of matches (may exceed cap).
int ref_distance_ratio_match(const float* featsI, const uint8_t* descI, int nI, const float* featsJ, const uint8_t* descJ, int nJ,
                             float ratio, uint32_t* out, int cap) {
    openMVG::features::AKAZE_Binary_Regions a, b;
    fill_regions(a, featsI, descI, nI);
    fill_regions(b, featsJ, descJ, nJ);
    openMVG::matching::IndMatches m;
    openMVG::matching::DistanceRatioMatch(ratio, openMVG::matching::BRUTE_FORCE_HAMMING, a, b, m);
    int n = 0;
    for (const auto& x : m) { if (n < cap) { out[2 * n] = x.i_; out[2 * n + 1] = x.j_; } ++n; }
    return n;
}

// the reference's own error functors (solver_fundamental_kernel.cpp:156-192, solver_homography_kernel.hpp:59-63); model row-major
void ref_model_residuals(int kind, const double* model, const double* x1, const double* x2, int n, double* out) {
    using namespace openMVG;
    Mat3 M;
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) M(r, c) = model[3 * r + c];
    for (int i = 0; i < n; ++i) {
        const Vec2 a(x1[2 * i], x1[2 * i + 1]), b(x2[2 * i], x2[2 * i + 1]);
        switch (kind) {
            case 0: out[i] = fundamental::kernel::EpipolarDistanceError::Error(M, a, b); break;
            case 1: out[i] = fundamental::kernel::SymmetricEpipolarDistanceError::Error(M, a, b); break;
            case 2: out[i] = fundamental::kernel::SampsonError::Error(M, a, b); break;
            default: out[i] = homography::kernel::AsymmetricError::Error(M, a, b); break;
        }
    }
}

int ref_hw_threads() { return (int)std::thread::hardware_concurrency(); }

}
