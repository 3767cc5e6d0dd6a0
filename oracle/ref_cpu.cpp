// ref_cpu.cpp -- harness that compiles the REFERENCE's own CPU code, unmodified, from where it
// lies under /root/reference (include paths only; no reference source is copied here).
// TEST INFRASTRUCTURE: built by oracle/Makefile into oracle/_ref/libcoloc_ref_cpu.so.
//   * KFAST<mt,true>            include/coloc/KFAST.h:503-540
//   * featureAngle/fastAtan2    include/coloc/FeatureAngle.h:160-247
//   * ArrayMatcherBruteForce<uchar,Hamming<uchar>> + NNdistanceRatio
//       lib/openMVG/src/openMVG/matching/matcher_brute_force.hpp:100-200, metric_hamming.hpp:80-99,
//       matching_filters.hpp:39-60  (what CPUMatcher.hpp:67-76 -> regions_matcher.hpp:137-184 runs)
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

int ref_hw_threads() { return (int)std::thread::hardware_concurrency(); }

}
