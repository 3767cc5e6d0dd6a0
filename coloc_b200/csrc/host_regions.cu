// host_regions.cu -- host-side pieces either side of the GPU path; no kernels here.
//   * region packing                 GPUDetector.hpp:167-182 (Keypoint -> SIOPointFeature)
//   * .feat / .desc files            Binary_Regions::Save / Load, binary_regions.hpp:51-66, descriptor.hpp:165-210, feature.hpp:148-182
//   * CPU-build match semantics      matching::DistanceRatioMatch (regions_matcher.hpp:137-184): ratio filter on the GPU 2-NN,
//                                    then IndMatch dedup (indMatch.hpp:43-52) and (x,y) dedup (indMatchDecoratorXY.hpp:43-70,131-146)
#include "matcher_priv.cuh"
#include <cmath>
#include <fstream>
#include <set>
#include <string>
#include <vector>

using namespace clb;

namespace {

struct Ind { uint32_t i, j; };
struct IndLess {
    bool operator()(const Ind& a, const Ind& b) const { return a.i < b.i || (a.i == b.i && a.j < b.j); }
};

// the decorator's ordering is NOT a strict weak order (it compares y1 whenever the x1 differ and nothing else), so the
// surviving set and its order depend on the tree insertion sequence; the same std::set range construction is used here
struct XYMatch { float x1, y1, x2, y2; Ind m; };
struct XYLess {
    bool operator()(const XYMatch& a, const XYMatch& b) const {
        if (a.x1 == b.x1 && a.y1 == b.y1 && a.x2 == b.x2 && a.y2 == b.y2) return false;
        if (a.x1 < b.x1 || a.x1 > b.x1) return a.y1 < b.y1;
        return false;
    }
};

int dedup(const int32_t* match, int nQ, const float* featsI, int strideI, const float* featsJ, int strideJ,
          uint32_t* pairs_out, int cap, int* n_out) {
    std::vector<Ind> put;
    for (int q = 0; q < nQ; ++q)
        if (match[q] >= 0) put.push_back(Ind{(uint32_t)match[q], (uint32_t)q});          // (database idx, query idx), regions_matcher.hpp:169
    const std::set<Ind, IndLess> uniq(put.begin(), put.end());
    std::vector<XYMatch> dec;
    dec.reserve(uniq.size());
    for (const Ind& m : uniq)
        dec.push_back(XYMatch{featsI[(size_t)m.i * strideI], featsI[(size_t)m.i * strideI + 1],
                              featsJ[(size_t)m.j * strideJ], featsJ[(size_t)m.j * strideJ + 1], m});
    const std::set<XYMatch, XYLess> kept(dec.begin(), dec.end());
    int n = 0;
    for (const XYMatch& k : kept) {
        if (n < cap) { pairs_out[2 * n] = k.m.i; pairs_out[2 * n + 1] = k.m.j; }
        ++n;
    }
    *n_out = n;
    return n > cap ? CLB_ECAPACITY : CLB_OK;
}

}  // namespace

extern "C" {

int clb_pack_regions(const clb_keypoint* kps, int n, float* feats_out) {
    if (n < 0 || (n > 0 && (!kps || !feats_out))) return CLB_EINVAL;
    for (int i = 0; i < n; ++i) {
        const float scale = static_cast<float>(std::pow(1.2f, kps[i].scale));       // the reference hard-codes 1.2f (:173)
        feats_out[4 * i] = scale * static_cast<float>(kps[i].x);
        feats_out[4 * i + 1] = scale * static_cast<float>(kps[i].y);
        feats_out[4 * i + 2] = 7.0f * scale;
        feats_out[4 * i + 3] = kps[i].angle;
    }
    return CLB_OK;
}

int clb_regions_save(const char* feat_path, const char* desc_path, const float* feats, const uint64_t* desc, int n) {
    if (n < 0 || (n > 0 && !desc) || (feat_path && n > 0 && !feats)) return CLB_EINVAL;
    if (feat_path) {
        std::ofstream f(feat_path);
        if (!f.is_open()) { snprintf(g_err, sizeof(g_err), "cannot open %s", feat_path); return CLB_EIO; }
        for (int i = 0; i < n; ++i) f << feats[4 * i] << " " << feats[4 * i + 1] << " " << feats[4 * i + 2] << " " << feats[4 * i + 3] << "\n";
        if (!f.good()) { snprintf(g_err, sizeof(g_err), "write failed: %s", feat_path); return CLB_EIO; }
    }
    if (desc_path) {
        std::ofstream f(desc_path, std::ios::out | std::ios::binary);
        if (!f.is_open()) { snprintf(g_err, sizeof(g_err), "cannot open %s", desc_path); return CLB_EIO; }
        const std::size_t count = (std::size_t)n;
        f.write(reinterpret_cast<const char*>(&count), sizeof(count));
        if (n > 0) f.write(reinterpret_cast<const char*>(desc), (std::streamsize)count * 64);
        if (!f.good()) { snprintf(g_err, sizeof(g_err), "write failed: %s", desc_path); return CLB_EIO; }
    }
    return CLB_OK;
}

int clb_regions_count(const char* desc_path, long long* n_out) {
    if (!desc_path || !n_out) return CLB_EINVAL;
    std::ifstream f(desc_path, std::ios::in | std::ios::binary);
    if (!f.is_open()) { snprintf(g_err, sizeof(g_err), "cannot open %s", desc_path); return CLB_EIO; }
    std::size_t count = 0;
    f.read(reinterpret_cast<char*>(&count), sizeof(count));
    if (!f.good()) { snprintf(g_err, sizeof(g_err), "short read: %s", desc_path); return CLB_EIO; }
    *n_out = (long long)count;
    return CLB_OK;
}

int clb_regions_load(const char* feat_path, const char* desc_path, float* feats_out, uint64_t* desc_out, int cap, int* n_out) {
    if (!n_out || cap < 0 || (!feat_path && !desc_path)) return CLB_EINVAL;
    long long nd = -1, nf = -1;
    if (desc_path) {
        std::ifstream f(desc_path, std::ios::in | std::ios::binary);
        if (!f.is_open()) { snprintf(g_err, sizeof(g_err), "cannot open %s", desc_path); return CLB_EIO; }
        std::size_t count = 0;
        f.read(reinterpret_cast<char*>(&count), sizeof(count));
        if (!f.good()) { snprintf(g_err, sizeof(g_err), "short read: %s", desc_path); return CLB_EIO; }
        nd = (long long)count;
        if (nd > cap) { *n_out = nd > 0x7FFFFFFF ? 0x7FFFFFFF : (int)nd; return CLB_ECAPACITY; }
        if (desc_out && nd > 0) {
            f.read(reinterpret_cast<char*>(desc_out), (std::streamsize)nd * 64);
            if (f.gcount() != (std::streamsize)nd * 64) { snprintf(g_err, sizeof(g_err), "short read: %s", desc_path); return CLB_EIO; }
        }
    }
    if (feat_path) {
        std::ifstream f(feat_path);
        if (!f.is_open()) { snprintf(g_err, sizeof(g_err), "cannot open %s", feat_path); return CLB_EIO; }
        float v[4];
        nf = 0;
        while (f >> v[0] >> v[1] >> v[2] >> v[3]) {
            if (nf < cap && feats_out) for (int k = 0; k < 4; ++k) feats_out[4 * nf + k] = v[k];
            ++nf;
        }
        if (nf > cap) { *n_out = nf > 0x7FFFFFFF ? 0x7FFFFFFF : (int)nf; return CLB_ECAPACITY; }
    }
    if (nd >= 0 && nf >= 0 && nd != nf) { snprintf(g_err, sizeof(g_err), "%lld features but %lld descriptors", nf, nd); *n_out = (int)(nd < nf ? nd : nf); return CLB_EINVAL; }
    *n_out = (int)(nd >= 0 ? nd : nf);
    return CLB_OK;
}

int clb_dedup_matches(const int32_t* match, int nQ, const float* featsI, int strideI, const float* featsJ, int strideJ,
                      uint32_t* pairs_out, int cap, int* n_out) {
    if (nQ < 0 || (nQ > 0 && (!match || !featsI || !featsJ)) || strideI < 2 || strideJ < 2 || cap < 0 || (cap > 0 && !pairs_out) || !n_out)
        return CLB_EINVAL;
    return dedup(match, nQ, featsI, strideI, featsJ, strideJ, pairs_out, cap, n_out);
}

int clb_distance_ratio_match(clb_matcher* m, const uint64_t* descI, const float* featsI, int nI,
                             const uint64_t* descJ, const float* featsJ, int nJ, float ratio,
                             uint32_t* pairs_out, int cap, int* n_out) {
    if (!m || !n_out || nI < 0 || nJ < 0) return CLB_EINVAL;
    *n_out = 0;
    if (nI == 0 || nJ == 0) return CLB_OK;                       // Match() returns before searching (regions_matcher.hpp:127-128, matcher_brute_force.hpp:107-112)
    std::vector<int32_t> match((size_t)nJ);
    if (nI < 2) return CLB_OK;                                    // SearchNeighbours refuses NN > rows: no matches
    int rc = clb_match_2nn_ratio(m, descJ, nJ, descI, nI, ratio, match.data(), nullptr, nullptr);
    if (rc) return rc;
    return clb_dedup_matches(match.data(), nJ, featsI, 4, featsJ, 4, pairs_out, cap, n_out);
}

}  // extern "C"
