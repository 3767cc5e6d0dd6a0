// coloc/GPUDetector.hpp -- drop-in for the reference's include/coloc/GPUDetector.hpp:29-292.
// Same class name, template shape, public members and method signatures, so that
//   FeatureDetector<bool, GPUDetector>   (FeatureDetector.hpp:21-32, coloc.hpp:64)
// compiles unchanged against it; every method forwards to the C ABI (include/coloc_b200.h).  No CUDA
// headers are needed by the host build.  Conventions kept from the reference: methods typed T=bool
// `return EXIT_SUCCESS` (=0) on success (GPUDetector.hpp:183, tested `== EXIT_SUCCESS` at coloc.hpp:241);
// errors print and exit like gpuErrchk (GPUMatcher.hpp:33-41).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include "coloc_b200.h"
// Three builds of the same shim:
//   (default)             inside the reference tree: real coloc/colocData.hpp (openMVG + SfM types) and cv::imread
//   CLB_SHIM_OPENMVG      the real openMVG boundary types (regions_factory.hpp, indMatch.hpp, types.hpp) without the SfM / ceres /
//                         OpenCV part of the tree: what this repo's compile test uses when /root/reference is present
//   CLB_SHIM_STANDALONE   no openMVG at all: the 60-line stand-ins of openmvg_min.hpp (tests on the GPU box, demos)
#if defined(CLB_SHIM_STANDALONE)
#include "openmvg_min.hpp"
#elif defined(CLB_SHIM_OPENMVG)
#include <map>
#include <memory>
#include "openMVG/types.hpp"
#include "openMVG/features/regions_factory.hpp"
#include "openMVG/matching/indMatch.hpp"
namespace coloc {      // include/coloc/colocData.hpp:19-42, the part of it the detector / matcher boundary touches
using FeatureMap = std::map<openMVG::IndexT, std::unique_ptr<openMVG::features::AKAZE_Binary_Regions>>;
struct DetectorOptions { float scale_factor; uint8_t scale_levels; unsigned int width, height, maxkp; uint8_t thresh; };
struct MatcherOptions { float distRatio; int thresh; unsigned int maxkp; };
}  // namespace coloc
#else
#include <opencv2/imgcodecs.hpp>
#include "coloc/colocData.hpp"
#endif

// include/coloc/Keypoint.h:155-164 (same layout as clb_keypoint; kept as its own type so host code that names
// `Keypoint` keeps compiling)
#ifndef CLB_SHIM_HAVE_KEYPOINT
#define CLB_SHIM_HAVE_KEYPOINT
struct Keypoint {
    int32_t x, y;
    uint8_t score;
    float angle;
    uint8_t scale;
    Keypoint() {}
    Keypoint(const int32_t _x, const int32_t _y, const uint8_t _score) : x(_x), y(_y), score(_score) {}
};
static_assert(sizeof(Keypoint) == sizeof(clb_keypoint), "Keypoint layout");
#endif

namespace coloc {

inline void clbCheck(int rc, const char* what) {
    if (rc != CLB_OK) {
        std::fprintf(stderr, "coloc_b200: %s failed: %s (%s)\n", what, clb_strerror(rc), clb_last_error());
        std::exit(rc);
    }
}

template <typename T>
class GPUDetector {
public:
    std::vector<Keypoint> kps;       // last frame's raw output, as GPUDetector.hpp:32-33
    std::vector<uint64_t> desc;
    bool receivedImg = false;

    GPUDetector(DetectorOptions opts) : thresh(opts.thresh) {
        clb_detector_opts o;
        o.scale_factor = opts.scale_factor; o.scale_levels = opts.scale_levels;
        o.width = (int32_t)opts.width; o.height = (int32_t)opts.height;
        // the reference sizes d_kps/d_desc to maxkp but never clamps (GPUDetector.hpp:135-136, SURVEY 3.1 note);
        // here the buffers grow to what a frame can produce so the latent overflow cannot happen
        o.maxkp = (int32_t)std::max<unsigned>(opts.maxkp, opts.width * opts.height / 16u + 1024u);
        o.thresh = opts.thresh;
        clbCheck(clb_detector_create(&o, 0, &det), "clb_detector_create");
        cap = o.maxkp;
    }
    ~GPUDetector() {}                         // the reference leaks on purpose; freeGPUMemory() is explicit
    void freeGPUMemory() {                    // GPUDetector.hpp:144-155
        clb_detector_destroy(det); det = nullptr;
        clb_host_free(h_kps); clb_host_free(h_desc); h_kps = nullptr; h_desc = nullptr;
    }

    // GPUDetector.hpp:158-184
    T detectFeaturesFile(uint8_t idx, coloc::FeatureMap& regions, std::string& imageName) {
        int w = 0, h = 0;
        std::vector<uint8_t> own;
        const uint8_t* data = loadGray(imageName, w, h, own);
        if (!data) { std::fprintf(stderr, "coloc_b200: cannot read %s\n", imageName.c_str()); return EXIT_FAILURE; }
        return detectFeaturesBuffer(idx, regions, data, w, h);
    }

    // the body of detectFeaturesFile after cv::imread -- also what detectFeaturesTopic (:188-212) needs
    T detectFeaturesBuffer(uint8_t idx, coloc::FeatureMap& regions, const uint8_t* gray, int w, int h) {
        detectAndDescribe(gray, (uint32_t)w, (uint32_t)h, thresh);
        using Regions = openMVG::features::AKAZE_Binary_Regions;
        regions[idx] = std::unique_ptr<Regions>(new Regions);
        regions[idx]->Features().resize(kps.size());
        regions[idx]->Descriptors().resize(kps.size());
        for (size_t i = 0; i < kps.size(); ++i) {
            const float scale = static_cast<float>(std::pow(1.2f, kps[i].scale));      // hard-coded 1.2f, :173
            regions[idx]->Features()[i] = {scale * static_cast<float>(kps[i].x), scale * static_cast<float>(kps[i].y),
                                           7.0f * scale, kps[i].angle};
            std::memcpy(&(regions[idx]->Descriptors()[i]), &(desc[i * 8]), 8 * sizeof(uint64_t));
        }
        receivedImg = true;
        return EXIT_SUCCESS;
    }

    // GPUDetector.hpp:188-212 (`#ifdef USE_STREAM`, a ROS topic frame).  The reference takes a cv_bridge::CvImagePtr; any pointer-like
    // object whose `->image` has `data`, `cols`, `rows` (cv::Mat's members) works here, so the shim needs neither ROS nor OpenCV.
    // The reference's copy of the packing loop writes `scale * x` into BOTH coordinates (:203, a typo: detectFeaturesFile :175-176
    // has x, y); this method packs x, y like detectFeaturesFile.
    template <class ImagePtr>
    void detectFeaturesTopic(uint8_t idx, coloc::FeatureMap& regions, ImagePtr imagePtr) {
        detectFeaturesBuffer(idx, regions, static_cast<const uint8_t*>(imagePtr->image.data), (int)imagePtr->image.cols,
                             (int)imagePtr->image.rows);
    }

private:
    // GPUDetector.hpp:216-291: H2D, pyramid, FAST+NMS+angle, LATCH, D2H -- one ABI call
    // The results land in page-locked scratch that lives as long as the detector (the describe kernel writes the descriptors
    // straight into it, clb_detect_describe); only the n entries of this frame are copied into the public vectors -- no
    // value-initialisation of `cap` entries (43 MB at 4K) per frame.
    void detectAndDescribe(const uint8_t* image, const uint32_t width, const uint32_t height, const uint8_t) {
        if (!h_kps) {
            void* p = nullptr;
            clbCheck(clb_host_alloc((size_t)cap * sizeof(clb_keypoint), &p), "clb_host_alloc"); h_kps = static_cast<clb_keypoint*>(p);
            clbCheck(clb_host_alloc((size_t)cap * 64, &p), "clb_host_alloc"); h_desc = static_cast<uint64_t*>(p);
        }
        int n = 0;
        clbCheck(clb_detect_describe(det, image, (int)width, (int)height, (int)width, h_kps, h_desc, cap, &n), "clb_detect_describe");
        kps.resize((size_t)n); desc.resize(8 * (size_t)n);
        if (n > 0) {
            std::memcpy(static_cast<void*>(kps.data()), h_kps, (size_t)n * sizeof(clb_keypoint));
            std::memcpy(desc.data(), h_desc, (size_t)n * 64);
        }
    }

    static const uint8_t* loadGray(const std::string& name, int& w, int& h, std::vector<uint8_t>& own) {
#if defined(CLB_SHIM_STANDALONE) || defined(CLB_SHIM_OPENMVG)
        // binary PGM (P5, maxval 255) only
        FILE* f = std::fopen(name.c_str(), "rb");
        if (!f) return nullptr;
        int maxv = 0;
        if (std::fscanf(f, "P5 %d %d %d", &w, &h, &maxv) != 3 || maxv != 255) { std::fclose(f); return nullptr; }
        std::fgetc(f);
        own.resize((size_t)w * h);
        const size_t got = std::fread(own.data(), 1, own.size(), f);
        std::fclose(f);
        return got == own.size() ? own.data() : nullptr;
#else
        static thread_local cv::Mat image;
        image = cv::imread(name, 0);                       // GPUDetector.hpp:161
        if (image.empty()) return nullptr;
        if (!image.isContinuous()) image = image.clone();
        w = image.cols; h = image.rows; (void)own;
        return image.data;
#endif
    }

    clb_detector* det = nullptr;
    clb_keypoint* h_kps = nullptr;   // page-locked result scratch, cap entries
    uint64_t* h_desc = nullptr;
    int cap = 0;
    const uint8_t thresh;
};

}  // namespace coloc
