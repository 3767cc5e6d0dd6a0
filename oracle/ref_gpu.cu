// ref_gpu.cu -- harness around the REFERENCE's three CUDA kernels, compiled unmodified from
// /root/reference/src/{CUDALERP,CLATCH,CUDAK2NN}.cu by oracle/Makefile into oracle/_ref/libcoloc_ref_gpu.so.
// TEST INFRASTRUCTURE: second oracle on the GPU box (pins the C restatement against the real kernels)
// and the "reference kernel on B200" column of BASELINE.md.  The texture / buffer set-up below follows
// what GPUDetector.hpp:70-138,216-291 and GPUMatcher.hpp:180-226 do around the kernels.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstring>
#include <cstdio>
#include "coloc/CUDALERP.h"
#include "coloc/CLATCH.h"
#include "coloc/CUDAK2NN.h"

#define CK(x) do { cudaError_t e__ = (x); if (e__ != cudaSuccess) { fprintf(stderr, "ref_gpu: %s -> %s\n", #x, cudaGetErrorString(e__)); return 1; } } while (0)

static_assert(sizeof(Keypoint) == 20, "Keypoint.h layout");

static cudaTextureDesc img_texdesc(cudaTextureReadMode mode) {
    cudaTextureDesc t; memset(&t, 0, sizeof(t));        // GPUDetector.hpp:96-100
    t.addressMode[0] = cudaAddressModeClamp; t.addressMode[1] = cudaAddressModeClamp;
    t.filterMode = cudaFilterModePoint; t.normalizedCoords = 0; t.readMode = mode;
    return t;
}

extern "C" {

// one pyramid level from level 0 (GPUDetector.hpp:232-255)
int ref_gpu_lerp(const uint8_t* img, int w, int h, float f, uint8_t* out, int neww, int newh) {
    cudaChannelFormatDesc cd = cudaCreateChannelDesc(8, 0, 0, 0, cudaChannelFormatKindUnsigned);
    cudaArray_t arr;
    CK(cudaMallocArray(&arr, &cd, w, h, cudaArrayTextureGather));
    CK(cudaMemcpy2DToArray(arr, 0, 0, img, w, w, h, cudaMemcpyHostToDevice));
    cudaResourceDesc rd; memset(&rd, 0, sizeof(rd)); rd.resType = cudaResourceTypeArray; rd.res.array.array = arr;
    cudaTextureDesc td = img_texdesc(cudaReadModeNormalizedFloat);
    cudaTextureObject_t tex;
    CK(cudaCreateTextureObject(&tex, &rd, &td, nullptr));
    uint8_t* d_out; size_t pitch;
    CK(cudaMallocPitch(&d_out, &pitch, neww, newh));
    CK(cudaMemset2D(d_out, pitch, 0, neww, newh));
    CUDALERP(tex, f, f, d_out, pitch, neww, newh, 0);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy2D(out, neww, d_out, pitch, neww, newh, cudaMemcpyDeviceToHost));
    cudaDestroyTextureObject(tex); cudaFree(d_out); cudaFreeArray(arr);
    return 0;
}

// CLATCH on caller-supplied level images (GPUDetector.hpp:70-94,239-244,280-288).  ms_out (optional): kernel time.
int ref_gpu_latch(const uint8_t* const* levels, const int* ws, const int* hs, int nlevels,
                  const Keypoint* kps, int n, uint64_t* desc, int iters, float* ms_out) {
    if (nlevels > 8 || n <= 0) return 1;
    // triplet table -> 1-D ushort4 array texture
    cudaChannelFormatDesc ct = cudaCreateChannelDesc(16, 16, 16, 16, cudaChannelFormatKindUnsigned);
    cudaArray_t trip_arr;
    CK(cudaMallocArray(&trip_arr, &ct, 512));
    CK(cudaMemcpy2DToArray(trip_arr, 0, 0, triplets, 2048 * sizeof(uint16_t), 2048 * sizeof(uint16_t), 1, cudaMemcpyHostToDevice));
    cudaResourceDesc rt; memset(&rt, 0, sizeof(rt)); rt.resType = cudaResourceTypeArray; rt.res.array.array = trip_arr;
    cudaTextureDesc tt; memset(&tt, 0, sizeof(tt));
    tt.addressMode[0] = cudaAddressModeClamp; tt.filterMode = cudaFilterModePoint; tt.readMode = cudaReadModeElementType; tt.normalizedCoords = 0;
    cudaTextureObject_t trip_tex;
    CK(cudaCreateTextureObject(&trip_tex, &rt, &tt, nullptr));
    // level textures: level 0 a cudaArray, the others pitched 2-D (element type)
    cudaChannelFormatDesc cd = cudaCreateChannelDesc(8, 0, 0, 0, cudaChannelFormatKindUnsigned);
    cudaTextureObject_t tex[8]; memset(tex, 0, sizeof(tex));
    cudaArray_t arr0 = nullptr; uint8_t* d_lv[8] = {nullptr};
    cudaTextureDesc td = img_texdesc(cudaReadModeElementType);
    for (int i = 0; i < nlevels; ++i) {
        if (!levels[i]) continue;
        cudaResourceDesc rd; memset(&rd, 0, sizeof(rd));
        if (i == 0) {
            CK(cudaMallocArray(&arr0, &cd, ws[0], hs[0], cudaArrayTextureGather));
            CK(cudaMemcpy2DToArray(arr0, 0, 0, levels[0], ws[0], ws[0], hs[0], cudaMemcpyHostToDevice));
            rd.resType = cudaResourceTypeArray; rd.res.array.array = arr0;
        } else {
            size_t pitch;
            CK(cudaMallocPitch(&d_lv[i], &pitch, ws[i], hs[i]));
            CK(cudaMemcpy2D(d_lv[i], pitch, levels[i], ws[i], ws[i], hs[i], cudaMemcpyHostToDevice));
            rd.resType = cudaResourceTypePitch2D; rd.res.pitch2D.desc = cd; rd.res.pitch2D.devPtr = d_lv[i];
            rd.res.pitch2D.height = hs[i]; rd.res.pitch2D.pitchInBytes = pitch; rd.res.pitch2D.width = ws[i];
        }
        CK(cudaCreateTextureObject(&tex[i], &rd, &td, nullptr));
    }
    cudaTextureObject_t* d_tex; Keypoint* d_kps; uint64_t* d_desc;
    CK(cudaMalloc(&d_tex, 8 * sizeof(cudaTextureObject_t)));
    CK(cudaMemcpy(d_tex, tex, 8 * sizeof(cudaTextureObject_t), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&d_kps, (size_t)n * sizeof(Keypoint)));
    CK(cudaMemcpy(d_kps, kps, (size_t)n * sizeof(Keypoint), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&d_desc, (size_t)n * 64));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    if (iters < 1) iters = 1;
    CLATCH(d_tex, trip_tex, d_kps, n, d_desc);            // warm-up + result
    CK(cudaDeviceSynchronize());
    cudaEventRecord(e0);
    for (int i = 0; i < iters; ++i) CLATCH(d_tex, trip_tex, d_kps, n, d_desc);
    cudaEventRecord(e1);
    CK(cudaDeviceSynchronize());
    if (ms_out) { cudaEventElapsedTime(ms_out, e0, e1); *ms_out /= iters; }
    CK(cudaMemcpy(desc, d_desc, (size_t)n * 64, cudaMemcpyDeviceToHost));
    for (int i = 0; i < nlevels; ++i) { if (tex[i]) cudaDestroyTextureObject(tex[i]); if (d_lv[i]) cudaFree(d_lv[i]); }
    if (arr0) cudaFreeArray(arr0);
    cudaDestroyTextureObject(trip_tex); cudaFreeArray(trip_arr);
    cudaFree(d_tex); cudaFree(d_kps); cudaFree(d_desc); cudaEventDestroy(e0); cudaEventDestroy(e1);
    return 0;
}

// CUDAK2NN as GPUMatcher::computeMatches drives it (GPUMatcher.hpp:180-226).  ms_out: kernel (+its device sync) time.
int ref_gpu_k2nn(const uint64_t* Q, int nQ, const uint64_t* T, int nT, int threshold, int* match, int iters, float* ms_out) {
    if (nQ <= 0 || nT <= 0) return 1;
    uint64_t *dQ, *dT; int* dM;
    CK(cudaMalloc(&dT, (size_t)(nT + 8) * 64)); CK(cudaMemset(dT, 0, (size_t)(nT + 8) * 64));
    CK(cudaMemcpy(dT, T, (size_t)nT * 64, cudaMemcpyHostToDevice));
    CK(cudaMalloc(&dQ, (size_t)nQ * 64)); CK(cudaMemcpy(dQ, Q, (size_t)nQ * 64, cudaMemcpyHostToDevice));
    CK(cudaMalloc(&dM, (size_t)nQ * 4));
    cudaResourceDesc rd; memset(&rd, 0, sizeof(rd));     // GPUMatcher.hpp:79-91,198-201
    rd.resType = cudaResourceTypeLinear; rd.res.linear.desc.f = cudaChannelFormatKindUnsigned;
    rd.res.linear.desc.x = 32; rd.res.linear.desc.y = 32; rd.res.linear.devPtr = dQ; rd.res.linear.sizeInBytes = (size_t)nQ * 64;
    cudaTextureDesc td; memset(&td, 0, sizeof(td));
    td.addressMode[0] = cudaAddressModeBorder; td.addressMode[1] = cudaAddressModeBorder;
    td.filterMode = cudaFilterModePoint; td.readMode = cudaReadModeElementType; td.normalizedCoords = 0;
    cudaTextureObject_t tex;
    CK(cudaCreateTextureObject(&tex, &rd, &td, nullptr));
    if (iters < 1) iters = 1;
    CUDAK2NN(dT, nT, tex, nQ, dM, threshold);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    for (int i = 0; i < iters; ++i) CUDAK2NN(dT, nT, tex, nQ, dM, threshold);
    cudaEventRecord(e1);
    CK(cudaDeviceSynchronize());
    if (ms_out) { cudaEventElapsedTime(ms_out, e0, e1); *ms_out /= iters; }
    CK(cudaMemcpy(match, dM, (size_t)nQ * 4, cudaMemcpyDeviceToHost));
    cudaDestroyTextureObject(tex); cudaFree(dQ); cudaFree(dT); cudaFree(dM); cudaEventDestroy(e0); cudaEventDestroy(e1);
    return 0;
}

}
