// tmap.cu -- TMA tensor maps over the pyramid levels (host side).
//
// The image kernels stage their tiles with cp.async.bulk.tensor.2d (SASS UTMALDG): one instruction moves a box of a level into
// shared memory, rows outside the image are zero-filled by the copy engine, and the 32 lanes that used to issue the loads do
// arithmetic instead.  cuTensorMapEncodeTiled is a driver entry point; it is looked up through the runtime
// (cudaGetDriverEntryPoint), so the library does not link libcuda.
#include "common.cuh"

namespace clb {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
        cudaGetLastError();
    }
    return fn;
}

// 2-D map over a pitched 8-bit image: elements of `elem` bytes (1 or 4), `w_elems` x `h` of them, rows `stride` bytes apart
// (a multiple of 16), box `box_w` x `box_h` elements (box_w * elem a multiple of 16, each dimension <= 256).
int make_tmap_2d(CUtensorMap* out, const void* base, int elem, uint64_t w_elems, uint64_t h, uint64_t stride, uint32_t box_w, uint32_t box_h) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) { snprintf(g_err, sizeof(g_err), "cuTensorMapEncodeTiled is not available from this driver"); return CLB_EUNSUPPORTED; }
    if ((reinterpret_cast<uintptr_t>(base) & 15) || (stride & 15) || box_w > 256 || box_h > 256 || ((box_w * elem) & 15) || w_elems == 0 || h == 0) {
        snprintf(g_err, sizeof(g_err), "tensor map: base %p stride %llu box %ux%u (elem %d) violates the TMA alignment rules", base,
                 (unsigned long long)stride, box_w, box_h, elem);
        return CLB_EINVAL;
    }
    const cuuint64_t dims[2] = {w_elems, h};
    const cuuint64_t strides[1] = {stride};
    const cuuint32_t box[2] = {box_w, box_h};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = fn(out, elem == 4 ? CU_TENSOR_MAP_DATA_TYPE_UINT32 : CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), dims, strides,
                          box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { snprintf(g_err, sizeof(g_err), "cuTensorMapEncodeTiled failed (%d)", (int)r); return CLB_ECUDA; }
    return CLB_OK;
}

}  // namespace clb
