// common.cuh -- shared device/host helpers for libcoloc_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda.h>        // CUtensorMap (types only; the driver entry point is looked up at run time, tmap.cu)
#include <stdint.h>
#include <stdio.h>
#include <atomic>
#include "../../include/coloc_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libcoloc_b200 is written for sm_100a (B200) only"
#endif

namespace clb {

extern thread_local char g_err[512];
extern std::atomic<int64_t> g_launches;
extern std::atomic<int> g_timing;
// process-wide planning knobs (clb_set_param; atomics, read once per call) and diagnostics of the last call
extern std::atomic<int> g_match_rq, g_match_chunks, g_match_tc, g_match_tc_chunks, g_match_slice_rows, g_match_balance;
extern std::atomic<int> g_last_match_path, g_last_match_balanced;
#ifdef CLB_LAB      // kernel-variant and timing knobs exist in the lab build only (make -C coloc_b200/csrc lab)
extern std::atomic<int> g_match_hs, g_match_tc_ver, g_match_tc_dbg, g_latch_tex, g_latch_var, g_fast_nms, g_fast_var;
#endif

int set_cuda_error(cudaError_t e, const char* what, const char* file, int line);

#define CLB_CUDA(call)                                                                         \
    do {                                                                                       \
        cudaError_t e__ = (call);                                                              \
        if (e__ != cudaSuccess) return ::clb::set_cuda_error(e__, #call, __FILE__, __LINE__);  \
    } while (0)

#define CLB_LAUNCHED()                                                                         \
    do {                                                                                       \
        ::clb::g_launches.fetch_add(1, std::memory_order_relaxed);                             \
        cudaError_t e__ = cudaGetLastError();                                                  \
        if (e__ != cudaSuccess) return ::clb::set_cuda_error(e__, "kernel launch", __FILE__, __LINE__); \
    } while (0)

// optional per-kernel timing with events on the launching stream (bench's roofline leg)
struct KernelTimer {
    cudaEvent_t a = nullptr, b = nullptr;
    float last_ms = 0.f;
    bool pending = false;
    void begin(cudaStream_t s) {
        if (!g_timing) return;
        if (!a) { cudaEventCreate(&a); cudaEventCreate(&b); }
        cudaEventRecord(a, s);
    }
    void end(cudaStream_t s) {
        if (!g_timing) return;
        cudaEventRecord(b, s); pending = true;
    }
    float ms() {
        if (pending) { cudaEventSynchronize(b); cudaEventElapsedTime(&last_ms, a, b); pending = false; }
        return last_ms;
    }
};
KernelTimer& timer(const char* which);
// event timeline of one call (clb_set_timing(2)): tl_mark records a named event on a stream, clb_timeline formats the
// offsets (ms from the first mark) after synchronising; for the step breakdowns under profiles/
void tl_mark(const char* name, cudaStream_t s);

// NVTX ranges around the stages (resize, fast, describe, expand, match, merge), behind COLOC_B200_NVTX=1 in the environment so
// that a timeline tool shows the host side of every launch; nvtx3 is header-only and binds the tool's library at run time
bool nvtx_on();
void nvtx_push(const char* name);
void nvtx_pop();
struct NvtxRange {
    bool on;
    explicit NvtxRange(const char* name) : on(nvtx_on()) { if (on) nvtx_push(name); }
    ~NvtxRange() { if (on) nvtx_pop(); }
};

// a growable device buffer
struct DevBuf {
    void* p = nullptr; size_t cap = 0;
    cudaError_t reserve(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 4 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

}  // namespace clb

// ---- TMA / mbarrier PTX shared by the image kernels ----------------------------------------------
namespace clb {
#ifdef __CUDACC__
__device__ __forceinline__ uint32_t ptx_smem(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void ptx_mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(ptx_smem(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void ptx_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(ptx_smem(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void ptx_mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\t"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(ptx_smem(bar)), "r"(parity) : "memory");
    } while (!ok);
}
// one box of a 2-D tensor map -> shared memory (SASS UTMALDG); rows / columns outside the tensor are zero-filled
__device__ __forceinline__ void ptx_tma_load_2d(void* dst, const CUtensorMap* map, int x, int y, uint64_t* bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(ptx_smem(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(x), "r"(y), "r"(ptx_smem(bar)) : "memory");
}
// programmatic dependent launch: wait for the grids this one depends on / let the dependent grid start its prologue
__device__ __forceinline__ void ptx_grid_dep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void ptx_grid_dep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#endif
#ifdef __CUDACC__
// kernel launch with the programmatic-dependent-launch attribute (the kernel calls griddepcontrol.wait before it touches what
// the grid in front of it wrote): its CTAs are scheduled while that grid drains instead of after it
template <class... KArgs, class... Args>
inline cudaError_t launch_ex(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, bool pdl, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}
#endif
int make_tmap_2d(CUtensorMap* out, const void* base, int elem, uint64_t w_elems, uint64_t h, uint64_t stride, uint32_t box_w, uint32_t box_h);
}  // namespace clb

// ---- stage launchers (defined in the .cu files) --------------------------------------------------
namespace clb {

struct LevelTable {            // passed by value to kernels (fits in param space)
    int n;
    int w[CLB_MAX_LEVELS], h[CLB_MAX_LEVELS], stride[CLB_MAX_LEVELS];
    float f[CLB_MAX_LEVELS];
    const uint8_t* img[CLB_MAX_LEVELS];
};

struct LevelTex {              // optional point-sampled texture views of the levels (LATCH gather)
    cudaTextureObject_t t[CLB_MAX_LEVELS];
};
int  make_level_textures(const LevelTable& lt, LevelTex& tx);   // levels need 512-B aligned bases, 32-B aligned strides
void destroy_level_textures(LevelTex& tx, int n);

// match.cu
size_t match_scratch_bytes(int nQ, int nT);
int launch_match(const uint64_t* dQ, int nQ, const uint64_t* dT, int nT, int64_t t_base,
                 void* scratch, size_t scratch_bytes, int sm_count,
                 int threshold, float ratio, int32_t* d_match, int32_t* d_best, int32_t* d_second,
                 clb_partial* d_partial, cudaStream_t s,
                 const void* Qx = nullptr, const void* Tx = nullptr);   // expanded operands -> tensor-core path
int launch_merge_partials(const clb_partial* parts, int nparts, int nQ, int threshold,
                          int32_t* d_match, int32_t* d_best, int32_t* d_second, cudaStream_t s);
int launch_fold_partials(const clb_partial* parts, int nparts, int nQ, clb_partial* out, cudaStream_t s);
// match_tc.cu -- tcgen05 path: +-1 E4M3 expansion of the packed rows, then a tensor-core distance sweep
bool   tc_wanted(int nQ, int nT);
size_t tc_q_bytes(int nQ);
size_t tc_t_bytes(int nT);
int launch_expand(const uint64_t* d_rows, int n, size_t out_bytes, void* d_out, cudaStream_t s);
void tc_plan(int nQ, int nT, int sm_count, int* qblocks_out, int* nchunks_out, int* tpc_out);
int  tc_version();
bool tc_needs_qx();      // v1 kernel (queries staged in shared memory) wants the expanded query image too
int launch_match_tc_kernel(const uint64_t* dQ, const void* Qx, int nQ, const void* Tx, int nT, int qblocks, int nchunks, int tpc,
                           uint2* part, cudaStream_t s);
// batched pair lists (api_fmap.cu): one v4 CTA per TcJob, one merge row block per MergeJob; operands are E2M1 images (expand4)
struct TcJob {
    const uint32_t* Q;        // packed query rows of the pair's slice (row 0 = first row of the slice)
    const uint8_t* Tx;        // expanded (E2M1, pre-swizzled) image of the pair's training set
    uint2* partial_chunk;     // partial[chunk][0 .. nQ)
    int nQ, qblock, tile0, ntiles, chunk_rows, _pad;
};
struct MergeJob {
    const uint2* partial;             // [piece slot][nQ]
    int32_t *match, *best, *second;   // any may be NULL
    clb_partial* part_out;            // optional per-query partial (row indices + t_base)
    long long t_base;
    // uniform chunks: `nchunks` pieces of `rows_per_chunk` rows for every query (pieces == NULL); balanced jobs: query block b
    // (256 queries) has pieces[b] pieces, piece k starting at DB row base_rows[b * maxp + k]
    const int* pieces;
    const int* base_rows;
    int nQ, nT, nchunks, rows_per_chunk, maxp, _pad;
};
size_t tc4_t_bytes(int nT);
double tc_cta_cost_us(int tiles);
double tc_balanced_cost_us(long long units, long long strips, int sm_count, long long* parts_out);   // modelled time of the balanced job layout
int launch_expand4(const uint64_t* d_rows, int n, void* d_out, cudaStream_t s);
int tc_plan_pairs(const int* nQ, const int* nT, int npairs, int sm_count, double* t_us_out = nullptr);   // -> DB tiles per chunk, shared by all pairs (+ modelled time)
int launch_match_tc_jobs(const TcJob* d_jobs, int njobs, cudaStream_t s);
int launch_merge_jobs(const MergeJob* d_jobs, int njobs, int max_nQ, int threshold, float ratio, cudaStream_t s);
// TMA tensor maps of one frame geometry (tmap.cu builds them, the detector caches them per (w, h)):
//   fast[L]    8-bit map over level L, box = the FAST tile + halo (80 x 40)
//   resize[L]  8-bit map over LEVEL 0, box = the source footprint of one output tile of level L
struct alignas(64) LevelMaps {
    CUtensorMap fast[CLB_MAX_LEVELS];
    CUtensorMap resize[CLB_MAX_LEVELS];
};
// pyramid.cu
struct ResizePlan {            // per-level source-box geometry + the column / row tables of every level, device-resident
    int tile_w[CLB_MAX_LEVELS], tile_h[CLB_MAX_LEVELS];   // output tile of a CTA (64 x 32 unless the scale is extreme)
    int tile_base[CLB_MAX_LEVELS + 1], tiles_x[CLB_MAX_LEVELS];   // first tile of each level, tiles per tile row
    uint32_t tiles_x_magic[CLB_MAX_LEVELS];                // tile / tiles_x == umulhi(tile, magic) (verified on the host)
    int box_w[CLB_MAX_LEVELS], box_h[CLB_MAX_LEVELS];     // bytes x rows staged per output tile (box_w a multiple of 16, <= 256)
    int col_off[CLB_MAX_LEVELS], row_off[CLB_MAX_LEVELS]; // first table entry of each level
    const uint2* tab;                                      // entries {origin (bit 31 = clamped at the edge), weight}
};
size_t resize_table_entries(const LevelTable& lt);
int resize_plan(const LevelTable& lt, ResizePlan& rp);                                  // box geometry (host)
int launch_resize_tables(const LevelTable& lt, const ResizePlan& rp, uint2* d_tab, cudaStream_t s);
int launch_resize(const LevelTable& lt, const ResizePlan& rp, const LevelMaps& maps, uint32_t* clear, int clear_words, cudaStream_t s, bool pdl = false);   // fills levels 1..n-1 from level 0; also zeroes clear[0..clear_words)
// fast.cu
size_t fast_scratch_bytes(const LevelTable& lt);
int fast_tile_box(int* w, int* h);
int launch_fast(const LevelTable& lt, const LevelMaps& maps, int thresh, int quirk, void* scratch, size_t scratch_bytes,
                clb_keypoint* d_kps, int cap, int* d_total, cudaStream_t s, bool pdl = false, bool standalone = true);
void fast_clear_region(const LevelTable& lt, void* scratch, uint32_t** ptr, int* words);
// latch.cu
int launch_describe(const LevelTable& lt, const clb_keypoint* d_kps, int n, const int* d_n, int cap,
                    uint64_t* d_desc, cudaStream_t s, const LevelTex* tex = nullptr, bool pdl = false);
int launch_sincos(const float* d_a, int n, float* d_s, float* d_c, cudaStream_t s);
int latch_init();   // uploads the triplet table
// microbench.cu
int microbench_pipes(double* out8);
#ifdef CLB_LAB
int tc_lab_tile_clocks(long long* out1024);      // lab build: per-tile clocks of CTA (0, 0) of the last match_tc_dbg = 7 sweep
#endif
int microbench_mma(double* out6);      // match_tc.cu: bare tcgen05.mma kind::mxf4 stream on every SM

}  // namespace clb
