// common.cuh -- shared device/host helpers for libcoloc_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
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
// pyramid.cu
int launch_resize(const LevelTable& lt, cudaStream_t s);   // fills levels 1..n-1 from level 0
// fast.cu
struct FastScratch { uint8_t* nms; int* row_count; int* row_off; int* total; };
size_t fast_scratch_bytes(const LevelTable& lt);
int launch_fast(const LevelTable& lt, int thresh, int quirk, void* scratch, size_t scratch_bytes,
                clb_keypoint* d_kps, int cap, int* d_total, cudaStream_t s);
// latch.cu
int launch_describe(const LevelTable& lt, const clb_keypoint* d_kps, int n, const int* d_n, int cap,
                    uint64_t* d_desc, cudaStream_t s, const LevelTex* tex = nullptr);
int launch_sincos(const float* d_a, int n, float* d_s, float* d_c, cudaStream_t s);
int latch_init();   // uploads the triplet table
// microbench.cu
int microbench_pipes(double* out8);
int microbench_mma(double* out6);      // match_tc.cu: bare tcgen05.mma kind::mxf4 stream on every SM

}  // namespace clb
