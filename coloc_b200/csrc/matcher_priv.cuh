// matcher_priv.cuh -- the matcher object behind the C ABI, shared by api_matcher.cu and api_fmap.cu
#pragma once
#include "common.cuh"

struct clb_matcher {
    clb_matcher_opts o;
    int device, sm_count;
    cudaStream_t stream;
    clb::DevBuf q, t, map, scratch, out;   // out: match | best | second (3 x nQ int32)
    clb::DevBuf qx, tx, mapx;              // +-1 expansions for the tensor-core path (match_tc.cu)
    int n_map;
    bool mapx_valid;
    int mapx_ver;                          // kernel version the cached map expansion was made for
    // pair lists (api_fmap.cu): job tables staged in pinned memory, per-chunk partials, host-call output
    clb::DevBuf jobs, pair_part, pair_out;
    void* jobs_host;
    size_t jobs_host_cap;
    cudaEvent_t jobs_ev;
    bool jobs_ev_set;
    // host-resident DB streamed in slices (clb_match_partial_hostdb): ping-pong raw / expanded slice buffers, per-slice partials
    cudaStream_t copy_stream;
    clb::DevBuf sraw[2], sx[2], sparts, sfold;
    clb::DevBuf geo;                       // clb_model_residuals staging (geometry.cu)
    cudaEvent_t ev_copied[2], ev_rawfree[2], ev_entry;
    bool stream_ready;
};

namespace clb {
// a (query rows, expanded training set) problem of a job list (api_fmap.cu)
struct PairView { const uint64_t* Q; int nQ; const uint8_t* Tx; int nT; };
extern int g_match_balance, g_last_match_balanced;
}
int run_views(clb_matcher* m, const clb::PairView* v, int nv, int threshold, float ratio,
              int32_t* d_match, int32_t* d_best, int32_t* d_second, clb_partial* d_part_out, long long t_base, cudaStream_t s);
