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
    cudaEvent_t ev_map;                    // recorded behind the copy of clb_set_map_dev / clb_set_map_from_fmap: map sweeps on any stream wait on it
    bool ev_map_set;
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
    // deferred head of a streamed slice's sweep: "wait for the slice copy, expand it" is enqueued by the sweep's launch path
    // AFTER its job-table upload, so that the small upload is not parked behind the wait (see match_partial_streamed)
    struct Pending { bool set; cudaEvent_t ev; const uint64_t* raw; int rows; size_t x_bytes; void* x; } pending;
    // an event the NEXT sweep launch must wait for (the query upload of clb_sharded_match_map on its own stream): waited for in
    // the launch path, not at call entry, so that the copy stream's entry event -- and with it the first DB slice -- is not
    // ordered behind that upload
    cudaEvent_t q_wait_ev;
    bool q_wait_set;
};

// enqueue a pending slice head on `s` (no-op when none is pending)
inline int clb_consume_pending(clb_matcher* m, cudaStream_t s) {
    if (m->q_wait_set) {
        m->q_wait_set = false;
        cudaError_t e = cudaStreamWaitEvent(s, m->q_wait_ev, 0);
        if (e != cudaSuccess) return clb::set_cuda_error(e, "cudaStreamWaitEvent", __FILE__, __LINE__);
    }
    if (!m->pending.set) return CLB_OK;
    m->pending.set = false;
    cudaError_t e = cudaStreamWaitEvent(s, m->pending.ev, 0);
    if (e != cudaSuccess) return clb::set_cuda_error(e, "cudaStreamWaitEvent", __FILE__, __LINE__);
    if (m->pending.x) return clb::launch_expand(m->pending.raw, m->pending.rows, m->pending.x_bytes, m->pending.x, s);
    return CLB_OK;
}

namespace clb {
// a (query rows, expanded training set) problem of a job list (api_fmap.cu)
struct PairView { const uint64_t* Q; int nQ; const uint8_t* Tx; int nT; };
}
int run_views(clb_matcher* m, const clb::PairView* v, int nv, int threshold, float ratio,
              int32_t* d_match, int32_t* d_best, int32_t* d_second, clb_partial* d_part_out, long long t_base, cudaStream_t s);
