// api_matcher.cu -- matcher half of the C ABI (include/coloc_b200.h).
// Mirrors GPUMatcher (include/coloc/GPUMatcher.hpp:45-272): the object owns d_descQ / d_descT /
// d_descM and the match buffer; host entry points are blocking (copy in, run, copy out, sync).
#include "common.cuh"
#include <string.h>
#include <new>

using namespace clb;

#include "matcher_priv.cuh"

// One (Q, T) problem: the tensor-core sweep goes through the job-list machinery (api_fmap.cu) when its balanced layout is
// modelled faster than whole waves of equal CTAs (mid-sized problems); everything else through launch_match.
static int match_dispatch(clb_matcher* m, const uint64_t* dQ, int nQ, const uint64_t* dT, int nT, int64_t t_base,
                          int threshold, float ratio, int32_t* d_match, int32_t* d_best, int32_t* d_second,
                          clb_partial* d_partial, cudaStream_t s, const void* Qx, const void* Tx) {
    if (Tx && nT > 0 && nQ > 0 && tc_version() == 4 && g_match_balance != 0) {
        const int sm = m->sm_count > 0 ? m->sm_count : 148;
        const long long qb = (nQ + 255) / 256, tiles = (nT + 127) / 128, units = qb * tiles;
        double t_uniform = 0.0;
        tc_plan_pairs(&nQ, &nT, 1, sm, &t_uniform);
        long long parts = sm;
        const double t_bal = tc_balanced_cost_us(units, qb, sm, &parts);
        if (g_match_balance == 1 || (units >= 4 * parts && t_bal < 0.97 * t_uniform)) {
            clb::PairView v; v.Q = dQ; v.nQ = nQ; v.Tx = static_cast<const uint8_t*>(Tx); v.nT = nT;
            return run_views(m, &v, 1, threshold, ratio, d_match, d_best, d_second, d_partial, t_base, s);
        }
    }
    g_last_match_balanced = 0;
    { const int prc = clb_consume_pending(m, s); if (prc) return prc; }
    return launch_match(dQ, nQ, dT, nT, t_base, m->scratch.p, m->scratch.cap, m->sm_count, threshold, ratio,
                        d_match, d_best, d_second, d_partial, s, Qx, Tx);
}

// Expand the operands when the tensor-core path applies; *Qx / *Tx stay NULL for the POPC path.
// `map` = T is the resident map (expanded once per clb_set_map, not per query set).
static int prepare_tc(clb_matcher* m, const uint64_t* dQ, int nQ, const uint64_t* dT, int nT, bool map, cudaStream_t s,
                      const void** Qx, const void** Tx) {
    *Qx = nullptr; *Tx = nullptr;
    if (!tc_wanted(nQ, nT)) {
        if (map && m->ev_map_set) CLB_CUDA(cudaStreamWaitEvent(s, m->ev_map, 0));
        return CLB_OK;
    }
    const size_t qb = tc_q_bytes(nQ), tb = tc_t_bytes(nT);
    int rc = CLB_OK;
    if (tc_needs_qx()) {
        CLB_CUDA(m->qx.reserve(qb));
        rc = launch_expand(dQ, nQ, qb, m->qx.p, s);
        if (rc) return rc;
        *Qx = m->qx.p;
    }
    if (map && m->ev_map_set) CLB_CUDA(cudaStreamWaitEvent(s, m->ev_map, 0));      // the map may have been set on another stream
    if (map) {
        if (!m->mapx_valid || m->mapx_ver != tc_version()) {
            CLB_CUDA(m->mapx.reserve(tb));
            rc = launch_expand(dT, nT, tb, m->mapx.p, s);
            if (rc) return rc;
            m->mapx_valid = true; m->mapx_ver = tc_version();
        }
        *Tx = m->mapx.p;
    } else {
        CLB_CUDA(m->tx.reserve(tb));
        rc = launch_expand(dT, nT, tb, m->tx.p, s);
        if (rc) return rc;
        *Tx = m->tx.p;
    }
    return CLB_OK;
}

static int ensure(clb_matcher* m, int nQ, int nT) {
    CLB_CUDA(m->scratch.reserve(match_scratch_bytes(nQ, nT)));
    CLB_CUDA(m->out.reserve((size_t)(nQ > 0 ? nQ : 1) * 12));
    return CLB_OK;
}

// shared tail of the host entry points: run on device buffers, copy results out
static int run_host(clb_matcher* m, const uint64_t* dQ, int nQ, const uint64_t* dT, int nT, int threshold, float ratio,
                    int32_t* match_out, int32_t* best_out, int32_t* second_out, bool map = false) {
    cudaStream_t s = m->stream;
    int rc = ensure(m, nQ, nT);
    if (rc) return rc;
    const void *Qx, *Tx;
    rc = prepare_tc(m, dQ, nQ, dT, nT, map, s, &Qx, &Tx);
    if (rc) return rc;
    int32_t* o = m->out.as<int32_t>();
    rc = match_dispatch(m, dQ, nQ, dT, nT, 0, threshold, ratio,
                      o, best_out ? o + nQ : nullptr, second_out ? o + 2 * (size_t)nQ : nullptr, nullptr, s, Qx, Tx);
    if (rc) return rc;
    if (match_out) CLB_CUDA(cudaMemcpyAsync(match_out, o, (size_t)nQ * 4, cudaMemcpyDeviceToHost, s));
    if (best_out) CLB_CUDA(cudaMemcpyAsync(best_out, o + nQ, (size_t)nQ * 4, cudaMemcpyDeviceToHost, s));
    if (second_out) CLB_CUDA(cudaMemcpyAsync(second_out, o + 2 * (size_t)nQ, (size_t)nQ * 4, cudaMemcpyDeviceToHost, s));
    CLB_CUDA(cudaStreamSynchronize(s));
    return CLB_OK;
}

namespace clb { std::atomic<int> g_match_slice_rows{131072}; }     // clb_set_param "match_slice_rows": rows of the FIRST H2D slice of the streamed host-DB path

// DB rows in HOST memory: slice k+1 is copied (copy stream) while slice k is expanded and swept (caller's stream); each slice
// leaves a per-query partial, the slices are folded in ascending row order (first-min-wins is preserved).  Slices grow
// geometrically (x4, capped at 4 Mi rows): only the small first copy is exposed, and a short sweep -- which pays the per-CTA
// prologue over fewer tiles -- happens once; a copy is ~5x faster than the sweep of the same rows at 50 k queries.
static constexpr int STREAM_MAX_SLICE = 4 << 20;
static int match_partial_streamed(clb_matcher* m, const uint64_t* dQ, int nQ, const uint64_t* T_host, int nT, int64_t t_base,
                                  clb_partial* d_partial_out, cudaStream_t s) {
    const int slice_knob = g_match_slice_rows;
    int first = slice_knob < 1024 ? 1024 : slice_knob;
    // a shard of a sharded map (250 k rows per GPU at 8 GPUs) is smaller than the default first slice and its whole copy would sit
    // in front of the sweep: first slice <= a quarter of the rows; shards below 256 Ki rows start with 32 Ki rows -- with 8 ranks
    // pulling from one host a rank sees ~21 GB/s, so every 16 Ki rows in front of the first sweep are 50 us of the step
    // (profiles/r2_e2e_timeline.md) (match_slice_rows >= 2^30: never cut)
    if (first > nT / 4 && slice_knob < (1 << 30)) first = nT / 4 > 65536 ? nT / 4 : (first < 32768 ? first : 32768);
    first = (first + 127) & ~127;
    int nsl = 0, rows_max = 0;
    for (long long start = 0, sz = first; start < nT; start += sz, sz = sz * 4 > STREAM_MAX_SLICE ? STREAM_MAX_SLICE : sz * 4) {
        const int rows = (int)(nT - start < sz ? nT - start : sz);
        if (rows > rows_max) rows_max = rows;
        ++nsl;
    }
    if (!m->stream_ready) {
        CLB_CUDA(cudaStreamCreateWithFlags(&m->copy_stream, cudaStreamNonBlocking));
        for (int b = 0; b < 2; ++b) {
            CLB_CUDA(cudaEventCreateWithFlags(&m->ev_copied[b], cudaEventDisableTiming));
            CLB_CUDA(cudaEventCreateWithFlags(&m->ev_rawfree[b], cudaEventDisableTiming));
        }
        CLB_CUDA(cudaEventCreateWithFlags(&m->ev_entry, cudaEventDisableTiming));
        m->stream_ready = true;
    }
    int rc = ensure(m, nQ, rows_max);
    if (rc) return rc;
    const bool tc = tc_wanted(nQ, first < nT ? first : nT);
    for (int b = 0; b < (nsl > 1 ? 2 : 1); ++b) {
        CLB_CUDA(m->sraw[b].reserve((size_t)rows_max * 64));
        if (tc) CLB_CUDA(m->sx[b].reserve(tc_t_bytes(rows_max)));
    }
    CLB_CUDA(m->sparts.reserve((size_t)nsl * nQ * sizeof(clb_partial)));
    const void* Qx = nullptr;
    if (tc && tc_needs_qx()) {
        CLB_CUDA(m->qx.reserve(tc_q_bytes(nQ)));
        rc = launch_expand(dQ, nQ, tc_q_bytes(nQ), m->qx.p, s);
        if (rc) return rc;
        Qx = m->qx.p;
    }
    // the slice buffers may still be in use by work queued earlier on `s`
    CLB_CUDA(cudaEventRecord(m->ev_entry, s));
    CLB_CUDA(cudaStreamWaitEvent(m->copy_stream, m->ev_entry, 0));
    clb_partial* parts = nsl > 1 ? m->sparts.as<clb_partial>() : d_partial_out;
    long long start = 0, sz = first;
    for (int k = 0; k < nsl; ++k) {
        const int b = k & 1;
        const int rows = (int)(nT - start < sz ? nT - start : sz);
        if (k >= 2) CLB_CUDA(cudaStreamWaitEvent(m->copy_stream, m->ev_rawfree[b], 0));
        tl_mark("slice_copy_begin", m->copy_stream);
        CLB_CUDA(cudaMemcpyAsync(m->sraw[b].p, T_host + (size_t)start * 8, (size_t)rows * 64, cudaMemcpyHostToDevice, m->copy_stream));
        CLB_CUDA(cudaEventRecord(m->ev_copied[b], m->copy_stream));
        tl_mark("slice_copy_end", m->copy_stream);
        // "wait for the copy, expand the slice" is handed to the sweep's launch path, which enqueues it AFTER its job-table
        // upload: parked behind the wait, that small upload became eligible only once the next slice's copy already held the
        // copy engine, and the sweep then sat out the whole copy (event timeline: a 0.85 ms sweep took 1.29 ms next to a
        // 0.44 ms copy; the streamed call cost copies + sweeps)
        const void* Tx = tc ? m->sx[b].p : nullptr;
        m->pending.set = true; m->pending.ev = m->ev_copied[b]; m->pending.raw = m->sraw[b].as<uint64_t>(); m->pending.rows = rows;
        m->pending.x_bytes = tc ? tc_t_bytes(rows) : 0; m->pending.x = tc ? m->sx[b].p : nullptr;
        rc = match_dispatch(m, dQ, nQ, m->sraw[b].as<uint64_t>(), rows, t_base + start, 0, 0.f, nullptr, nullptr, nullptr, parts + (size_t)k * nQ, s, Qx, Tx);
        if (rc) { m->pending.set = false; return rc; }
        if (m->pending.set || m->q_wait_set) { rc = clb_consume_pending(m, s); if (rc) return rc; }      // a path that launched nothing
        CLB_CUDA(cudaEventRecord(m->ev_rawfree[b], s));
        tl_mark("slice_swept", s);
        start += sz;
        sz = sz * 4 > STREAM_MAX_SLICE ? STREAM_MAX_SLICE : sz * 4;
    }
    if (nsl > 1) return launch_fold_partials(parts, nsl, nQ, d_partial_out, s);
    return CLB_OK;
}

extern "C" {

int clb_matcher_create(const clb_matcher_opts* opts, int device, clb_matcher** out) {
    if (!opts || !out || opts->maxkp < 1) { snprintf(g_err, sizeof(g_err), "bad matcher options"); return CLB_EINVAL; }
    CLB_CUDA(cudaSetDevice(device));
    clb_matcher* m = new (std::nothrow) clb_matcher();
    if (!m) return CLB_ENOMEM;
    m->o = *opts; m->device = device; m->n_map = 0; m->mapx_valid = false; m->mapx_ver = 0; m->stream = nullptr;
    m->jobs_host = nullptr; m->jobs_host_cap = 0; m->jobs_ev_set = false; m->stream_ready = false; m->copy_stream = nullptr; m->ev_map_set = false;
    m->pending.set = false; m->q_wait_set = false;
    cudaDeviceGetAttribute(&m->sm_count, cudaDevAttrMultiProcessorCount, device);
    cudaError_t e = cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = m->q.reserve((size_t)opts->maxkp * 64);
    if (e == cudaSuccess) e = m->t.reserve((size_t)opts->maxkp * 64);
    if (e != cudaSuccess) { int rc = set_cuda_error(e, "matcher allocation", __FILE__, __LINE__); clb_matcher_destroy(m); return rc; }
    *out = m;
    return CLB_OK;
}

void clb_matcher_destroy(clb_matcher* m) {
    if (!m) return;
    cudaSetDevice(m->device);
    m->q.release(); m->t.release(); m->map.release(); m->scratch.release(); m->out.release();
    m->qx.release(); m->tx.release(); m->mapx.release();
    m->jobs.release(); m->pair_part.release(); m->pair_out.release();
    if (m->jobs_host) cudaFreeHost(m->jobs_host);
    if (m->jobs_ev_set) cudaEventDestroy(m->jobs_ev);
    if (m->ev_map_set) cudaEventDestroy(m->ev_map);
    if (m->stream_ready) {
        for (int b = 0; b < 2; ++b) { cudaEventDestroy(m->ev_copied[b]); cudaEventDestroy(m->ev_rawfree[b]); }
        cudaEventDestroy(m->ev_entry);
        cudaStreamDestroy(m->copy_stream);
    }
    for (int b = 0; b < 2; ++b) { m->sraw[b].release(); m->sx[b].release(); }
    m->sparts.release(); m->sfold.release(); m->geo.release();
    if (m->stream) cudaStreamDestroy(m->stream);
    delete m;
}

static int check_sets(clb_matcher* m, const void* Q, int nQ, const void* T, int nT) {
    if (!m || nQ < 0 || nT < 0 || (nQ > 0 && !Q) || (nT > 0 && !T)) { snprintf(g_err, sizeof(g_err), "bad descriptor set"); return CLB_EINVAL; }
    return CLB_OK;
}

int clb_match_2nn(clb_matcher* m, const uint64_t* Q, int nQ, const uint64_t* T, int nT, int threshold,
                  int32_t* match_out, int32_t* best_d_out, int32_t* second_d_out) {
    int rc = check_sets(m, Q, nQ, T, nT);
    if (rc) return rc;
    if (nQ == 0) return CLB_OK;
    CLB_CUDA(cudaSetDevice(m->device));
    cudaStream_t s = m->stream;
    CLB_CUDA(m->q.reserve((size_t)nQ * 64));
    CLB_CUDA(cudaMemcpyAsync(m->q.p, Q, (size_t)nQ * 64, cudaMemcpyHostToDevice, s));
    if (nT > g_match_slice_rows && tc_wanted(nQ, g_match_slice_rows)) {
        // large DB: stream it in slices so that the H2D copy hides behind the sweep (results identical: slices fold in row order)
        CLB_CUDA(m->sfold.reserve((size_t)nQ * sizeof(clb_partial)));
        CLB_CUDA(m->out.reserve((size_t)nQ * 12));
        rc = match_partial_streamed(m, m->q.as<uint64_t>(), nQ, T, nT, 0, m->sfold.as<clb_partial>(), s);
        if (rc) return rc;
        int32_t* o = m->out.as<int32_t>();
        rc = launch_merge_partials(m->sfold.as<clb_partial>(), 1, nQ, threshold, o, best_d_out ? o + nQ : nullptr,
                                   second_d_out ? o + 2 * (size_t)nQ : nullptr, s);
        if (rc) return rc;
        if (match_out) CLB_CUDA(cudaMemcpyAsync(match_out, o, (size_t)nQ * 4, cudaMemcpyDeviceToHost, s));
        if (best_d_out) CLB_CUDA(cudaMemcpyAsync(best_d_out, o + nQ, (size_t)nQ * 4, cudaMemcpyDeviceToHost, s));
        if (second_d_out) CLB_CUDA(cudaMemcpyAsync(second_d_out, o + 2 * (size_t)nQ, (size_t)nQ * 4, cudaMemcpyDeviceToHost, s));
        CLB_CUDA(cudaStreamSynchronize(s));
        return CLB_OK;
    }
    CLB_CUDA(m->t.reserve((size_t)(nT > 0 ? nT : 1) * 64));
    if (nT > 0) CLB_CUDA(cudaMemcpyAsync(m->t.p, T, (size_t)nT * 64, cudaMemcpyHostToDevice, s));
    return run_host(m, m->q.as<uint64_t>(), nQ, m->t.as<uint64_t>(), nT, threshold, 0.f, match_out, best_d_out, second_d_out);
}

int clb_match_2nn_ratio(clb_matcher* m, const uint64_t* Q, int nQ, const uint64_t* T, int nT, float ratio,
                        int32_t* match_out, int32_t* best_d_out, int32_t* second_d_out) {
    int rc = check_sets(m, Q, nQ, T, nT);
    if (rc) return rc;
    if (!(ratio > 0.f)) return CLB_EINVAL;
    if (nQ == 0) return CLB_OK;
    CLB_CUDA(cudaSetDevice(m->device));
    cudaStream_t s = m->stream;
    CLB_CUDA(m->q.reserve((size_t)nQ * 64));
    CLB_CUDA(m->t.reserve((size_t)(nT > 0 ? nT : 1) * 64));
    CLB_CUDA(cudaMemcpyAsync(m->q.p, Q, (size_t)nQ * 64, cudaMemcpyHostToDevice, s));
    if (nT > 0) CLB_CUDA(cudaMemcpyAsync(m->t.p, T, (size_t)nT * 64, cudaMemcpyHostToDevice, s));
    return run_host(m, m->q.as<uint64_t>(), nQ, m->t.as<uint64_t>(), nT, 0, ratio, match_out, best_d_out, second_d_out);
}

int clb_match_2nn_dev(clb_matcher* m, const uint64_t* dQ, int nQ, const uint64_t* dT, int nT, int threshold,
                      int32_t* d_match_out, int32_t* d_best_out, int32_t* d_second_out, void* stream) {
    int rc = check_sets(m, dQ, nQ, dT, nT);
    if (rc) return rc;
    if (nQ == 0) return CLB_OK;
    CLB_CUDA(cudaSetDevice(m->device));
    rc = ensure(m, nQ, nT);
    if (rc) return rc;
    cudaStream_t s = stream ? (cudaStream_t)stream : m->stream;
    const void *Qx, *Tx;
    rc = prepare_tc(m, dQ, nQ, dT, nT, false, s, &Qx, &Tx);
    if (rc) return rc;
    return match_dispatch(m, dQ, nQ, dT, nT, 0, threshold, 0.f,
                        d_match_out, d_best_out, d_second_out, nullptr, s, Qx, Tx);
}

int clb_set_map(clb_matcher* m, const uint64_t* M, int nM) {
    if (!m || nM < 0 || (nM > 0 && !M)) return CLB_EINVAL;
    CLB_CUDA(cudaSetDevice(m->device));
    CLB_CUDA(m->map.reserve((size_t)(nM > 0 ? nM : 1) * 64));
    if (nM > 0) CLB_CUDA(cudaMemcpyAsync(m->map.p, M, (size_t)nM * 64, cudaMemcpyHostToDevice, m->stream));
    CLB_CUDA(cudaStreamSynchronize(m->stream));
    m->n_map = nM;
    m->mapx_valid = false;      // re-expanded lazily by the next clb_match_map that takes the tensor-core path
    return CLB_OK;
}

int clb_match_map(clb_matcher* m, const uint64_t* Q, int nQ, int threshold,
                  int32_t* match_out, int32_t* best_d_out, int32_t* second_d_out) {
    if (!m || nQ < 0 || (nQ > 0 && !Q)) return CLB_EINVAL;
    if (nQ == 0) return CLB_OK;
    CLB_CUDA(cudaSetDevice(m->device));
    CLB_CUDA(m->q.reserve((size_t)nQ * 64));
    CLB_CUDA(cudaMemcpyAsync(m->q.p, Q, (size_t)nQ * 64, cudaMemcpyHostToDevice, m->stream));
    return run_host(m, m->q.as<uint64_t>(), nQ, m->map.as<uint64_t>(), m->n_map, threshold < 0 ? m->o.thresh : threshold, 0.f,
                    match_out, best_d_out, second_d_out, true);
}

int clb_set_map_dev(clb_matcher* m, const uint64_t* dM, int nM, void* stream) {
    if (!m || nM < 0 || (nM > 0 && !dM)) return CLB_EINVAL;
    CLB_CUDA(cudaSetDevice(m->device));
    cudaStream_t s = stream ? (cudaStream_t)stream : m->stream;
    CLB_CUDA(m->map.reserve((size_t)(nM > 0 ? nM : 1) * 64));
    if (nM > 0) CLB_CUDA(cudaMemcpyAsync(m->map.p, dM, (size_t)nM * 64, cudaMemcpyDeviceToDevice, s));
    if (!m->ev_map_set) { CLB_CUDA(cudaEventCreateWithFlags(&m->ev_map, cudaEventDisableTiming)); m->ev_map_set = true; }
    CLB_CUDA(cudaEventRecord(m->ev_map, s));      // later map sweeps, on whatever stream, are ordered behind this copy
    m->n_map = nM;
    m->mapx_valid = false;
    return CLB_OK;
}

int clb_match_map_dev(clb_matcher* m, const uint64_t* dQ, int nQ, int threshold,
                      int32_t* d_match_out, int32_t* d_best_out, int32_t* d_second_out, void* stream) {
    if (!m || nQ < 0 || (nQ > 0 && !dQ)) return CLB_EINVAL;
    if (nQ == 0) return CLB_OK;
    CLB_CUDA(cudaSetDevice(m->device));
    int rc = ensure(m, nQ, m->n_map);
    if (rc) return rc;
    cudaStream_t s = stream ? (cudaStream_t)stream : m->stream;
    const void *Qx, *Tx;
    rc = prepare_tc(m, dQ, nQ, m->map.as<uint64_t>(), m->n_map, true, s, &Qx, &Tx);
    if (rc) return rc;
    return match_dispatch(m, dQ, nQ, m->map.as<uint64_t>(), m->n_map, 0, threshold < 0 ? m->o.thresh : threshold, 0.f, d_match_out, d_best_out, d_second_out, nullptr, s, Qx, Tx);
}

int clb_match_map_partial_dev(clb_matcher* m, const uint64_t* dQ, int nQ, int64_t t_base,
                              clb_partial* d_partial_out, void* stream) {
    if (!m || nQ < 0 || (nQ > 0 && (!dQ || !d_partial_out))) return CLB_EINVAL;
    if (nQ == 0) return CLB_OK;
    CLB_CUDA(cudaSetDevice(m->device));
    int rc = ensure(m, nQ, m->n_map);
    if (rc) return rc;
    cudaStream_t s = stream ? (cudaStream_t)stream : m->stream;
    const void *Qx, *Tx;
    rc = prepare_tc(m, dQ, nQ, m->map.as<uint64_t>(), m->n_map, true, s, &Qx, &Tx);
    if (rc) return rc;
    return match_dispatch(m, dQ, nQ, m->map.as<uint64_t>(), m->n_map, t_base, 0, 0.f,
                        nullptr, nullptr, nullptr, d_partial_out, s, Qx, Tx);
}

int clb_match_partial_dev(clb_matcher* m, const uint64_t* dQ, int nQ, const uint64_t* dT, int nT,
                          int64_t t_base, clb_partial* d_partial_out, void* stream) {
    int rc = check_sets(m, dQ, nQ, dT, nT);
    if (rc) return rc;
    if (!d_partial_out && nQ > 0) return CLB_EINVAL;
    if (nQ == 0) return CLB_OK;
    CLB_CUDA(cudaSetDevice(m->device));
    rc = ensure(m, nQ, nT);
    if (rc) return rc;
    cudaStream_t s = stream ? (cudaStream_t)stream : m->stream;
    const void *Qx, *Tx;
    rc = prepare_tc(m, dQ, nQ, dT, nT, false, s, &Qx, &Tx);
    if (rc) return rc;
    return match_dispatch(m, dQ, nQ, dT, nT, t_base, 0, 0.f,
                        nullptr, nullptr, nullptr, d_partial_out, s, Qx, Tx);
}

int clb_match_partial_hostdb(clb_matcher* m, const uint64_t* dQ, int nQ, const uint64_t* T_host, int nT,
                             int64_t t_base, clb_partial* d_partial_out, void* stream) {
    int rc = check_sets(m, dQ, nQ, T_host, nT);
    if (rc) return rc;
    if (!d_partial_out && nQ > 0) return CLB_EINVAL;
    if (nQ == 0) return CLB_OK;
    CLB_CUDA(cudaSetDevice(m->device));
    cudaStream_t s = stream ? (cudaStream_t)stream : m->stream;
    if (nT == 0) {
        rc = ensure(m, nQ, 0);
        if (rc) return rc;
        return match_dispatch(m, dQ, nQ, nullptr, 0, t_base, 0, 0.f,
                              nullptr, nullptr, nullptr, d_partial_out, s, nullptr, nullptr);
    }
    return match_partial_streamed(m, dQ, nQ, T_host, nT, t_base, d_partial_out, s);
}

int clb_merge_partials_dev(clb_matcher* m, const clb_partial* d_parts, int nparts, int nQ, int threshold,
                           int32_t* d_match_out, int32_t* d_best_out, int32_t* d_second_out, void* stream) {
    if (!m || nparts < 0 || nQ < 0 || (nQ > 0 && nparts > 0 && !d_parts)) return CLB_EINVAL;
    CLB_CUDA(cudaSetDevice(m->device));
    return launch_merge_partials(d_parts, nparts, nQ, threshold, d_match_out, d_best_out, d_second_out,
                                 stream ? (cudaStream_t)stream : m->stream);
}

}  // extern "C"
