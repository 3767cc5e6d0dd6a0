// api_fmap.cu -- device-resident FeatureMap, batched pair matching and the map-database builder (C ABI, include/coloc_b200.h).
//
// The reference keeps every frame's descriptors in host `Regions` (coloc::FeatureMap, idx -> unique_ptr<AKAZE_Binary_Regions>,
// GPUDetector.hpp:167-182) and re-uploads both sets for every pair (GPUMatcher.hpp:165-172 -> :180-226) and the gathered landmark
// rows for every map update (colocData.hpp:89-121 -> GPUMatcher.hpp:110-117).  Here a clb_fmap holds each set once on the device,
// packed rows plus the tensor-core operand image, and
//   * clb_match_pairs[_dev]   runs a whole pair list (GPUMatcher::computeMatches(FeatureMap&, PairWiseMatches&), :143-155)
//                             as ONE sweep launch + ONE merge launch,
//   * clb_set_map_from_fmap   is setupMapDatabase + setMapData as a device gather.
#include "matcher_priv.cuh"
#include <map>
#include <vector>
#include <algorithm>
#include <new>
#include <string.h>

using namespace clb;

namespace {

struct FSet {
    DevBuf desc, kps, tx;
    int n = 0;
    bool has_kps = false;
    ~FSet() { desc.release(); kps.release(); tx.release(); }
};

}  // namespace

struct clb_fmap {
    int device;
    cudaStream_t stream;
    std::map<uint32_t, FSet*> sets;
    DevBuf gather_tab;           // clb_set_map_from_fmap: (slot, feat) list + slot base pointers
    void* stage;                 // pinned staging for that table
    size_t stage_cap;
};

static FSet* find_set(const clb_fmap* f, uint32_t idx) {
    auto it = f->sets.find(idx);
    return it == f->sets.end() ? nullptr : it->second;
}

static int pinned_reserve(void** p, size_t* cap, size_t bytes) {
    if (bytes <= *cap) return CLB_OK;
    if (*p) cudaFreeHost(*p);
    *p = nullptr; *cap = 0;
    const size_t want = bytes + bytes / 2 + 4096;
    CLB_CUDA(cudaMallocHost(p, want));
    *cap = want;
    return CLB_OK;
}

// (re)place a set: rows come from host or device memory; the operand image is built right away
static int put_set(clb_fmap* f, uint32_t idx, const clb_keypoint* kps, const uint64_t* desc, int n, cudaMemcpyKind kind, cudaStream_t s) {
    if (!f || n < 0 || (n > 0 && !desc)) { snprintf(g_err, sizeof(g_err), "bad feature set"); return CLB_EINVAL; }
    CLB_CUDA(cudaSetDevice(f->device));
    FSet* set = find_set(f, idx);
    if (!set) {
        set = new (std::nothrow) FSet();
        if (!set) return CLB_ENOMEM;
        f->sets[idx] = set;
    }
    set->n = 0; set->has_kps = false;
    CLB_CUDA(set->desc.reserve((size_t)(n > 0 ? n : 1) * 64));
    CLB_CUDA(set->tx.reserve(tc4_t_bytes(n) + 16));
    if (n > 0) {
        CLB_CUDA(cudaMemcpyAsync(set->desc.p, desc, (size_t)n * 64, kind, s));
        int rc = launch_expand4(set->desc.as<uint64_t>(), n, set->tx.p, s);
        if (rc) return rc;
        if (kps) {
            CLB_CUDA(set->kps.reserve((size_t)n * sizeof(clb_keypoint)));
            CLB_CUDA(cudaMemcpyAsync(set->kps.p, kps, (size_t)n * sizeof(clb_keypoint), kind, s));
            set->has_kps = true;
        }
    }
    set->n = n;
    return CLB_OK;
}

extern "C" {

int clb_fmap_create(int device, clb_fmap** out) {
    if (!out) return CLB_EINVAL;
    CLB_CUDA(cudaSetDevice(device));
    clb_fmap* f = new (std::nothrow) clb_fmap();
    if (!f) return CLB_ENOMEM;
    f->device = device; f->stream = nullptr; f->stage = nullptr; f->stage_cap = 0;
    cudaError_t e = cudaStreamCreateWithFlags(&f->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) { delete f; return set_cuda_error(e, "fmap stream", __FILE__, __LINE__); }
    *out = f;
    return CLB_OK;
}

void clb_fmap_destroy(clb_fmap* f) {
    if (!f) return;
    cudaSetDevice(f->device);
    for (auto& kv : f->sets) delete kv.second;
    f->gather_tab.release();
    if (f->stage) cudaFreeHost(f->stage);
    if (f->stream) cudaStreamDestroy(f->stream);
    delete f;
}

int clb_fmap_put(clb_fmap* f, uint32_t idx, const clb_keypoint* kps, const uint64_t* desc, int n) {
    if (!f) return CLB_EINVAL;
    int rc = put_set(f, idx, kps, desc, n, cudaMemcpyHostToDevice, f->stream);
    if (rc) return rc;
    CLB_CUDA(cudaStreamSynchronize(f->stream));
    return CLB_OK;
}

int clb_fmap_put_dev(clb_fmap* f, uint32_t idx, const clb_keypoint* d_kps, const uint64_t* d_desc, int n, void* stream) {
    if (!f) return CLB_EINVAL;
    return put_set(f, idx, d_kps, d_desc, n, cudaMemcpyDeviceToDevice, stream ? (cudaStream_t)stream : f->stream);
}

int clb_fmap_count(const clb_fmap* f, uint32_t idx, int* n_out) {
    if (!f || !n_out) return CLB_EINVAL;
    const FSet* set = find_set(f, idx);
    if (!set) { snprintf(g_err, sizeof(g_err), "feature set %u is not in the map", idx); return CLB_ENOTFOUND; }
    *n_out = set->n;
    return CLB_OK;
}

int clb_fmap_size(const clb_fmap* f) { return f ? (int)f->sets.size() : 0; }

int clb_fmap_erase(clb_fmap* f, uint32_t idx) {
    if (!f) return CLB_EINVAL;
    auto it = f->sets.find(idx);
    if (it == f->sets.end()) { snprintf(g_err, sizeof(g_err), "feature set %u is not in the map", idx); return CLB_ENOTFOUND; }
    CLB_CUDA(cudaSetDevice(f->device));
    delete it->second;
    f->sets.erase(it);
    return CLB_OK;
}

int clb_fmap_get(clb_fmap* f, uint32_t idx, clb_keypoint* kps_out, uint64_t* desc_out, int cap, int* n_out) {
    if (!f || !n_out) return CLB_EINVAL;
    const FSet* set = find_set(f, idx);
    if (!set) { snprintf(g_err, sizeof(g_err), "feature set %u is not in the map", idx); return CLB_ENOTFOUND; }
    *n_out = set->n;
    if (set->n > cap) return CLB_ECAPACITY;
    CLB_CUDA(cudaSetDevice(f->device));
    if (kps_out && !set->has_kps && set->n > 0) { snprintf(g_err, sizeof(g_err), "feature set %u holds no keypoints", idx); return CLB_EINVAL; }
    if (desc_out && set->n > 0) CLB_CUDA(cudaMemcpyAsync(desc_out, set->desc.p, (size_t)set->n * 64, cudaMemcpyDeviceToHost, f->stream));
    if (kps_out && set->n > 0) CLB_CUDA(cudaMemcpyAsync(kps_out, set->kps.p, (size_t)set->n * sizeof(clb_keypoint), cudaMemcpyDeviceToHost, f->stream));
    CLB_CUDA(cudaStreamSynchronize(f->stream));
    return CLB_OK;
}

int clb_fmap_desc_dev(const clb_fmap* f, uint32_t idx, const uint64_t** d_desc_out, const clb_keypoint** d_kps_out, int* n_out) {
    if (!f) return CLB_EINVAL;
    const FSet* set = find_set(f, idx);
    if (!set) { snprintf(g_err, sizeof(g_err), "feature set %u is not in the map", idx); return CLB_ENOTFOUND; }
    if (d_desc_out) *d_desc_out = set->desc.as<uint64_t>();
    if (d_kps_out) *d_kps_out = set->has_kps ? set->kps.as<clb_keypoint>() : nullptr;
    if (n_out) *n_out = set->n;
    return CLB_OK;
}

}  // extern "C"

// ---- batched pair matching --------------------------------------------------------------------
namespace {

// gather of landmark rows: 4 threads per 64-byte row
__global__ void __launch_bounds__(256)
gather_rows_kernel(const uint4* const* __restrict__ bases, const int2* __restrict__ list, int n, uint4* __restrict__ out) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= 4ll * n) return;
    const int2 e = list[t >> 2];
    out[t] = __ldg(bases[e.x] + 4ll * e.y + (t & 3));
}

}  // namespace

namespace clb { std::atomic<int> g_match_balance{-1}, g_last_match_balanced{0}; }
// g_match_balance: clb_set_param "match_balance": -1 choose by the time model, 0 uniform chunks, 1 balanced jobs

// One sweep launch + one merge launch over a list of (query rows, expanded training set) views.
// Two job layouts, chosen by the CTA time model (tc_cta_cost_us):
//   uniform   every view cut into chunks of one common length (tc_plan_pairs): CTA = (view, 256-query block, chunk)
//   balanced  the (view, query block) strips laid end to end and cut into one contiguous range per SM (stream-K style): a strip
//             that straddles a range border becomes two jobs, so an SM runs ~1.5 jobs of the same total length instead of
//             whole waves of equal CTAs -- what mid-sized problems (one 20 k x 20 k frame pair) need to fill 148 SMs.
// Jobs are sorted longest first (blockIdx.x order = LPT packing).
int run_views(clb_matcher* m, const clb::PairView* v, int nv, int threshold, float ratio,
              int32_t* d_match, int32_t* d_best, int32_t* d_second, clb_partial* d_part_out, long long t_base, cudaStream_t s) {
    const int sm = m->sm_count > 0 ? m->sm_count : 148;
    std::vector<int> nQ(nv), nT(nv);
    long long units = 0, strips = 0;
    int max_nQ = 0;
    for (int p = 0; p < nv; ++p) {
        // the sweep reads query rows and operand images as 128-bit words: a misaligned pointer is an argument error, not a fault
        if (v[p].nQ > 0 && v[p].nT > 0 && ((reinterpret_cast<uintptr_t>(v[p].Q) | reinterpret_cast<uintptr_t>(v[p].Tx)) & 15)) {
            snprintf(g_err, sizeof(g_err), "descriptor pointers must be 16-byte aligned"); return CLB_EINVAL;
        }
        nQ[p] = v[p].nQ > 0 ? v[p].nQ : 0; nT[p] = v[p].nT;
        if (nQ[p] > max_nQ) max_nQ = nQ[p];
        if (nQ[p] <= 0 || nT[p] <= 0) continue;
        const long long qb = (nQ[p] + 255) / 256, tiles = (nT[p] + 127) / 128;
        strips += qb; units += qb * tiles;
    }
    double t_uniform = 0.0;
    const int tpc = tc_plan_pairs(nQ.data(), nT.data(), nv, sm, &t_uniform);
    // balanced layout: one range per SM (more when a range would exceed the 2^22 rows a packed key can index)
    long long parts = sm;
    const double t_balanced = tc_balanced_cost_us(units, strips, sm, &parts);
    const double quota = units > 0 ? (double)units / (double)parts : 1.0;
    bool balanced = units >= 4 * parts && t_balanced < 0.97 * t_uniform;
    if (g_match_balance == 0) balanced = false;
    if (g_match_balance == 1 && units > 0) balanced = true;
    constexpr int MINP = 4;             // no piece shorter than this many tiles (cuts snap to the strip ends)

    // pass 1: count jobs / partial rows / piece-table entries; pass 2 fills the tables
    struct Cut { int tile0, ntiles; };
    auto cut_strip = [&](long long u0, int tiles, std::vector<Cut>& out) {
        out.clear();
        int t0 = 0;
        if (balanced) {
            // cut points = multiples of quota inside (u0, u0 + tiles)
            long long k = (long long)((double)u0 / quota) + 1;
            for (;; ++k) {
                const long long cut = (long long)((double)k * quota + 0.5) - u0;
                if (cut >= tiles) break;
                if (cut - t0 < MINP || tiles - cut < MINP) continue;
                out.push_back(Cut{t0, (int)(cut - t0)});
                t0 = (int)cut;
            }
            out.push_back(Cut{t0, tiles - t0});
        } else {
            for (; t0 < tiles; t0 += tpc) out.push_back(Cut{t0, tiles - t0 < tpc ? tiles - t0 : tpc});
        }
    };
    std::vector<Cut> cuts;
    size_t njobs = 0, part_rows = 0, tab_ints = 0;
    std::vector<int> maxp(nv, 0);
    {
        long long u = 0;
        for (int p = 0; p < nv; ++p) {
            if (nQ[p] <= 0 || nT[p] <= 0) continue;
            const int qb = (nQ[p] + 255) / 256, tiles = (nT[p] + 127) / 128;
            for (int b = 0; b < qb; ++b, u += tiles) {
                cut_strip(u, tiles, cuts);
                njobs += cuts.size();
                if ((int)cuts.size() > maxp[p]) maxp[p] = (int)cuts.size();
            }
            part_rows += (size_t)maxp[p] * nQ[p];
            if (balanced) tab_ints += (size_t)qb * (1 + maxp[p]);
        }
    }
    if (njobs > 0x7FFFFFFFull) { snprintf(g_err, sizeof(g_err), "pair list too large for one launch"); return CLB_EINVAL; }
    const size_t jobs_bytes = njobs * sizeof(TcJob), merge_bytes = (size_t)nv * sizeof(MergeJob);
    const size_t tab_bytes = jobs_bytes + merge_bytes + tab_ints * sizeof(int);
    // the previous call's table copy must have left the pinned staging area before it is rewritten
    if (m->jobs_ev_set) CLB_CUDA(cudaEventSynchronize(m->jobs_ev));
    int rc = pinned_reserve(&m->jobs_host, &m->jobs_host_cap, tab_bytes);
    if (rc) return rc;
    CLB_CUDA(m->jobs.reserve(tab_bytes));
    CLB_CUDA(m->pair_part.reserve((part_rows > 0 ? part_rows : 1) * sizeof(uint2)));
    TcJob* tj = reinterpret_cast<TcJob*>(m->jobs_host);
    MergeJob* mj = reinterpret_cast<MergeJob*>(tj + njobs);
    int* tab = reinterpret_cast<int*>(mj + nv);
    const int* d_tab = reinterpret_cast<const int*>(static_cast<char*>(m->jobs.p) + jobs_bytes + merge_bytes);
    uint2* part = m->pair_part.as<uint2>();
    size_t j = 0, part_off = 0, tab_off = 0;
    long long out_off = 0, u = 0;
    for (int p = 0; p < nv; ++p) {
        const int tiles = nT[p] > 0 ? (nT[p] + 127) / 128 : 0;
        const int qb = nQ[p] > 0 && tiles > 0 ? (nQ[p] + 255) / 256 : 0;
        MergeJob& mg = mj[p];
        mg.partial = part + part_off;
        mg.match = d_match ? d_match + out_off : nullptr;
        mg.best = d_best ? d_best + out_off : nullptr;
        mg.second = d_second ? d_second + out_off : nullptr;
        mg.part_out = d_part_out ? d_part_out + out_off : nullptr;
        mg.t_base = t_base;
        mg.nQ = nQ[p]; mg.nT = nT[p]; mg.nchunks = qb > 0 ? maxp[p] : 0; mg.rows_per_chunk = tpc * 128; mg.maxp = maxp[p]; mg._pad = 0;
        mg.pieces = nullptr; mg.base_rows = nullptr;
        int* np_tab = nullptr; int* base_tab = nullptr;
        if (balanced && qb > 0) {
            np_tab = tab + tab_off; base_tab = np_tab + qb;
            mg.pieces = d_tab + tab_off; mg.base_rows = mg.pieces + qb;
            tab_off += (size_t)qb * (1 + maxp[p]);
        }
        for (int b = 0; b < qb; ++b, u += tiles) {
            cut_strip(u, tiles, cuts);
            if (np_tab) np_tab[b] = (int)cuts.size();
            for (size_t c = 0; c < cuts.size(); ++c) {
                if (base_tab) base_tab[(size_t)b * maxp[p] + c] = cuts[c].tile0 * 128;
                TcJob& job = tj[j++];
                job.Q = reinterpret_cast<const uint32_t*>(v[p].Q);
                job.Tx = v[p].Tx;
                job.partial_chunk = part + part_off + c * (size_t)nQ[p];
                job.nQ = nQ[p]; job.qblock = b; job.tile0 = cuts[c].tile0; job.ntiles = cuts[c].ntiles;
                job.chunk_rows = std::min(cuts[c].ntiles * 128, nT[p] - cuts[c].tile0 * 128);
                job._pad = 0;
            }
        }
        part_off += (size_t)maxp[p] * nQ[p];
        out_off += nQ[p];
    }
    // longest CTAs first: the hardware hands out blockIdx.x in order, so this is longest-processing-time-first packing
    std::stable_sort(tj, tj + njobs, [](const TcJob& a, const TcJob& b) { return a.ntiles > b.ntiles; });
    if (tab_bytes > 0) {
        CLB_CUDA(cudaMemcpyAsync(m->jobs.p, m->jobs_host, tab_bytes, cudaMemcpyHostToDevice, s));
        if (!m->jobs_ev_set) { CLB_CUDA(cudaEventCreateWithFlags(&m->jobs_ev, cudaEventDisableTiming)); m->jobs_ev_set = true; }
        CLB_CUDA(cudaEventRecord(m->jobs_ev, s));
    }
    rc = clb_consume_pending(m, s);        // a streamed slice: its copy is waited for (and expanded) here, behind the table upload
    if (rc) return rc;
    rc = launch_match_tc_jobs(m->jobs.as<TcJob>(), (int)njobs, s);
    if (rc) return rc;
    g_last_match_path = 4;
    g_last_match_balanced = balanced ? 1 : 0;
    return launch_merge_jobs(reinterpret_cast<const MergeJob*>(m->jobs.as<TcJob>() + njobs), nv, max_nQ, threshold, ratio, s);
}

static int run_pairs(clb_matcher* m, clb_fmap* f, const clb_pair* pairs, int npairs, int threshold, float ratio,
                     int32_t* d_match, int32_t* d_best, int32_t* d_second, cudaStream_t s) {
    std::vector<PairView> views((size_t)npairs);
    for (int p = 0; p < npairs; ++p) {
        const FSet* q = find_set(f, pairs[p].query_idx);
        const FSet* t = find_set(f, pairs[p].train_idx);
        const int hi = pairs[p].q_hi < 0 ? q->n : pairs[p].q_hi;
        views[p].Q = q->desc.as<uint64_t>() + 8ll * pairs[p].q_lo;
        views[p].nQ = hi - pairs[p].q_lo;
        views[p].Tx = t->tx.as<uint8_t>();
        views[p].nT = t->n;
    }
    return run_views(m, views.data(), npairs, threshold, ratio, d_match, d_best, d_second, nullptr, 0, s);
}

static int check_pairs(clb_matcher* m, clb_fmap* f, const clb_pair* pairs, int npairs, long long* total_rows) {
    if (!m || !f || npairs < 0 || (npairs > 0 && !pairs)) { snprintf(g_err, sizeof(g_err), "bad pair list"); return CLB_EINVAL; }
    if (m->device != f->device) { snprintf(g_err, sizeof(g_err), "matcher and feature map live on different devices"); return CLB_EINVAL; }
    long long tot = 0;
    for (int p = 0; p < npairs; ++p) {
        const FSet* q = find_set(f, pairs[p].query_idx);
        const FSet* t = find_set(f, pairs[p].train_idx);
        if (!q || !t) { snprintf(g_err, sizeof(g_err), "pair %d: feature set %u is not in the map", p, !q ? pairs[p].query_idx : pairs[p].train_idx); return CLB_ENOTFOUND; }
        const int hi = pairs[p].q_hi < 0 ? q->n : pairs[p].q_hi;
        if (pairs[p].q_lo < 0 || hi < pairs[p].q_lo || hi > q->n) { snprintf(g_err, sizeof(g_err), "pair %d: query rows [%d, %d) outside the set's %d rows", p, pairs[p].q_lo, hi, q->n); return CLB_EINVAL; }
        tot += hi - pairs[p].q_lo;
    }
    *total_rows = tot;
    return CLB_OK;
}

extern "C" {

int clb_match_pairs_rows(clb_matcher* m, clb_fmap* f, const clb_pair* pairs, int npairs, long long* rows_out) {
    long long tot = 0;
    int rc = check_pairs(m, f, pairs, npairs, &tot);
    if (rc) return rc;
    if (rows_out) *rows_out = tot;
    return CLB_OK;
}

int clb_match_pairs_dev(clb_matcher* m, clb_fmap* f, const clb_pair* pairs, int npairs, int threshold, float ratio,
                        int32_t* d_match_out, int32_t* d_best_out, int32_t* d_second_out, void* stream) {
    long long tot = 0;
    int rc = check_pairs(m, f, pairs, npairs, &tot);
    if (rc) return rc;
    if (tot == 0) return CLB_OK;
    if (!d_match_out) return CLB_EINVAL;
    CLB_CUDA(cudaSetDevice(m->device));
    return run_pairs(m, f, pairs, npairs, threshold < 0 ? m->o.thresh : threshold, ratio, d_match_out, d_best_out, d_second_out,
                     stream ? (cudaStream_t)stream : m->stream);
}

int clb_match_pairs(clb_matcher* m, clb_fmap* f, const clb_pair* pairs, int npairs, int threshold, float ratio,
                    int32_t* match_out, int32_t* best_d_out, int32_t* second_d_out) {
    long long tot = 0;
    int rc = check_pairs(m, f, pairs, npairs, &tot);
    if (rc) return rc;
    if (tot == 0) return CLB_OK;
    if (!match_out) return CLB_EINVAL;
    CLB_CUDA(cudaSetDevice(m->device));
    CLB_CUDA(m->pair_out.reserve((size_t)tot * 12));
    int32_t* o = m->pair_out.as<int32_t>();
    cudaStream_t s = m->stream;
    rc = run_pairs(m, f, pairs, npairs, threshold < 0 ? m->o.thresh : threshold, ratio, o, best_d_out ? o + tot : nullptr,
                   second_d_out ? o + 2 * tot : nullptr, s);
    if (rc) return rc;
    CLB_CUDA(cudaMemcpyAsync(match_out, o, (size_t)tot * 4, cudaMemcpyDeviceToHost, s));
    if (best_d_out) CLB_CUDA(cudaMemcpyAsync(best_d_out, o + tot, (size_t)tot * 4, cudaMemcpyDeviceToHost, s));
    if (second_d_out) CLB_CUDA(cudaMemcpyAsync(second_d_out, o + 2 * tot, (size_t)tot * 4, cudaMemcpyDeviceToHost, s));
    CLB_CUDA(cudaStreamSynchronize(s));
    return CLB_OK;
}

// ---- matching::Filter (matching_filters.hpp:118-213): symmetric (cross-check) and / or distance-ratio filtering of 2-NN lists ----
}  // extern "C"
namespace {
__global__ void __launch_bounds__(256)
filter_flags_kernel(const int32_t* __restrict__ m01, const int32_t* __restrict__ best01, const int32_t* __restrict__ second01,
                    const int32_t* __restrict__ m10, int nA, int nB, float ratio, uint8_t* __restrict__ flags) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nA) return;
    const int j = m01[i];
    // SymmetricMatches (:76-94): index == vec_reversematches[vec_matches[i * NN] * NN]
    const bool sym = j >= 0 && j < nB && m10[j] == i;
    // NNdistanceRatio (:39-60): *iter < fratio * *(iter + 1), on the float distances
    const bool rat = (float)best01[i] < ratio * (float)second01[i];
    flags[i] = (uint8_t)((sym ? 1 : 0) | (rat ? 2 : 0));
}
}  // namespace
extern "C" {

int clb_match_filtered(clb_matcher* m, const uint64_t* A, int nA, const uint64_t* B, int nB, int filter, float ratio,
                       uint32_t* pairs_out, int cap, int* n_out) {
    if (!m || !A || !B || !n_out || nA < 1 || nB < 1 || filter < 1 || filter > 3 || cap < 0 || (cap > 0 && !pairs_out)) {
        snprintf(g_err, sizeof(g_err), "bad arguments (filter 1 = symmetric, 2 = distance ratio, 3 = both; both sets non-empty)"); return CLB_EINVAL;
    }
    if ((filter & 2) && (nA < 2 || nB < 2)) { snprintf(g_err, sizeof(g_err), "the distance-ratio filter needs two neighbours: at least two rows per set"); return CLB_EINVAL; }
    CLB_CUDA(cudaSetDevice(m->device));
    cudaStream_t s = m->stream;
    const size_t tot = (size_t)nA + nB;
    CLB_CUDA(m->q.reserve((size_t)nA * 64));
    CLB_CUDA(m->t.reserve((size_t)nB * 64));
    CLB_CUDA(m->qx.reserve(tc4_t_bytes(nA) + 16));
    CLB_CUDA(m->tx.reserve(tc4_t_bytes(nB) + 16));
    CLB_CUDA(m->pair_out.reserve(tot * 12 + (size_t)nA + 256));
    CLB_CUDA(cudaMemcpyAsync(m->q.p, A, (size_t)nA * 64, cudaMemcpyHostToDevice, s));
    CLB_CUDA(cudaMemcpyAsync(m->t.p, B, (size_t)nB * 64, cudaMemcpyHostToDevice, s));
    int rc = launch_expand4(m->q.as<uint64_t>(), nA, m->qx.p, s);
    if (rc) return rc;
    rc = launch_expand4(m->t.as<uint64_t>(), nB, m->tx.p, s);
    if (rc) return rc;
    // both directions in ONE sweep launch: (A rows -> B image) and (B rows -> A image); threshold -1 keeps every best index
    PairView v[2];
    v[0].Q = m->q.as<uint64_t>(); v[0].nQ = nA; v[0].Tx = m->tx.as<uint8_t>(); v[0].nT = nB;
    v[1].Q = m->t.as<uint64_t>(); v[1].nQ = nB; v[1].Tx = m->qx.as<uint8_t>(); v[1].nT = nA;
    int32_t* o = m->pair_out.as<int32_t>();
    rc = run_views(m, v, 2, -1, 0.f, o, o + tot, o + 2 * tot, nullptr, 0, s);
    if (rc) return rc;
    uint8_t* d_flags = reinterpret_cast<uint8_t*>(o + 3 * tot);
    filter_flags_kernel<<<(nA + 255) / 256, 256, 0, s>>>(o, o + tot, o + 2 * tot, o + nA, nA, nB, ratio, d_flags);
    CLB_LAUNCHED();
    std::vector<int32_t> m01((size_t)nA);
    std::vector<uint8_t> fl((size_t)nA);
    CLB_CUDA(cudaMemcpyAsync(m01.data(), o, (size_t)nA * 4, cudaMemcpyDeviceToHost, s));
    CLB_CUDA(cudaMemcpyAsync(fl.data(), d_flags, (size_t)nA, cudaMemcpyDeviceToHost, s));
    CLB_CUDA(cudaStreamSynchronize(s));
    // index lists as Filter() builds them; its output loops stop one short of each list's end (`i < size() - 1`,
    // matching_filters.hpp:175,184,199): the LAST surviving index is dropped, and so it is here
    const uint8_t need = (uint8_t)filter;
    std::vector<int> idx;
    for (int i = 0; i < nA; ++i) if ((fl[(size_t)i] & need) == need) idx.push_back(i);
    const int n = idx.empty() ? 0 : (int)idx.size() - 1;
    *n_out = n;
    if (n > cap) return CLB_ECAPACITY;
    for (int k = 0; k < n; ++k) { pairs_out[2 * k] = (uint32_t)idx[(size_t)k]; pairs_out[2 * k + 1] = (uint32_t)m01[(size_t)idx[(size_t)k]]; }
    return CLB_OK;      // ascending in i_ and unique by construction = the state after Filter()'s sort + unique
}

// ---- map database: setupMapDatabase (colocData.hpp:89-121) + setMapData (GPUMatcher.hpp:110-117) on the device ----
int clb_set_map_from_fmap(clb_matcher* m, clb_fmap* f, const clb_obs* obs, int n, int* n_map_out, void* stream) {
    if (!m || !f || n < 0 || (n > 0 && !obs)) { snprintf(g_err, sizeof(g_err), "bad observation list"); return CLB_EINVAL; }
    if (m->device != f->device) { snprintf(g_err, sizeof(g_err), "matcher and feature map live on different devices"); return CLB_EINVAL; }
    CLB_CUDA(cudaSetDevice(m->device));
    cudaStream_t s = stream ? (cudaStream_t)stream : m->stream;
    std::map<uint32_t, int> slot_of;
    std::vector<const uint64_t*> bases;
    int kept = 0;
    for (int i = 0; i < n; ++i) {
        if (obs[i].feat == 0xFFFFFFFFu) continue;                        // UndefinedIndexT: landmark without a feature, skipped
        const FSet* set = find_set(f, obs[i].idx);
        if (!set) { snprintf(g_err, sizeof(g_err), "observation %d: feature set %u is not in the map", i, obs[i].idx); return CLB_ENOTFOUND; }
        if (obs[i].feat >= (uint32_t)set->n) { snprintf(g_err, sizeof(g_err), "observation %d: feature %u outside set %u (%d rows)", i, obs[i].feat, obs[i].idx, set->n); return CLB_EINVAL; }
        if (slot_of.emplace(obs[i].idx, (int)bases.size()).second) bases.push_back(set->desc.as<uint64_t>());
        ++kept;
    }
    const size_t base_bytes = (bases.size() * sizeof(void*) + 15) & ~(size_t)15;
    const size_t bytes = base_bytes + (size_t)kept * sizeof(int2);
    int rc = pinned_reserve(&f->stage, &f->stage_cap, bytes);
    if (rc) return rc;
    if (!bases.empty()) memcpy(f->stage, bases.data(), bases.size() * sizeof(void*));
    int2* list = reinterpret_cast<int2*>(static_cast<char*>(f->stage) + base_bytes);
    for (int i = 0, k = 0; i < n; ++i) {
        if (obs[i].feat == 0xFFFFFFFFu) continue;
        list[k++] = make_int2(slot_of[obs[i].idx], (int)obs[i].feat);
    }
    CLB_CUDA(f->gather_tab.reserve(bytes > 0 ? bytes : 16));
    CLB_CUDA(m->map.reserve((size_t)(kept > 0 ? kept : 1) * 64));
    if (kept > 0) {
        CLB_CUDA(cudaMemcpyAsync(f->gather_tab.p, f->stage, bytes, cudaMemcpyHostToDevice, s));
        const long long threads = 4ll * kept;
        gather_rows_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, s>>>(
            reinterpret_cast<const uint4* const*>(f->gather_tab.p),
            reinterpret_cast<const int2*>(static_cast<char*>(f->gather_tab.p) + base_bytes), kept, m->map.as<uint4>());
        CLB_LAUNCHED();
    }
    m->n_map = kept;
    m->mapx_valid = false;
    if (n_map_out) *n_map_out = kept;
    CLB_CUDA(cudaStreamSynchronize(s));       // the pinned table may be rewritten by the next call
    return CLB_OK;
}

}  // extern "C"
