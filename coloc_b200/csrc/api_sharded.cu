// api_sharded.cu -- DB-row-sharded map matching behind the C ABI (include/coloc_b200.h, clb_sharded_*).
//
// The reference matches a frame against ONE resident map on one GPU: GPUMatcher::setMapData (GPUMatcher.hpp:110-117) uploads
// it, matchSceneWithMap / matchFeaturesWithMap (:174-178, :252-271) sweep it.  Here the map rows are cut into contiguous
// shards, one per GPU; every GPU sweeps its shard for all queries (clb_match_map_partial_dev -> per-query
// (best_d, best_i_global, second_d)), the 16-byte-per-query partials are exchanged, and merge_partials_kernel folds them in
// ascending shard order -- which preserves the reference's first-min-wins (CUDAK2NN.cu:67-72) -- and applies the margin test.
//
// Two ways to own the GPUs, same data path:
//   device list   ONE process drives N devices (clb_sharded_create): a C++ host that binds GPUMatcher can shard a 2 M-row map
//                 without any launcher.  Partials travel to the first device by peer copies (NVLink DMA, cudaMemcpyPeerAsync).
//   rank          one process per GPU (clb_sharded_create_rank, what torchrun / MPI launch): the exchange is ONE ncclAllGather
//                 on the caller's stream.  NCCL is bound at run time (dlopen of libnccl.so.2 -- inside a torch process that is
//                 the copy torch already loaded), so the library itself has no link-time NCCL dependency.
#include "matcher_priv.cuh"
#include <dlfcn.h>
#include <string.h>
#include <new>
#include <vector>

using namespace clb;

namespace {

// ---- NCCL, bound at run time.  Only ABI-stable pieces of nccl.h are restated: the 128-byte unique id, opaque comm,
// ncclChar = 0, ncclSuccess = 0.
struct NcclId { char internal[128]; };
typedef void* NcclComm;
struct NcclApi {
    void* h = nullptr;
    int (*GetUniqueId)(NcclId*) = nullptr;
    int (*CommInitRank)(NcclComm*, int, NcclId, int) = nullptr;
    int (*CommDestroy)(NcclComm) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, NcclComm, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int*) = nullptr;
};
NcclApi g_nccl;

int nccl_load() {
    if (g_nccl.h) return CLB_OK;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) { snprintf(g_err, sizeof(g_err), "libnccl.so.2 not found: %s", dlerror()); return CLB_EUNSUPPORTED; }
    NcclApi a; a.h = h;
    *(void**)&a.GetUniqueId = dlsym(h, "ncclGetUniqueId");
    *(void**)&a.CommInitRank = dlsym(h, "ncclCommInitRank");
    *(void**)&a.CommDestroy = dlsym(h, "ncclCommDestroy");
    *(void**)&a.AllGather = dlsym(h, "ncclAllGather");
    *(void**)&a.GetErrorString = dlsym(h, "ncclGetErrorString");
    *(void**)&a.GetVersion = dlsym(h, "ncclGetVersion");
    if (!a.GetUniqueId || !a.CommInitRank || !a.CommDestroy || !a.AllGather || !a.GetErrorString) {
        snprintf(g_err, sizeof(g_err), "libnccl.so.2 lacks a required symbol"); return CLB_EUNSUPPORTED;
    }
    g_nccl = a;
    return CLB_OK;
}

int nccl_fail(int rc, const char* what) {
    snprintf(g_err, sizeof(g_err), "%s: %s", what, g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "nccl error");
    return CLB_ECUDA;
}
#define CLB_NCCL(call) do { int r__ = (call); if (r__ != 0) return nccl_fail(r__, #call); } while (0)

struct Shard {
    int device = 0;
    clb_matcher* m = nullptr;
    cudaStream_t stream = nullptr;        // device-list mode: this device's work stream (= m->stream)
    DevBuf q, part;                       // queries and this shard's partials on this device
    cudaEvent_t ev_q = nullptr, ev_done = nullptr;
    long long lo = 0, hi = 0;             // global map rows [lo, hi)
    const uint64_t* host_rows = nullptr;  // clb_sharded_set_map_host: shard rows stay in (pinned) host memory, streamed per call
};

}  // namespace

struct clb_sharded {
    clb_matcher_opts o;
    bool rank_mode = false;
    int n = 1, rank = 0;                  // shards (devices or ranks); this process's rank (rank mode)
    std::vector<Shard> sh;                // device-list mode: n entries; rank mode: 1
    long long n_map = 0;
    NcclComm comm = nullptr;
    DevBuf gath, out;                     // on the first device / this rank's device: [n][nQ] partials, 3 x nQ results
    void* h_q = nullptr; size_t h_q_cap = 0;          // pinned staging (pageable callers)
    int32_t* h_out = nullptr; size_t h_out_cap = 0;
    cudaStream_t up_stream = nullptr;     // query upload next to the first DB slice (host-DB path)
    cudaEvent_t ev_up = nullptr;
};

static void shard_range(long long n, int r, int world, long long* lo, long long* hi) {
    const long long base = n / world, rem = n % world;     // as coloc_b200/shard.py: the first n % world shards get one more row
    *lo = r * base + (r < rem ? r : rem);
    *hi = *lo + base + (r < rem ? 1 : 0);
}

static int pinned_reserve_any(void** p, size_t* cap, size_t bytes) {
    if (bytes <= *cap) return CLB_OK;
    if (*p) cudaFreeHost(*p);
    *p = nullptr; *cap = 0;
    const size_t want = bytes + bytes / 4 + 4096;
    CLB_CUDA(cudaMallocHost(p, want));
    *cap = want;
    return CLB_OK;
}

static bool is_pinned(const void* p) {
    cudaPointerAttributes pa;
    const bool pinned = cudaPointerGetAttributes(&pa, p) == cudaSuccess && pa.type == cudaMemoryTypeHost;
    cudaGetLastError();
    return pinned;
}

extern "C" {

void clb_sharded_destroy(clb_sharded* s) {
    if (!s) return;
    for (Shard& x : s->sh) {
        cudaSetDevice(x.device);
        if (x.m) cudaStreamSynchronize(x.m->stream);
        x.q.release(); x.part.release();
        if (x.ev_q) cudaEventDestroy(x.ev_q);
        if (x.ev_done) cudaEventDestroy(x.ev_done);
        if (x.m) clb_matcher_destroy(x.m);
    }
    if (!s->sh.empty()) cudaSetDevice(s->sh[0].device);
    if (s->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(s->comm);
    s->gath.release(); s->out.release();
    if (s->h_q) cudaFreeHost(s->h_q);
    if (s->h_out) cudaFreeHost(s->h_out);
    if (s->ev_up) cudaEventDestroy(s->ev_up);
    if (s->up_stream) cudaStreamDestroy(s->up_stream);
    delete s;
}

static int make_shard(clb_sharded* s, Shard& x, int device) {
    x.device = device;
    int rc = clb_matcher_create(&s->o, device, &x.m);
    if (rc) return rc;
    x.stream = x.m->stream;
    CLB_CUDA(cudaEventCreateWithFlags(&x.ev_q, cudaEventDisableTiming));
    CLB_CUDA(cudaEventCreateWithFlags(&x.ev_done, cudaEventDisableTiming));
    return CLB_OK;
}

int clb_sharded_create(const clb_matcher_opts* opts, const int* devices, int ndev, clb_sharded** out) {
    if (!opts || !out || ndev < 1 || ndev > 64) { snprintf(g_err, sizeof(g_err), "bad device list"); return CLB_EINVAL; }
    int have = 0;
    CLB_CUDA(cudaGetDeviceCount(&have));
    clb_sharded* s = new (std::nothrow) clb_sharded();
    if (!s) return CLB_ENOMEM;
    s->o = *opts; s->n = ndev; s->sh.resize(ndev);
    for (int i = 0; i < ndev; ++i) {
        const int dev = devices ? devices[i] : i;
        if (dev < 0 || dev >= have) { snprintf(g_err, sizeof(g_err), "device %d of the list does not exist (%d visible)", dev, have); clb_sharded_destroy(s); return CLB_EINVAL; }
        for (int k = 0; k < i; ++k) if (s->sh[k].device == dev) { snprintf(g_err, sizeof(g_err), "device %d listed twice", dev); clb_sharded_destroy(s); return CLB_EINVAL; }
        const int rc = make_shard(s, s->sh[i], dev);
        if (rc) { clb_sharded_destroy(s); return rc; }
    }
    // peer access lets the partial copies run as NVLink DMA without a host bounce; absent it cudaMemcpyPeerAsync still works
    for (int i = 1; i < ndev; ++i) {
        int can = 0;
        cudaDeviceCanAccessPeer(&can, s->sh[0].device, s->sh[i].device);
        if (can) { cudaSetDevice(s->sh[0].device); cudaDeviceEnablePeerAccess(s->sh[i].device, 0); cudaGetLastError(); }
        cudaDeviceCanAccessPeer(&can, s->sh[i].device, s->sh[0].device);
        if (can) { cudaSetDevice(s->sh[i].device); cudaDeviceEnablePeerAccess(s->sh[0].device, 0); cudaGetLastError(); }
    }
    cudaSetDevice(s->sh[0].device);
    *out = s;
    return CLB_OK;
}

int clb_sharded_unique_id(void* id128_out) {
    if (!id128_out) return CLB_EINVAL;
    int rc = nccl_load();
    if (rc) return rc;
    NcclId id;
    CLB_NCCL(g_nccl.GetUniqueId(&id));
    memcpy(id128_out, &id, sizeof(id));
    return CLB_OK;
}

int clb_sharded_create_rank(const clb_matcher_opts* opts, int device, int rank, int nranks, const void* id128, clb_sharded** out) {
    if (!opts || !out || nranks < 1 || rank < 0 || rank >= nranks || (nranks > 1 && !id128)) { snprintf(g_err, sizeof(g_err), "bad rank / id"); return CLB_EINVAL; }
    clb_sharded* s = new (std::nothrow) clb_sharded();
    if (!s) return CLB_ENOMEM;
    s->o = *opts; s->rank_mode = true; s->n = nranks; s->rank = rank; s->sh.resize(1);
    int rc = make_shard(s, s->sh[0], device);
    if (rc) { clb_sharded_destroy(s); return rc; }
    if (nranks > 1) {
        rc = nccl_load();
        if (rc) { clb_sharded_destroy(s); return rc; }
        NcclId id; memcpy(&id, id128, sizeof(id));
        cudaSetDevice(device);
        const int nrc = g_nccl.CommInitRank(&s->comm, nranks, id, rank);
        if (nrc != 0) { s->comm = nullptr; clb_sharded_destroy(s); return nccl_fail(nrc, "ncclCommInitRank"); }
    }
    *out = s;
    return CLB_OK;
}

int clb_sharded_shards(const clb_sharded* s) { return s ? s->n : 0; }

int clb_sharded_range(const clb_sharded* s, long long n_rows, int shard, long long* lo_out, long long* hi_out) {
    if (!s || shard < 0 || shard >= s->n || n_rows < 0 || !lo_out || !hi_out) return CLB_EINVAL;
    shard_range(n_rows, shard, s->n, lo_out, hi_out);
    return CLB_OK;
}

// setMapData for the whole map (HOST pointer, nM rows): every shard copies its own rows and keeps them resident.
// In rank mode every rank passes the same whole map (or use clb_sharded_set_map_shard with its rows only).
int clb_sharded_set_map(clb_sharded* s, const uint64_t* M, long long nM) {
    if (!s || nM < 0 || nM > 0x7FFFFFFFll * 64 || (nM > 0 && !M)) return CLB_EINVAL;
    for (int i = 0; i < (int)s->sh.size(); ++i) {
        Shard& x = s->sh[i];
        shard_range(nM, s->rank_mode ? s->rank : i, s->n, &x.lo, &x.hi);
        x.host_rows = nullptr;
        if (x.hi - x.lo > 0x7FFFFFFFll) { snprintf(g_err, sizeof(g_err), "shard of %lld rows is too large", x.hi - x.lo); return CLB_EINVAL; }
        const int rc = clb_set_map(x.m, M + 8 * x.lo, (int)(x.hi - x.lo));      // blocking H2D, as setMapData
        if (rc) return rc;
    }
    s->n_map = nM;
    return CLB_OK;
}

// The whole map stays in HOST memory (pinned for overlap; the caller keeps it alive): every shard streams its rows in slices
// behind its sweep on every match call (clb_match_partial_hostdb) -- the "map lives in host memory" boundary number of SURVEY 8(d).
int clb_sharded_set_map_host(clb_sharded* s, const uint64_t* M, long long nM) {
    if (!s || nM < 0 || (nM > 0 && !M)) return CLB_EINVAL;
    for (int i = 0; i < (int)s->sh.size(); ++i) {
        Shard& x = s->sh[i];
        shard_range(nM, s->rank_mode ? s->rank : i, s->n, &x.lo, &x.hi);
        if (x.hi - x.lo > 0x7FFFFFFFll) { snprintf(g_err, sizeof(g_err), "shard of %lld rows is too large", x.hi - x.lo); return CLB_EINVAL; }
        x.host_rows = M + 8 * x.lo;
    }
    s->n_map = nM;
    return CLB_OK;
}

// rank mode: this rank's rows only -- global rows [row_lo, row_lo + rows) of a map of n_total rows.  `where`: 0 = HOST rows copied
// to the device now and kept resident; 1 = DEVICE rows (copied device-to-device on the shard's stream); 2 = HOST rows that STAY
// in host memory (pinned for overlap) and are streamed in slices behind the sweep on every clb_sharded_match_map* call -- the
// "map lives in host memory" boundary number of SURVEY 8(d).
int clb_sharded_set_map_shard(clb_sharded* s, const uint64_t* rows_ptr, long long row_lo, int rows, long long n_total, int where) {
    if (!s || !s->rank_mode || rows < 0 || row_lo < 0 || row_lo + rows > n_total || (rows > 0 && !rows_ptr) || where < 0 || where > 2) return CLB_EINVAL;
    Shard& x = s->sh[0];
    x.lo = row_lo; x.hi = row_lo + rows; x.host_rows = nullptr;
    s->n_map = n_total;
    if (where == 0) return clb_set_map(x.m, rows_ptr, rows);
    if (where == 1) return clb_set_map_dev(x.m, rows_ptr, rows, nullptr);
    x.host_rows = rows_ptr;
    return CLB_OK;
}

// enqueue one shard's sweep: queries already on the shard's device in x.q
static int shard_sweep(Shard& x, int nQ, cudaStream_t st) {
    CLB_CUDA(x.part.reserve((size_t)nQ * sizeof(clb_partial)));
    if (x.host_rows) return clb_match_partial_hostdb(x.m, x.q.as<uint64_t>(), nQ, x.host_rows, (int)(x.hi - x.lo), x.lo, x.part.as<clb_partial>(), st);
    return clb_match_map_partial_dev(x.m, x.q.as<uint64_t>(), nQ, x.lo, x.part.as<clb_partial>(), st);
}

// rank mode, DEVICE queries and outputs on this rank's device, asynchronous on `stream` (NULL = the shard's own stream):
// local sweep -> ncclAllGather of nQ x 16 B per rank -> merge.  Every rank gets the full result.
int clb_sharded_match_map_dev(clb_sharded* s, const uint64_t* dQ, int nQ, int threshold,
                              int32_t* d_match_out, int32_t* d_best_out, int32_t* d_second_out, void* stream) {
    if (!s || !s->rank_mode || nQ < 0 || (nQ > 0 && (!dQ || !d_match_out))) return CLB_EINVAL;
    if (nQ == 0) return CLB_OK;
    Shard& x = s->sh[0];
    CLB_CUDA(cudaSetDevice(x.device));
    cudaStream_t st = stream ? (cudaStream_t)stream : x.stream;
    const int thr = threshold < 0 ? s->o.thresh : threshold;
    CLB_CUDA(x.part.reserve((size_t)nQ * sizeof(clb_partial)));
    int rc;
    tl_mark("sweep_begin", st);
    if (x.host_rows) rc = clb_match_partial_hostdb(x.m, dQ, nQ, x.host_rows, (int)(x.hi - x.lo), x.lo, x.part.as<clb_partial>(), st);
    else             rc = clb_match_map_partial_dev(x.m, dQ, nQ, x.lo, x.part.as<clb_partial>(), st);
    if (rc) return rc;
    const clb_partial* parts = x.part.as<clb_partial>();
    tl_mark("sweep_end", st);
    if (s->n > 1) {
        CLB_CUDA(s->gath.reserve((size_t)s->n * nQ * sizeof(clb_partial)));
        CLB_NCCL(g_nccl.AllGather(x.part.p, s->gath.p, (size_t)nQ * sizeof(clb_partial), 0 /* ncclChar */, s->comm, st));
        parts = s->gath.as<clb_partial>();
        tl_mark("allgather_end", st);
    }
    rc = clb_merge_partials_dev(x.m, parts, s->n, nQ, thr, d_match_out, d_best_out, d_second_out, st);
    tl_mark("merge_end", st);
    return rc;
}

// matchFeaturesWithMap (GPUMatcher.hpp:252-271) over all shards, HOST queries in, HOST results out, blocking.
// device-list mode: the queries go to every device, the sweeps run concurrently, the partials meet on the first device.
// rank mode: collective -- every rank calls it with the same queries and gets the same result.
int clb_sharded_match_map(clb_sharded* s, const uint64_t* Q, int nQ, int threshold,
                          int32_t* match_out, int32_t* best_d_out, int32_t* second_d_out) {
    if (!s || nQ < 0 || (nQ > 0 && (!Q || !match_out))) return CLB_EINVAL;
    if (nQ == 0) return CLB_OK;
    const int thr = threshold < 0 ? s->o.thresh : threshold;
    const size_t qbytes = (size_t)nQ * 64;
    Shard& x0 = s->sh[0];
    CLB_CUDA(cudaSetDevice(x0.device));
    // pageable queries are staged once in pinned memory so that the uploads (N of them in device-list mode) are asynchronous
    const uint64_t* src = Q;
    if (!is_pinned(Q)) {
        int rc = pinned_reserve_any(&s->h_q, &s->h_q_cap, qbytes);
        if (rc) return rc;
        memcpy(s->h_q, Q, qbytes);
        src = static_cast<const uint64_t*>(s->h_q);
    }
    CLB_CUDA(s->out.reserve((size_t)nQ * 12));
    int32_t* o = s->out.as<int32_t>();
    int rc;
    if (s->rank_mode) {
        cudaStream_t st = x0.stream;
        CLB_CUDA(x0.q.reserve(qbytes));
        tl_mark("call_begin", st);
        if (x0.host_rows) {
            // the first DB slice is as urgent as the queries: give the query upload its own stream so that both DMA queues start at once
            if (!s->up_stream) { CLB_CUDA(cudaStreamCreateWithFlags(&s->up_stream, cudaStreamNonBlocking)); CLB_CUDA(cudaEventCreateWithFlags(&s->ev_up, cudaEventDisableTiming)); }
            CLB_CUDA(cudaMemcpyAsync(x0.q.p, src, qbytes, cudaMemcpyHostToDevice, s->up_stream));
            CLB_CUDA(cudaEventRecord(s->ev_up, s->up_stream));
            tl_mark("queries_uploaded", s->up_stream);
            x0.m->q_wait_ev = s->ev_up; x0.m->q_wait_set = true;      // the first sweep launch waits for it (matcher_priv.cuh)
        } else {
            CLB_CUDA(cudaMemcpyAsync(x0.q.p, src, qbytes, cudaMemcpyHostToDevice, st));
        }
        rc = clb_sharded_match_map_dev(s, x0.q.as<uint64_t>(), nQ, thr, o, best_d_out ? o + nQ : nullptr, second_d_out ? o + 2 * (size_t)nQ : nullptr, st);
        if (rc) return rc;
    } else {
        CLB_CUDA(s->gath.reserve((size_t)s->n * nQ * sizeof(clb_partial)));
        for (int i = 0; i < s->n; ++i) {
            Shard& x = s->sh[i];
            CLB_CUDA(cudaSetDevice(x.device));
            CLB_CUDA(x.q.reserve(qbytes));
            CLB_CUDA(cudaMemcpyAsync(x.q.p, src, qbytes, cudaMemcpyHostToDevice, x.stream));
            rc = shard_sweep(x, nQ, x.stream);
            if (rc) return rc;
            // partials of shard i -> slot i of the gather buffer on the first device
            clb_partial* dst = s->gath.as<clb_partial>() + (size_t)i * nQ;
            if (i == 0) CLB_CUDA(cudaMemcpyAsync(dst, x.part.p, (size_t)nQ * sizeof(clb_partial), cudaMemcpyDeviceToDevice, x.stream));
            else        CLB_CUDA(cudaMemcpyPeerAsync(dst, x0.device, x.part.p, x.device, (size_t)nQ * sizeof(clb_partial), x.stream));
            CLB_CUDA(cudaEventRecord(x.ev_done, x.stream));
        }
        CLB_CUDA(cudaSetDevice(x0.device));
        for (int i = 1; i < s->n; ++i) CLB_CUDA(cudaStreamWaitEvent(x0.stream, s->sh[i].ev_done, 0));
        rc = clb_merge_partials_dev(x0.m, s->gath.as<clb_partial>(), s->n, nQ, thr, o, best_d_out ? o + nQ : nullptr,
                                    second_d_out ? o + 2 * (size_t)nQ : nullptr, x0.stream);
        if (rc) return rc;
    }
    cudaStream_t st = x0.stream;
    CLB_CUDA(cudaMemcpyAsync(match_out, o, (size_t)nQ * 4, cudaMemcpyDeviceToHost, st));
    if (best_d_out) CLB_CUDA(cudaMemcpyAsync(best_d_out, o + nQ, (size_t)nQ * 4, cudaMemcpyDeviceToHost, st));
    if (second_d_out) CLB_CUDA(cudaMemcpyAsync(second_d_out, o + 2 * (size_t)nQ, (size_t)nQ * 4, cudaMemcpyDeviceToHost, st));
    tl_mark("results_on_host", st);
    CLB_CUDA(cudaStreamSynchronize(st));
    return CLB_OK;
}

}  // extern "C"
