// api.cu -- the C ABI (include/coloc_b200.h): detector / matcher objects that own device
// buffers for their lifetime, as the reference's GPUDetector / GPUMatcher do
// (include/coloc/GPUDetector.hpp:70-155, GPUMatcher.hpp:70-108).
#include "common.cuh"
#include <string.h>
#include <stdlib.h>
#include <nvtx3/nvToolsExt.h>
#include <map>
#include <mutex>
#include <string>
#include <vector>
#include <new>

namespace clb {
thread_local char g_err[512] = "";
std::atomic<int64_t> g_launches{0};
std::atomic<int> g_timing{0};

int set_cuda_error(cudaError_t e, const char* what, const char* file, int line) {
    snprintf(g_err, sizeof(g_err), "%s: %s (%s:%d)", what, cudaGetErrorString(e), file, line);
    cudaGetLastError();
    return e == cudaErrorMemoryAllocation ? CLB_ENOMEM : CLB_ECUDA;
}

static std::mutex g_timer_mu;
static std::map<std::string, KernelTimer> g_timers;
KernelTimer& timer(const char* which) {
    std::lock_guard<std::mutex> lk(g_timer_mu);
    return g_timers[which];
}
bool nvtx_on() { static const bool on = [] { const char* e = getenv("COLOC_B200_NVTX"); return e && *e && *e != '0'; }(); return on; }
void nvtx_push(const char* name) { nvtxRangePushA(name); }
void nvtx_pop() { nvtxRangePop(); }

struct TlMark { std::string name; cudaEvent_t ev; };
static std::vector<TlMark> g_tl;
void tl_mark(const char* name, cudaStream_t s) {
    if (g_timing.load() != 2) return;
    std::lock_guard<std::mutex> lk(g_timer_mu);
    if (g_tl.size() >= 512) return;
    TlMark m; m.name = name;
    if (cudaEventCreate(&m.ev) != cudaSuccess) return;
    cudaEventRecord(m.ev, s);
    g_tl.push_back(m);
}
}  // namespace clb

using namespace clb;

static inline int align_up(int v, int a) { return (v + a - 1) / a * a; }

// ------------------------------------------------------------------------------ detector
struct clb_detector {
    clb_detector_opts o;
    int device, sm_count;
    cudaStream_t stream;
    DevBuf pyr;          // all levels, packed, 128-B aligned strides
    DevBuf fast_scratch;
    DevBuf kps, desc, misc;   // misc: [0] total count
    DevBuf stage;        // staging for stage-level entry points
    int* h_total;        // pinned
    size_t level_off[CLB_MAX_LEVELS];
    LevelTex tex;        // texture views of the pyramid levels for the frame geometry tex_w x tex_h
    int tex_w, tex_h, tex_n;
    bool tex_fresh;      // texture objects created since the last describe launch: that launch must be a plain one (see run_pipeline)
    cudaStream_t copy_stream;   // keypoint D2H of the blocking per-frame call, behind the describe kernel
    cudaEvent_t ev_count, ev_fast, ev_done;
    bool copy_ready;
    // per frame geometry (rebuilt when w x h changes): TMA tensor maps of the levels, the resize kernel's source-box plan and
    // its column / row tables
    LevelMaps maps;
    ResizePlan rplan;
    DevBuf rtab;
    int geo_w, geo_h;
    int fast_quirk;      // 1 = the compiled reference's KFAST column walk (KFAST.h:243 `1 << 32`), 0 = every column scanned
};

static void level_dims(const clb_detector_opts& o, int w, int h, int* ws, int* hs, float* fs) {
    // GPUDetector.hpp:107-114
    float f = 1.0f;
    ws[0] = w; hs[0] = h; fs[0] = 1.0f;
    for (int i = 1; i < o.scale_levels; ++i) {
        f *= o.scale_factor;
        ws[i] = (int)(uint32_t)((float)w / f + 0.5f);
        hs[i] = (int)(uint32_t)((float)h / f + 0.5f);
        fs[i] = f;
    }
}

// lay the levels of a w x h frame out inside d->pyr; returns bytes needed
static size_t make_table(clb_detector* d, int w, int h, LevelTable& lt) {
    float fs[CLB_MAX_LEVELS];
    lt.n = d->o.scale_levels;
    level_dims(d->o, w, h, lt.w, lt.h, fs);
    size_t off = 0;
    for (int i = 0; i < lt.n; ++i) {
        lt.f[i] = fs[i];
        lt.stride[i] = align_up(lt.w[i] > 0 ? lt.w[i] : 1, 128);
        d->level_off[i] = off;
        off += (size_t)lt.stride[i] * (lt.h[i] > 0 ? lt.h[i] : 1) + 256;
        off = (off + 511) & ~(size_t)511;          // 512 B: cudaDeviceProp::textureAlignment
    }
    for (int i = 0; i < lt.n; ++i) lt.img[i] = d->pyr.as<uint8_t>() ? d->pyr.as<uint8_t>() + d->level_off[i] : nullptr;
    return off;
}

// tensor maps of the FAST tiles, one per level (the only maps a single-level call needs)
static int make_fast_maps(const LevelTable& lt, LevelMaps& maps) {
    int bw = 0, bh = 0;
    fast_tile_box(&bw, &bh);
    for (int L = 0; L < lt.n; ++L) {
        if (!lt.img[L] || lt.w[L] < 1 || lt.h[L] < 1) continue;
        const int rc = make_tmap_2d(&maps.fast[L], lt.img[L], 1, (uint64_t)lt.w[L], (uint64_t)lt.h[L], (uint64_t)lt.stride[L], (uint32_t)bw, (uint32_t)bh);
        if (rc) return rc;
    }
    return CLB_OK;
}

// everything that depends on the frame geometry only: tensor maps, resize plan + tables (the pyramid buffer itself never moves)
static int ensure_geometry(clb_detector* d, const LevelTable& lt, cudaStream_t s) {
    if (d->geo_w == lt.w[0] && d->geo_h == lt.h[0]) return CLB_OK;
    d->geo_w = d->geo_h = 0;
    memset(&d->maps, 0, sizeof(d->maps));
    int rc = make_fast_maps(lt, d->maps);
    if (rc) return rc;
    rc = resize_plan(lt, d->rplan);
    if (rc) return rc;
    for (int L = 1; L < lt.n; ++L) {
        if (lt.w[L] < 1 || lt.h[L] < 1) continue;
        rc = make_tmap_2d(&d->maps.resize[L], lt.img[0], 1, (uint64_t)lt.w[0], (uint64_t)lt.h[0], (uint64_t)lt.stride[0],
                          (uint32_t)d->rplan.box_w[L], (uint32_t)d->rplan.box_h[L]);
        if (rc) return rc;
    }
    // the previous geometry's tables may still be read by work queued on another stream of this detector
    CLB_CUDA(cudaDeviceSynchronize());
    CLB_CUDA(d->rtab.reserve(resize_table_entries(lt) * sizeof(uint2)));
    d->rplan.tab = d->rtab.as<uint2>();
    rc = launch_resize_tables(lt, d->rplan, d->rtab.as<uint2>(), s);
    if (rc) return rc;
    d->geo_w = lt.w[0]; d->geo_h = lt.h[0];
    return CLB_OK;
}

static int resize_frame(clb_detector* d, const LevelTable& lt, cudaStream_t s, bool pdl) {
    uint32_t* clear = nullptr; int words = 0;
    fast_clear_region(lt, d->fast_scratch.p, &clear, &words);
    return launch_resize(lt, d->rplan, d->maps, clear, words, s, pdl);
}

// texture views follow the frame geometry (the pyramid buffer itself never moves)
static int ensure_textures(clb_detector* d, const LevelTable& lt, int w, int h) {
    if (d->tex_n == lt.n && d->tex_w == w && d->tex_h == h) return CLB_OK;
    destroy_level_textures(d->tex, d->tex_n);
    d->tex_n = 0;
    int rc = make_level_textures(lt, d->tex);
    if (rc) return rc;
    d->tex_n = lt.n; d->tex_w = w; d->tex_h = h; d->tex_fresh = true;
    return CLB_OK;
}

// The describe kernel joins the frame's programmatic-dependent-launch chain EXCEPT on the first launch after its texture objects
// were (re)created: a programmatically launched grid can start on SMs whose texture-header cache still holds the header a
// destroyed object left in the same slot (measured: the first frame of a detector created after another one was destroyed had
// 1/7 of its descriptors -- one GPC's CTAs -- computed from the old header); a plain launch goes through the full kernel
// boundary that applies the invalidation.
static bool describe_pdl(clb_detector* d) { const bool pdl = !d->tex_fresh; d->tex_fresh = false; return pdl; }

extern "C" {

const char* clb_strerror(int code) {
    switch (code) {
        case CLB_OK: return "ok";
        case CLB_EINVAL: return "invalid argument";
        case CLB_ECUDA: return "CUDA runtime error";
        case CLB_ECAPACITY: return "output capacity too small";
        case CLB_ENOMEM: return "out of device memory";
        case CLB_EUNSUPPORTED: return "unsupported";
        case CLB_ENOTFOUND: return "feature set not found";
        case CLB_EIO: return "file i/o failed";
        default: return "unknown error";
    }
}
const char* clb_last_error(void) { return g_err; }
int clb_version(void) { return 100; }
int clb_device_count(int* n_out) { if (!n_out) return CLB_EINVAL; CLB_CUDA(cudaGetDeviceCount(n_out)); return CLB_OK; }
int64_t clb_launch_count(void) { return g_launches.load(); }
int clb_set_timing(int on) { g_timing = on; return CLB_OK; }
// "name:ms name:ms ..." of the marks recorded since the last call (timing level 2), offsets from the first mark; clears them
int clb_timeline(char* buf, int cap) {
    if (!buf || cap < 2) return CLB_EINVAL;
    std::lock_guard<std::mutex> lk(g_timer_mu);
    buf[0] = 0;
    int off = 0;
    for (size_t i = 0; i < g_tl.size(); ++i) {
        cudaEventSynchronize(g_tl[i].ev);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, g_tl[0].ev, g_tl[i].ev);
        const int n = snprintf(buf + off, (size_t)(cap - off), "%s%s:%.4f", i ? " " : "", g_tl[i].name.c_str(), ms);
        if (n < 0 || n >= cap - off) break;
        off += n;
    }
    for (auto& m : g_tl) cudaEventDestroy(m.ev);
    g_tl.clear();
    cudaGetLastError();
    return CLB_OK;
}
float clb_last_kernel_ms(const char* which) { return which ? timer(which).ms() : 0.f; }
// Planning knobs are process-wide atomics (set them before the calls they should affect); kernel-variant and timing knobs
// exist in the lab build only and answer CLB_EUNSUPPORTED here otherwise, so a stray call cannot change what the product runs.
static bool lab_knob(const char* name) {
    static const char* const k[] = {"match_hs", "match_tc_ver", "match_tc_dbg", "latch_tex", "latch_var", "fast_nms", "fast_var"};
    for (const char* n : k) if (!strcmp(name, n)) return true;
    return false;
}
int clb_set_param(const char* name, int value) {
    if (!name) return CLB_EINVAL;
    if (!strcmp(name, "match_rq")) { if (value != 0 && value != 1 && value != 2 && value != 4) return CLB_EINVAL; g_match_rq = value; }
    else if (!strcmp(name, "match_chunks")) { if (value < 0) return CLB_EINVAL; g_match_chunks = value; }
    else if (!strcmp(name, "match_tc")) { if (value < -1 || value > 1) return CLB_EINVAL; g_match_tc = value; }
    else if (!strcmp(name, "match_slice_rows")) g_match_slice_rows = value < 1024 ? 1024 : value;
    else if (!strcmp(name, "match_balance")) { if (value < -1 || value > 1) return CLB_EINVAL; g_match_balance = value; }
    else if (!strcmp(name, "match_tc_chunks")) { if (value < 0) return CLB_EINVAL; g_match_tc_chunks = value; }
#ifdef CLB_LAB
    else if (!strcmp(name, "match_hs")) { if (value < 0 || value > 3) return CLB_EINVAL; g_match_hs = value; }
    else if (!strcmp(name, "match_tc_ver")) { if (value < 1 || value > 4) return CLB_EINVAL; g_match_tc_ver = value; }
    else if (!strcmp(name, "match_tc_dbg")) g_match_tc_dbg = value;
    else if (!strcmp(name, "latch_tex")) g_latch_tex = value ? 1 : 0;
    else if (!strcmp(name, "latch_var")) g_latch_var = value & 7;
    else if (!strcmp(name, "fast_nms")) g_fast_nms = value ? 1 : 0;
    else if (!strcmp(name, "fast_var")) g_fast_var = value & 63;
#else
    else if (lab_knob(name)) { snprintf(g_err, sizeof(g_err), "%s exists in the lab build only (make -C coloc_b200/csrc lab)", name); return CLB_EUNSUPPORTED; }
#endif
    else return CLB_EINVAL;
    return CLB_OK;
}
int clb_get_param(const char* name, int* value) {
    if (!name || !value) return CLB_EINVAL;
    if (!strcmp(name, "match_rq")) *value = g_match_rq;
    else if (!strcmp(name, "match_chunks")) *value = g_match_chunks;
    else if (!strcmp(name, "match_tc")) *value = g_match_tc;
    else if (!strcmp(name, "match_tc_chunks")) *value = g_match_tc_chunks;
    else if (!strcmp(name, "match_slice_rows")) *value = g_match_slice_rows;
    else if (!strcmp(name, "last_match_path")) *value = g_last_match_path;
    else if (!strcmp(name, "last_match_balanced")) *value = g_last_match_balanced;
    else if (!strcmp(name, "match_balance")) *value = g_match_balance;
#ifdef CLB_LAB
    else if (!strcmp(name, "lab")) *value = 1;
    else if (!strcmp(name, "match_hs")) *value = g_match_hs;
    else if (!strcmp(name, "match_tc_ver")) *value = g_match_tc_ver;
    else if (!strcmp(name, "fast_var")) *value = g_fast_var;
    else if (!strcmp(name, "latch_var")) *value = g_latch_var;
#else
    else if (!strcmp(name, "lab")) *value = 0;
    else if (!strcmp(name, "match_hs")) *value = 2;
    else if (!strcmp(name, "match_tc_ver")) *value = 4;
    else if (!strcmp(name, "fast_var")) *value = 55;
    else if (!strcmp(name, "latch_var")) *value = 3;
#endif
    else return CLB_EINVAL;
    return CLB_OK;
}
int clb_microbench_pipes(int device, double* out8) {
    if (!out8) return CLB_EINVAL;
    CLB_CUDA(cudaSetDevice(device));
    return microbench_pipes(out8);
}

int clb_host_alloc(size_t bytes, void** out) {
    if (!out || bytes == 0) return CLB_EINVAL;
    CLB_CUDA(cudaMallocHost(out, bytes));
    return CLB_OK;
}
void clb_host_free(void* p) { if (p) cudaFreeHost(p); }

int clb_microbench_mma(int device, double* out6) {
    if (!out6) return CLB_EINVAL;
    CLB_CUDA(cudaSetDevice(device));
    return microbench_mma(out6);
}

#ifdef CLB_LAB
// lab library only (not in include/coloc_b200.h): per-tile clocks of CTA (0, 0) of the last sweep run with match_tc_dbg = 7
int clb_lab_tile_clocks(long long* out1024) { return out1024 ? tc_lab_tile_clocks(out1024) : CLB_EINVAL; }
#endif

int clb_detector_create(const clb_detector_opts* opts, int device, clb_detector** out) {
    if (!opts || !out) return CLB_EINVAL;
    if (opts->scale_levels < 1 || opts->scale_levels > CLB_MAX_LEVELS || opts->width < 1 || opts->height < 1 ||
        opts->maxkp < 1 || opts->thresh < 0 || opts->thresh > 255 || !(opts->scale_factor >= 1.0f)) {
        snprintf(g_err, sizeof(g_err), "bad detector options"); return CLB_EINVAL;
    }
    CLB_CUDA(cudaSetDevice(device));
    clb_detector* d = new (std::nothrow) clb_detector();
    if (!d) return CLB_ENOMEM;
    d->o = *opts; d->device = device; d->h_total = nullptr; d->fast_quirk = 1; d->geo_w = d->geo_h = 0;
    memset(&d->tex, 0, sizeof(d->tex)); d->tex_w = d->tex_h = d->tex_n = 0; d->tex_fresh = false;
    cudaDeviceGetAttribute(&d->sm_count, cudaDevAttrMultiProcessorCount, device);
    cudaError_t e = cudaStreamCreateWithFlags(&d->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaMallocHost(&d->h_total, 64);
    LevelTable lt;
    size_t bytes = make_table(d, opts->width, opts->height, lt);
    if (e == cudaSuccess) e = d->pyr.reserve(bytes);
    make_table(d, opts->width, opts->height, lt);
    if (e == cudaSuccess) e = d->fast_scratch.reserve(fast_scratch_bytes(lt));
    if (e == cudaSuccess) e = d->kps.reserve((size_t)opts->maxkp * sizeof(clb_keypoint));
    if (e == cudaSuccess) e = d->desc.reserve((size_t)opts->maxkp * 64);
    if (e == cudaSuccess) e = d->misc.reserve(256);
    if (e != cudaSuccess) { int rc = set_cuda_error(e, "detector allocation", __FILE__, __LINE__); clb_detector_destroy(d); return rc; }
    int rc = latch_init();
    if (rc) { clb_detector_destroy(d); return rc; }
    *out = d;
    return CLB_OK;
}

void clb_detector_destroy(clb_detector* d) {
    if (!d) return;
    cudaSetDevice(d->device);
    destroy_level_textures(d->tex, d->tex_n);
    d->pyr.release(); d->fast_scratch.release(); d->kps.release(); d->desc.release(); d->misc.release(); d->stage.release(); d->rtab.release();
    if (d->h_total) cudaFreeHost(d->h_total);
    if (d->copy_ready) { cudaEventDestroy(d->ev_count); cudaEventDestroy(d->ev_fast); cudaEventDestroy(d->ev_done); cudaStreamDestroy(d->copy_stream); }
    if (d->stream) cudaStreamDestroy(d->stream);
    delete d;
}

int clb_detector_set_param(clb_detector* d, const char* name, int value) {
    if (!d || !name) return CLB_EINVAL;
    if (!strcmp(name, "fast_quirk")) { if (value != 0 && value != 1) return CLB_EINVAL; d->fast_quirk = value; return CLB_OK; }
    snprintf(g_err, sizeof(g_err), "unknown detector parameter %s", name);
    return CLB_EINVAL;
}
int clb_detector_get_param(const clb_detector* d, const char* name, int* value_out) {
    if (!d || !name || !value_out) return CLB_EINVAL;
    if (!strcmp(name, "fast_quirk")) { *value_out = d->fast_quirk; return CLB_OK; }
    return CLB_EINVAL;
}

int clb_detector_level_dims(const clb_detector* d, int w, int h, int32_t* ws, int32_t* hs, float* fs) {
    if (!d || !ws || !hs) return CLB_EINVAL;
    float tmp[CLB_MAX_LEVELS];
    level_dims(d->o, w, h, ws, hs, fs ? fs : tmp);
    return CLB_OK;
}

// the blocking per-frame call waits by polling the event: a sleeping wait (cudaEventSynchronize / cudaStreamSynchronize) is
// woken 10-20 us late, a visible share of a 0.13 - 0.3 ms call; the reference's caller blocks here as well (GPUDetector.hpp:290)
static int spin_wait(cudaEvent_t ev) {
    for (;;) {
        const cudaError_t e = cudaEventQuery(ev);
        if (e == cudaSuccess) return CLB_OK;
        if (e != cudaErrorNotReady) return set_cuda_error(e, "cudaEventQuery", __FILE__, __LINE__);
    }
}

static int check_frame(clb_detector* d, int w, int h, int stride) {
    if (!d || w < 1 || h < 1 || stride < w) { snprintf(g_err, sizeof(g_err), "bad frame geometry"); return CLB_EINVAL; }
    if (w > d->o.width || h > d->o.height) { snprintf(g_err, sizeof(g_err), "frame %dx%d exceeds detector size %dx%d", w, h, d->o.width, d->o.height); return CLB_EINVAL; }
    return CLB_OK;
}

// device pipeline on a level-0 image already inside d->pyr (table lt)
// resize -> [scan ends] -> FAST score + NMS -> compaction + orientation -> LATCH, chained by programmatic dependent launch:
// each kernel's CTAs are scheduled while the one in front drains and wait (griddepcontrol.wait) before they read its output.
// The resize kernel itself is a PLAIN launch: a programmatically launched grid does not get the L1 / texture-cache
// invalidation of a kernel boundary (measured: with the whole chain programmatic, the describe kernel of frame 2 read
// frame 1's keypoints out of L1 through its non-coherent loads), so every frame starts with one full boundary and the
// kernels behind it only ever see lines that were fetched after their producer finished.
static int run_pipeline(clb_detector* d, const LevelTable& lt, cudaStream_t s) {
    int rc = ensure_textures(d, lt, lt.w[0], lt.h[0]);
    if (rc) return rc;
    rc = resize_frame(d, lt, s, false);
    if (rc) return rc;
    int* d_total = d->misc.as<int>();
    rc = launch_fast(lt, d->maps, d->o.thresh, d->fast_quirk, d->fast_scratch.p, d->fast_scratch.cap, d->kps.as<clb_keypoint>(), d->o.maxkp, d_total, s, true, false);
    if (rc) return rc;
    return launch_describe(lt, d->kps.as<clb_keypoint>(), 0, d_total, d->o.maxkp, d->desc.as<uint64_t>(), s, &d->tex, describe_pdl(d));
}

int clb_detect_describe_dev(clb_detector* d, const uint8_t* d_gray, int w, int h, int stride,
                            const clb_keypoint** d_kps, const uint64_t** d_desc, int* n_out, void* stream) {
    int rc = check_frame(d, w, h, stride);
    if (rc) return rc;
    if (!d_gray || !n_out) return CLB_EINVAL;
    CLB_CUDA(cudaSetDevice(d->device));
    cudaStream_t s = stream ? (cudaStream_t)stream : d->stream;
    LevelTable lt;
    make_table(d, w, h, lt);
    rc = ensure_geometry(d, lt, s);
    if (rc) return rc;
    CLB_CUDA(cudaMemcpy2DAsync(const_cast<uint8_t*>(lt.img[0]), lt.stride[0], d_gray, stride, w, h, cudaMemcpyDeviceToDevice, s));
    rc = run_pipeline(d, lt, s);
    if (rc) return rc;
    CLB_CUDA(cudaMemcpyAsync(d->h_total, d->misc.p, sizeof(int), cudaMemcpyDeviceToHost, s));
    CLB_CUDA(cudaStreamSynchronize(s));
    *n_out = *d->h_total;
    if (d_kps) *d_kps = d->kps.as<clb_keypoint>();
    if (d_desc) *d_desc = d->desc.as<uint64_t>();
    if (*n_out > d->o.maxkp) { snprintf(g_err, sizeof(g_err), "%d keypoints exceed maxkp %d", *n_out, d->o.maxkp); return CLB_ECAPACITY; }
    return CLB_OK;
}

int clb_detect_describe(clb_detector* d, const uint8_t* gray, int w, int h, int stride,
                        clb_keypoint* kps_out, uint64_t* desc_out, int cap, int* n_out) {
    int rc = check_frame(d, w, h, stride);
    if (rc) return rc;
    if (!gray || !n_out || cap < 0) return CLB_EINVAL;
    CLB_CUDA(cudaSetDevice(d->device));
    cudaStream_t s = d->stream;
    LevelTable lt;
    make_table(d, w, h, lt);
    rc = ensure_geometry(d, lt, s);
    if (rc) return rc;
    CLB_CUDA(cudaMemcpy2DAsync(const_cast<uint8_t*>(lt.img[0]), lt.stride[0], gray, stride, w, h, cudaMemcpyHostToDevice, s));
    if (!d->copy_ready) {
        CLB_CUDA(cudaStreamCreateWithFlags(&d->copy_stream, cudaStreamNonBlocking));
        CLB_CUDA(cudaEventCreateWithFlags(&d->ev_count, cudaEventDisableTiming));
        CLB_CUDA(cudaEventCreateWithFlags(&d->ev_fast, cudaEventDisableTiming));
        CLB_CUDA(cudaEventCreateWithFlags(&d->ev_done, cudaEventDisableTiming));
        d->copy_ready = true;
    }
    // the keypoints are final after the FAST stage: their count comes back on a second stream while the describe kernel runs
    // (so the kernels of the frame stay one programmatic-dependent-launch chain), and the keypoint D2H copy follows it there;
    // only the descriptor copy comes behind the describe kernel
    rc = ensure_textures(d, lt, lt.w[0], lt.h[0]);
    if (rc) return rc;
    rc = resize_frame(d, lt, s, false);               // plain launch: the frame's one full kernel boundary (see run_pipeline)
    if (rc) return rc;
    int* d_total = d->misc.as<int>();
    rc = launch_fast(lt, d->maps, d->o.thresh, d->fast_quirk, d->fast_scratch.p, d->fast_scratch.cap, d->kps.as<clb_keypoint>(), d->o.maxkp, d_total, s, true, false);
    if (rc) return rc;
    CLB_CUDA(cudaEventRecord(d->ev_fast, s));
    // a pinned (page-locked) descriptor buffer is written by the describe kernel itself, row by row over PCIe while it runs:
    // the 64 n-byte D2H copy behind the kernel disappears from the call's latency.  Pageable buffers take the copy.
    uint64_t* desc_dev = d->desc.as<uint64_t>();
    bool desc_direct = false;
    if (desc_out && cap > 0) {
        cudaPointerAttributes pa;
        if (cudaPointerGetAttributes(&pa, desc_out) == cudaSuccess && pa.type == cudaMemoryTypeHost && pa.devicePointer) {
            desc_dev = static_cast<uint64_t*>(pa.devicePointer); desc_direct = true;
        }
        cudaGetLastError();
    }
    rc = launch_describe(lt, d->kps.as<clb_keypoint>(), 0, d_total, desc_direct ? (cap < d->o.maxkp ? cap : d->o.maxkp) : d->o.maxkp, desc_dev, s, &d->tex, describe_pdl(d));
    if (rc) return rc;
    CLB_CUDA(cudaEventRecord(d->ev_done, s));
    CLB_CUDA(cudaStreamWaitEvent(d->copy_stream, d->ev_fast, 0));
    CLB_CUDA(cudaMemcpyAsync(d->h_total, d_total, sizeof(int), cudaMemcpyDeviceToHost, d->copy_stream));
    CLB_CUDA(cudaEventRecord(d->ev_count, d->copy_stream));
    rc = spin_wait(d->ev_count);
    if (rc) return rc;
    const int n = *d->h_total;
    *n_out = n;
    if (n > d->o.maxkp || n > cap) {
        cudaStreamSynchronize(s);
        snprintf(g_err, sizeof(g_err), "%d keypoints exceed capacity (cap %d, maxkp %d)", n, cap, d->o.maxkp);
        return CLB_ECAPACITY;
    }
    if (n > 0) {
        if (kps_out) {
            CLB_CUDA(cudaMemcpyAsync(kps_out, d->kps.p, (size_t)n * sizeof(clb_keypoint), cudaMemcpyDeviceToHost, d->copy_stream));
            CLB_CUDA(cudaEventRecord(d->ev_count, d->copy_stream));
        }
        if (desc_out && !desc_direct) {
            CLB_CUDA(cudaMemcpyAsync(desc_out, d->desc.p, (size_t)n * 64, cudaMemcpyDeviceToHost, s));
            CLB_CUDA(cudaEventRecord(d->ev_done, s));
        }
        if (kps_out) { rc = spin_wait(d->ev_count); if (rc) return rc; }
    }
    return spin_wait(d->ev_done);
}

int clb_pyramid(clb_detector* d, const uint8_t* gray, int w, int h, int stride, uint8_t* levels_out) {
    int rc = check_frame(d, w, h, stride);
    if (rc) return rc;
    if (!gray || !levels_out) return CLB_EINVAL;
    CLB_CUDA(cudaSetDevice(d->device));
    cudaStream_t s = d->stream;
    LevelTable lt;
    make_table(d, w, h, lt);
    rc = ensure_geometry(d, lt, s);
    if (rc) return rc;
    CLB_CUDA(cudaMemcpy2DAsync(const_cast<uint8_t*>(lt.img[0]), lt.stride[0], gray, stride, w, h, cudaMemcpyHostToDevice, s));
    rc = resize_frame(d, lt, s, false);
    if (rc) return rc;
    size_t off = 0;
    for (int i = 0; i < lt.n; ++i) {
        if (lt.w[i] > 0 && lt.h[i] > 0)
            CLB_CUDA(cudaMemcpy2DAsync(levels_out + off, lt.w[i], lt.img[i], lt.stride[i], lt.w[i], lt.h[i], cudaMemcpyDeviceToHost, s));
        off += (size_t)lt.w[i] * lt.h[i];
    }
    CLB_CUDA(cudaStreamSynchronize(s));
    return CLB_OK;
}

int clb_fast_nms_angle(clb_detector* d, const uint8_t* level, int w, int h, int stride, int level_idx,
                       clb_keypoint* kps_out, int cap, int* n_out) {
    int rc = check_frame(d, w, h, stride);
    if (rc) return rc;
    if (!level || !n_out || level_idx < 0 || level_idx > 255 || cap < 0) return CLB_EINVAL;
    CLB_CUDA(cudaSetDevice(d->device));
    cudaStream_t s = d->stream;
    // a one-level table: the image is staged as "level 0" of the detector's pyramid buffer
    LevelTable lt; memset(&lt, 0, sizeof(lt));
    lt.n = 1; lt.w[0] = w; lt.h[0] = h; lt.stride[0] = align_up(w, 128); lt.f[0] = 1.f;
    lt.img[0] = d->pyr.as<uint8_t>();
    CLB_CUDA(cudaMemcpy2DAsync(d->pyr.p, lt.stride[0], level, stride, w, h, cudaMemcpyHostToDevice, s));
    int* d_total = d->misc.as<int>();
    LevelMaps one; memset(&one, 0, sizeof(one));
    rc = make_fast_maps(lt, one);
    if (rc) return rc;
    rc = launch_fast(lt, one, d->o.thresh, d->fast_quirk, d->fast_scratch.p, d->fast_scratch.cap, d->kps.as<clb_keypoint>(), d->o.maxkp, d_total, s, false, true);
    if (rc) return rc;
    CLB_CUDA(cudaMemcpyAsync(d->h_total, d_total, sizeof(int), cudaMemcpyDeviceToHost, s));
    CLB_CUDA(cudaStreamSynchronize(s));
    const int n = *d->h_total;
    *n_out = n;
    if (n > d->o.maxkp || n > cap) { snprintf(g_err, sizeof(g_err), "%d keypoints exceed capacity", n); return CLB_ECAPACITY; }
    if (n > 0 && kps_out) {
        CLB_CUDA(cudaMemcpyAsync(kps_out, d->kps.p, (size_t)n * sizeof(clb_keypoint), cudaMemcpyDeviceToHost, s));
        CLB_CUDA(cudaStreamSynchronize(s));
        for (int i = 0; i < n; ++i) kps_out[i].scale = (uint8_t)level_idx;
    }
    return CLB_OK;
}

int clb_describe(clb_detector* d, const uint8_t* const* levels, const int32_t* ws, const int32_t* hs,
                 const int32_t* strides, int nlevels, const clb_keypoint* kps, int n, uint64_t* desc_out) {
    if (!d || !levels || !ws || !hs || !strides || nlevels < 1 || nlevels > CLB_MAX_LEVELS || n < 0 || (n > 0 && (!kps || !desc_out))) return CLB_EINVAL;
    if (n > d->o.maxkp) { snprintf(g_err, sizeof(g_err), "%d keypoints exceed maxkp %d", n, d->o.maxkp); return CLB_ECAPACITY; }
    for (int i = 0; i < n; ++i) if (kps[i].scale >= nlevels || !levels[kps[i].scale]) { snprintf(g_err, sizeof(g_err), "keypoint %d names missing level %d", i, kps[i].scale); return CLB_EINVAL; }
    if (n == 0) return CLB_OK;
    CLB_CUDA(cudaSetDevice(d->device));
    cudaStream_t s = d->stream;
    LevelTable lt; memset(&lt, 0, sizeof(lt));
    lt.n = nlevels;
    size_t off = 0;
    for (int i = 0; i < nlevels; ++i) {
        lt.w[i] = ws[i]; lt.h[i] = hs[i]; lt.stride[i] = align_up(ws[i] > 0 ? ws[i] : 1, 128); lt.f[i] = 1.f;
        off += (size_t)lt.stride[i] * (hs[i] > 0 ? hs[i] : 1) + 256;
        off = (off + 511) & ~(size_t)511;
    }
    CLB_CUDA(d->stage.reserve(off));
    off = 0;
    for (int i = 0; i < nlevels; ++i) {
        lt.img[i] = d->stage.as<uint8_t>() + off;
        if (levels[i] && ws[i] > 0 && hs[i] > 0)
            CLB_CUDA(cudaMemcpy2DAsync(d->stage.as<uint8_t>() + off, lt.stride[i], levels[i], strides[i], ws[i], hs[i], cudaMemcpyHostToDevice, s));
        else lt.img[i] = nullptr;
        off += (size_t)lt.stride[i] * (hs[i] > 0 ? hs[i] : 1) + 256;
        off = (off + 511) & ~(size_t)511;
    }
    LevelTex tx;
    int trc = make_level_textures(lt, tx);
    if (trc) return trc;
    CLB_CUDA(cudaMemcpyAsync(d->kps.p, kps, (size_t)n * sizeof(clb_keypoint), cudaMemcpyHostToDevice, s));
    int rc = launch_describe(lt, d->kps.as<clb_keypoint>(), n, nullptr, d->o.maxkp, d->desc.as<uint64_t>(), s, &tx);
    cudaError_t e1 = rc ? cudaSuccess : cudaMemcpyAsync(desc_out, d->desc.p, (size_t)n * 64, cudaMemcpyDeviceToHost, s);
    cudaError_t e2 = cudaStreamSynchronize(s);
    destroy_level_textures(tx, nlevels);
    if (rc) return rc;
    CLB_CUDA(e1); CLB_CUDA(e2);
    return CLB_OK;
}

int clb_sincos(clb_detector* d, const float* angles, int n, float* sin_out, float* cos_out) {
    if (!d || n < 0 || (n > 0 && (!angles || !sin_out || !cos_out))) return CLB_EINVAL;
    if (n == 0) return CLB_OK;
    CLB_CUDA(cudaSetDevice(d->device));
    cudaStream_t s = d->stream;
    CLB_CUDA(d->stage.reserve((size_t)n * 12));
    float* a = d->stage.as<float>();
    CLB_CUDA(cudaMemcpyAsync(a, angles, (size_t)n * 4, cudaMemcpyHostToDevice, s));
    int rc = launch_sincos(a, n, a + n, a + 2 * (size_t)n, s);
    if (rc) return rc;
    CLB_CUDA(cudaMemcpyAsync(sin_out, a + n, (size_t)n * 4, cudaMemcpyDeviceToHost, s));
    CLB_CUDA(cudaMemcpyAsync(cos_out, a + 2 * (size_t)n, (size_t)n * 4, cudaMemcpyDeviceToHost, s));
    CLB_CUDA(cudaStreamSynchronize(s));
    return CLB_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------ frame stream
// `depth` independent detector contexts used round-robin; see include/coloc_b200.h.
struct clb_stream_slot {
    clb_detector* det;
    uint8_t* h_img; size_t h_img_cap;      // pinned staging for pageable inputs
    clb_keypoint* h_kps; uint64_t* h_desc; // pinned results (maxkp entries)
    uint64_t tag;
};
struct clb_stream {
    int depth, head, tail, inflight, device;
    clb_detector_opts o;
    clb_stream_slot* slots;
};

extern "C" {

void clb_stream_destroy(clb_stream* s) {
    if (!s) return;
    cudaSetDevice(s->device);
    for (int i = 0; s->slots && i < s->depth; ++i) {
        if (s->slots[i].det) { cudaStreamSynchronize(s->slots[i].det->stream); clb_detector_destroy(s->slots[i].det); }
        if (s->slots[i].h_img) cudaFreeHost(s->slots[i].h_img);
        if (s->slots[i].h_kps) cudaFreeHost(s->slots[i].h_kps);
        if (s->slots[i].h_desc) cudaFreeHost(s->slots[i].h_desc);
    }
    delete[] s->slots;
    delete s;
}

int clb_stream_create(const clb_detector_opts* opts, int device, int depth, clb_stream** out) {
    if (!opts || !out || depth < 1 || depth > 16) { snprintf(g_err, sizeof(g_err), "bad stream options (depth 1..16)"); return CLB_EINVAL; }
    clb_stream* s = new (std::nothrow) clb_stream();
    if (!s) return CLB_ENOMEM;
    s->depth = depth; s->head = s->tail = s->inflight = 0; s->device = device; s->o = *opts;
    s->slots = new (std::nothrow) clb_stream_slot[depth]();
    if (!s->slots) { delete s; return CLB_ENOMEM; }
    for (int i = 0; i < depth; ++i) {
        int rc = clb_detector_create(opts, device, &s->slots[i].det);
        cudaError_t e = cudaSuccess;
        if (!rc) e = cudaMallocHost(&s->slots[i].h_kps, (size_t)opts->maxkp * sizeof(clb_keypoint));
        if (!rc && e == cudaSuccess) e = cudaMallocHost(&s->slots[i].h_desc, (size_t)opts->maxkp * 64);
        if (!rc && e != cudaSuccess) rc = set_cuda_error(e, "stream pinned allocation", __FILE__, __LINE__);
        if (rc) { clb_stream_destroy(s); return rc; }
    }
    *out = s;
    return CLB_OK;
}

int clb_stream_in_flight(const clb_stream* s) { return s ? s->inflight : 0; }

int clb_stream_set_param(clb_stream* s, const char* name, int value) {
    if (!s) return CLB_EINVAL;
    for (int i = 0; i < s->depth; ++i) { const int rc = clb_detector_set_param(s->slots[i].det, name, value); if (rc) return rc; }
    return CLB_OK;
}

int clb_stream_submit(clb_stream* s, const uint8_t* gray, int w, int h, int stride, uint64_t tag) {
    if (!s || !gray) return CLB_EINVAL;
    if (s->inflight == s->depth) { snprintf(g_err, sizeof(g_err), "%d frames already in flight", s->depth); return CLB_ECAPACITY; }
    clb_stream_slot& sl = s->slots[s->tail];
    clb_detector* d = sl.det;
    int rc = check_frame(d, w, h, stride);
    if (rc) return rc;
    CLB_CUDA(cudaSetDevice(d->device));
    cudaStream_t st = d->stream;
    LevelTable lt;
    make_table(d, w, h, lt);
    // pinned input: DMA straight from the caller's buffer; pageable input: stage it so the copy is truly asynchronous
    cudaPointerAttributes pa;
    const bool pinned = cudaPointerGetAttributes(&pa, gray) == cudaSuccess && pa.type == cudaMemoryTypeHost;
    cudaGetLastError();
    const uint8_t* src = gray; int src_stride = stride;
    if (!pinned) {
        const size_t need = (size_t)w * h;
        if (need > sl.h_img_cap) {
            if (sl.h_img) cudaFreeHost(sl.h_img);
            sl.h_img = nullptr; sl.h_img_cap = 0;
            CLB_CUDA(cudaMallocHost(&sl.h_img, need));
            sl.h_img_cap = need;
        }
        for (int y = 0; y < h; ++y) memcpy(sl.h_img + (size_t)y * w, gray + (size_t)y * stride, (size_t)w);
        src = sl.h_img; src_stride = w;
    }
    rc = ensure_geometry(d, lt, st);
    if (rc) return rc;
    CLB_CUDA(cudaMemcpy2DAsync(const_cast<uint8_t*>(lt.img[0]), lt.stride[0], src, src_stride, w, h, cudaMemcpyHostToDevice, st));
    rc = run_pipeline(d, lt, st);
    if (rc) return rc;
    CLB_CUDA(cudaMemcpyAsync(d->h_total, d->misc.p, sizeof(int), cudaMemcpyDeviceToHost, st));
    sl.tag = tag;
    s->tail = (s->tail + 1) % s->depth;
    ++s->inflight;
    return CLB_OK;
}

int clb_stream_next(clb_stream* s, uint64_t* tag_out, const clb_keypoint** kps_out, const uint64_t** desc_out,
                    const clb_keypoint** d_kps_out, const uint64_t** d_desc_out, int* n_out) {
    if (!s || !n_out) return CLB_EINVAL;
    if (s->inflight == 0) { snprintf(g_err, sizeof(g_err), "no frame in flight"); return CLB_EINVAL; }
    clb_stream_slot& sl = s->slots[s->head];
    clb_detector* d = sl.det;
    s->head = (s->head + 1) % s->depth;
    --s->inflight;
    CLB_CUDA(cudaSetDevice(d->device));
    cudaStream_t st = d->stream;
    CLB_CUDA(cudaStreamSynchronize(st));
    const int n = *d->h_total;
    *n_out = n;
    if (tag_out) *tag_out = sl.tag;
    if (n > d->o.maxkp) { snprintf(g_err, sizeof(g_err), "%d keypoints exceed maxkp %d", n, d->o.maxkp); return CLB_ECAPACITY; }
    if (n > 0 && (kps_out || desc_out)) {
        if (kps_out) CLB_CUDA(cudaMemcpyAsync(sl.h_kps, d->kps.p, (size_t)n * sizeof(clb_keypoint), cudaMemcpyDeviceToHost, st));
        if (desc_out) CLB_CUDA(cudaMemcpyAsync(sl.h_desc, d->desc.p, (size_t)n * 64, cudaMemcpyDeviceToHost, st));
        CLB_CUDA(cudaStreamSynchronize(st));
    }
    if (kps_out) *kps_out = sl.h_kps;
    if (desc_out) *desc_out = sl.h_desc;
    if (d_kps_out) *d_kps_out = d->kps.as<clb_keypoint>();
    if (d_desc_out) *d_desc_out = d->desc.as<uint64_t>();
    return CLB_OK;
}

}  // extern "C"
