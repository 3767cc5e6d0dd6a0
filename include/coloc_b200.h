/* coloc_b200.h -- C ABI of the B200-native CoLoC front end (libcoloc_b200.so).
 *
 * Drop-in boundary for the reference's GPU detector / matcher policies.  Every entry point names
 * the reference interface it replaces (paths relative to the reference repo root).  Plain
 * pointers and sizes only; no C++/torch types.  All functions return 0 (CLB_OK) or a CLB_E*
 * code; clb_strerror() explains it and clb_last_error() returns the CUDA detail of the last
 * failure on the calling thread.
 *
 * Host-pointer entry points (suffix-less) have the reference's blocking semantics: they copy
 * inputs H2D, run, copy results D2H and return after a stream sync (GPUDetector.hpp:290,
 * CUDAK2NN.cu:80).  `_dev` entry points take DEVICE pointers plus a cudaStream_t (passed as
 * void*) and only enqueue work; a NULL stream means the object's own (non-blocking) stream, pass
 * cudaStreamLegacy / cudaStreamPerThread explicitly to use a default stream.
 *
 * Concurrency contract: an object (detector, matcher, fmap, sharded, stream) owns ONE set of device scratch buffers, so one
 * call per object may be in flight at a time -- consecutive `_dev` calls on the SAME stream are ordered by it; before using
 * an object from a different stream, synchronise the previous one.  (The map copy of clb_set_map_dev is the exception: later
 * map sweeps on any stream wait for it by event.)  Objects are not thread-safe; different objects are independent.
 */
#ifndef COLOC_B200_H
#define COLOC_B200_H
#include <stdint.h>
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

#define CLB_OK            0
#define CLB_EINVAL        1   /* bad argument */
#define CLB_ECUDA         2   /* a CUDA runtime call failed (see clb_last_error) */
#define CLB_ECAPACITY     3   /* caller buffer / maxkp too small; *n_out holds the needed count */
#define CLB_ENOMEM        4
#define CLB_EUNSUPPORTED  5
#define CLB_ENOTFOUND     6   /* a feature-set index that is not in the clb_fmap */
#define CLB_EIO           7   /* a .feat / .desc file could not be opened, read or written */

#define CLB_MAX_LEVELS    16
#define CLB_NO_MATCH      (-1)
#define CLB_BEST_NONE     100000   /* CUDAK2NN.cu:54 initial best   */
#define CLB_SECOND_NONE   200000   /* CUDAK2NN.cu:54 initial second */

/* include/coloc/Keypoint.h:155-164 -- same 20-byte layout, copied raw by GPUDetector.hpp:281 */
typedef struct clb_keypoint {
    int32_t x, y;            /* level coordinates */
    uint8_t score; uint8_t _pad0[3];
    float   angle;           /* (-pi, pi] */
    uint8_t scale; uint8_t _pad1[3];   /* pyramid level */
} clb_keypoint;

/* include/coloc/colocData.hpp:29-36 DetectorOptions, field for field */
typedef struct clb_detector_opts {
    float    scale_factor;   /* 1.2 in coloc_node.cpp:78 */
    int32_t  scale_levels;   /* 8 */
    int32_t  width, height;  /* maximum frame size (buffers are sized for it) */
    int32_t  maxkp;          /* capacity of the keypoint / descriptor buffers */
    int32_t  thresh;         /* FAST threshold, 40 in coloc_node.cpp:81 */
} clb_detector_opts;

/* include/coloc/colocData.hpp:38-42 MatcherOptions */
typedef struct clb_matcher_opts {
    float    distRatio;      /* used by the ratio-test mode only (CPUMatcher.hpp:58-89) */
    int32_t  thresh;         /* absolute margin for map matching, 60 in coloc_node.cpp:85 */
    int32_t  maxkp;          /* initial capacity in descriptors (grows on demand) */
} clb_matcher_opts;

typedef struct clb_detector clb_detector;
typedef struct clb_matcher  clb_matcher;

/* per-query partial 2-NN state exchanged between DB shards (multi-GPU) */
typedef struct clb_partial {
    int32_t best_d;     /* CLB_BEST_NONE when the shard was empty */
    int32_t best_i;     /* GLOBAL row index of the first row attaining best_d, -1 if none */
    int32_t second_d;   /* second-smallest distance in the shard (== best_d on ties) */
    int32_t _pad;
} clb_partial;

const char* clb_strerror(int code);
const char* clb_last_error(void);
int  clb_version(void);
int  clb_device_count(int* n_out);

/* ---------------------------------------------------------------- detector --------------- */
/* GPUDetector::GPUDetector(DetectorOptions)  include/coloc/GPUDetector.hpp:70-138 */
int  clb_detector_create(const clb_detector_opts* opts, int device, clb_detector** out);
/* GPUDetector::freeGPUMemory  GPUDetector.hpp:144-155 */
void clb_detector_destroy(clb_detector* d);

/* per-detector parameters.  "fast_quirk" (default 1): 1 = FAST scans exactly the columns the COMPILED reference scans --
 * on levels whose width has (w - 38) % 16 == 0 (214 = level 6 of a 640-wide frame, 1286 = level 6 of a 3840-wide one) the
 * `1 << 32` at KFAST.h:243 evaluates to 1 on x86 and the last 32 columns of a row are dropped when the block walk of
 * KFAST.h:451-459 / :256-262 lands on column w - 35 (data dependent, per row); 0 = the intended semantics, every column
 * 3..w-4 (a deliberate deviation from the reference: different keypoints on those levels). */
int  clb_detector_set_param(clb_detector* d, const char* name, int value);
int  clb_detector_get_param(const clb_detector* d, const char* name, int* value_out);

/* level geometry of GPUDetector.hpp:107-114 for a w x h frame; arrays of scale_levels entries */
int  clb_detector_level_dims(const clb_detector* d, int w, int h, int32_t* ws, int32_t* hs, float* fs);

/* GPUDetector::detectAndDescribe  GPUDetector.hpp:216-291 (everything after cv::imread):
 * gray (HOST, 8-bit, `stride` bytes per row) -> keypoints in level-major / row-major order
 * (:262-276) and 8 x uint64 descriptors per keypoint (:282-288).  If the frame has more than
 * `cap` (or maxkp) keypoints returns CLB_ECAPACITY with *n_out = needed. */
int  clb_detect_describe(clb_detector* d, const uint8_t* gray, int w, int h, int stride,
                         clb_keypoint* kps_out, uint64_t* desc_out, int cap, int* n_out);
/* page-locked host memory for callers that want the copies of the calls above to overlap the kernels (any cudaMallocHost /
 * cudaHostRegister memory works the same).  With a page-locked desc_out, clb_detect_describe lets the describe kernel write
 * the descriptors straight into it while it runs (no D2H copy behind the kernel); a page-locked gray is DMA-ed from directly. */
int  clb_host_alloc(size_t bytes, void** out);
void clb_host_free(void* p);
/* same with a DEVICE image; results stay on the device (d_kps / d_desc owned by the detector,
 * valid until the next call); *n_out is read back (one 4-byte D2H + stream sync). */
int  clb_detect_describe_dev(clb_detector* d, const uint8_t* d_gray, int w, int h, int stride,
                             const clb_keypoint** d_kps, const uint64_t** d_desc, int* n_out, void* stream);

/* stage-level entry points (parity tests) */
/* CUDALERP x (levels-1)  src/CUDALERP.cu:153-183 as driven by GPUDetector.hpp:249-255.
 * levels_out (HOST): levels packed back to back, level i is ws[i] x hs[i] with stride ws[i];
 * level 0 is a copy of the input. */
int  clb_pyramid(clb_detector* d, const uint8_t* gray, int w, int h, int stride, uint8_t* levels_out);
/* KFAST<true,true> + featureAngle on one level image  include/coloc/KFAST.h:503,
 * FeatureAngle.h:197; caller GPUDetector.hpp:268-274.  level_idx is stored in kp.scale. */
int  clb_fast_nms_angle(clb_detector* d, const uint8_t* level, int w, int h, int stride, int level_idx,
                        clb_keypoint* kps_out, int cap, int* n_out);
/* CLATCH  src/CLATCH.cu:153-193 on caller-supplied level images (HOST pointers). */
int  clb_describe(clb_detector* d, const uint8_t* const* levels, const int32_t* ws, const int32_t* hs,
                  const int32_t* strides, int nlevels, const clb_keypoint* kps, int n, uint64_t* desc_out);
/* device sin/cos used by the describe kernel (CLATCH.cu:161), for ulp-level parity tests */
int  clb_sincos(clb_detector* d, const float* angles, int n, float* sin_out, float* cos_out);

/* ---------------------------------------------------------------- frame stream ----------- */
/* Throughput mode for a stream of frames (BASELINE config 5).  The reference calls
 * detectAndDescribe once per frame and blocks (GPUDetector.hpp:290); a stream keeps `depth` frames in
 * flight, each on its own CUDA stream with its own pyramid / keypoint / descriptor buffers and pinned
 * result staging, so the H2D copy of frame k+1 and the D2H copy of frame k-1 overlap the kernels of
 * frame k.  Results are bit-identical to clb_detect_describe and come back in submission order. */
typedef struct clb_stream clb_stream;
int  clb_stream_create(const clb_detector_opts* opts, int device, int depth, clb_stream** out);
void clb_stream_destroy(clb_stream* s);
/* enqueue one frame (HOST pointer; pinned memory is copied from directly, pageable memory is staged).
 * Returns CLB_ECAPACITY when `depth` frames are already in flight (call clb_stream_next first). */
int  clb_stream_submit(clb_stream* s, const uint8_t* gray, int w, int h, int stride, uint64_t tag);
/* wait for the OLDEST frame in flight.  kps_out / desc_out point into pinned HOST memory, d_kps_out /
 * d_desc_out at the DEVICE copies (any of them may be NULL); all stay valid until `depth` - 1 further
 * frames have been submitted.  CLB_EINVAL when nothing is in flight. */
int  clb_stream_next(clb_stream* s, uint64_t* tag_out, const clb_keypoint** kps_out, const uint64_t** desc_out,
                     const clb_keypoint** d_kps_out, const uint64_t** d_desc_out, int* n_out);
int  clb_stream_in_flight(const clb_stream* s);
int  clb_stream_set_param(clb_stream* s, const char* name, int value);   /* clb_detector_set_param on every slot */

/* ---------------------------------------------------------------- matcher ---------------- */
/* GPUMatcher::GPUMatcher(MatcherOptions)  include/coloc/GPUMatcher.hpp:70-95 */
int  clb_matcher_create(const clb_matcher_opts* opts, int device, clb_matcher** out);
/* GPUMatcher::freeGPUMemory  GPUMatcher.hpp:102-108 */
void clb_matcher_destroy(clb_matcher* m);

/* GPUMatcher::computeMatches(void* hQ, void* hT, int nQ, int nT, uint8_t threshold)
 * GPUMatcher.hpp:180-226 -> CUDAK2NN  src/CUDAK2NN.cu:42-81.  HOST pointers.
 * match_out[q] = first T row attaining the minimum distance if second - best > threshold,
 * else -1.  best_d_out / second_d_out optional (may be NULL). */
int  clb_match_2nn(clb_matcher* m, const uint64_t* Q, int nQ, const uint64_t* T, int nT, int threshold,
                   int32_t* match_out, int32_t* best_d_out, int32_t* second_d_out);
/* same on DEVICE pointers (16-byte aligned), enqueue only */
int  clb_match_2nn_dev(clb_matcher* m, const uint64_t* dQ, int nQ, const uint64_t* dT, int nT, int threshold,
                       int32_t* d_match_out, int32_t* d_best_out, int32_t* d_second_out, void* stream);

/* Allocation-free 2-NN.  No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same result as clb_match_2nn_dev
 * (match[], best, second; margin test `second - best > threshold`), computed by the POPC sweep + chunk merge with the per-chunk
 * scratch in a 4 MiB file-scope device array that every GPU loading the library holds until exit -- no matcher object, no
 * cudaMalloc / cudaFree or other allocating call (the first call on a device only looks the array's address up).  Limits:
 * nQ <= 32 768 and nT <= 2^26 rows; larger sizes return CLB_EINVAL (clb_last_error names the limit).  DEVICE pointers, 16-byte
 * aligned; best / second outputs may be NULL; enqueue only, on `stream` (NULL = the legacy default stream).  Every call shares
 * the one array (it keeps the last call's contents; every slot a call reads it has written itself, nothing needs clearing):
 * calls that overlap on the GPU -- different streams, different threads -- race; order them on one stream or by events. */
int  clb_match_2nn_noalloc(const uint64_t* dQ, int nQ, const uint64_t* dT, int nT, int threshold,
                           int32_t* d_match_out, int32_t* d_best_out, int32_t* d_second_out, void* stream);

/* GPUMatcher::setMapData  GPUMatcher.hpp:110-117 (HOST pointer; H2D once per map) */
int  clb_set_map(clb_matcher* m, const uint64_t* M, int nM);
/* GPUMatcher::setQueryImage + matchFeaturesWithMap  GPUMatcher.hpp:130-141,252-271
 * threshold < 0 -> opts.thresh.  match_out[q] = MAP row or -1 (the caller flips it into
 * IndMatch(map, frame) as :263-266 does). */
int  clb_match_map(clb_matcher* m, const uint64_t* Q, int nQ, int threshold,
                   int32_t* match_out, int32_t* best_d_out, int32_t* second_d_out);

/* device-resident variants: the map / DB shard is copied device-to-device once (and expanded once for
 * the tensor-core path), every later call only brings queries.  dM / dQ are DEVICE pointers, 16-byte aligned; enqueue only.
 * clb_match_map_partial_dev is the DB-row-sharding step on the resident rows (global indices start at t_base). */
int  clb_set_map_dev(clb_matcher* m, const uint64_t* dM, int nM, void* stream);
int  clb_match_map_dev(clb_matcher* m, const uint64_t* dQ, int nQ, int threshold,
                       int32_t* d_match_out, int32_t* d_best_out, int32_t* d_second_out, void* stream);
int  clb_match_map_partial_dev(clb_matcher* m, const uint64_t* dQ, int nQ, int64_t t_base,
                               clb_partial* d_partial_out, void* stream);

/* ratio-test mode mirroring the CPU build (CPUMatcher.hpp:67-76 -> matching_filters.hpp:39-60):
 * keep iff best < ratio * second (float compare).  HOST pointers. */
int  clb_match_2nn_ratio(clb_matcher* m, const uint64_t* Q, int nQ, const uint64_t* T, int nT, float ratio,
                         int32_t* match_out, int32_t* best_d_out, int32_t* second_d_out);

/* DB-row sharding: local 2-NN of every query against DEVICE rows dT[0..nT) whose
 * global indices start at t_base; d_partial_out[nQ] (device).  Enqueue only. */
int  clb_match_partial_dev(clb_matcher* m, const uint64_t* dQ, int nQ, const uint64_t* dT, int nT,
                           int64_t t_base, clb_partial* d_partial_out, void* stream);
/* As clb_match_partial_dev with the DB rows in HOST memory: they are copied in slices (the first of "match_slice_rows"
 * rows, default 524 288, each next one 4x larger) on a second stream while earlier slices are expanded and swept, so the H2D copy hides behind the
 * sweep; pinned host memory overlaps, pageable memory works but copies synchronously.  Queries and output on the DEVICE;
 * asynchronous on `stream`: T_host must stay valid until that stream is synchronised.  clb_match_2nn (all host) takes the
 * same path for DBs larger than one slice.  Bit-identical to the unsliced scan (slices fold in ascending row order). */
int  clb_match_partial_hostdb(clb_matcher* m, const uint64_t* dQ, int nQ, const uint64_t* T_host, int nT,
                              int64_t t_base, clb_partial* d_partial_out, void* stream);
/* merge `nparts` partial arrays laid out [part][nQ] (ascending t_base order) and apply the margin
 * test: first-min-wins across parts, global second = 2nd smallest of the union.  DEVICE pointers. */
int  clb_merge_partials_dev(clb_matcher* m, const clb_partial* d_parts, int nparts, int nQ, int threshold,
                            int32_t* d_match_out, int32_t* d_best_out, int32_t* d_second_out, void* stream);

/* ---------------------------------------------------------------- sharded map matching --- */
/* GPUMatcher::setMapData (GPUMatcher.hpp:110-117) + matchSceneWithMap / matchFeaturesWithMap
 * (:174-178, :252-271) with the map rows cut into contiguous shards, one per GPU.  Shard r of n holds rows
 * [r*base + min(r, rem), ...) with base = nM / n, rem = nM % n (clb_sharded_range).  Every GPU sweeps its shard for all
 * queries, the 16-byte-per-query clb_partial arrays are exchanged and merged in ascending shard order: bit-identical to the
 * single-GPU scan (first row attaining the minimum wins across shard borders).
 *   clb_sharded_create       ONE process drives `ndev` devices (devices == NULL: 0..ndev-1); partials reach the first device
 *                            by peer copies over NVLink.
 *   clb_sharded_create_rank  one process per GPU (torchrun / MPI): rank 0 calls clb_sharded_unique_id, the caller's own
 *                            transport broadcasts the 128 bytes, every rank creates its handle; the exchange is one
 *                            ncclAllGather per call.  NCCL is bound at run time (libnccl.so.2); CLB_EUNSUPPORTED without it. */
typedef struct clb_sharded clb_sharded;
int  clb_sharded_create(const clb_matcher_opts* opts, const int* devices, int ndev, clb_sharded** out);
int  clb_sharded_unique_id(void* id128_out);
int  clb_sharded_create_rank(const clb_matcher_opts* opts, int device, int rank, int nranks, const void* id128, clb_sharded** out);
void clb_sharded_destroy(clb_sharded* s);
int  clb_sharded_shards(const clb_sharded* s);
int  clb_sharded_range(const clb_sharded* s, long long n_rows, int shard, long long* lo_out, long long* hi_out);
/* setMapData: M = the WHOLE map (HOST); every shard uploads its own rows and keeps them resident (expanded once). */
int  clb_sharded_set_map(clb_sharded* s, const uint64_t* M, long long nM);
/* the map STAYS in host memory (pinned for overlap; kept alive by the caller): each match call streams every shard's rows
 * in slices behind its sweep (clb_match_partial_hostdb). */
int  clb_sharded_set_map_host(clb_sharded* s, const uint64_t* M, long long nM);
/* rank mode, this rank's rows only: global rows [row_lo, row_lo + rows) of a map of n_total rows.
 * where: 0 HOST rows -> resident, 1 DEVICE rows -> resident (copied), 2 HOST rows streamed per call. */
int  clb_sharded_set_map_shard(clb_sharded* s, const uint64_t* rows_ptr, long long row_lo, int rows, long long n_total, int where);
/* matchFeaturesWithMap over all shards: HOST queries in, HOST match[] (global MAP row or -1) out, blocking.  threshold < 0 ->
 * opts.thresh.  Rank mode: collective, every rank passes the same queries and gets the same result. */
int  clb_sharded_match_map(clb_sharded* s, const uint64_t* Q, int nQ, int threshold,
                           int32_t* match_out, int32_t* best_d_out, int32_t* second_d_out);
/* rank mode, DEVICE queries / outputs on this rank's device, asynchronous on `stream`: sweep -> ncclAllGather -> merge. */
int  clb_sharded_match_map_dev(clb_sharded* s, const uint64_t* dQ, int nQ, int threshold,
                               int32_t* d_match_out, int32_t* d_best_out, int32_t* d_second_out, void* stream);

/* ---------------------------------------------------------------- device FeatureMap ------ */
/* SURVEY 8(f)-1/2.  The reference keeps each frame's descriptors in host Regions (coloc::FeatureMap = idx ->
 * unique_ptr<AKAZE_Binary_Regions>, GPUDetector.hpp:167-182), uploads BOTH sets again for every pair
 * (GPUMatcher.hpp:165-172 -> :180-226) and the gathered landmark rows for every map update (colocData.hpp:89-121 ->
 * GPUMatcher.hpp:110-117).  A clb_fmap keeps every set on the device once (packed rows, optional keypoints and the
 * tensor-core operand image); pair lists and the map database are then built from device memory. */
typedef struct clb_fmap clb_fmap;
int  clb_fmap_create(int device, clb_fmap** out);
void clb_fmap_destroy(clb_fmap* f);
/* regions[idx] = (kps, desc): replaces an existing entry (GPUDetector.hpp:182 `regions[idx] = std::move(...)`).
 * kps may be NULL.  Host pointers, blocking. */
int  clb_fmap_put(clb_fmap* f, uint32_t idx, const clb_keypoint* kps, const uint64_t* desc, int n);
/* same from DEVICE buffers (the outputs of clb_detect_describe_dev / clb_stream_next), asynchronous on `stream`;
 * later calls that use the entry must be ordered after it (same stream, or synchronise). */
int  clb_fmap_put_dev(clb_fmap* f, uint32_t idx, const clb_keypoint* d_kps, const uint64_t* d_desc, int n, void* stream);
int  clb_fmap_count(const clb_fmap* f, uint32_t idx, int* n_out);          /* CLB_ENOTFOUND when absent */
int  clb_fmap_size(const clb_fmap* f);                                     /* number of entries */
int  clb_fmap_erase(clb_fmap* f, uint32_t idx);
/* copy an entry back to the host (kps_out / desc_out may be NULL); CLB_ECAPACITY with *n_out = rows when cap is short */
int  clb_fmap_get(clb_fmap* f, uint32_t idx, clb_keypoint* kps_out, uint64_t* desc_out, int cap, int* n_out);
/* device pointers of an entry (valid until the entry is replaced or erased) */
int  clb_fmap_desc_dev(const clb_fmap* f, uint32_t idx, const uint64_t** d_desc_out, const clb_keypoint** d_kps_out, int* n_out);

/* One pair of a pair list: query rows [q_lo, q_hi) of set query_idx against all rows of set train_idx (q_hi < 0 = to
 * the end of the set).  Row slices are what SURVEY 8(e) shards across GPUs. */
typedef struct clb_pair {
    uint32_t query_idx, train_idx;
    int32_t  q_lo, q_hi;
} clb_pair;
/* GPUMatcher::computeMatches(FeatureMap&, PairWiseMatches&) (GPUMatcher.hpp:143-155) for a whole pair list in ONE sweep
 * launch + ONE merge launch.  Outputs are the per-pair match[] arrays of clb_match_2nn concatenated in list order
 * (clb_match_pairs_rows gives the total length); match values index rows of the pair's training set.  ratio > 0
 * applies the NNdistanceRatio test (clb_match_2nn_ratio) instead of the margin test; threshold < 0 = the matcher's
 * own.  best / second outputs may be NULL.  Always the tensor-core kernel (match_tc_ver 4), any size. */
int  clb_match_pairs_rows(clb_matcher* m, clb_fmap* f, const clb_pair* pairs, int npairs, long long* rows_out);
int  clb_match_pairs(clb_matcher* m, clb_fmap* f, const clb_pair* pairs, int npairs, int threshold, float ratio,
                     int32_t* match_out, int32_t* best_d_out, int32_t* second_d_out);
/* DEVICE outputs, asynchronous on `stream` (NULL = the matcher's stream); one call in flight per matcher */
int  clb_match_pairs_dev(clb_matcher* m, clb_fmap* f, const clb_pair* pairs, int npairs, int threshold, float ratio,
                         int32_t* d_match_out, int32_t* d_best_out, int32_t* d_second_out, void* stream);

/* One landmark observation: feature `feat` of set `idx` (openMVG Observation::id_feat of view idx). */
typedef struct clb_obs {
    uint32_t idx, feat;
} clb_obs;
/* colocData::setupMapDatabase (colocData.hpp:89-121) + GPUMatcher::setMapData (GPUMatcher.hpp:110-117) without the host
 * round trip: map row k = descriptor `feat` of set `idx` of the k-th observation whose feat != 0xFFFFFFFF
 * (UndefinedIndexT entries are skipped, as the reference does); *n_map_out = rows kept.  The rows become the matcher's
 * resident map (clb_match_map*).  CLB_ENOTFOUND / CLB_EINVAL for an unknown set / a feature outside its set. */
int  clb_set_map_from_fmap(clb_matcher* m, clb_fmap* f, const clb_obs* obs, int n, int* n_map_out, void* stream);

/* ---------------------------------------------------------------- host side of the path -- */
/* Region packing of GPUDetector::detectFeaturesFile (GPUDetector.hpp:167-182): feats_out[4k..4k+3] =
 * SIOPointFeature{1.2^s * x, 1.2^s * y, 7 * 1.2^s, angle} of keypoint k (1.2f hard-coded there, float pow).  HOST. */
int  clb_pack_regions(const clb_keypoint* kps, int n, float* feats_out);
/* Binary_Regions<SIOPointFeature,64>::Save / Load (binary_regions.hpp:51-66): `.feat` = one "x y scale orientation" text line
 * per region (feature.hpp:168-182, default ostream float format), `.desc` = size_t count + count x 64 raw bytes
 * (descriptor.hpp:165-210).  Either path may be NULL to skip that file.  HOST. */
int  clb_regions_save(const char* feat_path, const char* desc_path, const float* feats, const uint64_t* desc, int n);
int  clb_regions_count(const char* desc_path, long long* n_out);
/* CLB_ECAPACITY with *n_out = rows in the file when cap is short; CLB_EINVAL when the two files disagree in length */
int  clb_regions_load(const char* feat_path, const char* desc_path, float* feats_out, uint64_t* desc_out, int cap, int* n_out);
/* Tail of matching::DistanceRatioMatch (regions_matcher.hpp:166-181) on a match[] array (match[q] = database row or -1):
 * (database idx, query idx) pairs -> IndMatch::getDeduplicated -> IndMatchDecorator<float>::getDeduplicated over the
 * (x, y) of both sides; feats* = positions with `stride*` floats per region (x, y first).  Output order = the reference's
 * (iteration order of its std::set).  HOST, no GPU.  CLB_ECAPACITY with *n_out = needed pairs. */
int  clb_dedup_matches(const int32_t* match, int nQ, const float* featsI, int strideI, const float* featsJ, int strideJ,
                       uint32_t* pairs_out, int cap, int* n_out);
/* matching::DistanceRatioMatch(ratio, BRUTE_FORCE_HAMMING, regions_I, regions_J, out) as the CPU build calls it
 * (CPUMatcher.hpp:58-89): regions_I is the database, regions_J the queries; GPU 2-NN + ratio test, host dedup.
 * feats = 4 floats per region (clb_pack_regions layout).  pairs_out = (i_, j_) pairs. */
int  clb_distance_ratio_match(clb_matcher* m, const uint64_t* descI, const float* featsI, int nI,
                              const uint64_t* descJ, const float* featsJ, int nJ, float ratio,
                              uint32_t* pairs_out, int cap, int* n_out);

/* matching::Filter (matching_filters.hpp:118-213) over brute-force Hamming 2-NN lists in BOTH directions:
 * filter 1 = MATCHFILTER_SYMMETRIC (keep i iff the nearest row of B to A[i] has A[i] as ITS nearest row, SymmetricMatches :76-94),
 * 2 = MATCHFILTER_NNDISTANCERATIO (:39-60), 3 = both.  One GPU sweep launch covers both directions.  pairs_out = (i_, j_) =
 * (row of A, row of B), ascending, as Filter() leaves them -- including its loops that stop one short of the list's end (the last
 * surviving index is dropped).  Nearest rows resolve distance ties to the lowest index (CUDAK2NN.cu:67-72); openMVG's
 * partial_sort leaves ties unspecified.  HOST pointers.  CLB_ECAPACITY with *n_out = needed pairs. */
int  clb_match_filtered(clb_matcher* m, const uint64_t* A, int nA, const uint64_t* B, int nB, int filter, float ratio,
                        uint32_t* pairs_out, int cap, int* n_out);

/* ---------------------------------------------------------------- geometric pre-filter --- */
/* SURVEY 8(f)-4.  The Errors() pass of the AC-RANSAC kernel adaptors (robust_estimator_ACRansacKernelAdaptator.hpp:93-99,
 * 244-252) that RobustMatcher::filterFundamental / filterEssential / filterHomography (RobustMatcher.hpp:128-239) run per
 * hypothesis, for a batch of hypotheses.  models: n_models x 9 doubles, ROW-major 3x3; x1, x2: n x 2 doubles (the matched
 * positions of view 1 / view 2); residuals_out: n_models x n doubles or NULL; inliers_out[h] = #{residual < threshold} or NULL.
 * kind: 0 EpipolarDistanceError, 1 SymmetricEpipolarDistanceError, 2 SampsonError (solver_fundamental_kernel.cpp:156-192),
 * 3 homography AsymmetricError (solver_homography_kernel.hpp:59-63).  fp64; agrees with the reference's functors to 1e-9
 * relative (FMA contraction differs).  Hypothesis solvers and the a-contrario selection stay on the host. */
#define CLB_RES_EPIPOLAR            0
#define CLB_RES_SYMMETRIC_EPIPOLAR  1
#define CLB_RES_SAMPSON             2
#define CLB_RES_HOMOGRAPHY          3
int  clb_model_residuals(clb_matcher* m, int kind, const double* models, int n_models, const double* x1, const double* x2, int n,
                         double threshold, double* residuals_out, int32_t* inliers_out);
int  clb_model_residuals_dev(clb_matcher* m, int kind, const double* d_models, int n_models, const double* d_x1, const double* d_x2,
                             int n, double threshold, double* d_residuals_out, int32_t* d_inliers_out, void* stream);

/* ---------------------------------------------------------------- measurement ------------ */
/* pipe micro-benchmarks that back the match roofline: dependent-free POPC / LOP3
 * streams and a mixed stream; results in ops per clock per SM and the SM clock seen.
 * out[0]=popc/clk/SM out[1]=lop3/clk/SM out[2]=mixed popc/clk/SM out[3]=mixed lop3/clk/SM
 * out[4]=SM MHz during the run  out[5]=SM count
 * out[6]=chip G POPC32/s (wall time)  out[7]=chip G LOP3/s (wall time) */
int  clb_microbench_pipes(int device, double* out8);
/* measured tensor-pipe peak behind match_tc4_kernel's roofline: a bare stream of the sweep's own instruction (tcgen05.mma
 * kind::mxf4.block_scale M128 N128 K64, A in TMEM, B in shared memory, +-1 operands) on every SM, ~10 ms.
 * out[0]=cycles per MMA  out[1]=MAC/clk/SM  out[2]=SM MHz while it ran  out[3]=chip TFLOP/s (CUDA events)  out[4]=SMs  out[5]=MMAs per SM */
int  clb_microbench_mma(int device, double* out6);
/* number of kernels this library launched on the calling process so far (bench's gpu_launches) */
int64_t clb_launch_count(void);
/* time of the last match kernel (ms, CUDA events on its own stream); 0 when timing is off */
int  clb_set_timing(int on);   /* 0 off, 1 per-kernel event pairs (clb_last_kernel_ms), 2 also named marks for clb_timeline */
/* event timeline of the calls since the last clb_timeline (timing level 2): "name:ms name:ms ..." offsets from the first mark */
int  clb_timeline(char* buf, int cap);
/* Process-wide planning knobs (atomics; set them before the calls they should affect): "match_tc" (-1 auto: tensor-core
 * path from 2^24 comparisons up, 0 never, 1 always), "match_tc_chunks" / "match_chunks" (0 = auto), "match_rq" (0 auto, 1, 2, 4
 * queries per thread of the POPC kernel), "match_balance" (-1 the time model chooses, 0 uniform chunks, 1 balanced job list),
 * "match_slice_rows" (first H2D slice of the streamed host-DB path).  clb_get_param also answers "last_match_path" (0 = POPC
 * kernel, 4 = tensor-core kernel), "last_match_balanced" and "lab".
 * Kernel-variant and timing knobs ("match_hs", "match_tc_ver", "match_tc_dbg", "fast_var", "fast_nms", "latch_var", "latch_tex")
 * exist only in the lab build (make -C coloc_b200/csrc lab -> libcoloc_b200_lab.so, the same sources with -DCLB_LAB, which also
 * carries the superseded kernel generations for A/B); the product library answers CLB_EUNSUPPORTED to them, so nothing a caller
 * sets can change which kernels it runs.  Thresholds: the reference passes uint8_t (CUDAK2NN.cu:46); this ABI takes int and
 * compares `second - best > threshold` in int, so values above 255 simply never match rather than wrapping. */
int  clb_set_param(const char* name, int value);
int  clb_get_param(const char* name, int* value_out);
float clb_last_kernel_ms(const char* which);   /* "match", "describe", "fast", "resize", "merge" */

#ifdef __cplusplus
}
#endif
#endif
