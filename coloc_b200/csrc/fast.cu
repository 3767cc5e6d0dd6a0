// fast.cu -- FAST-9/16 corner score + 3x3 non-max suppression + intensity-centroid orientation,
// all pyramid levels in one pass, keypoints emitted in the reference's order.
//
// Replaces the CPU stage of the reference detector: KFAST<true,true> (include/coloc/KFAST.h:164-540)
// and featureAngle/fastAtan2 (include/coloc/FeatureAngle.h:160-247) as called per level by
// GPUDetector.hpp:262-276 -- and with them the 7 D2H pyramid copies + the H2D keypoint copy that
// bracket that stage.  Semantics (bit-exact set, scores, order):
//   corner  : >= 9 contiguous ring pixels all > p+t or all < p-t (strict)        KFAST.h:272-290
//   scanned : rows 3..h-4, cols 3..w-4                                          KFAST.h:431,451,243
//   score   : max over the 16 arcs of 9 of max(min(p-x), min(x-p))              KFAST.h:304-374
//   NMS     : score strictly greater than the 8 neighbours' scores              KFAST.h:484-495
//   order   : level-major, then row-major (y, then x)                           GPUDetector.hpp:262-276
//   angle   : fastAtan2(sum dy*I, sum dx*I) over the 37-px disc                 FeatureAngle.h:197-247
// The `1 << 32` edge at KFAST.h:243: on x86 the shift count wraps and the mask becomes 0, so on levels whose width has
// (w - 38) % 16 == 0 (214 = level 6 of 640, 1286 = level 6 of 3840) the compiled reference drops the last 32 columns of every
// row whose block walk (KFAST.h:451-459, 16-px retreat at :256-262) lands exactly on column w - 35.  `quirk` = 1 (the default,
// clb_detector_set_param "fast_quirk") reproduces that: fast_xend_kernel replays the walk per row on the pre-test mask and the
// score kernel stops each row at its xend; quirk = 0 gives the intended semantics (every column 3..w-4).
//
// Kernels: (1) score+NMS per 64x16 tile staged in shared memory, NMS'd score map + per-row counts
// out; (2) one-block exclusive scan of the row counts; (3) warp-per-row ordered compaction that
// also computes the angle.  Integer work, HBM/L2-bound: pyramid read once, 1 B/px map written.
#include "common.cuh"

namespace clb {

static constexpr int FT_TW = 64, FT_TH = 16;           // output tile
static constexpr int FT_SW = FT_TW + 8, FT_SH = FT_TH + 8;   // pixels needed: +-4 (3 ring + 1 NMS)
static constexpr int FT_SSTRIDE = 80;                  // smem row stride (bytes)
static constexpr int FT_CW = FT_TW + 2, FT_CH = FT_TH + 2;   // score region: +-1
static constexpr int FT_CSTRIDE = 68;

#ifdef CLB_LAB
std::atomic<int> g_fast_nms{1};   // debug knob: 0 = emit every corner (raw score map)
std::atomic<int> g_fast_var{55};   // clb_set_param "fast_var": bit 0 = queued score kernel (fast_nms2_kernel), bit 1 = fused scan + compaction +
                      // orientation (compact2_kernel), bit 2 = 64x32 tiles instead of 64x16, bit 3 = score kernel compiled for
                      // 8 CTAs per SM (32 registers), bit 4 = pre-test on four positions per thread in 16-bit lanes, bit 5 = (row, tile)
                      // occupancy bitmap: the tail visits occupied tiles only; 0 = the first-generation path
#else
static constexpr int g_fast_nms = 1, g_fast_var = 55;    // the product build carries the default variant only
#endif

struct RowMap {   // global row index <-> (level, y); for the v2 score kernel also tile index <-> (level, tile)
    int n; int row_base[CLB_MAX_LEVELS + 1];
    int tile_base[CLB_MAX_LEVELS + 1];      // first tile of each level for the tile height of this launch (host-computed:
    int tiles_x[CLB_MAX_LEVELS];            // the kernel's level search is compares only, no divisions per thread)
    uint32_t tiles_x_magic[CLB_MAX_LEVELS]; // tile / tiles_x == umulhi(tile, magic) (0: divide); entries of tile_base above n are INT_MAX
    int bm_words;                           // words per row of the (row, 64-px tile) occupancy bitmap
    uint32_t quirk_mask;                    // bit L: level L takes its per-row scan end from xend[] (KFAST.h:243 edge)
};

// rows of the quirk levels laid end to end: one warp of fast_xend_kernel per row
struct QuirkRows { int n; int level[CLB_MAX_LEVELS]; int row_base[CLB_MAX_LEVELS + 1]; };
static constexpr int FQ_WORDS = 32;         // block bits per row: 32 words x 32 blocks x 16 px = 16 384 px wide at most

__device__ __forceinline__ bool fast_pretest(const uint8_t* __restrict__ p, int stride, int t) {
    // cardinal early-out of KFAST.h:186-235: two adjacent compass points beyond the threshold = one vertical and one horizontal
    const int c = p[0];
    const int pn = p[-3 * stride], ps = p[3 * stride], pe = p[3], pw = p[-3];
    return (min(max(pn, ps), max(pe, pw)) > c + t) | (max(min(pn, ps), min(pe, pw)) < c - t);
}

// Per-row scan end of the reference's column walk (KFAST.h:451-459 with the retreat of :256-262 and the mask of :243).
// The walk starts at column 3 and looks at 32 columns: if some pass the pre-test but none of the first 16 does it advances by
// 16, otherwise by 32.  With B = (w - 38) / 16 blocks of 16 columns in front of column w - 35, the walk either lands on block B
// (32 columns left -> `(1 << 32) - 1` == 0 on x86 -> the tail is dropped, xend = w - 35) or steps over it (xend = w - 3).
// One warp per row, four positions per lane and step (columns 3 + 128 i + 4 lane ..+3: one quarter of a 16-column block, read
// as aligned 32-bit words of the three rows the pre-test touches, the loads of four steps in flight together -- the first
// version walked 32 positions per step with byte loads and spent 22 us on a 1286 x 723 level waiting for them); two ballots
// per step give the "any pre-test pass" bit of the step's eight blocks, lane 0 replays the walk.
__device__ __forceinline__ uint32_t xend_pass4(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t n0, uint32_t n1, uint32_t s0, uint32_t s1, int t) {
    // positions x .. x+3 with x = 4k + 3: centre / north / south bytes = (word0.b3, word1.b0..b2), west = word0.b0..b3,
    // east = (word1.b2, word1.b3, word2.b0, word2.b1)
    const uint32_t c = __byte_perm(w0, w1, 0x6543), n = __byte_perm(n0, n1, 0x6543), s = __byte_perm(s0, s1, 0x6543);
    const uint32_t e = __byte_perm(w1, w2, 0x5432), wv = w0;
    uint32_t any = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int cc = (c >> (8 * k)) & 255, nn = (n >> (8 * k)) & 255, ss = (s >> (8 * k)) & 255, ee = (e >> (8 * k)) & 255, ww = (wv >> (8 * k)) & 255;
        const bool pass = (min(max(nn, ss), max(ee, ww)) > cc + t) | (max(min(nn, ss), min(ee, ww)) < cc - t);
        any |= (uint32_t)pass << k;
    }
    return any;
}

__global__ void __launch_bounds__(128)
fast_xend_kernel(const LevelTable lt, const RowMap rm, const QuirkRows qr, int thresh, int* __restrict__ xend) {
    __shared__ uint32_t sbits[4][FQ_WORDS];
    ptx_grid_dep_wait();                // every CTA: the level comes from the resize kernel, and the score kernel behind this
    ptx_grid_dep_launch();              // one relies on this grid not finishing before that one has
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int qrow = blockIdx.x * 4 + wid;
    if (qrow >= qr.row_base[qr.n]) return;
    int k = 0;
    while (k + 1 < qr.n && qrow >= qr.row_base[k + 1]) ++k;
    const int L = qr.level[k], y = qrow - qr.row_base[k];
    const int w = lt.w[L], h = lt.h[L], stride = lt.stride[L];
    int* out = xend + rm.row_base[L] + y;
    if (y < 3 || y >= h - 3) { if (lane == 0) *out = w - 3; return; }
    // rows are 128-byte aligned and at least w + 1 bytes of every row are addressable (stride = align_up(w, 128) > w - 12)
    const uint32_t* __restrict__ rc = reinterpret_cast<const uint32_t*>(lt.img[L] + (size_t)y * stride);
    const uint32_t* __restrict__ rn = reinterpret_cast<const uint32_t*>(lt.img[L] + (size_t)(y - 3) * stride);
    const uint32_t* __restrict__ rs = reinterpret_cast<const uint32_t*>(lt.img[L] + (size_t)(y + 3) * stride);
    const int B = (w - 38) >> 4;                                  // blocks 0..B are looked at: columns 3 .. w - 20
    const int nsteps = (B + 1 + 7) >> 3;                          // 8 blocks (128 columns) per step
    const int wmax = (w - 20 + 8) >> 2;                           // last word index any valid position needs (x + 8 <= w - 12)
    constexpr int XU = 10;                                        // steps whose loads are in flight together (1286 px = 10 steps)
    for (int i0 = 0; i0 < nsteps; i0 += XU) {
        uint32_t v[XU][7];
#pragma unroll
        for (int u = 0; u < XU; ++u) {
            const int wi = 32 * (i0 + u) + lane;                  // word of columns 4 wi .. 4 wi + 3; this lane's positions start at 4 wi + 3
            const bool ok = (i0 + u) < nsteps && 4 * wi + 3 <= w - 20;
            const int a = ok ? wi : 0;
            v[u][0] = __ldg(rc + a); v[u][1] = __ldg(rc + min(a + 1, wmax)); v[u][2] = __ldg(rc + min(a + 2, wmax));
            v[u][3] = __ldg(rn + a); v[u][4] = __ldg(rn + min(a + 1, wmax));
            v[u][5] = __ldg(rs + a); v[u][6] = __ldg(rs + min(a + 1, wmax));
        }
#pragma unroll
        for (int u = 0; u < XU; ++u) {
            if (i0 + u >= nsteps) break;
            const int x = 4 * (32 * (i0 + u) + lane) + 3;
            uint32_t any = xend_pass4(v[u][0], v[u][1], v[u][2], v[u][3], v[u][4], v[u][5], v[u][6], thresh);
            // positions beyond column w - 20 belong to no block the walk looks at
            const int nvalid = min(max(w - 20 - x + 1, 0), 4);
            any &= (1u << nvalid) - 1u;
            const uint32_t bal = __ballot_sync(0xFFFFFFFFu, any != 0u);
            const uint32_t blk = __ballot_sync(0xFFFFFFFFu, lane < 8 && ((bal >> (4 * lane)) & 0xFu) != 0u) & 0xFFu;
            if (lane == 0) {
                const int i = i0 + u;                              // blocks 8 i .. 8 i + 7
                uint32_t* wp = &sbits[wid][i >> 2];
                if ((i & 3) == 0) *wp = blk; else *wp |= blk << (8 * (i & 3));
            }
        }
    }
    __syncwarp();
    if (lane == 0) {
        int b = 0;
        while (b < B) {
            const bool lo = (sbits[wid][b >> 5] >> (b & 31)) & 1u, hi = (sbits[wid][(b + 1) >> 5] >> ((b + 1) & 31)) & 1u;
            b += (!lo && hi) ? 1 : 2;
        }
        *out = b == B ? w - 35 : w - 3;
    }
}

#ifdef CLB_LAB      // first-generation kernels, kept for A/B only (clb_set_param "fast_var")
__device__ __forceinline__ int fast_score_px(const uint8_t* __restrict__ p, int stride, int t) {
    const int c = p[0];
    const int hi = c + t, lo = c - t;
    // cardinal early-out (KFAST.h:186-235): any 9-arc contains two adjacent compass points
    const int p9 = p[3 * stride], p5 = p[3], p1 = p[-3 * stride], p13 = p[-3];
    const bool b = ((p9 > hi) & (p5 > hi)) | ((p5 > hi) & (p1 > hi)) | ((p1 > hi) & (p13 > hi)) | ((p13 > hi) & (p9 > hi));
    const bool d = ((p9 < lo) & (p5 < lo)) | ((p5 < lo) & (p1 < lo)) | ((p1 < lo) & (p13 < lo)) | ((p13 < lo) & (p9 < lo));
    if (!(b | d)) return 0;
    int ring[16];
    ring[0] = p9;                 ring[1] = p[3 * stride + 1];  ring[2] = p[2 * stride + 2];  ring[3] = p[stride + 3];
    ring[4] = p5;                 ring[5] = p[-stride + 3];     ring[6] = p[-2 * stride + 2]; ring[7] = p[-3 * stride + 1];
    ring[8] = p1;                 ring[9] = p[-3 * stride - 1]; ring[10] = p[-2 * stride - 2]; ring[11] = p[-stride - 3];
    ring[12] = p13;               ring[13] = p[stride - 3];     ring[14] = p[2 * stride - 2]; ring[15] = p[3 * stride - 1];
    uint32_t mb = 0, md = 0;
#pragma unroll
    for (int k = 0; k < 16; ++k) { mb |= (uint32_t)(ring[k] > hi) << k; md |= (uint32_t)(ring[k] < lo) << k; }
    // 9 contiguous set bits in the circular 16-bit mask
    auto run9 = [](uint32_t m) {
        m |= m << 16;
        uint32_t a = m & (m >> 1);
        a &= a >> 2;
        a &= a >> 4;
        a &= m >> 8;
        return (a & 0xFFFFu) != 0;
    };
    if (!(run9(mb) | run9(md))) return 0;
    // score = max over the 16 arcs of 9 of max(min_k(c - x_k), min_k(x_k - c))
    //       = max(c - min_arcs(max_k x_k), max_arcs(min_k x_k) - c).
    // Sliding max/min of width 9 on the raw ring values (all in [0,255], non-negative: ptxas 12.9
    // mis-packs *signed* 32-bit min/max chains into VIMNMX3.U16x2 at -O1+):
    // width 8 by three doubling steps, then one more element.
    int mn1[16], mx1[16], mn2[16], mx2[16], mn4[16], mx4[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) { mn1[k] = min(ring[k], ring[(k + 1) & 15]); mx1[k] = max(ring[k], ring[(k + 1) & 15]); }
#pragma unroll
    for (int k = 0; k < 16; ++k) { mn2[k] = min(mn1[k], mn1[(k + 2) & 15]); mx2[k] = max(mx1[k], mx1[(k + 2) & 15]); }
#pragma unroll
    for (int k = 0; k < 16; ++k) { mn4[k] = min(mn2[k], mn2[(k + 4) & 15]); mx4[k] = max(mx2[k], mx2[(k + 4) & 15]); }
    int lo_of_max = 255, hi_of_min = 0;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        lo_of_max = min(lo_of_max, max(mx4[k], ring[(k + 8) & 15]));
        hi_of_min = max(hi_of_min, min(mn4[k], ring[(k + 8) & 15]));
    }
    const int dark = c - lo_of_max, bright = hi_of_min - c;
    const int best = dark > bright ? dark : bright;
    return best & 0xFF;
}

__global__ void __launch_bounds__(256)
fast_nms_kernel(const LevelTable lt, const RowMap rm, int thresh, uint8_t* __restrict__ nms_map,
                int do_nms, int* __restrict__ row_count, const int* __restrict__ xend) {
    __shared__ __align__(16) uint8_t px[FT_SH * FT_SSTRIDE];
    __shared__ uint8_t sc[FT_CH * FT_CSTRIDE];

    int tile = blockIdx.x, L = 0, tx_n = 0;
    for (; L < lt.n; ++L) {
        tx_n = (lt.w[L] + FT_TW - 1) / FT_TW;
        const int cnt = tx_n * ((lt.h[L] + FT_TH - 1) / FT_TH);
        if (tile < cnt) break;
        tile -= cnt;
    }
    if (L >= lt.n) return;
    const int w = lt.w[L], h = lt.h[L], stride = lt.stride[L];
    const uint8_t* __restrict__ img = lt.img[L];
    const int x0 = (tile % tx_n) * FT_TW, y0 = (tile / tx_n) * FT_TH;
    const int tid = threadIdx.x;

    // stage (TW+8) x (TH+8) pixels, clamped at the image border (border values are never scored)
    for (int i = tid; i < FT_SH * (FT_SW / 4); i += 256) {
        const int r = i / (FT_SW / 4), cq = i % (FT_SW / 4);
        const int gy = min(max(y0 - 4 + r, 0), h - 1);
        const int gx = x0 - 4 + cq * 4;
        const uint8_t* row = img + (size_t)gy * stride;
        uint32_t v;
        if (gx >= 0 && gx + 3 < w && ((((uintptr_t)(row + gx)) & 3) == 0)) {
            v = __ldg(reinterpret_cast<const uint32_t*>(row + gx));
        } else {
            v = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) v |= (uint32_t)__ldg(row + min(max(gx + k, 0), w - 1)) << (8 * k);
        }
        *reinterpret_cast<uint32_t*>(&px[r * FT_SSTRIDE + cq * 4]) = v;
    }
    __syncthreads();

    // scores on the (TW+2) x (TH+2) region
    for (int i = tid; i < FT_CH * FT_CW; i += 256) {
        const int r = i / FT_CW, c = i % FT_CW;
        const int gx = x0 - 1 + c, gy = y0 - 1 + r;
        int s = 0;
        if (gx >= 3 && gx < w - 3 && gy >= 3 && gy < h - 3 && (!((rm.quirk_mask >> L) & 1u) || gx < xend[rm.row_base[L] + gy]))
            s = fast_score_px(&px[(r + 3) * FT_SSTRIDE + (c + 3)], FT_SSTRIDE, thresh);
        sc[r * FT_CSTRIDE + c] = (uint8_t)s;
    }
    __syncthreads();

    // NMS + write the map; per-row counts via ballot (each warp covers 32 px of one tile row)
    uint8_t* map = nms_map;   // per-level maps are addressed through the row map (stride = level stride)
    const int grow0 = rm.row_base[L];
#pragma unroll
    for (int k = 0; k < (FT_TW * FT_TH) / 256; ++k) {
        const int i = k * 256 + tid;
        const int r = i / FT_TW, c = i % FT_TW;
        const int gx = x0 + c, gy = y0 + r;
        const uint8_t* s1 = &sc[(r + 1) * FT_CSTRIDE + (c + 1)];
        const int s = s1[0];
        bool keep = false;
        if (s && !do_nms) keep = true;
        else if (s) {
            keep = (s > s1[-1]) & (s > s1[1]) &
                   (s > s1[-FT_CSTRIDE - 1]) & (s > s1[-FT_CSTRIDE]) & (s > s1[-FT_CSTRIDE + 1]) &
                   (s > s1[FT_CSTRIDE - 1]) & (s > s1[FT_CSTRIDE]) & (s > s1[FT_CSTRIDE + 1]);
        }
        const bool inside = gx < w && gy < h;
        keep = keep && inside;
        if (inside) map[(size_t)(grow0 + gy) * lt.stride[0] + gx] = keep ? (uint8_t)s : (uint8_t)0;
        const uint32_t bal = __ballot_sync(0xFFFFFFFFu, keep);
        if ((tid & 31) == 0 && bal) { atomicAdd(&row_count[grow0 + gy], __popc(bal)); atomicAdd(&row_count[rm.row_base[rm.n] + ((grow0 + gy) >> 6)], __popc(bal)); }
    }
}

// exclusive scan of row counts (one block); also writes the total
__global__ void __launch_bounds__(1024)
row_scan_kernel(const int* __restrict__ row_count, int nrows, int* __restrict__ row_off, int* __restrict__ total) {
    __shared__ int warp_sum[32];
    __shared__ int carry;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < nrows; base += 1024) {
        const int i = base + tid;
        const int v = i < nrows ? row_count[i] : 0;
        int x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xFFFFFFFFu, x, o); if (lane >= o) x += y; }
        if (lane == 31) warp_sum[wid] = x;
        __syncthreads();
        if (wid == 0) {
            int s = warp_sum[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xFFFFFFFFu, s, o); if (lane >= o) s += y; }
            warp_sum[lane] = s;
        }
        __syncthreads();
        const int prefix = carry + (wid ? warp_sum[wid - 1] : 0) + x - v;
        if (i < nrows) row_off[i] = prefix;
        __syncthreads();
        if (tid == 1023) carry = prefix + v;
        __syncthreads();
    }
    if (tid == 0) *total = carry;
}

#endif   // CLB_LAB

// FeatureAngle.h:160-177 with GCC's -O3 contraction written out (oracle-checked exhaustively)
__device__ __forceinline__ float fast_atan2_ref(float y, float x) {
    const float PI = 3.1415927f;
    const float ax = fabsf(x), ay = fabsf(y);
    float a;
    if (ax >= ay) {
        const float c = __fdiv_rn(ay, __fadd_rn(ax, 1.17549435e-38f));
        const float cc = __fmul_rn(c, c);
        a = __fmul_rn(__fmaf_rn(__fmaf_rn(__fmaf_rn(-0.0443265555479f, cc, 0.1555786518f), cc, -0.325808397f), cc, 0.9997878412f), c);
    } else {
        const float c = __fdiv_rn(ax, __fadd_rn(ay, 1.17549435e-38f));
        const float cc = __fmul_rn(c, c);
        const float p = __fmaf_rn(__fmaf_rn(__fmaf_rn(-0.0443265555479f, cc, 0.1555786518f), cc, -0.325808397f), cc, 0.9997878412f);
        a = __fmaf_rn(-p, c, 1.57079635f);
    }
    if (x < 0.0f) a = __fsub_rn(PI, a);
    if (y < 0.0f) a = -a;
    return a;
}

__device__ __forceinline__ float feature_angle(const uint8_t* __restrict__ img, int x, int y, int stride) {
    int xs = 0, ys = 0;
#pragma unroll
    for (int dy = -3; dy <= 3; ++dy) {
        const int hw = (dy == -3 || dy == 3) ? 1 : ((dy == -2 || dy == 2) ? 2 : 3);
        const uint8_t* r = img + (size_t)(y + dy) * stride + x;
#pragma unroll
        for (int dx = -3; dx <= 3; ++dx) {
            if (dx < -hw || dx > hw) continue;
            const int v = __ldg(r + dx);
            xs += dx * v; ys += dy * v;
        }
    }
    return fast_atan2_ref((float)ys, (float)xs);
}

#ifdef CLB_LAB
// one warp per global row: ordered compaction, 128 map bytes per step (one 32-bit load per lane, 4 ballots)
__global__ void __launch_bounds__(256)
compact_kernel(const LevelTable lt, const RowMap rm, const uint8_t* __restrict__ nms_map,
               const int* __restrict__ row_count, const int* __restrict__ row_off,
               clb_keypoint* __restrict__ kps, int cap) {
    const int grow = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (grow >= rm.row_base[rm.n]) return;
    if (row_count[grow] == 0) return;
    int L = 0;
    while (L + 1 < rm.n && grow >= rm.row_base[L + 1]) ++L;
    const int y = grow - rm.row_base[L];
    const int w = lt.w[L];
    const uint32_t* row = reinterpret_cast<const uint32_t*>(nms_map + (size_t)grow * lt.stride[0]);   // stride0 is 128-B aligned
    int out = row_off[grow];
    for (int xb = 0; xb < w; xb += 128) {
        const int x0 = xb + lane * 4;
        uint32_t v = x0 < w ? __ldg(row + (x0 >> 2)) : 0u;      // bytes beyond w inside the padded stride are never written: mask them
        if (x0 + 3 >= w) {
            const int keep = w - x0;                             // 1..3 valid bytes (or <= 0)
            v = keep > 0 ? (v & (0xFFFFFFFFu >> (8 * (4 - keep)))) : 0u;
        }
        // order inside the 128-px step is x = xb + 4*lane + k: count lanes below (all 4 bytes) + own lower bytes
        const uint32_t nz = ((v & 0xFFu) != 0) + ((v & 0xFF00u) != 0) + ((v & 0xFF0000u) != 0) + ((v >> 24) != 0);
        uint32_t pre = nz;                                       // inclusive warp scan of nz
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, pre, o); if (lane >= o) pre += t; }
        int pos = out + (int)(pre - nz);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint32_t s = (v >> (8 * k)) & 0xFFu;
            if (s) {
                if (pos < cap) {
                    clb_keypoint kp;
                    kp.x = x0 + k; kp.y = y; kp.score = (uint8_t)s; kp._pad0[0] = kp._pad0[1] = kp._pad0[2] = 0;
                    kp.angle = 0.f;
                    kp.scale = (uint8_t)L; kp._pad1[0] = kp._pad1[1] = kp._pad1[2] = 0;
                    kps[pos] = kp;
                }
                ++pos;
            }
        }
        out += (int)__shfl_sync(0xFFFFFFFFu, pre, 31);
    }
}

// orientation, one thread per keypoint (FeatureAngle.h:197-247)
__global__ void __launch_bounds__(256)
angle_kernel(const LevelTable lt, clb_keypoint* __restrict__ kps, const int* __restrict__ d_total, int cap) {
    const int n = min(*d_total, cap);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int L = kps[i].scale;
        kps[i].angle = feature_angle(lt.img[L], kps[i].x, kps[i].y, lt.stride[L]);
    }
}

#endif   // CLB_LAB

// ---------------------------------------------------------------------------------------------------------------
// v2 score kernel: the same arithmetic on DENSE warps.  On textured frames ~10 % of the positions pass the cardinal
// pre-test and ~3 % are corners, so in the one-pass kernel nearly every warp pays for the ring masks and the sliding
// min/max of a few lanes.  Here each stage compacts its survivors into a shared-memory queue (ballot + one shared
// atomic per warp) and the next stage runs one survivor per thread:
//   stage    (TH+8) x 80 pixels, 8-byte loads on interior tiles
//   pre-test every position of the (TW+2) x (TH+2) score region: min(max(N,S), max(E,W)) > p+t or the dark twin
//            (two adjacent compass points = one vertical and one horizontal one)            -> queue 1
//   ring     queue 1: 16-bit bright/dark masks, 9-run test                                   -> queue 2
//   score    queue 2: sliding min/max of width 9, score byte into the shared score map
//   nms      queue 2 again: 8 neighbours of each corner of the output tile, survivors into the output map + row counts
//   output   one 32-bit word (4 px) per thread to the global map
template <int TH, bool TMA = false>
struct Fast2 {
    static constexpr int SH = TH + 8;              // staged rows: gy = y0 - 4 + r
    static constexpr int XO = TMA ? 16 : 8;        // staged cols: gx = x0 - XO + j (a TMA box starts on a 16-byte boundary)
    static constexpr int SSTR = TMA ? 96 : 80;
    static constexpr int CW = FT_TW + 2, CH = TH + 2;   // score region: gx = x0 - 1 + c, gy = y0 - 1 + r
    static constexpr int CSTR = 72;                // score byte (r, c) at r * CSTR + c + 3: output col 0 is word aligned
    static constexpr int NPOS = CH * CW;
};

__device__ __forceinline__ void fast_ring(const uint8_t* __restrict__ p, int stride, int (&ring)[16]) {
    ring[0] = p[3 * stride];      ring[1] = p[3 * stride + 1];  ring[2] = p[2 * stride + 2];   ring[3] = p[stride + 3];
    ring[4] = p[3];               ring[5] = p[-stride + 3];     ring[6] = p[-2 * stride + 2];  ring[7] = p[-3 * stride + 1];
    ring[8] = p[-3 * stride];     ring[9] = p[-3 * stride - 1]; ring[10] = p[-2 * stride - 2]; ring[11] = p[-stride - 3];
    ring[12] = p[-3];             ring[13] = p[stride - 3];     ring[14] = p[2 * stride - 2];  ring[15] = p[3 * stride - 1];
}

// TMA = the tile + halo arrives by ONE cp.async.bulk.tensor.2d (SASS UTMALDG) from the level's tensor map: box 96 x (TH + 8)
// at (x0 - 16, y0 - 4) (a box starts on a 16-byte boundary), rows / columns outside the level zero-filled by the copy engine (border pixels are never scored, so
// zero is as good as the clamped values the load loop staged); the staging loop and its per-lane address arithmetic -- a
// fifth of the kernel's instructions in the ncu source view -- are gone.
template <int TH, int MINB, bool SWAR, bool TMA>
__global__ void __launch_bounds__(256, MINB)
fast_nms2_kernel(const LevelTable lt, const RowMap rm, const __grid_constant__ LevelMaps maps, int thresh, uint8_t* __restrict__ nms_map,
                 int do_nms, int* __restrict__ row_count, uint32_t* __restrict__ bitmap, const int* __restrict__ xend) {
    typedef Fast2<TH, TMA> F;
    __shared__ __align__(128) uint8_t px[F::SH * F::SSTR];
    __shared__ __align__(8) uint64_t bar;
    __shared__ __align__(16) uint8_t sc[F::CH * F::CSTR];
    __shared__ __align__(16) uint8_t so[TH * FT_TW];
    __shared__ uint16_t q1[F::NPOS], q2[F::NPOS];
    __shared__ int n1, n2, rowcnt[TH], s_xlim[F::CH];

    // tile -> (level, tile row, tile column): four-step search of the largest L with tile_base[L] <= blockIdx.x (entries above the
    // level count are INT_MAX) and a multiply-shift for the division -- this decode was 8 % of the kernel's instructions as a loop
    // plus an integer division, executed by ~12 700 short-lived CTAs per 4K frame
    const int bx = (int)blockIdx.x;
    int L = bx >= rm.tile_base[8] ? 8 : 0;
    L += bx >= rm.tile_base[L + 4] ? 4 : 0;
    L += bx >= rm.tile_base[L + 2] ? 2 : 0;
    L += bx >= rm.tile_base[L + 1] ? 1 : 0;
    if (L >= rm.n) return;
    const int tile = bx - rm.tile_base[L], tx_n = rm.tiles_x[L];
    const uint32_t magic = rm.tiles_x_magic[L];
    const int w = lt.w[L], h = lt.h[L], stride = lt.stride[L];
    const uint8_t* __restrict__ img = lt.img[L];
    const int tyi = magic ? (int)__umulhi((uint32_t)tile, magic) : tile / tx_n;
    const int x0 = (tile - tyi * tx_n) * FT_TW, y0 = tyi * TH;
    const int tid = threadIdx.x, lane = tid & 31;

    // Which grid is in front of this one?  Without quirk levels it is the resize kernel: wait for it.  With quirk levels it is
    // fast_xend_kernel, every CTA of which sat out the resize kernel (griddepcontrol.wait) BEFORE releasing its dependents: by the
    // time this grid exists the levels are complete and visible, so only the tiles of a quirk level -- which read xend[] -- have to
    // wait for that grid itself.  The ~11 us latency-bound scan-end pass then runs next to the level-0 tiles instead of in front.
    const bool lvl_quirk = (rm.quirk_mask >> L) & 1u;
    const bool need_wait = rm.quirk_mask == 0u || lvl_quirk;
    if (TMA && tid == 0) {
        ptx_mbar_init(&bar, 1);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        if (need_wait) ptx_grid_dep_wait();            // the level is written by the resize kernel in front of this one
        ptx_mbar_expect_tx(&bar, (uint32_t)(F::SH * F::SSTR));
        ptx_tma_load_2d(px, &maps.fast[L], x0 - F::XO, y0 - 4, &bar);
    }
    ptx_grid_dep_launch();
    if (tid == 0) { n1 = 0; n2 = 0; }
    if (tid < TH) rowcnt[tid] = 0;
    if (!TMA || lvl_quirk) ptx_grid_dep_wait();      // xend[] comes from the kernel in front as well
    if (tid < F::CH) {                 // scan end of each score-region row: w - 3, or the reference's column walk (fast_xend_kernel)
        const int gy = y0 - 1 + tid;
        s_xlim[tid] = (((rm.quirk_mask >> L) & 1u) && gy >= 3 && gy < h - 3) ? __ldcg(xend + rm.row_base[L] + gy) : w - 3;
    }
    for (int i = tid; i < F::CH * F::CSTR / 4; i += 256) reinterpret_cast<uint32_t*>(sc)[i] = 0u;
    for (int i = tid; i < TH * FT_TW / 4; i += 256) reinterpret_cast<uint32_t*>(so)[i] = 0u;

    // stage
    const bool wide = x0 >= 8 && x0 + F::SSTR - 8 <= w && (stride & 7) == 0 && (reinterpret_cast<uintptr_t>(img) & 7) == 0;
    if (TMA) {
        // nothing to do: the box is in flight
    } else if (wide) {
        for (int i = tid; i < F::SH * 10; i += 256) {          // x0 - 8 is 8-byte aligned
            const int r = i / 10, q = i - r * 10;
            const int gy = min(max(y0 - 4 + r, 0), h - 1);
            const uint2 v = __ldg(reinterpret_cast<const uint2*>(img + (size_t)gy * stride + (x0 - 8)) + q);
            *reinterpret_cast<uint2*>(&px[r * F::SSTR + q * 8]) = v;
        }
    } else {
        for (int i = tid; i < F::SH * (F::SSTR / 4); i += 256) {
            const int r = i / (F::SSTR / 4), cq = i - r * (F::SSTR / 4);
            const int gy = min(max(y0 - 4 + r, 0), h - 1);
            const int gx = x0 - 8 + cq * 4;
            const uint8_t* row = img + (size_t)gy * stride;
            uint32_t v;
            if (gx >= 0 && gx + 3 < w && ((reinterpret_cast<uintptr_t>(row + gx)) & 3) == 0) {
                v = __ldg(reinterpret_cast<const uint32_t*>(row + gx));
            } else {
                v = 0;
#pragma unroll
                for (int k = 0; k < 4; ++k) v |= (uint32_t)__ldg(row + min(max(gx + k, 0), w - 1)) << (8 * k);
            }
            *reinterpret_cast<uint32_t*>(&px[r * F::SSTR + cq * 4]) = v;
        }
    }
    __syncthreads();
    if (TMA) ptx_mbar_wait(&bar, 0);

    // pre-test -> q1
    if (SWAR) {
        // four positions per thread, two per register in 16-bit lanes: VIMNMX.U16x2 / VIADD.16x2 are native on sm_100a (the
        // byte-wise video intrinsics are not).  Item (r, wq) = the four score-region columns 4wq-3 .. 4wq of row r, an aligned
        // word of the staged pixels; east / west neighbours come from the adjacent words by byte permute.
        //   bright > p + t  <=>  bright + ~p - t      has its lane's sign bit clear   (~p = -p - 1 in a 16-bit lane)
        //   dark   < p - t  <=>  dark   + ~p + t + 1  has its lane's sign bit set
        // Thread -> (word column wq = tid % 18, row tid / 18): a pass covers 14 rows and only the row changes between passes,
        // so the column terms (first valid position, staged-pixel address) are computed once per thread (the first version
        // re-derived (row, column) from a flat item index with a division per pass: 30 of its 110 instructions per item).
        constexpr int WPR = 18, RPP = 256 / WPR, NPASS = (F::CH + RPP - 1) / RPP;
        const uint32_t k_b = ((uint32_t)(-thresh) & 0xFFFFu) * 0x10001u, k_d = ((uint32_t)(thresh + 1) & 0xFFFFu) * 0x10001u;
        const int xlo = max(3, x0 - 1);                                        // valid gx range of this tile's score region: [xlo, xhi)
        const int wq = tid % WPR, rr = tid / WPR;
        const int c0 = 4 * wq - 3, gxb = x0 - 1 + c0;
        const uint32_t lowmask = 0xFu << min(max(xlo - gxb, 0), 4);
        const int xcap = x0 + F::CW - 1 - gxb;
        const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
        for (int p = 0; p < NPASS; ++p) {
            const int r = rr + RPP * p;
            uint32_t cm = 0;
            if (rr < RPP && r < F::CH) {
                const int gy = y0 - 1 + r;
                const int nval = min(max(min(s_xlim[r] - gxb, xcap), 0), 4);
                const uint32_t vmask = lowmask & ((1u << nval) - 1u);
                if (gy >= 3 && gy < h - 3 && vmask) {
                    const uint32_t* wp = reinterpret_cast<const uint32_t*>(&px[(r + 3) * F::SSTR + (F::XO - 4) + 4 * wq]);
                    const uint32_t cw = wp[0], cl = wp[-1], cr = wp[1];
                    const uint32_t nw = wp[-3 * F::SSTR / 4], sw = wp[3 * F::SSTR / 4];
                    const uint32_t ew = __byte_perm(cw, cr, 0x6543), ww = __byte_perm(cl, cw, 0x4321);
                    uint32_t fl[2];
#pragma unroll
                    for (int hf = 0; hf < 2; ++hf) {
                        const uint32_t sel = hf ? 0x4341u : 0x4240u;      // positions 1,3 / 0,2 -> 16-bit lanes
                        const uint32_t c2 = __byte_perm(cw, 0u, sel), n2 = __byte_perm(nw, 0u, sel), s2 = __byte_perm(sw, 0u, sel);
                        const uint32_t e2 = __byte_perm(ew, 0u, sel), w2 = __byte_perm(ww, 0u, sel);
                        const uint32_t bright = __vminu2(__vmaxu2(n2, s2), __vmaxu2(e2, w2));
                        const uint32_t dark = __vmaxu2(__vminu2(n2, s2), __vminu2(e2, w2));
                        const uint32_t nc = ~c2;
                        const uint32_t eb = __vadd2(bright, __vadd2(nc, k_b));
                        const uint32_t ed = __vadd2(dark, __vadd2(nc, k_d));
                        fl[hf] = (~eb | ed) & 0x80008000u;
                    }
                    const uint32_t t2 = ((fl[0] >> 15) | (fl[1] >> 14)) & 0x00030003u;   // bit 0/16: positions 0/2, bit 1/17: 1/3
                    cm = (t2 | (t2 >> 14)) & vmask;
                }
            }
            // queue slots: a lane has 0..4 candidates; the warp prefix of such small counts is three ballots (one per bit of
            // the count) and six population counts instead of a five-step shuffle scan
            const uint32_t cnt = __popc(cm);
            const uint32_t b0 = __ballot_sync(0xFFFFFFFFu, cnt & 1u), b1 = __ballot_sync(0xFFFFFFFFu, cnt & 2u), b2 = __ballot_sync(0xFFFFFFFFu, cnt & 4u);
            const int total = __popc(b0) + 2 * __popc(b1) + 4 * __popc(b2);
            if (total) {
                int b = 0;
                if (lane == 0) b = atomicAdd(&n1, total);
                b = __shfl_sync(0xFFFFFFFFu, b, 0);
                int pos = b + __popc(b0 & lt_mask) + 2 * __popc(b1 & lt_mask) + 4 * __popc(b2 & lt_mask);
                const int i0 = r * F::CW + c0;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (cm & (1u << k)) q1[pos++] = (uint16_t)(i0 + k);
            }
        }
    } else
    for (int base = 0; base < F::NPOS; base += 256) {
        const int i = base + tid;
        bool cand = false;
        if (i < F::NPOS) {
            const int r = i / F::CW, c = i - r * F::CW;
            const int gx = x0 - 1 + c, gy = y0 - 1 + r;
            if (gx >= 3 && gx < s_xlim[r] && gy >= 3 && gy < h - 3) {
                const uint8_t* p = &px[(r + 3) * F::SSTR + (c + F::XO - 1)];
                const int cv = p[0];
                const int pn = p[-3 * F::SSTR], ps = p[3 * F::SSTR], pe = p[3], pw = p[-3];
                const int bright = min(max(pn, ps), max(pe, pw));
                const int dark = max(min(pn, ps), min(pe, pw));
                cand = (bright > cv + thresh) | (dark < cv - thresh);
            }
        }
        const uint32_t bal = __ballot_sync(0xFFFFFFFFu, cand);
        if (bal) {
            int b = 0;
            if (lane == 0) b = atomicAdd(&n1, __popc(bal));
            b = __shfl_sync(0xFFFFFFFFu, b, 0);
            if (cand) q1[b + __popc(bal & ((1u << lane) - 1u))] = (uint16_t)i;
        }
    }
    __syncthreads();

    // ring masks + 9-run test -> q2
    const int c1 = n1;
    for (int base = 0; base < c1; base += 256) {
        const int k = base + tid;
        bool corner = false;
        int i = 0;
        if (k < c1) {
            i = q1[k];
            const int r = i / F::CW, c = i - r * F::CW;
            const uint8_t* p = &px[(r + 3) * F::SSTR + (c + F::XO - 1)];
            const int cv = p[0];
            const int hi = cv + thresh, lo = cv - thresh;
            int ring[16];
            fast_ring(p, F::SSTR, ring);
            uint32_t mb = 0, md = 0;
#pragma unroll
            for (int j = 0; j < 16; ++j) { mb |= (uint32_t)(ring[j] > hi) << j; md |= (uint32_t)(ring[j] < lo) << j; }
            auto run9 = [](uint32_t m) {
                m |= m << 16;
                uint32_t a = m & (m >> 1);
                a &= a >> 2;
                a &= a >> 4;
                a &= m >> 8;
                return (a & 0xFFFFu) != 0;
            };
            corner = run9(mb) | run9(md);
        }
        const uint32_t bal = __ballot_sync(0xFFFFFFFFu, corner);
        if (bal) {
            int b = 0;
            if (lane == 0) b = atomicAdd(&n2, __popc(bal));
            b = __shfl_sync(0xFFFFFFFFu, b, 0);
            if (corner) q2[b + __popc(bal & ((1u << lane) - 1u))] = (uint16_t)i;
        }
    }
    __syncthreads();

    // score (KFAST.h:304-374): max(c - min_arcs(max_k x_k), max_arcs(min_k x_k) - c), sliding windows of 9 on the ring
    const int c2 = n2;
    for (int k = tid; k < c2; k += 256) {
        const int i = q2[k];
        const int r = i / F::CW, c = i - r * F::CW;
        const uint8_t* p = &px[(r + 3) * F::SSTR + (c + F::XO - 1)];
        const int cv = p[0];
        int ring[16];
        fast_ring(p, F::SSTR, ring);
        int mn1[16], mx1[16], mn2[16], mx2[16], mn4[16], mx4[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) { mn1[j] = min(ring[j], ring[(j + 1) & 15]); mx1[j] = max(ring[j], ring[(j + 1) & 15]); }
#pragma unroll
        for (int j = 0; j < 16; ++j) { mn2[j] = min(mn1[j], mn1[(j + 2) & 15]); mx2[j] = max(mx1[j], mx1[(j + 2) & 15]); }
#pragma unroll
        for (int j = 0; j < 16; ++j) { mn4[j] = min(mn2[j], mn2[(j + 4) & 15]); mx4[j] = max(mx2[j], mx2[(j + 4) & 15]); }
        int lo_of_max = 255, hi_of_min = 0;
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            lo_of_max = min(lo_of_max, max(mx4[j], ring[(j + 8) & 15]));
            hi_of_min = max(hi_of_min, min(mn4[j], ring[(j + 8) & 15]));
        }
        const int dark = cv - lo_of_max, bright = hi_of_min - cv;
        sc[r * F::CSTR + c + 3] = (uint8_t)((dark > bright ? dark : bright) & 0xFF);
    }
    __syncthreads();

    // NMS (KFAST.h:484-495) over the corners of the output tile: strictly greater than the 8 neighbours' scores
    for (int k = tid; k < c2; k += 256) {
        const int i = q2[k];
        const int r = i / F::CW, c = i - r * F::CW;
        if (r < 1 || r > TH || c < 1 || c > FT_TW) continue;            // halo of the score region
        const uint8_t* s1 = &sc[r * F::CSTR + c + 3];
        const int s = s1[0];
        bool keep = s != 0;
        if (keep && do_nms)
            keep = (s > s1[-1]) & (s > s1[1]) &
                   (s > s1[-F::CSTR - 1]) & (s > s1[-F::CSTR]) & (s > s1[-F::CSTR + 1]) &
                   (s > s1[F::CSTR - 1]) & (s > s1[F::CSTR]) & (s > s1[F::CSTR + 1]);
        if (keep && x0 + c - 1 < w && y0 + r - 1 < h) {
            so[(r - 1) * FT_TW + (c - 1)] = (uint8_t)s;
            atomicAdd(&rowcnt[r - 1], 1);
        }
    }
    __syncthreads();

    // output: 4 px per thread (the map's row stride is stride[0], 128-byte aligned, >= every tile's right edge)
    const int grow0 = rm.row_base[L];
    const int mstride = lt.stride[0];
    for (int i = tid; i < TH * (FT_TW / 4); i += 256) {
        const int r = i / (FT_TW / 4), cw = i - r * (FT_TW / 4);
        const int gx = x0 + cw * 4, gy = y0 + r;
        if (gy < h && gx < w) {
            const uint32_t v = reinterpret_cast<const uint32_t*>(so)[i];
            uint8_t* dst = nms_map + (size_t)(grow0 + gy) * mstride + gx;
            if (gx + 3 < mstride) *reinterpret_cast<uint32_t*>(dst) = v;
            else for (int b = 0; b < 4 && gx + b < w; ++b) dst[b] = (uint8_t)(v >> (8 * b));
        }
    }
    if (tid < TH && rowcnt[tid]) {
        atomicAdd(&row_count[grow0 + y0 + tid], rowcnt[tid]);
        const int tx = tile % tx_n;          // which 64-px tiles of a row hold keypoints: the tail only visits those
        if (bitmap) atomicOr(&bitmap[(size_t)(grow0 + y0 + tid) * rm.bm_words + (tx >> 5)], 1u << (tx & 31));
    }
    if (tid < 32 && TH <= 32) {
        // keypoints per chunk of 64 global rows (stored behind the row counters): the tail's row offsets then cost two loads per
        // thread instead of a sum over every earlier row.  A tile's rows lie in at most two chunks.
        const int grow = grow0 + y0 + tid;
        const int c = tid < TH ? rowcnt[tid] : 0;
        const int first = (grow0 + y0) >> 6;
        int a = (grow >> 6) == first ? c : 0, b = c - a;
#pragma unroll
        for (int o = 16; o; o >>= 1) { a += __shfl_xor_sync(0xFFFFFFFFu, a, o); b += __shfl_xor_sync(0xFFFFFFFFu, b, o); }
        if (tid == 0 && a) atomicAdd(&row_count[rm.row_base[rm.n] + first], a);
        if (tid == 1 && b) atomicAdd(&row_count[rm.row_base[rm.n] + first + 1], b);
    }
}

// v2 tail: row offsets (sum of the counts of all earlier rows, computed by every block for itself -- a few
// thousand ints out of L2), ordered compaction, orientation of the row's keypoints, total.  One launch instead of three.
__global__ void __launch_bounds__(256)
compact2_kernel(const LevelTable lt, const RowMap rm, const uint8_t* __restrict__ nms_map,
                const int* __restrict__ row_count, clb_keypoint* __restrict__ kps, int cap, int* __restrict__ total,
                const uint32_t* __restrict__ bitmap) {
    __shared__ int wsum[8];
    __shared__ int rcnt[8];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int nrows = rm.row_base[rm.n];
    const int grow0 = blockIdx.x * 8, grow = grow0 + wid;
    ptx_grid_dep_wait();                // the score map, row counts and tile bitmap of the kernel in front
    ptx_grid_dep_launch();
    // keypoints in front of this block's rows = whole chunks of 64 rows (counted by the score kernel) + the rows of the last,
    // partial chunk
    int acc = 0;
    const int nchunk = grow0 >> 6;
    for (int i = tid; i < nchunk; i += 256) acc += __ldcg(row_count + nrows + i);
    for (int i = (nchunk << 6) + tid; i < grow0; i += 256) acc += __ldcg(row_count + i);
#pragma unroll
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, o);
    if (lane == 0) { wsum[wid] = acc; rcnt[wid] = grow < nrows ? __ldcg(row_count + grow) : 0; }
    __syncthreads();
    int out = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) { out += wsum[k]; if (k < wid) out += rcnt[k]; }
    const int mine = rcnt[wid];
    if (grow == nrows - 1 && lane == 0) *total = out + mine;
    if (grow >= nrows || mine == 0) return;
    int L = 0;
    while (L + 1 < rm.n && grow >= rm.row_base[L + 1]) ++L;
    const int y = grow - rm.row_base[L];
    const int w = lt.w[L];
    const uint32_t* row = reinterpret_cast<const uint32_t*>(nms_map + (size_t)grow * lt.stride[0]);
    const int first = out;
    // one step = 128 map bytes, 4 per lane, in x order: ordered emission through a warp scan of the per-lane counts
    auto emit = [&](uint32_t v, int x0) {
        if (x0 + 3 >= w) {                                           // bytes beyond w inside the padded stride are masked
            const int keep = w - x0;
            v = keep > 0 ? (v & (0xFFFFFFFFu >> (8 * (4 - min(keep, 4))))) : 0u;
        }
        if (!__ballot_sync(0xFFFFFFFFu, v != 0u)) return;
        const uint32_t nz = ((v & 0xFFu) != 0) + ((v & 0xFF00u) != 0) + ((v & 0xFF0000u) != 0) + ((v >> 24) != 0);
        uint32_t pre = nz;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, pre, o); if (lane >= o) pre += t; }
        int pos = out + (int)(pre - nz);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint32_t sv = (v >> (8 * k)) & 0xFFu;
            if (sv) {
                if (pos < cap) {
                    clb_keypoint kp;
                    kp.x = x0 + k; kp.y = y; kp.score = (uint8_t)sv; kp._pad0[0] = kp._pad0[1] = kp._pad0[2] = 0;
                    kp.angle = 0.f;
                    kp.scale = (uint8_t)L; kp._pad1[0] = kp._pad1[1] = kp._pad1[2] = 0;
                    kps[pos] = kp;
                }
                ++pos;
            }
        }
        out += (int)__shfl_sync(0xFFFFFFFFu, pre, 31);
    };
    bool sparse = false;
    const int tiles = rm.tiles_x[L];
    if (bitmap) {                     // few occupied tiles (the usual case): visit only those, two 64-px tiles per step
        const uint32_t* bm = bitmap + (size_t)grow * rm.bm_words;
        const int nw = (tiles + 31) >> 5;
        int occupied = 0;
        for (int wi = 0; wi < nw; ++wi) occupied += __popc(__ldg(bm + wi));
        sparse = 2 * occupied <= tiles;
        if (sparse) {
            int pend = -1;
            const int sub = (lane & 15) * 4;
            for (int wi = 0; wi < nw; ++wi) {
                uint32_t bits = __ldg(bm + wi);                       // warp-uniform
                while (bits) {
                    const int t = wi * 32 + __ffs(bits) - 1;
                    bits &= bits - 1;
                    if (pend < 0) { pend = t; continue; }
                    const int x0 = (lane < 16 ? pend : t) * FT_TW + sub;
                    emit(x0 < w ? __ldg(row + (x0 >> 2)) : 0u, x0);
                    pend = -1;
                }
            }
            if (pend >= 0) {
                const int x0 = pend * FT_TW + sub;
                emit(lane < 16 && x0 < w ? __ldg(row + (x0 >> 2)) : 0u, x0);
            }
        }
    }
    if (!sparse) {
        // dense rows: 128 px per step, the loads of CB steps are issued together -- the steps themselves are serialised by
        // their ballots and scans, and one L2 round trip per step was most of this kernel's time
        constexpr int CB = 16;
        for (int xc = 0; xc < w; xc += 128 * CB) {
            uint32_t vv[CB];
#pragma unroll
            for (int j = 0; j < CB; ++j) {
                const int x0 = xc + j * 128 + lane * 4;
                vv[j] = x0 < w ? __ldg(row + (x0 >> 2)) : 0u;
            }
#pragma unroll
            for (int j = 0; j < CB; ++j) emit(vv[j], xc + j * 128 + lane * 4);
        }
    }
    __syncwarp();
    // orientation (FeatureAngle.h:197-247), one lane per keypoint of this row
    const uint8_t* __restrict__ img = lt.img[L];
    const int stride = lt.stride[L];
    for (int j = first + lane; j < first + mine && j < cap; j += 32) {
        const int x = *reinterpret_cast<volatile int*>(&kps[j].x);
        kps[j].angle = feature_angle(img, x, y, stride);
    }
}

static RowMap make_rowmap(const LevelTable& lt) {
    RowMap rm; rm.n = lt.n; rm.row_base[0] = 0; rm.bm_words = 0; rm.quirk_mask = 0;
    for (int i = 0; i < lt.n; ++i) rm.row_base[i + 1] = rm.row_base[i] + lt.h[i];
    for (int i = 0; i <= CLB_MAX_LEVELS; ++i) rm.tile_base[i] = 0;
    for (int i = 0; i < CLB_MAX_LEVELS; ++i) { rm.tiles_x[i] = 0; rm.tiles_x_magic[i] = 0u; }
    return rm;
}

// tile grid of the v2 score kernel for tile height `th`: tile_base[] (INT_MAX above the level count: the kernel's level search
// reads all 17 entries), tiles per row and the multiply-shift constant of the division by it.  floor(t * ceil(2^32 / d) / 2^32) ==
// t / d for every t with t * d < 2^32 (the error term t * (d - 2^32 mod d) / (d * 2^32) stays below 1 / d); otherwise the kernel divides.
static void fill_tile_grid(RowMap& rm, const LevelTable& lt, int th) {
    rm.tile_base[0] = 0;
    for (int L = 0; L < lt.n; ++L) {
        const int txn = (lt.w[L] + FT_TW - 1) / FT_TW, tyn = (lt.h[L] + th - 1) / th;
        rm.tiles_x[L] = txn;
        rm.tile_base[L + 1] = rm.tile_base[L] + txn * tyn;
        const unsigned long long tiles = (unsigned long long)txn * (unsigned long long)tyn;
        rm.tiles_x_magic[L] = (txn > 1 && tiles * (unsigned long long)txn < 0x100000000ull) ? (uint32_t)(0x100000000ull / (unsigned long long)txn) + 1u : 0u;
    }
    for (int L = lt.n + 1; L <= CLB_MAX_LEVELS; ++L) rm.tile_base[L] = 0x7FFFFFFF;
}

// widths on which the compiled reference's `1 << 32` (KFAST.h:243) can drop a row's last 32 columns
static inline bool fast_quirk_width(int w) { return w >= 38 && ((w - 38) & 15) == 0; }

// scratch layout: [nms map: total_rows x stride0][row_count: total_rows ints + one per chunk of 64 rows][tile bitmap][row_off][xend]
struct FastLayout { size_t rows, map_bytes, cnt_bytes, bm_words, bm_bytes; };
static FastLayout fast_layout(const LevelTable& lt) {
    FastLayout f; f.rows = 0;
    for (int i = 0; i < lt.n; ++i) f.rows += lt.h[i];
    f.map_bytes = (f.rows * (size_t)lt.stride[0] + 255) & ~(size_t)255;
    f.cnt_bytes = ((f.rows + f.rows / 64 + 2) * sizeof(int) + 255) & ~(size_t)255;
    f.bm_words = ((size_t)(lt.n > 0 ? (lt.w[0] + FT_TW - 1) / FT_TW : 1) + 31) / 32;
    f.bm_bytes = (f.rows * f.bm_words * 4 + 255) & ~(size_t)255;
    return f;
}
size_t fast_scratch_bytes(const LevelTable& lt) {
    const FastLayout f = fast_layout(lt);
    return f.map_bytes + 3 * f.cnt_bytes + f.bm_bytes + 256;
}

int fast_tile_box(int* w, int* h) { *w = Fast2<32, true>::SSTR; *h = Fast2<32, true>::SH; return CLB_OK; }

// standalone = no resize kernel ran in front (stage-level entry point): the row counters are cleared here
int launch_fast(const LevelTable& lt, const LevelMaps& maps, int thresh, int quirk, void* scratch, size_t scratch_bytes,
                clb_keypoint* d_kps, int cap, int* d_total, cudaStream_t s, bool pdl, bool standalone) {
    NvtxRange nvtx_range("clb:fast");
    if (fast_scratch_bytes(lt) > scratch_bytes) { snprintf(g_err, sizeof(g_err), "fast scratch too small"); return CLB_EINVAL; }
    RowMap rm = make_rowmap(lt);
    const FastLayout lay = fast_layout(lt);
    const size_t rows = lay.rows, map_bytes = lay.map_bytes, cnt_bytes = lay.cnt_bytes, bm_bytes = lay.bm_bytes;
    uint8_t* nms = reinterpret_cast<uint8_t*>(scratch);
    // [map][row_count + chunk counts][tile bitmap][row_off][xend]: the counts and the bitmap are cleared together (by the resize kernel)
    rm.bm_words = (int)lay.bm_words;
    int* row_count = reinterpret_cast<int*>(nms + map_bytes);
    uint32_t* bitmap = reinterpret_cast<uint32_t*>(nms + map_bytes + cnt_bytes);
    int* xend = reinterpret_cast<int*>(nms + map_bytes + 2 * cnt_bytes + bm_bytes);
    const int fvar = g_fast_var, fnms = g_fast_nms;
    (void)fnms;
    const bool use_bm = (fvar & 35) == 35;       // bit 5: the tail visits occupied tiles only (needs both v2 kernels)
    if (standalone) CLB_CUDA(cudaMemsetAsync(row_count, 0, cnt_bytes + bm_bytes, s));
    timer("fast").begin(s);
    // the reference's per-row scan end on the widths its `1 << 32` bites (quirk = 1): one small launch over those levels' rows
    QuirkRows qr; qr.n = 0; qr.row_base[0] = 0;
    if (quirk) {
        for (int L = 0; L < lt.n; ++L) {
            if (!fast_quirk_width(lt.w[L]) || lt.h[L] < 7) continue;
            if (lt.w[L] > 38 + 16 * (32 * FQ_WORDS - 1)) { snprintf(g_err, sizeof(g_err), "fast_quirk: level %d is wider than %d px", L, 38 + 16 * (32 * FQ_WORDS - 1)); return CLB_EUNSUPPORTED; }
            rm.quirk_mask |= 1u << L;
            qr.level[qr.n] = L; qr.row_base[qr.n + 1] = qr.row_base[qr.n] + lt.h[L]; ++qr.n;
        }
        if (qr.n > 0) {
            CLB_CUDA(launch_ex(fast_xend_kernel, dim3((qr.row_base[qr.n] + 3) / 4), dim3(128), 0, s, pdl, lt, rm, qr, thresh, xend));
            CLB_LAUNCHED();
        }
    }
#ifdef CLB_LAB
    int* row_off = reinterpret_cast<int*>(nms + map_bytes + cnt_bytes + bm_bytes);
    if (fvar & 1) {
        typedef void (*k2_t)(const LevelTable, const RowMap, const LevelMaps, int, uint8_t*, int, int*, uint32_t*, const int*);
        static const k2_t k2[2][2][2] = {      // [64x32 tiles][8 CTAs per SM][16-bit-lane pre-test]
            {{fast_nms2_kernel<16, 6, false, false>, fast_nms2_kernel<16, 6, true, false>}, {fast_nms2_kernel<16, 8, false, false>, fast_nms2_kernel<16, 8, true, false>}},
            {{fast_nms2_kernel<32, 6, false, false>, fast_nms2_kernel<32, 6, true, false>}, {fast_nms2_kernel<32, 8, false, false>, fast_nms2_kernel<32, 8, true, false>}}};
        const int big = (fvar & 4) ? 1 : 0, th = big ? 32 : 16;
        fill_tile_grid(rm, lt, th);
        const int ntiles = rm.tile_base[lt.n];
        k2_t kern = fvar == 55 ? (k2_t)fast_nms2_kernel<32, 6, true, true> : k2[big][(fvar & 8) ? 1 : 0][(fvar & 16) ? 1 : 0];   // 55 = the product kernel (TMA-staged)
        CLB_CUDA(launch_ex(kern, dim3(ntiles), dim3(256), 0, s, pdl, lt, rm, maps, thresh, nms, fnms, row_count, use_bm ? bitmap : (uint32_t*)nullptr, (const int*)xend));
    } else {
        int tiles = 0;
        for (int L = 0; L < lt.n; ++L) tiles += ((lt.w[L] + FT_TW - 1) / FT_TW) * ((lt.h[L] + FT_TH - 1) / FT_TH);
        fast_nms_kernel<<<tiles, 256, 0, s>>>(lt, rm, thresh, nms, fnms, row_count, xend);
    }
    CLB_LAUNCHED();
    if (fvar & 2) {
        CLB_CUDA(launch_ex(compact2_kernel, dim3(((int)rows + 7) / 8), dim3(256), 0, s, pdl, lt, rm, (const uint8_t*)nms, (const int*)row_count, d_kps, cap, d_total,
                           use_bm ? (const uint32_t*)bitmap : (const uint32_t*)nullptr));
        CLB_LAUNCHED();
    } else {
        row_scan_kernel<<<1, 1024, 0, s>>>(row_count, (int)rows, row_off, d_total);
        CLB_LAUNCHED();
        compact_kernel<<<((int)rows + 7) / 8, 256, 0, s>>>(lt, rm, nms, row_count, row_off, d_kps, cap);
        CLB_LAUNCHED();
        angle_kernel<<<296, 256, 0, s>>>(lt, d_kps, d_total, cap);
        CLB_LAUNCHED();
    }
#else
    fill_tile_grid(rm, lt, 32);
    CLB_CUDA(launch_ex(fast_nms2_kernel<32, 6, true, true>, dim3(rm.tile_base[lt.n]), dim3(256), 0, s, pdl, lt, rm, maps, thresh, nms, 1, row_count, bitmap, (const int*)xend));
    CLB_LAUNCHED();
    CLB_CUDA(launch_ex(compact2_kernel, dim3(((int)rows + 7) / 8), dim3(256), 0, s, pdl, lt, rm, (const uint8_t*)nms, (const int*)row_count, d_kps, cap, d_total, (const uint32_t*)bitmap));
    CLB_LAUNCHED();
#endif
    timer("fast").end(s);
    return CLB_OK;
}

// offset / length (32-bit words) of the row counters + tile bitmap inside the scratch buffer: the resize kernel clears them
void fast_clear_region(const LevelTable& lt, void* scratch, uint32_t** ptr, int* words) {
    const FastLayout f = fast_layout(lt);
    *ptr = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(scratch) + f.map_bytes);
    *words = (int)((f.cnt_bytes + f.bm_bytes) / 4);
}

}  // namespace clb
