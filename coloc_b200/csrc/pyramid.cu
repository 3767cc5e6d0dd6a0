// pyramid.cu -- image pyramid: every level i >= 1 is a bilinear down-scale of LEVEL 0.
//
// Replaces CUDALERP (reference src/CUDALERP.cu:153-183) as driven by GPUDetector.hpp:249-255:
// 7 launches on 7 streams + a zero-fill each become ONE launch covering all levels.  The float
// expression tree is the one nvcc emits for the reference kernel (PTX-checked):
//   fx = fma(x+.5, f, -.5); wt = fx - floor(fx); xa = fma(1-wtx, p00, wtx*p10); ...
//   res = fma(fma(1-wty, xa, wty*xb), 255, .5); store (uint8) truncation
// with texels read as p/255.f (cudaReadModeNormalizedFloat) and clamp-to-edge addressing.  The 2x2
// footprint is the one the texture unit picks for tex2Dgather at u = fx+.5: u rounded to 8 fractional
// bits, origin floor(u_fix - .5) (measured on B200 against the reference kernel: bit-exact; plain
// floor(fx) differs at ~0.1 % of the pixels by up to 118 grey levels).
//
// Design (not a port).  The first version read its four texels per pixel with byte-wide LDG and normalised them through a
// 256-entry shared-memory table; ncu put it at 88 % issue-active with the LSU data pipe at 73 % -- 32 lanes with 32 random table
// indices are a ~3.5-way bank conflict on every lookup.  This version:
//   * column / row terms (footprint origin, weight, edge flag) depend on (level, x) or (level, y) only: resize_tables_kernel
//     writes them once per frame geometry, the detector caches them, a pixel costs one table load instead of ~19 instructions;
//   * the level-0 footprint of a 64 x 16 output tile is staged by ONE TMA box copy (cp.async.bulk.tensor.2d, SASS UTMALDG; the
//     box starts on a 16-byte boundary, which the copy engine insists on), so a texel is a byte-wide shared load whose address
//     is "row base + column offset" -- no clamps, no 64-bit address arithmetic, 4 loads per pixel instead of 4 + 4;
//   * texel bytes become floats as 0x4B0000vv = 2^23 + v and the exact v/255 Newton step runs on PACKED fp32
//     (add / mul / fma.f32x2, SASS FADD2 / FMUL2 / FFMA2: two pixels per instruction), as do the lerps;
//   * the kernel also clears the FAST stage's row counters (one memset node less per frame) and takes part in the frame's
//     programmatic-dependent-launch chain.
#include "common.cuh"

namespace clb {

static constexpr int RZ_TW = 64, RZ_TH = 32;          // output tile: 32 lanes x 2 px, 4 warps x 8 rows
static constexpr int RZ_THREADS = 128;
static constexpr int RZ_MAX_BOX = 256 * 120;          // staged bytes per tile at most (level 7: 256 x 115)

//   unorm8       RN(v / 255.f) for v in 0..255 as q = v*r, q' = fma(fma(-255, q, v), r, q) with r = RN(1/255): one Newton
//                correction is exactly the correctly rounded quotient for all 256 inputs (checked exhaustively in exact
//                rational arithmetic, tools/check_unorm8.py).
//   floor_pos    floor of 0 <= x < 2^22 by a round-down add of 2^23 (the ulp there is 1).
//   round_int    cvt.rni (ties to even) of |x| < 2^22 by a round-to-nearest add of 1.5 * 2^23.
__device__ __forceinline__ float floor_pos(float x) { return __fsub_rn(__fadd_rd(x, 8388608.0f), 8388608.0f); }
__device__ __forceinline__ int round_int(float x) { return (int)__float_as_uint(__fadd_rn(x, 12582912.0f)) - 0x4B400000; }
__device__ __forceinline__ int gather_origin(float u) { return (round_int(__fmul_rn(u, 256.0f)) - 128) >> 8; }

// one table entry per output column / row of every level >= 1: x = clamped footprint origin in level 0 (bit 31: the +1
// neighbour is clamped onto it, i.e. the origin is the last column / row), y = weight of the +1 neighbour (float bits)
__global__ void __launch_bounds__(256)
resize_tables_kernel(const LevelTable lt, const ResizePlan rp, uint2* __restrict__ tab) {
    const int L = blockIdx.y + 1;
    if (L >= lt.n) return;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int dw = lt.w[L], dh = lt.h[L];
    if (i >= dw + dh) return;
    const bool is_col = i < dw;
    const int k = is_col ? i : i - dw;
    const int n0 = is_col ? lt.w[0] : lt.h[0];
    const float f = lt.f[L];
    const float fx = __fmaf_rn((float)(uint32_t)k + 0.5f, f, -0.5f);      // >= 0.5 f - 0.5 >= 0 for f >= 1
    const float wt = __fsub_rn(fx, floor_pos(fx));
    const int g = gather_origin(__fadd_rn(fx, 0.5f));
    const int g0 = min(max(g, 0), n0 - 1), g1 = min(max(g + 1, 0), n0 - 1);
    tab[(is_col ? rp.col_off[L] : rp.row_off[L]) + k] = make_uint2((uint32_t)g0 | (g1 == g0 ? 0x80000000u : 0u), __float_as_uint(wt));
}

__device__ __forceinline__ float2 unorm8x2(uint32_t a, uint32_t b) {      // a, b = 0x4B0000vv
    const float2 fv = __fadd2_rn(make_float2(__uint_as_float(a), __uint_as_float(b)), make_float2(-8388608.0f, -8388608.0f));
    const float r = __uint_as_float(0x3B808081u);
    const float2 q = __fmul2_rn(fv, make_float2(r, r));
    return __ffma2_rn(__ffma2_rn(make_float2(-255.0f, -255.0f), q, fv), make_float2(r, r), q);
}

__global__ void __launch_bounds__(RZ_THREADS)
resize_levels_kernel(const LevelTable lt, const ResizePlan rp, const __grid_constant__ LevelMaps maps,
                     uint32_t* __restrict__ clear, int clear_words) {
    extern __shared__ __align__(128) uint8_t rz_smem[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ __align__(16) uint4 rowterm[RZ_TH];      // per output row of the tile: {row offset, +1-row offset, wty, 1 - wty}

    // nothing below may run before the grid in front has finished: level 0 comes from it, and on the first frame of a geometry
    // so do the column / row tables (resize_tables_kernel) -- every thread waits, not only the one that issues the TMA copy
    ptx_grid_dep_wait();
    // the FAST stage's row counters and tile bitmap start every frame at zero (no kernel of this frame has touched them yet)
    for (int i = blockIdx.x * RZ_THREADS + threadIdx.x; i < clear_words; i += gridDim.x * RZ_THREADS) clear[i] = 0u;

    // which level / tile does this CTA own?  (host-computed bases and an exact multiply-shift division: no divisions here)
    int L = 1;
    while (L + 1 < lt.n && (int)blockIdx.x >= rp.tile_base[L + 1]) ++L;
    const int tile = blockIdx.x - rp.tile_base[L];
    const int tyi = rp.tiles_x[L] == 1 ? tile : (int)__umulhi((uint32_t)tile, rp.tiles_x_magic[L]), txi = tile - tyi * rp.tiles_x[L];
    const int dw = lt.w[L], dh = lt.h[L], ds = lt.stride[L];
    uint8_t* __restrict__ dst = const_cast<uint8_t*>(lt.img[L]);
    const int tw = rp.tile_w[L], th = rp.tile_h[L];              // 64 x 32 unless the level's scale makes that footprint too large
    const int tx0 = txi * tw, ty0 = tyi * th;
    const int bw = rp.box_w[L], bh = rp.box_h[L];
    const uint2* __restrict__ ctab = rp.tab + rp.col_off[L];
    const uint2* __restrict__ rtab = rp.tab + rp.row_off[L];
    // box origin: the footprint origin of the tile's first column / row, the column rounded down to 16 bytes (TMA start alignment)
    const int bx0 = (int)(__ldg(&ctab[tx0]).x & 0x7FFFFFFFu) & ~15;
    const int by0 = (int)(__ldg(&rtab[ty0]).x & 0x7FFFFFFFu);
    if (threadIdx.x == 0) {
        ptx_mbar_init(&bar, 1);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        ptx_mbar_expect_tx(&bar, (uint32_t)(bw * bh));
        ptx_tma_load_2d(rz_smem, &maps.resize[L], bx0, by0, &bar);      // rz_smem[r][j] = level0[by0 + r][bx0 + j]
    }
    ptx_grid_dep_launch();
    if (threadIdx.x < RZ_TH) {
        const int y = min(ty0 + (int)threadIdx.x, dh - 1);
        const uint2 re = __ldg(&rtab[y]);
        const uint32_t r0 = ((re.x & 0x7FFFFFFFu) - (uint32_t)by0) * (uint32_t)bw;
        const float wty = __uint_as_float(re.y);
        rowterm[threadIdx.x] = make_uint4(r0, r0 + ((re.x >> 31) ? 0u : (uint32_t)bw), re.y, __float_as_uint(__fsub_rn(1.0f, wty)));
    }

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int x = tx0 + 2 * lane;                 // this thread's two columns x, x + 1
    constexpr int RPW = RZ_TH / (RZ_THREADS / 32);   // rows per warp
    // column terms: box-relative offsets of the texel pair (the +1 neighbour collapses onto the origin at the right edge), weights
    uint32_t c0[2], c1[2];
    float wtx[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const uint2 e = __ldg(&ctab[min(x + k, dw - 1)]);
        c0[k] = (e.x & 0x7FFFFFFFu) - (uint32_t)bx0;
        c1[k] = c0[k] + ((e.x >> 31) ? 0u : 1u);
        wtx[k] = __uint_as_float(e.y);
    }
    const float2 wx = make_float2(wtx[0], wtx[1]);
    const float2 ix = __fadd2_rn(make_float2(1.0f, 1.0f), make_float2(-wtx[0], -wtx[1]));
    __syncthreads();                              // barrier initialised, row terms written
    ptx_mbar_wait(&bar, 0);

    if (x >= dw || 2 * lane >= tw) return;
    const bool pair = x + 1 < dw && 2 * lane + 1 < tw;
    const int rows = min(min(RPW, th - RPW * warp), dh - (ty0 + RPW * warp));
    uint8_t* o = dst + (size_t)(ty0 + RPW * warp) * ds + x;
#pragma unroll 2
    for (int rr = 0; rr < rows; ++rr, o += ds) {
        const uint4 rt = rowterm[RPW * warp + rr];
        const uint8_t* r0 = rz_smem + rt.x;
        const uint8_t* r1 = rz_smem + rt.y;
        const float wty = __uint_as_float(rt.z), ivy = __uint_as_float(rt.w);
        const uint32_t bias = 0x4B000000u;
        const float2 p00 = unorm8x2(bias | r0[c0[0]], bias | r0[c0[1]]);
        const float2 p10 = unorm8x2(bias | r0[c1[0]], bias | r0[c1[1]]);
        const float2 p01 = unorm8x2(bias | r1[c0[0]], bias | r1[c0[1]]);
        const float2 p11 = unorm8x2(bias | r1[c1[0]], bias | r1[c1[1]]);
        const float2 xa = __ffma2_rn(ix, p00, __fmul2_rn(wx, p10));
        const float2 xb = __ffma2_rn(ix, p01, __fmul2_rn(wx, p11));
        const float2 v = __ffma2_rn(make_float2(ivy, ivy), xa, __fmul2_rn(make_float2(wty, wty), xb));
        const float2 res = __ffma2_rn(v, make_float2(255.0f, 255.0f), make_float2(0.5f, 0.5f));
        const float2 fl = __fadd2_rd(res, make_float2(8388608.0f, 8388608.0f));        // res in [0.5, 255.5]: trunc = floor
        const uint32_t b01 = __byte_perm(__float_as_uint(fl.x), __float_as_uint(fl.y), 0x7740);   // byte 0 of each
        if (pair) *reinterpret_cast<uint16_t*>(o) = (uint16_t)b01;      // x is even, rows are 128-B aligned
        else *o = (uint8_t)b01;
    }
}

size_t resize_table_entries(const LevelTable& lt) {
    size_t n = 0;
    for (int L = 1; L < lt.n; ++L) n += (size_t)lt.w[L] + lt.h[L];
    return n + 8;
}

// source-box geometry of every level: the widest / tallest footprint a tile can have, from the same float formulas.  A tile's
// footprint spans (T - 1) * f + 2 texels, + 1 for the origin's rounding, + 15 for the box start rounded down to 16 bytes; the
// tile shrinks on levels whose scale would push that past what one TMA box (256 wide) or the staging buffer holds.
int resize_plan(const LevelTable& lt, ResizePlan& rp) {
    int off = 0;
    for (int L = 0; L < CLB_MAX_LEVELS; ++L) { rp.box_w[L] = rp.box_h[L] = 16; rp.col_off[L] = rp.row_off[L] = 0; rp.tile_w[L] = RZ_TW; rp.tile_h[L] = RZ_TH; rp.tiles_x[L] = 1; rp.tiles_x_magic[L] = 0xFFFFFFFFu; }
    for (int L = 0; L <= CLB_MAX_LEVELS; ++L) rp.tile_base[L] = 0;
    for (int L = 1; L < lt.n; ++L) {
        const double f = lt.f[L];
        int tw = RZ_TW, th = RZ_TH;
        while (tw > 2 && (int)((tw - 1) * f) + 19 > 256) tw -= 2;
        while (th > 1 && (((int)((tw - 1) * f) + 19 + 15) & ~15) * ((int)((th - 1) * f) + 4) > RZ_MAX_BOX) --th;
        const int bw = ((int)((tw - 1) * f) + 19 + 15) & ~15, bh = (int)((th - 1) * f) + 4;
        if (bw > 256 || bh > 256 || bw * bh > RZ_MAX_BOX) {
            snprintf(g_err, sizeof(g_err), "pyramid level %d: scale %.3f needs a %d x %d source box for a 2 x 1 tile", L, f, bw, bh);
            return CLB_EUNSUPPORTED;
        }
        rp.tile_w[L] = tw; rp.tile_h[L] = th;
        rp.box_w[L] = bw; rp.box_h[L] = bh;
        // tile index -> (row, column) of the level's tile grid by multiply-shift, checked exhaustively here
        const int txn = lt.w[L] > 0 ? (lt.w[L] + tw - 1) / tw : 1, tyn = lt.h[L] > 0 ? (lt.h[L] + th - 1) / th : 0;
        const uint32_t magic = txn == 1 ? 0xFFFFFFFFu : (uint32_t)(0x100000000ull / (uint64_t)txn) + 1u;
        for (long long t = 0; t < (long long)txn * tyn; ++t) {
            const uint32_t q = txn == 1 ? (uint32_t)t : (uint32_t)(((uint64_t)t * magic) >> 32);
            if (q != (uint32_t)(t / txn)) { snprintf(g_err, sizeof(g_err), "pyramid level %d: tile grid %d x %d is too large for the tile decode", L, txn, tyn); return CLB_EUNSUPPORTED; }
        }
        rp.tiles_x[L] = txn; rp.tiles_x_magic[L] = magic;
        rp.tile_base[L + 1] = rp.tile_base[L] + txn * tyn;
        rp.col_off[L] = off; off += lt.w[L];
        rp.row_off[L] = off; off += lt.h[L];
    }
    return CLB_OK;
}

int launch_resize_tables(const LevelTable& lt, const ResizePlan& rp, uint2* d_tab, cudaStream_t s) {
    if (lt.n < 2) return CLB_OK;
    int most = 1;
    for (int L = 1; L < lt.n; ++L) most = lt.w[L] + lt.h[L] > most ? lt.w[L] + lt.h[L] : most;
    resize_tables_kernel<<<dim3((most + 255) / 256, lt.n - 1), 256, 0, s>>>(lt, rp, d_tab);
    CLB_LAUNCHED();
    return CLB_OK;
}

int launch_resize(const LevelTable& lt, const ResizePlan& rp, const LevelMaps& maps, uint32_t* clear, int clear_words, cudaStream_t s, bool pdl) {
    NvtxRange nvtx_range("clb:resize");
    int smem = 0;
    const int tiles = lt.n > 1 ? rp.tile_base[lt.n] : 0;
    for (int L = 1; L < lt.n; ++L) {
        const int need = (rp.box_w[L] * rp.box_h[L] + 127) & ~127;
        smem = need > smem ? need : smem;
    }
    if (tiles == 0) {
        if (clear_words > 0) CLB_CUDA(cudaMemsetAsync(clear, 0, (size_t)clear_words * 4, s));
        return CLB_OK;
    }
    static bool attr_done[64] = {false};
    int dev = 0;
    CLB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_done[dev]) {
        CLB_CUDA(cudaFuncSetAttribute(resize_levels_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, RZ_MAX_BOX + 256));
        if (dev >= 0 && dev < 64) attr_done[dev] = true;
    }
    timer("resize").begin(s);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(tiles); cfg.blockDim = dim3(RZ_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    CLB_CUDA(cudaLaunchKernelEx(&cfg, resize_levels_kernel, lt, rp, maps, clear, clear_words));
    timer("resize").end(s);
    CLB_LAUNCHED();
    return CLB_OK;
}

}  // namespace clb
