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
#include "common.cuh"

namespace clb {

static constexpr int RZ_TX = 32, RZ_TY = 4, RZ_PX = 4, RZ_ROWS = 2;   // thread block 32x4, 4 px x 2 rows per thread
static constexpr int RZ_TILE_W = RZ_TX * RZ_PX;                       // 128 x 8 output pixels per CTA
static constexpr int RZ_TILE_H = RZ_TY * RZ_ROWS;
static_assert(RZ_TX * RZ_TY == 128, "the texel table is filled 2 entries per thread");

// The conversions below stay off the XU pipe (I2F / F2I / FRND / the FCHK of an IEEE division issue at 16 lanes per
// clock per SM and were ~12 per output pixel: the first version of this kernel was bound by them, not by memory).
//   unorm8       RN(v / 255.f) for v in 0..255 as q = v*r, q' = fma(fma(-255, q, v), r, q) with r = RN(1/255): one Newton
//                correction is exactly the correctly rounded quotient for all 256 inputs (checked exhaustively in exact
//                rational arithmetic, tools/check_unorm8.py); v -> float through the 2^23 exponent trick.
//   floor_pos    floor of 0 <= x < 2^22 by a round-down add of 2^23 (the ulp there is 1).
//   round_int    cvt.rni (ties to even) of |x| < 2^22 by a round-to-nearest add of 1.5 * 2^23.
__device__ __forceinline__ float unorm8(uint32_t v) {
    const float fv = __fsub_rn(__uint_as_float(0x4B000000u | v), 8388608.0f);
    const float r = __uint_as_float(0x3B808081u);
    const float q = __fmul_rn(fv, r);
    return __fmaf_rn(__fmaf_rn(-255.0f, q, fv), r, q);
}
__device__ __forceinline__ float floor_pos(float x) { return __fsub_rn(__fadd_rd(x, 8388608.0f), 8388608.0f); }
__device__ __forceinline__ int round_int(float x) { return (int)__float_as_uint(__fadd_rn(x, 12582912.0f)) - 0x4B400000; }
__device__ __forceinline__ int gather_origin(float u) { return (round_int(__fmul_rn(u, 256.0f)) - 128) >> 8; }

__global__ void __launch_bounds__(RZ_TX * RZ_TY)
resize_levels_kernel(const LevelTable lt) {
    // which level / tile does this CTA own?
    int tile = blockIdx.x, L = 1, tx_n = 0;
    for (; L < lt.n; ++L) {
        tx_n = (lt.w[L] + RZ_TILE_W - 1) / RZ_TILE_W;
        const int cnt = tx_n * ((lt.h[L] + RZ_TILE_H - 1) / RZ_TILE_H);
        if (tile < cnt) break;
        tile -= cnt;
    }
    if (L >= lt.n) return;
    const int w0 = lt.w[0], h0 = lt.h[0], s0 = lt.stride[0];
    const uint8_t* __restrict__ src = lt.img[0];
    uint8_t* __restrict__ dst = const_cast<uint8_t*>(lt.img[L]);
    const int dw = lt.w[L], dh = lt.h[L], ds = lt.stride[L];
    const float f = lt.f[L];

    // texel -> float table: the exact unorm8() of every byte value, built once per CTA (2 entries per thread); a lookup is one
    // shared load on the otherwise idle LSU pipe instead of 5 issue slots on the ALU / FMA pipes this kernel is bound by
    __shared__ float lut[256];
    {
        const int t = threadIdx.y * RZ_TX + threadIdx.x;
        lut[t] = unorm8((uint32_t)t);
        lut[t + RZ_TX * RZ_TY] = unorm8((uint32_t)(t + RZ_TX * RZ_TY));
    }
    __syncthreads();
    const int yb = (tile / tx_n) * RZ_TILE_H + threadIdx.y * RZ_ROWS;
    const int xb = (tile % tx_n) * RZ_TILE_W + threadIdx.x * RZ_PX;
    if (yb >= dh || xb >= dw) return;

    // the column terms (footprint origin, weights) are computed once and serve both rows of the thread
    uint32_t cx0[RZ_PX], cx1[RZ_PX];
    float wtx[RZ_PX], ivx[RZ_PX];
#pragma unroll
    for (int k = 0; k < RZ_PX; ++k) {
        const int x = xb + k;
        const float fx = __fmaf_rn((float)(uint32_t)x + 0.5f, f, -0.5f);
        const float flx = floor_pos(fx);          // fy, fx >= 0.5 * f - 0.5 >= 0 for f >= 1
        wtx[k] = __fsub_rn(fx, flx); ivx[k] = __fsub_rn(1.0f, wtx[k]);
        const int gx = gather_origin(__fadd_rn(fx, 0.5f));
        cx0[k] = (uint32_t)min(max(gx, 0), w0 - 1); cx1[k] = (uint32_t)min(max(gx + 1, 0), w0 - 1);
    }
#pragma unroll
    for (int rr = 0; rr < RZ_ROWS; ++rr) {
        const int y = yb + rr;
        if (y >= dh) break;
        const float fy = __fmaf_rn((float)(uint32_t)y + 0.5f, f, -0.5f);
        const float fly = floor_pos(fy);
        const float wty = __fsub_rn(fy, fly), ivy = __fsub_rn(1.0f, wty);
        const int gy = gather_origin(__fadd_rn(fy, 0.5f));
        // 32-bit offsets into level 0 (a level is far below 4 GiB): one IMAD.WIDE.U32 per load instead of 64-bit chains
        const uint32_t o0 = (uint32_t)min(max(gy, 0), h0 - 1) * (uint32_t)s0, o1 = (uint32_t)min(max(gy + 1, 0), h0 - 1) * (uint32_t)s0;
        uint32_t packed = 0;
#pragma unroll
        for (int k = 0; k < RZ_PX; ++k) {
            const float p00 = lut[__ldg(src + (o0 + cx0[k]))], p10 = lut[__ldg(src + (o0 + cx1[k]))];
            const float p01 = lut[__ldg(src + (o1 + cx0[k]))], p11 = lut[__ldg(src + (o1 + cx1[k]))];
            const float xa = __fmaf_rn(ivx[k], p00, __fmul_rn(wtx[k], p10));
            const float xbv = __fmaf_rn(ivx[k], p01, __fmul_rn(wtx[k], p11));
            const float res = __fmaf_rn(__fmaf_rn(ivy, xa, __fmul_rn(wty, xbv)), 255.0f, 0.5f);
            packed |= (__float_as_uint(__fadd_rd(res, 8388608.0f)) & 0xFFu) << (8 * k);   // res in [0.5, 255.5]: trunc = floor
        }
        uint8_t* o = dst + (size_t)y * ds + xb;
        if (xb + RZ_PX <= dw && (((uintptr_t)o) & 3) == 0) {
            *reinterpret_cast<uint32_t*>(o) = packed;
        } else {
            for (int k = 0; k < RZ_PX && xb + k < dw; ++k) o[k] = (uint8_t)(packed >> (8 * k));
        }
    }
}

int launch_resize(const LevelTable& lt, cudaStream_t s) {
    int tiles = 0;
    for (int L = 1; L < lt.n; ++L)
        tiles += ((lt.w[L] + RZ_TILE_W - 1) / RZ_TILE_W) * ((lt.h[L] + RZ_TILE_H - 1) / RZ_TILE_H);
    if (tiles == 0) return CLB_OK;
    timer("resize").begin(s);
    resize_levels_kernel<<<tiles, dim3(RZ_TX, RZ_TY), 0, s>>>(lt);
    timer("resize").end(s);
    CLB_LAUNCHED();
    return CLB_OK;
}

}  // namespace clb
