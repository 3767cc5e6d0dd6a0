// match_tc.cu -- brute-force 512-bit Hamming 2-NN on the 5th-gen tensor cores (tcgen05 + TMEM).
//
// Same contract as match.cu (reference src/CUDAK2NN.cu:42-81: ascending scan, first row attaining
// the minimum wins, `second` = second smallest distance, == best on ties); this file only changes
// HOW the nQ x nT distance matrix is produced.  sm_100a has no 1-bit MMA, but
//     hamming(q, t) = (512 - <q', t'>) / 2     with q', t' in {-1, +1}^512
// and <q', t'> is a K = 512 dot product that tcgen05.mma kind::f8f6f4 computes exactly with E4M3
// operands (+-1.0) and fp32 accumulation (every partial sum is an integer of magnitude <= 512).
// One SM then produces 16 distances per clock (8192 MAC/clk) against ~1.4 on the POPC/LOP3 pipes.
//
// Pipeline (one CTA per SM, 10 warps, no clusters):
//   expand_kernel   packed 64-B rows -> +-1 E4M3 bytes, written ONCE in the exact shared-memory image
//                   the MMA wants: tiles of 128 rows, 4 K-blocks of [128 rows x 128 B], 128-B swizzle
//                   (16-B chunk c of row r stored at chunk c ^ (r & 7)).  A tile K-block is 16 KB
//                   contiguous, so plain 1-D bulk copies (cp.async.bulk, SASS UBLKCP) stage it -- no
//                   tensor maps.
//   match_tc_kernel warp 0 lane 0: producer -- 2 query tiles (128 KB, resident for the whole sweep),
//                   then a 6-stage ring of DB K-blocks.  warp 1 lane 0: 32 tcgen05.mma (M=128, N=128,
//                   K=32) per DB tile into one of two 256-column TMEM buffers; tcgen05.commit frees
//                   ring slots / publishes the buffer.  warps 2-9: epilogue, one query per thread:
//                   tcgen05.ld 32 columns at a time; a group of 8 distances is skipped with 4 FMNMX +
//                   1 compare unless one of them beats the thread's current second best (after the
//                   first few tiles that is ~1 % of the groups), otherwise the packed (distance, row)
//                   keys go through the same 3-op best/second update as match.cu.
//   The per-(chunk, query) key pairs land in the same `partial` array match.cu's merge kernel folds.
#include "common.cuh"
#include <vector>
#include <algorithm>

namespace clb {

std::atomic<int> g_match_tc{-1};        // -1 auto, 0 never, 1 always (clb_set_param "match_tc")
std::atomic<int> g_match_tc_chunks{0};  // 0 = heuristic
#ifdef CLB_LAB      // the lab build (make lab) keeps the superseded kernel generations and the timing modes for A/B
std::atomic<int> g_match_tc_dbg{0};     // timing experiments: 1 no TMEM reads, 2 TMEM reads only, 3 no MMAs, ... (see tc4_cta)
std::atomic<int> g_match_tc_ver{4};     // 1 / 2 = E4M3 operands, queries in smem / TMEM; 3 / 4 = E2M1 operands (kind::mxf4), queries in smem / TMEM
#else
static constexpr int g_match_tc_ver = 4;   // the product build carries the v4 kernel only
#endif

static constexpr int TC_TILE = 128;                      // rows per operand tile (UMMA M and N)
static constexpr int TC_KB = 4;                          // K-blocks of 128 B per row (512 E4M3)
static constexpr int TC_KB_BYTES = TC_TILE * 128;        // 16 KB
static constexpr int TC_TILE_BYTES = TC_KB * TC_KB_BYTES;   // 64 KB
static constexpr int TC_QTILES = 2;                      // query tiles per CTA
static constexpr int TC_QROWS = TC_QTILES * TC_TILE;     // 256 queries per CTA
static constexpr int TC_STAGES = 6;
static constexpr int TC_EPI_WARPS = 8;
static constexpr int TC_THREADS = 32 * (2 + TC_EPI_WARPS);
static constexpr int TC_KEY_SHIFT = 22;                  // same key as match.cu: dist << 22 | row-in-chunk
static constexpr int TC_MAX_TILES_PER_CHUNK = (1 << TC_KEY_SHIFT) / TC_TILE;
static constexpr int TC_SMEM_A = TC_QTILES * TC_TILE_BYTES;          // 131072
static constexpr int TC_SMEM_B = TC_STAGES * TC_KB_BYTES;            //  98304
static constexpr int TC_SMEM_BAR = 256;
static constexpr int TC_SMEM_TOTAL = TC_SMEM_A + TC_SMEM_B + TC_SMEM_BAR + 1024;   // + alignment slack
// kind::f8f6f4 instruction descriptor: D=F32 (bits 4-5 = 1), A=B=E4M3 (0), K-major both, N>>3 at 17, M>>4 at 24
static constexpr uint32_t TC_IDESC = (1u << 4) | ((uint32_t)(TC_TILE >> 3) << 17) | ((uint32_t)(TC_TILE >> 4) << 24);

// ---- PTX helpers ----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t tc_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void tc_mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void tc_mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tc_mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// non-blocking phase test
__device__ __forceinline__ bool tc_mbar_test(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\t"
                 "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                 "selp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
// bounded wait: a protocol bug traps (launch error) instead of hanging the GPU
__device__ __forceinline__ void tc_mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok, spins = 0;
    long long t0 = 0;
    for (;;) {
        asm volatile("{\n\t.reg .pred p;\n\t"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
        if (ok) return;
        if ((++spins & 1023u) == 0) {
            const long long now = clock64();
            if (t0 == 0) t0 = now;
            else if (now - t0 > 8000000000ll) __trap();
        }
    }
}
// latency-critical wait: spins on the non-suspending test (a suspended try_wait is woken late)
__device__ __forceinline__ void tc_mbar_spin(uint32_t bar, uint32_t parity) {
    uint32_t spins = 0;
    long long t0 = 0;
    while (!tc_mbar_test(bar, parity)) {
        if ((++spins & 4095u) == 0) {
            const long long now = clock64();
            if (t0 == 0) t0 = now;
            else if (now - t0 > 8000000000ll) __trap();
        }
    }
}
__device__ __forceinline__ void tc_bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
// one lane of a converged warp (the same one every time for a full mask); lets ptxas keep the
// uniform-datapath instructions (UTCQMMA, UBLKCP, UTCBAR) out of per-lane ELECT/BRA loops
__device__ __forceinline__ bool tc_elect_one() {
    uint32_t pred = 0;
    asm volatile("{\n\t.reg .pred px;\n\telect.sync _|px, 0xFFFFFFFF;\n\t@px mov.s32 %0, 1;\n\t}" : "+r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// K-major, 128-B swizzle shared-memory matrix descriptor: rows of 128 B, 8-row groups 1024 B apart
__device__ __forceinline__ uint64_t tc_desc(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(1024u >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ void tc_mma(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t accumulate, uint32_t idesc = TC_IDESC) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// 32 lanes x 32 consecutive 32-bit columns: thread t gets lane (base lane + t)
__device__ __forceinline__ void tc_ld32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                   "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                   "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                 : "r"(taddr) : "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ float tc_max3(float a, float b, float c) { return fmaxf(fmaxf(a, b), c); }

#ifdef CLB_LAB
// ---- operand expansion ------------------------------------------------------------------------
// 4 descriptor bits -> 4 E4M3 bytes: bit 1 -> 0x38 (+1.0), bit 0 -> 0xB8 (-1.0)
__device__ __forceinline__ uint32_t tc_nib(uint32_t x) {
    return 0xB8B8B8B8u ^ (((x * 0x00204081u) & 0x01010101u) << 7);
}
// One thread per 16-byte output chunk, in output order (coalesced stores).  n_pad rows (a multiple
// of 128 x `tile_multiple`); rows >= n are zero (E4M3 0.0) and masked out again by the consumer.
__global__ void __launch_bounds__(256)
expand_kernel(const uint32_t* __restrict__ rows, int n, size_t n_chunks16, uint4* __restrict__ out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_chunks16) return;
    const int cpos = (int)(i & 7), r = (int)((i >> 3) & 127), kb = (int)((i >> 10) & 3);
    const size_t tile = i >> 12;
    const int c = cpos ^ (r & 7);
    const size_t row = tile * TC_TILE + r;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (row < (size_t)n) {
        const uint32_t w = __ldg(rows + row * 16 + kb * 4 + (c >> 1));
        const uint32_t h = (w >> ((c & 1) * 16)) & 0xFFFFu;
        v.x = tc_nib(h & 15u); v.y = tc_nib((h >> 4) & 15u); v.z = tc_nib((h >> 8) & 15u); v.w = tc_nib(h >> 12);
    }
    out[i] = v;
}

// ---- main kernel ------------------------------------------------------------------------------
// per-thread epilogue state
struct TcState { uint32_t best, second; float thr; };

// 32 accumulator columns (DB rows row0 .. row0+31 of the chunk) for this thread's query, in two groups of 16:
// a group costs 8 three-input max ops + 1 compare + 1 branch unless one of its distances beats the current second best
__device__ __forceinline__ void tc_process32(const float (&v)[32], uint32_t row0, TcState& st) {
#pragma unroll
    for (int g = 0; g < 2; ++g) {
        const float* x = &v[16 * g];
        const float m0 = tc_max3(x[0], x[1], x[2]), m1 = tc_max3(x[3], x[4], x[5]), m2 = tc_max3(x[6], x[7], x[8]);
        const float m3 = tc_max3(x[9], x[10], x[11]), m4 = tc_max3(x[12], x[13], x[14]);
        const float m = tc_max3(tc_max3(m0, m1, m2), tc_max3(m3, m4, x[15]), -1e30f);
        if (m > st.thr) {
            // rare after the first tiles, and then usually one element of one lane: descend by sub-groups of 4 so the
            // warp does not replay 16 predicated updates for it
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float m4 = fmaxf(tc_max3(x[4 * q], x[4 * q + 1], x[4 * q + 2]), x[4 * q + 3]);
                if (m4 > st.thr) {
#pragma unroll
                    for (int e = 4 * q; e < 4 * q + 4; ++e) {
                        const float xe = x[e];
                        if (xe > st.thr) {   // thr only rises, so testing against the group-entry value is exact
                            const uint32_t d = (uint32_t)(512 - __float2int_rn(xe)) >> 1;
                            const uint32_t key = (d << TC_KEY_SHIFT) | (row0 + 16 * g + e);
                            const uint32_t mx = max(key, st.best);
                            st.second = min(st.second, mx);
                            st.best = min(st.best, key);
                        }
                    }
                }
            }
            st.thr = st.second == 0xFFFFFFFFu ? -1e30f : (float)(512 - 2 * (int)(st.second >> TC_KEY_SHIFT));
        }
    }
}

#endif   // CLB_LAB (v1-v3 expansion and packed-key epilogue state)

// v4 epilogue state in the accumulator's own domain: the two largest dot products seen (= the two smallest distances,
// d = (512 - x) / 2) and the row of the largest.  No conversion and no key packing inside the sweep: an update is
// 2 FMNMX + FSETP + SEL + FMNMX per element, branch-free inside a sub-group of 4.  (The packed-key version converted
// every updating element with F2I on the XU pipe; while the per-query thresholds are still open -- the first ~100
// tiles of a sweep, i.e. most of a frame-pair-sized call -- that path was the long pole: 4 800 cycles for the first
// tile, ~3 200 for each of the next seven, against 1 148 in steady state.)
struct TcStateF { float b, s; uint32_t idx; };

__device__ __forceinline__ void tc_process32f(const float (&v)[32], uint32_t row0, TcStateF& st) {
    // maxima of the eight sub-groups of 4 first (independent: they pipeline), then one test per group of 16 and, behind it,
    // one per sub-group: the update path re-uses the sub-group maxima instead of recomputing them
    float mq[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) mq[q] = fmaxf(tc_max3(v[4 * q], v[4 * q + 1], v[4 * q + 2]), v[4 * q + 3]);
#pragma unroll
    for (int g = 0; g < 2; ++g) {
        const float m = fmaxf(tc_max3(mq[4 * g], mq[4 * g + 1], mq[4 * g + 2]), mq[4 * g + 3]);
        if (m > st.s) {
#pragma unroll
            for (int q = 4 * g; q < 4 * g + 4; ++q) {
                if (mq[q] > st.s) {       // s only rises: elements that do not beat it change nothing
#pragma unroll
                    for (int e = 4 * q; e < 4 * q + 4; ++e) {
                        const float xe = v[e];
                        st.s = fmaxf(st.s, fminf(st.b, xe));
                        st.idx = xe > st.b ? row0 + (uint32_t)e : st.idx;     // strict: the first row attaining the minimum wins
                        st.b = fmaxf(st.b, xe);
                    }
                }
            }
        }
    }
}

// packed keys of a finished sweep (the format the merge kernels take): distance << 22 | row; the second key's row bits
// are all ones so that second >= best also holds on equal distances
__device__ __forceinline__ uint2 tc_statef_keys(const TcStateF& st) {
    const uint32_t kb = st.b < -600.f ? 0xFFFFFFFFu : ((((uint32_t)(512 - (int)st.b)) >> 1) << TC_KEY_SHIFT) | st.idx;
    const uint32_t ks = st.s < -600.f ? 0xFFFFFFFFu : ((((uint32_t)(512 - (int)st.s)) >> 1) << TC_KEY_SHIFT) | ((1u << TC_KEY_SHIFT) - 1u);
    return make_uint2(kb, ks);
}

#ifdef CLB_LAB
// One 128-column accumulator tile: four 32-column TMEM loads, the next one in flight while the
// current one is scanned; the buffer is handed back to the MMA warp as soon as the last load landed.
template <bool MASKED>
__device__ __forceinline__ void tc_epilogue_tile(uint32_t t, uint32_t row0, int valid, uint32_t bar_release, int lane, TcState& st) {
    float va[32], vb[32];
    tc_ld32(t, va);
    tc_wait_ld();
    tc_ld32(t + 32, vb);
    if (MASKED) {
#pragma unroll
        for (int i = 0; i < 32; ++i) if (i >= valid) va[i] = -1e30f;
    }
    tc_process32(va, row0, st);
    tc_wait_ld();
    tc_ld32(t + 64, va);
    if (MASKED) {
#pragma unroll
        for (int i = 0; i < 32; ++i) if (32 + i >= valid) vb[i] = -1e30f;
    }
    tc_process32(vb, row0 + 32, st);
    tc_wait_ld();
    tc_ld32(t + 96, vb);
    if (MASKED) {
#pragma unroll
        for (int i = 0; i < 32; ++i) if (64 + i >= valid) va[i] = -1e30f;
    }
    tc_process32(va, row0 + 64, st);
    tc_wait_ld();
    tc_fence_before();
    __syncwarp();
    if (lane == 0) tc_mbar_arrive(bar_release);
    if (MASKED) {
#pragma unroll
        for (int i = 0; i < 32; ++i) if (96 + i >= valid) vb[i] = -1e30f;
    }
    tc_process32(vb, row0 + 96, st);
}

// Same tile, all four loads issued at once: the accumulator buffer goes back to the MMA warp after ONE TMEM round trip
// and the scan of the 128 distances overlaps the next MMA pass (used where the tensor pipe outruns a serial epilogue).
template <bool MASKED>
__device__ __forceinline__ void tc_epilogue_tile_wide(uint32_t t, uint32_t row0, int valid, uint32_t bar_release, int lane, TcState& st) {
    float v0[32], v1[32], v2[32], v3[32];
    tc_ld32(t, v0); tc_ld32(t + 32, v1); tc_ld32(t + 64, v2); tc_ld32(t + 96, v3);
    tc_wait_ld();
    tc_fence_before();
    __syncwarp();
    if (lane == 0) tc_mbar_arrive(bar_release);
    if (MASKED) {
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            if (i >= valid) v0[i] = -1e30f;
            if (32 + i >= valid) v1[i] = -1e30f;
            if (64 + i >= valid) v2[i] = -1e30f;
            if (96 + i >= valid) v3[i] = -1e30f;
        }
    }
    tc_process32(v0, row0, st);
    tc_process32(v1, row0 + 32, st);
    tc_process32(v2, row0 + 64, st);
    tc_process32(v3, row0 + 96, st);
}

__global__ void __launch_bounds__(TC_THREADS, 1)
match_tc_kernel(const uint8_t* __restrict__ Qx, int nQ, const uint8_t* __restrict__ Tx, int nT, int tiles_per_chunk,
                uint2* __restrict__ partial, int dbg) {
    extern __shared__ uint8_t tc_smem_raw[];
    const uint32_t raw = tc_smem_u32(tc_smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;               // 128-B swizzle atoms need 1024-B alignment
    const uint32_t sA = base, sB = base + TC_SMEM_A, sBar = sB + TC_SMEM_B;
    const uint32_t bar_full = sBar, bar_empty = sBar + 8 * TC_STAGES, bar_a = sBar + 16 * TC_STAGES;
    const uint32_t bar_tfull = bar_a + 8, bar_tempty = bar_tfull + 16, tmem_slot = bar_tempty + 16;
    uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(tc_smem_raw + (tmem_slot - raw));

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int total_tiles = (nT + TC_TILE - 1) / TC_TILE;
    const int tile0 = blockIdx.y * tiles_per_chunk;
    const int ntiles = min(tiles_per_chunk, total_tiles - tile0);
    const int chunk_rows = min(ntiles * TC_TILE, nT - tile0 * TC_TILE);

    if (tid == 0) {
        for (int s = 0; s < TC_STAGES; ++s) { tc_mbar_init(bar_full + 8 * s, 1); tc_mbar_init(bar_empty + 8 * s, 1); }
        tc_mbar_init(bar_a, 1);
        for (int b = 0; b < 2; ++b) { tc_mbar_init(bar_tfull + 8 * b, 1); tc_mbar_init(bar_tempty + 8 * b, TC_EPI_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 1) {   // TMEM: all 512 columns (two 256-column accumulator buffers), one CTA per SM
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(tmem_slot) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot_ptr;

    if (warp == 0) {
        // ===== producer (whole warp converged, one elected lane issues) =====
        const uint8_t* qsrc = Qx + (size_t)blockIdx.x * TC_SMEM_A;
        if (tc_elect_one()) {
            tc_mbar_expect_tx(bar_a, TC_SMEM_A);
            for (int i = 0; i < TC_QTILES * TC_KB; ++i)
                tc_bulk_load(sA + i * TC_KB_BYTES, qsrc + (size_t)i * TC_KB_BYTES, TC_KB_BYTES, bar_a);
        }
        __syncwarp();
        const uint8_t* tsrc = Tx + (size_t)tile0 * TC_TILE_BYTES;
        const int nit = ntiles * TC_KB;
        for (int it = 0; it < nit; ++it) {
            const int s = it % TC_STAGES;
            tc_mbar_wait(bar_empty + 8 * s, (uint32_t)(((it / TC_STAGES) & 1) ^ 1));
            if (tc_elect_one()) {
                tc_mbar_expect_tx(bar_full + 8 * s, TC_KB_BYTES);
                tc_bulk_load(sB + s * TC_KB_BYTES, tsrc + (size_t)it * TC_KB_BYTES, TC_KB_BYTES, bar_full + 8 * s);
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        // ===== MMA issuer (whole warp converged, one elected lane issues) =====
        tc_mbar_wait(bar_a, 0);
        tc_fence_after();
        const uint64_t desc_hi = ((uint64_t)(1024u >> 4) << 32) | (1ull << 46) | (2ull << 61);
        int it = 0;
        for (int j = 0; j < ntiles; ++j) {
            const int buf = j & 1;
            tc_mbar_wait(bar_tempty + 8 * buf, (uint32_t)(((j >> 1) & 1) ^ 1));
            tc_fence_after();
            for (int kb = 0; kb < TC_KB; ++kb, ++it) {
                const int s = it % TC_STAGES;
                tc_mbar_wait(bar_full + 8 * s, (uint32_t)((it / TC_STAGES) & 1));
                tc_fence_after();
                if (tc_elect_one()) {
                    const uint64_t b0 = desc_hi | (uint64_t)(((sB + s * TC_KB_BYTES) & 0x3FFFFu) >> 4);
                    const uint64_t a0 = desc_hi | (uint64_t)(((sA + kb * TC_KB_BYTES) & 0x3FFFFu) >> 4);
                    const uint32_t d0 = tmem + (uint32_t)(buf * 256);
                    const uint32_t idesc = dbg == 4 ? ((1u << 4) | (8u << 17) | (8u << 24)) : TC_IDESC;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
#pragma unroll
                        for (int m = 0; m < TC_QTILES; ++m) {
                            if (dbg != 3) tc_mma(d0 + (uint32_t)(m * TC_TILE), a0 + (uint64_t)(m * (TC_TILE_BYTES >> 4) + k * 2),
                                                 b0 + (uint64_t)(k * 2), (uint32_t)((kb | k) != 0), idesc);
                        }
                    }
                    tc_commit(bar_empty + 8 * s);          // ring slot free once these MMAs have read it
                    if (kb == TC_KB - 1) tc_commit(bar_tfull + 8 * buf);   // accumulators of tile j complete
                }
                __syncwarp();
            }
        }
    } else {
        // ===== epilogue: one query per thread =====
        const int quarter = warp & 3;                      // TMEM lane quarter this warp may read
        const int m = (warp - 2) >> 2;                     // which query tile (column half of the buffer)
        const int qi = blockIdx.x * TC_QROWS + m * TC_TILE + quarter * 32 + lane;
        const uint32_t tbase = tmem + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(m * TC_TILE);
        TcState st; st.best = 0xFFFFFFFFu; st.second = 0xFFFFFFFFu; st.thr = -1e30f;
        const int full_tiles = chunk_rows / TC_TILE;         // a ragged DB leaves one masked tile at the end
        for (int j = 0; j < ntiles; ++j) {
            const int buf = j & 1;
            tc_mbar_wait(bar_tfull + 8 * buf, (uint32_t)((j >> 1) & 1));
            tc_fence_after();
            const uint32_t t = tbase + (uint32_t)(buf * 256);
            if (dbg == 1 || dbg == 2) {      // timing experiments only (results are garbage)
                if (dbg == 2) {
                    float va[32];
                    tc_ld32(t, va); tc_ld32(t + 32, va); tc_ld32(t + 64, va); tc_ld32(t + 96, va);
                    tc_wait_ld();
                    if (va[0] == 12345.f) st.best = 0;
                }
                tc_fence_before(); __syncwarp();
                if (lane == 0) tc_mbar_arrive(bar_tempty + 8 * buf);
                continue;
            }
            if (j < full_tiles) tc_epilogue_tile<false>(t, (uint32_t)(j * TC_TILE), TC_TILE, bar_tempty + 8 * buf, lane, st);
            else                tc_epilogue_tile<true>(t, (uint32_t)(j * TC_TILE), chunk_rows - j * TC_TILE, bar_tempty + 8 * buf, lane, st);
        }
        if (qi < nQ) partial[(size_t)blockIdx.y * nQ + qi] = make_uint2(st.best, st.second);
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
    }
}

// ---- v2: query tiles resident in TMEM (A operand from tensor memory) --------------------------------
// ncu on the kernel above: tensor pipe 62 % active with the shared-memory tensor-read path equally
// busy -- an SS-mode M=128 x N=128 x K=32 MMA reads 8 KB of shared memory per 64 math cycles, all the
// bandwidth an SM has, and the bulk copies write into the same banks.  Here the two query tiles are
// expanded in registers by their epilogue threads and parked in TMEM (columns 0-255: 128 lanes x 128
// packed words per tile) with tcgen05.st, so each MMA only reads its 4 KB B slice; shared memory is all
// DB ring (14 K-block stages, 224 KB in flight per SM).  The two 128-column accumulators (columns 256-511)
// alternate: while the epilogue warps of query tile 0 scan D0, the MMA warp fills D1 from the same DB
// tile, so there is no second buffer per tile and none is needed.
#endif   // CLB_LAB (v1 kernel)
static constexpr int TC2_STAGES = 14;
static constexpr int TC2_SMEM_B = TC2_STAGES * TC_KB_BYTES;          // 229376
static constexpr int TC2_SMEM_BAR = 512;
static constexpr int TC2_SMEM_TOTAL = TC2_SMEM_B + TC2_SMEM_BAR + 1024;

__device__ __forceinline__ void tc_mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t accumulate, uint32_t idesc = TC_IDESC) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], [%1], %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tc_st32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
                 "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
                 "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
                 ::"r"(taddr),
                   "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
                   "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]),
                   "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]),
                   "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
                 : "memory");
}
__device__ __forceinline__ void tc_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

#ifdef CLB_LAB
__global__ void __launch_bounds__(TC_THREADS, 1)
match_tc2_kernel(const uint32_t* __restrict__ Q, int nQ, const uint8_t* __restrict__ Tx, int nT, int tiles_per_chunk,
                 uint2* __restrict__ partial, int dbg) {
    extern __shared__ uint8_t tc_smem_raw[];
    const uint32_t raw = tc_smem_u32(tc_smem_raw);
    const uint32_t sB = (raw + 1023u) & ~1023u;
    const uint32_t sBar = sB + TC2_SMEM_B;
    const uint32_t bar_full = sBar, bar_empty = sBar + 8 * TC2_STAGES, bar_a = sBar + 16 * TC2_STAGES;
    const uint32_t bar_tfull = bar_a + 8, bar_tempty = bar_tfull + 16, tmem_slot = bar_tempty + 16;
    uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(tc_smem_raw + (tmem_slot - raw));

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int total_tiles = (nT + TC_TILE - 1) / TC_TILE;
    const int tile0 = blockIdx.y * tiles_per_chunk;
    const int ntiles = min(tiles_per_chunk, total_tiles - tile0);
    const int chunk_rows = min(ntiles * TC_TILE, nT - tile0 * TC_TILE);

    if (tid == 0) {
        for (int s = 0; s < TC2_STAGES; ++s) { tc_mbar_init(bar_full + 8 * s, 1); tc_mbar_init(bar_empty + 8 * s, 1); }
        tc_mbar_init(bar_a, TC_EPI_WARPS);
        for (int b = 0; b < 2; ++b) { tc_mbar_init(bar_tfull + 8 * b, 1); tc_mbar_init(bar_tempty + 8 * b, TC_EPI_WARPS / 2); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(tmem_slot) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot_ptr;

    if (warp == 0) {
        // ===== producer: DB K-blocks, 16 KB each, straight from the pre-swizzled image =====
        const uint8_t* tsrc = Tx + (size_t)tile0 * TC_TILE_BYTES;
        const int nit = ntiles * TC_KB;
        for (int it = 0; it < nit; ++it) {
            const int s = it % TC2_STAGES;
            tc_mbar_wait(bar_empty + 8 * s, (uint32_t)(((it / TC2_STAGES) & 1) ^ 1));
            if (tc_elect_one()) {
                tc_mbar_expect_tx(bar_full + 8 * s, TC_KB_BYTES);
                tc_bulk_load(sB + s * TC_KB_BYTES, tsrc + (size_t)it * TC_KB_BYTES, TC_KB_BYTES, bar_full + 8 * s);
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        // ===== MMA issuer: D_m (+)= A_m[tmem] x B[smem], m alternating per DB tile =====
        tc_mbar_wait(bar_a, 0);
        tc_fence_after();
        const uint64_t desc_hi = ((uint64_t)(1024u >> 4) << 32) | (1ull << 46) | (2ull << 61);
        for (int j = 0; j < ntiles; ++j) {
#pragma unroll
            for (int m = 0; m < TC_QTILES; ++m) {
                tc_mbar_wait(bar_tempty + 8 * m, (uint32_t)((j & 1) ^ 1));
                tc_fence_after();
                for (int kb = 0; kb < TC_KB; ++kb) {
                    const int it = j * TC_KB + kb;
                    const int s = it % TC2_STAGES;
                    if (m == 0) {
                        tc_mbar_wait(bar_full + 8 * s, (uint32_t)((it / TC2_STAGES) & 1));
                        tc_fence_after();
                    }
                    if (tc_elect_one()) {
                        const uint64_t b0 = desc_hi | (uint64_t)(((sB + s * TC_KB_BYTES) & 0x3FFFFu) >> 4);
                        const uint32_t a0 = tmem + (uint32_t)(m * TC_TILE + kb * 32);
                        const uint32_t d0 = tmem + (uint32_t)(256 + m * TC_TILE);
                        if (dbg == 6) {      // timing experiment: same MACs as 2 x N=256 instructions (results are garbage)
                            for (int k = 0; k < 2; ++k)
                                tc_mma_ts(tmem + 256u, a0 + (uint32_t)(k * 8), (desc_hi | (uint64_t)((sB & 0x3FFFFu) >> 4)) + (uint64_t)(k * 2), (uint32_t)((kb | k) != 0),
                                          (1u << 4) | (32u << 17) | (8u << 24));
                        } else
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            tc_mma_ts(d0, a0 + (uint32_t)(k * 8), b0 + (uint64_t)(k * 2), (uint32_t)((kb | k) != 0));
                        if (m == TC_QTILES - 1) tc_commit(bar_empty + 8 * s);
                        if (kb == TC_KB - 1) tc_commit(bar_tfull + 8 * m);
                    }
                    __syncwarp();
                }
            }
        }
    } else {
        // ===== epilogue warps: first park this thread's query in TMEM, then scan its accumulator row =====
        const int quarter = warp & 3;
        const int m = (warp - 2) >> 2;
        const int qi = blockIdx.x * TC_QROWS + m * TC_TILE + quarter * 32 + lane;
        const uint32_t lane_base = tmem + ((uint32_t)(quarter * 32) << 16);
        {
            uint4 q4[4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
                q4[i] = qi < nQ ? __ldg(reinterpret_cast<const uint4*>(Q + (size_t)qi * 16) + i) : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
            for (int i = 0; i < 4; ++i) {                      // 4 packed words -> 32 TMEM columns (128 E4M3)
                const uint32_t w[4] = {q4[i].x, q4[i].y, q4[i].z, q4[i].w};
                uint32_t r[32];
#pragma unroll
                for (int c = 0; c < 32; ++c) r[c] = qi < nQ ? tc_nib((w[c >> 3] >> (4 * (c & 7))) & 15u) : 0u;
                tc_st32(lane_base + (uint32_t)(m * TC_TILE + 32 * i), r);
            }
            tc_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) tc_mbar_arrive(bar_a);
        }
        const uint32_t tbase = lane_base + (uint32_t)(256 + m * TC_TILE);
        TcState st; st.best = 0xFFFFFFFFu; st.second = 0xFFFFFFFFu; st.thr = -1e30f;
        const int full_tiles = chunk_rows / TC_TILE;
        for (int j = 0; j < ntiles; ++j) {
            tc_mbar_wait(bar_tfull + 8 * m, (uint32_t)(j & 1));
            tc_fence_after();
            if (dbg == 1 || dbg == 2) {      // timing experiments only (results are garbage)
                if (dbg == 2) {
                    float va[32];
                    tc_ld32(tbase, va); tc_ld32(tbase + 32, va); tc_ld32(tbase + 64, va); tc_ld32(tbase + 96, va);
                    tc_wait_ld();
                    if (va[0] == 12345.f) st.best = 0;
                }
                tc_fence_before(); __syncwarp();
                if (lane == 0) tc_mbar_arrive(bar_tempty + 8 * m);
                continue;
            }
            if (j < full_tiles) tc_epilogue_tile<false>(tbase, (uint32_t)(j * TC_TILE), TC_TILE, bar_tempty + 8 * m, lane, st);
            else                tc_epilogue_tile<true>(tbase, (uint32_t)(j * TC_TILE), chunk_rows - j * TC_TILE, bar_tempty + 8 * m, lane, st);
        }
        if (qi < nQ) partial[(size_t)blockIdx.y * nQ + qi] = make_uint2(st.best, st.second);
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
    }
}

#endif   // CLB_LAB (v2 kernel)

// ---- v3: E2M1 (fp4) operands, kind::mxf4 block-scaled MMA at twice the fp8 rate ------------------------
// +-1 is exact in E2M1 too (0x2 / 0xA), and kind::mxf4 multiplies K = 64 elements per instruction at 16 384 MAC/clk/SM:
// 32 comparisons per clock per SM.  The block scales the instruction insists on are UE8M0 bytes; 0x7F = 2^0, so a TMEM
// region filled with 0x7F7F7F7F serves as scale factor A and B whatever the layout rules say.  Operands are packed two
// per byte: a row is 256 B = 2 K-blocks of 128 B, a 128-row tile 32 KB (half the L2 -> smem traffic per comparison).
// Both operands come from shared memory (SS mode); the two accumulators alternate as in v2.
static constexpr int TC3_KB = 2;                                    // K-blocks (128 B = 256 E2M1) per row
static constexpr int TC3_TILE_BYTES = TC3_KB * TC_KB_BYTES;         // 32 KB
static constexpr int TC3_SMEM_A = TC_QTILES * TC3_TILE_BYTES;       // 64 KB
static constexpr int TC3_STAGES = 10;
static constexpr int TC3_SMEM_B = TC3_STAGES * TC_KB_BYTES;         // 160 KB
static constexpr int TC3_SMEM_TOTAL = TC3_SMEM_A + TC3_SMEM_B + TC2_SMEM_BAR + 1024;
// block-scaled instruction descriptor: A = B = E2M1 (MXF4Format 1) at bits 7 / 10, N>>3 at 17, scale format UE8M0 at 23, M>>4 at 24
static constexpr uint32_t TC3_IDESC = (1u << 7) | (1u << 10) | ((uint32_t)(TC_TILE >> 3) << 17) | (1u << 23) | ((uint32_t)(TC_TILE >> 4) << 24);

__device__ __forceinline__ void tc_mma_mxf4(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t accumulate, uint32_t sfa, uint32_t sfb) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], %1, %2, %3, [%5], [%6], p;\n\t}"
                 ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(TC3_IDESC), "r"(accumulate), "r"(sfa), "r"(sfb) : "memory");
}

// 8 descriptor bits -> 8 E2M1 nibbles (one 32-bit word): bit 1 -> 0x2 (+1.0), bit 0 -> 0xA (-1.0); bit i lands in nibble i
__device__ __forceinline__ uint32_t tc_nib4(uint32_t x) {
    uint32_t y = (x | (x << 12)) & 0x000F000Fu;
    y = (y | (y << 6)) & 0x03030303u;
    y = (y | (y << 3)) & 0x11111111u;
    return 0xAAAAAAAAu ^ (y << 3);
}
// One thread per 16-byte output chunk (= one packed 32-bit word of the row), output order.
// Layout: [tile of 128 rows][2 K-blocks][128 rows][128 B], 16-B chunk c of row r at chunk c ^ (r & 7).
__global__ void __launch_bounds__(256)
expand4_kernel(const uint32_t* __restrict__ rows, int n, size_t n_chunks16, uint4* __restrict__ out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_chunks16) return;
    const int cpos = (int)(i & 7), r = (int)((i >> 3) & 127), kb = (int)((i >> 10) & 1);
    const size_t tile = i >> 11;
    const int c = cpos ^ (r & 7);
    const size_t row = tile * TC_TILE + r;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);                        // E2M1 0x0 = +0.0 for the padding rows
    if (row < (size_t)n) {
        const uint32_t w = __ldg(rows + row * 16 + kb * 8 + c);
        v.x = tc_nib4(w & 255u); v.y = tc_nib4((w >> 8) & 255u); v.z = tc_nib4((w >> 16) & 255u); v.w = tc_nib4(w >> 24);
    }
    out[i] = v;
}

#ifdef CLB_LAB
__global__ void __launch_bounds__(TC_THREADS, 1)
match_tc3_kernel(const uint8_t* __restrict__ Qx, int nQ, const uint8_t* __restrict__ Tx, int nT, int tiles_per_chunk,
                 uint2* __restrict__ partial, int dbg) {
    extern __shared__ uint8_t tc_smem_raw[];
    const uint32_t raw = tc_smem_u32(tc_smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    const uint32_t sA = base, sB = base + TC3_SMEM_A, sBar = sB + TC3_SMEM_B;
    const uint32_t bar_full = sBar, bar_empty = sBar + 8 * TC3_STAGES, bar_a = sBar + 16 * TC3_STAGES;
    const uint32_t bar_tfull = bar_a + 8, bar_tempty = bar_tfull + 16, bar_sf = bar_tempty + 16, tmem_slot = bar_sf + 8;
    uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(tc_smem_raw + (tmem_slot - raw));

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int total_tiles = (nT + TC_TILE - 1) / TC_TILE;
    const int tile0 = blockIdx.y * tiles_per_chunk;
    const int ntiles = min(tiles_per_chunk, total_tiles - tile0);
    const int chunk_rows = min(ntiles * TC_TILE, nT - tile0 * TC_TILE);

    if (tid == 0) {
        for (int s = 0; s < TC3_STAGES; ++s) { tc_mbar_init(bar_full + 8 * s, 1); tc_mbar_init(bar_empty + 8 * s, 1); }
        tc_mbar_init(bar_a, 1);
        tc_mbar_init(bar_sf, TC_EPI_WARPS / 2);
        for (int b = 0; b < 2; ++b) { tc_mbar_init(bar_tfull + 8 * b, 1); tc_mbar_init(bar_tempty + 8 * b, TC_EPI_WARPS / 2); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(tmem_slot) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot_ptr;
    // TMEM map: columns 0-127 / 128-255 accumulators D_0 / D_1, columns 256-319 scale factors (all 1.0)
    const uint32_t tm_sf = tmem + 256u;

    if (warp == 0) {
        // ===== producer =====
        const uint8_t* qsrc = Qx + (size_t)blockIdx.x * TC3_SMEM_A;
        if (tc_elect_one()) {
            tc_mbar_expect_tx(bar_a, TC3_SMEM_A);
            for (int i = 0; i < TC_QTILES * TC3_KB; ++i)
                tc_bulk_load(sA + i * TC_KB_BYTES, qsrc + (size_t)i * TC_KB_BYTES, TC_KB_BYTES, bar_a);
        }
        __syncwarp();
        const uint8_t* tsrc = Tx + (size_t)tile0 * TC3_TILE_BYTES;
        const int nit = ntiles * TC3_KB;
        for (int it = 0; it < nit; ++it) {
            const int s = it % TC3_STAGES;
            tc_mbar_wait(bar_empty + 8 * s, (uint32_t)(((it / TC3_STAGES) & 1) ^ 1));
            if (tc_elect_one()) {
                tc_mbar_expect_tx(bar_full + 8 * s, TC_KB_BYTES);
                tc_bulk_load(sB + s * TC_KB_BYTES, tsrc + (size_t)it * TC_KB_BYTES, TC_KB_BYTES, bar_full + 8 * s);
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        tc_mbar_wait(bar_a, 0);
        tc_mbar_wait(bar_sf, 0);
        tc_fence_after();
        const uint64_t desc_hi = ((uint64_t)(1024u >> 4) << 32) | (1ull << 46) | (2ull << 61);
        for (int j = 0; j < ntiles; ++j) {
#pragma unroll
            for (int m = 0; m < TC_QTILES; ++m) {
                tc_mbar_wait(bar_tempty + 8 * m, (uint32_t)((j & 1) ^ 1));
                tc_fence_after();
#pragma unroll
                for (int kb = 0; kb < TC3_KB; ++kb) {
                    const int it = j * TC3_KB + kb;
                    const int s = it % TC3_STAGES;
                    if (m == 0) {
                        tc_mbar_wait(bar_full + 8 * s, (uint32_t)((it / TC3_STAGES) & 1));
                        tc_fence_after();
                    }
                    if (tc_elect_one()) {
                        const uint64_t b0 = desc_hi | (uint64_t)(((sB + s * TC_KB_BYTES) & 0x3FFFFu) >> 4);
                        const uint64_t a0 = desc_hi | (uint64_t)(((sA + m * TC3_TILE_BYTES + kb * TC_KB_BYTES) & 0x3FFFFu) >> 4);
                        const uint32_t d0 = tmem + (uint32_t)(m * TC_TILE);
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            tc_mma_mxf4(d0, a0 + (uint64_t)(k * 2), b0 + (uint64_t)(k * 2), (uint32_t)((kb | k) != 0), tm_sf, tm_sf + 32u);
                        if (m == TC_QTILES - 1) tc_commit(bar_empty + 8 * s);
                        if (kb == TC3_KB - 1) tc_commit(bar_tfull + 8 * m);
                    }
                    __syncwarp();
                }
            }
        }
    } else {
        // ===== epilogue warps =====
        const int quarter = warp & 3;
        const int m = (warp - 2) >> 2;
        const int qi = blockIdx.x * TC_QROWS + m * TC_TILE + quarter * 32 + lane;
        const uint32_t lane_base = tmem + ((uint32_t)(quarter * 32) << 16);
        if (m == 0) {   // scale-factor region: every byte UE8M0 1.0, all 128 lanes x 64 columns
            uint32_t r[32];
#pragma unroll
            for (int c = 0; c < 32; ++c) r[c] = 0x7F7F7F7Fu;
            tc_st32(lane_base + 256u, r);
            tc_st32(lane_base + 288u, r);
            tc_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) tc_mbar_arrive(bar_sf);
        }
        const uint32_t tbase = lane_base + (uint32_t)(m * TC_TILE);
        TcState st; st.best = 0xFFFFFFFFu; st.second = 0xFFFFFFFFu; st.thr = -1e30f;
        const int full_tiles = chunk_rows / TC_TILE;
        for (int j = 0; j < ntiles; ++j) {
            tc_mbar_wait(bar_tfull + 8 * m, (uint32_t)(j & 1));
            tc_fence_after();
            if (dbg == 1 || dbg == 4 || dbg == 5) {
                tc_fence_before(); __syncwarp();
                if (lane == 0) tc_mbar_arrive(bar_tempty + 8 * m);
                continue;
            }
            if (j < full_tiles) tc_epilogue_tile_wide<false>(tbase, (uint32_t)(j * TC_TILE), TC_TILE, bar_tempty + 8 * m, lane, st);
            else                tc_epilogue_tile_wide<true>(tbase, (uint32_t)(j * TC_TILE), chunk_rows - j * TC_TILE, bar_tempty + 8 * m, lane, st);
        }
        if (qi < nQ) partial[(size_t)blockIdx.y * nQ + qi] = make_uint2(st.best, st.second);
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
    }
}

#endif   // CLB_LAB (v3 kernel)

#ifdef CLB_LAB
// match_tc_dbg = 7: per-tile clocks of CTA (0, 0) -- [0][j] DB K-block pair of tile j issued by the TMA warp, [1][j] MMAs of tile j
// issued (query tile 0), [2][j] accumulator j seen by epilogue warp 0, [3][j] epilogue warp 0 done with tile j (clb_lab_tile_clocks)
__device__ long long g_tc_trace[6][256];     // [4][j] TMEM loads of tile j landed, [5][j] first 32 columns processed
#endif

// ---- v4: v3 with the query tiles in TMEM (TS mode) and 16 epilogue warps --------------------------------
// tools/tc_dbg.py on v3: 5.4 T cmp/s, 6.1 T with the epilogue's TMEM reads disabled -- the scan of 128 distances per
// thread per tile had become the long pole.  Here every accumulator row is shared by two threads (64 columns each),
// both loads are issued at once and the buffer is handed back after one TMEM round trip; the two halves of a query are
// merged by packed key at the end (exact: min of the bests, second smallest of the four).  The queries sit in TMEM
// (columns 320-447, 8 E2M1 per 32-bit column), so shared memory is a 14-stage DB ring again.
static constexpr int TC4_EPI_WARPS = 16;
static constexpr int TC4_MMA1_WARP = 2 + TC4_EPI_WARPS;                 // second MMA-issuing warp comes after the epilogue warps
static constexpr int TC4_THREADS = 32 * (3 + TC4_EPI_WARPS);
static constexpr int TC4_STAGES = 14;
static constexpr int TC4_SMEM_B = TC4_STAGES * TC_KB_BYTES;
static constexpr int TC4_SMEM_TOTAL = TC4_SMEM_B + TC2_SMEM_BAR + 1024;

__device__ __forceinline__ void tc_mma_mxf4_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t accumulate, uint32_t sfa, uint32_t sfb) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], [%1], %2, %3, [%5], [%6], p;\n\t}"
                 ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(TC3_IDESC), "r"(accumulate), "r"(sfa), "r"(sfb) : "memory");
}

// same instruction with N = 64 (half an accumulator: DB rows 0-63 or 64-127 of the tile)
__device__ __forceinline__ void tc_mma_mxf4_ts_n64(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t accumulate, uint32_t sfa, uint32_t sfb) {
    constexpr uint32_t idesc = (1u << 7) | (1u << 10) | ((uint32_t)(64 >> 3) << 17) | (1u << 23) | ((uint32_t)(TC_TILE >> 4) << 24);
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], [%1], %2, %3, [%5], [%6], p;\n\t}"
                 ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(sfa), "r"(sfb) : "memory");
}

template <bool MASKED>
__device__ __forceinline__ void tc_epilogue_half_tile(uint32_t t, uint32_t row0, int valid, uint32_t bar_release, int lane, TcStateF& st,
                                                      long long* tr = nullptr) {      // tr: lab trace (clocks after the TMEM loads landed / after the first 32 columns)
    float v0[32], v1[32];
    tc_ld32(t, v0); tc_ld32(t + 32, v1);
    tc_wait_ld();
    if (tr) tr[0] = clock64();
    tc_fence_before();
    __syncwarp();
    if (lane == 0) tc_mbar_arrive(bar_release);
    if (MASKED) {
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            if (i >= valid) v0[i] = -1e30f;
            if (32 + i >= valid) v1[i] = -1e30f;
        }
    }
    tc_process32f(v0, row0, st);
    if (tr) tr[256] = clock64();
    tc_process32f(v1, row0 + 32, st);
}

// One CTA of the v4 sweep: query rows [qblock*256, +256) of Q against DB tiles [tile0, tile0+ntiles) of the expanded image Tx
// (chunk_rows of them real); the per-query (best, second) packed keys of this chunk go to partial_chunk[query row].
//
// Warp roles: 0 = TMA producer; 1 and 18 = MMA issue for query tile 0 / 1 (each accumulator's chain has its own issuing
// warp: tcgen05.mma issue is nearly synchronous -- the queue behind the tensor pipe is ~2 instructions deep -- so a single
// issuing warp lets the pipe drain at every barrier wait; with two, one warp's waits are covered by the other's MMAs:
// 1 450 -> see profiles/r1_ncu_summary.md cycles per DB tile against the 1 029 of a bare MMA stream); 2..17 = epilogue.
// dbg (timing experiments, results are garbage unless 0 or 7): 1 no TMEM reads, 2 TMEM reads without the reduction,
// 3 no MMAs, 7 normal sweep + real SM clock printed by a few CTAs.
// LAB = false (the product instantiation) folds dbg to 0: none of the timing modes, clocks or printfs is compiled in.
template <bool LAB>
__device__ __forceinline__ void
tc4_cta(const uint32_t* __restrict__ Q, int nQ, const uint8_t* __restrict__ Tx, int qblock, int tile0, int ntiles, int chunk_rows,
        uint2* __restrict__ partial_chunk, int dbg_rt) {
    const int dbg = LAB ? dbg_rt : 0;
    extern __shared__ uint8_t tc_smem_raw[];
    const uint32_t raw = tc_smem_u32(tc_smem_raw);
    const uint32_t sB = (raw + 1023u) & ~1023u;
    const uint32_t sBar = sB + TC4_SMEM_B;
    const uint32_t bar_full = sBar, bar_empty = sBar + 8 * TC4_STAGES, bar_a = sBar + 16 * TC4_STAGES;
    const uint32_t bar_tfull = bar_a + 8, bar_tempty = bar_tfull + 32, tmem_slot = bar_tempty + 32;   // [query tile][column half]
    uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(tc_smem_raw + (tmem_slot - raw));
    uint2* smem_res = reinterpret_cast<uint2*>(tc_smem_raw + (sB - raw));      // ring reused after the sweep
    // split: each accumulator is handed over in two 64-column halves with their own barriers (four half-accumulators rotate)
    const bool split = dbg == 21;                              // experiment (match_tc_dbg = 21): N = 64 MMAs cost ~44 cycles instead of 32, 1 426 vs 1 148 cycles per tile

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    long long dbg_c0 = 0; unsigned long long dbg_t0 = 0;
    if ((dbg == 7 || (dbg >= 10 && dbg != 20)) && tid == 0) { dbg_c0 = clock64(); asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(dbg_t0)); }
    long long dbg_c1 = 0;

    if (tid == 0) {
        for (int s = 0; s < TC4_STAGES; ++s) { tc_mbar_init(bar_full + 8 * s, 1); tc_mbar_init(bar_empty + 8 * s, TC_QTILES); }
        tc_mbar_init(bar_a, TC4_EPI_WARPS);
        for (int b = 0; b < 4; ++b) { tc_mbar_init(bar_tfull + 8 * b, 1); tc_mbar_init(bar_tempty + 8 * b, split ? TC4_EPI_WARPS / 4 : TC4_EPI_WARPS / 2); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(tmem_slot) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (dbg == 7 && tid == 0) dbg_c1 = clock64();
    const uint32_t tmem = *tmem_slot_ptr;
    // TMEM map: 0-127 / 128-255 accumulators D_0 / D_1; 256-319 unit scale factors; 320-383 / 384-447 query tiles A_0 / A_1
    const uint32_t tm_sf = tmem + 256u, tm_a = tmem + 320u;

    if (warp == 0) {
        const uint8_t* tsrc = Tx + (size_t)tile0 * TC3_TILE_BYTES;
        if (dbg == 18 || dbg == 19) {      // timing experiment: TMA fill of the ring at the sweep's rate, paced by the clock, nobody waits on it
            const long long t0 = clock64();
            for (int it = 0; it < ntiles * TC3_KB; ++it) {
                while (clock64() - t0 < 515ll * it) { }
                const int s = it % TC4_STAGES;
                if (tc_elect_one()) {
                    tc_mbar_expect_tx(bar_full + 8 * s, TC_KB_BYTES);
                    tc_bulk_load(sB + s * TC_KB_BYTES, tsrc + (size_t)it * TC_KB_BYTES, TC_KB_BYTES, bar_full + 8 * s);
                }
                __syncwarp();
            }
        }
        const int nit = dbg >= 10 && dbg != 20 ? 0 : ntiles * TC3_KB;
        for (int it = 0; it < nit; ++it) {
            const int s = it % TC4_STAGES;
            tc_mbar_wait(bar_empty + 8 * s, (uint32_t)(((it / TC4_STAGES) & 1) ^ 1));
            if (tc_elect_one()) {
                tc_mbar_expect_tx(bar_full + 8 * s, TC_KB_BYTES);
                tc_bulk_load(sB + s * TC_KB_BYTES, tsrc + (size_t)it * TC_KB_BYTES, TC_KB_BYTES, bar_full + 8 * s);
#ifdef CLB_LAB
                if (LAB && dbg == 7 && qblock == 0 && tile0 == 0 && (it & 1) && (it >> 1) < 256) g_tc_trace[0][it >> 1] = clock64();
#endif
            }
            __syncwarp();
        }
    } else if (warp == 1 || warp == TC4_MMA1_WARP) {
        const int m = warp == 1 ? 0 : 1;
        tc_mbar_wait(bar_a, 0);
        tc_fence_after();
        const uint64_t desc_hi = ((uint64_t)(1024u >> 4) << 32) | (1ull << 46) | (2ull << 61);
        const uint32_t d0 = tmem + (uint32_t)(m * TC_TILE);
        if (dbg >= 10 && dbg != 20) {     // timing experiments on a bare MMA stream (warp 1 only): 10 nothing else, 11 + a commit, 12 + a wait on a
                             // completed barrier, 13 + fence::after_thread_sync, 14 + elect per MMA group of 4
            if (m == 0 && (dbg < 15 || dbg >= 17)) {
                for (int j = 0; j < ntiles; ++j) {
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        if (dbg == 12) tc_mbar_wait(bar_a, 0);
                        if (dbg == 13) tc_fence_after();
                        if (tc_elect_one()) {
                            const uint64_t b0 = desc_hi | (uint64_t)(((sB + ((2 * j + (g & 1)) % TC4_STAGES) * TC_KB_BYTES) & 0x3FFFFu) >> 4);
#pragma unroll
                            if (dbg == 17 || dbg == 18) {
                                const uint64_t a0 = desc_hi | (uint64_t)(((sB + ((2 * j + (g & 1) + 7) % TC4_STAGES) * TC_KB_BYTES) & 0x3FFFFu) >> 4);
#pragma unroll
                                for (int k = 0; k < 4; ++k)
                                    tc_mma_mxf4(tmem + (uint32_t)((g >> 1) * TC_TILE), a0 + (uint64_t)(k * 2), b0 + (uint64_t)(k * 2), (uint32_t)(((g & 1) | k) != 0), tm_sf, tm_sf + 32u);
                            } else {
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                tc_mma_mxf4_ts(tmem + (uint32_t)((g >> 1) * TC_TILE), tm_a + (uint32_t)(g * 32 + k * 8), b0 + (uint64_t)(k * 2), (uint32_t)(((g & 1) | k) != 0), tm_sf, tm_sf + 32u);
                            }
                            if (dbg == 11) tc_commit(bar_tfull + 8);
                        }
                        __syncwarp();
                    }
                }
                if (tc_elect_one()) tc_commit(bar_tfull);
                __syncwarp();
                tc_mbar_wait(bar_tfull, 0);
            }
        } else {
        // The issuing thread is blocked while its MMAs execute and nothing is queued behind them, so every instruction
        // between two MMAs is tensor-pipe idle time: the operand barriers are checked BEFORE the accumulator hand-off wait
        // (which has to be sat out anyway) and the MMAs of a hand-off unit then go out back to back.
        const uint32_t a0 = tm_a + (uint32_t)(m * 64);
        const bool no_mma = dbg == 3;
        for (int j = 0; j < ntiles; ++j) {
            const int it0 = j * TC3_KB, s0 = it0 % TC4_STAGES, s1 = (it0 + 1) % TC4_STAGES;
            tc_mbar_wait(bar_full + 8 * s0, (uint32_t)((it0 / TC4_STAGES) & 1));
            tc_mbar_wait(bar_full + 8 * s1, (uint32_t)(((it0 + 1) / TC4_STAGES) & 1));
            const uint64_t b0 = desc_hi | (uint64_t)(((sB + s0 * TC_KB_BYTES) & 0x3FFFFu) >> 4);
            const uint64_t b1 = desc_hi | (uint64_t)(((sB + s1 * TC_KB_BYTES) & 0x3FFFFu) >> 4);
            if (!split) {
                tc_mbar_wait(bar_tempty + 16 * m, (uint32_t)((j & 1) ^ 1));
                tc_fence_after();
                if (tc_elect_one()) {
                    if (!no_mma) {
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            tc_mma_mxf4_ts(d0, a0 + (uint32_t)(k * 8), b0 + (uint64_t)(k * 2), (uint32_t)(k != 0), tm_sf, tm_sf + 32u);
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            tc_mma_mxf4_ts(d0, a0 + (uint32_t)(32 + k * 8), b1 + (uint64_t)(k * 2), 1u, tm_sf, tm_sf + 32u);
                    }
                    tc_commit(bar_tfull + 16 * m);
                    tc_commit(bar_empty + 8 * s0);
                    tc_commit(bar_empty + 8 * s1);
#ifdef CLB_LAB
                    if (LAB && dbg == 7 && qblock == 0 && tile0 == 0 && m == 0 && j < 256) g_tc_trace[1][j] = clock64();
#endif
                }
                __syncwarp();
            } else {
#pragma unroll
                for (int hf = 0; hf < 2; ++hf) {         // DB rows 0-63 / 64-127 of the tile = accumulator columns 0-63 / 64-127
                    tc_mbar_wait(bar_tempty + 16 * m + 8 * hf, (uint32_t)((j & 1) ^ 1));
                    tc_fence_after();
                    if (tc_elect_one()) {
                        const uint64_t r0 = b0 + (uint64_t)(hf * 512), r1 = b1 + (uint64_t)(hf * 512);   // 64 rows x 128 B = 8 KB further
                        const uint32_t dh = d0 + (uint32_t)(hf * 64);
                        if (!no_mma) {
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                tc_mma_mxf4_ts_n64(dh, a0 + (uint32_t)(k * 8), r0 + (uint64_t)(k * 2), (uint32_t)(k != 0), tm_sf, tm_sf + 32u);
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                tc_mma_mxf4_ts_n64(dh, a0 + (uint32_t)(32 + k * 8), r1 + (uint64_t)(k * 2), 1u, tm_sf, tm_sf + 32u);
                        }
                        tc_commit(bar_tfull + 16 * m + 8 * hf);
                        if (hf == 1) { tc_commit(bar_empty + 8 * s0); tc_commit(bar_empty + 8 * s1); }
                    }
                    __syncwarp();
                }
            }
        }
        }
    } else {
        const int e = warp - 2;
        const int quarter = warp & 3, m = (e >> 2) & 1, half = e >> 3;
        const int row = quarter * 32 + lane;
        long long pc[6] = {0, 0, 0, 0, 0, 0};                    // dbg 7: phase clocks of epilogue warp 0
        const bool prof = LAB && dbg == 7 && e == 0 && lane == 0 && qblock == 0 && tile0 == 0;
        if (prof) pc[0] = clock64();
        const int qi = qblock * TC_QROWS + m * TC_TILE + row;
        const uint32_t lane_base = tmem + ((uint32_t)(quarter * 32) << 16);
        {
            uint32_t r[32];
            if (e < 4) {      // unit scale factors, 128 lanes x 64 columns
#pragma unroll
                for (int c = 0; c < 32; ++c) r[c] = 0x7F7F7F7Fu;
                tc_st32(lane_base + 256u, r);
                tc_st32(lane_base + 288u, r);
            }
            // this thread's half of its query: 8 packed words -> 32 TMEM columns of 8 E2M1 each
            uint4 q4[2];
#pragma unroll
            for (int i = 0; i < 2; ++i)
                q4[i] = qi < nQ ? __ldg(reinterpret_cast<const uint4*>(Q + (size_t)qi * 16 + half * 8) + i) : make_uint4(0u, 0u, 0u, 0u);
            const uint32_t w[8] = {q4[0].x, q4[0].y, q4[0].z, q4[0].w, q4[1].x, q4[1].y, q4[1].z, q4[1].w};
#pragma unroll
            for (int c = 0; c < 32; ++c) r[c] = qi < nQ ? tc_nib4((w[c >> 2] >> (8 * (c & 3))) & 255u) : 0u;
            tc_st32(lane_base + 320u + (uint32_t)(m * 64 + half * 32), r);
            tc_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) tc_mbar_arrive(bar_a);
        }
        if (prof) pc[1] = clock64();
        const uint32_t tbase = lane_base + (uint32_t)(m * TC_TILE + half * 64);
        const uint32_t hb = (uint32_t)(16 * m + (split ? 8 * half : 0));      // this warp's hand-off barriers
        TcStateF st; st.b = -1e30f; st.s = -1e30f; st.idx = 0u;
        if (dbg == 15 || dbg == 16) {      // timing experiment: bare TMEM read stream (16: only query tile 0's warps)
            float acc = 0.f;
            if (dbg == 15 || m == 0)
                for (int j = 0; j < ntiles; ++j) {
                    float v0[32], v1[32];
                    tc_ld32(tbase, v0); tc_ld32(tbase + 32, v1);
                    tc_wait_ld();
#pragma unroll
                    for (int i = 0; i < 32; ++i) acc += v0[i] + v1[i];
                }
            if (acc == 12345.678f) st.idx = 1u;
        }
        for (int j = 0; j < (dbg >= 10 && dbg != 20 ? 0 : ntiles); ++j) {
            if (dbg == 20) tc_mbar_spin(bar_tfull + hb, (uint32_t)(j & 1));
            else           tc_mbar_wait(bar_tfull + hb, (uint32_t)(j & 1));
            tc_fence_after();
            if (dbg == 1) {
                tc_fence_before(); __syncwarp();
                if (lane == 0) tc_mbar_arrive(bar_tempty + hb);
                continue;
            }
            if (dbg == 2) {
                float v0[32], v1[32];
                tc_ld32(tbase, v0); tc_ld32(tbase + 32, v1);
                tc_wait_ld();
                tc_fence_before(); __syncwarp();
                if (lane == 0) tc_mbar_arrive(bar_tempty + hb);
                float acc = 0.f;
#pragma unroll
                for (int i = 0; i < 32; ++i) acc += v0[i] + v1[i];
                if (acc == 12345.678f) st.idx = 1u;
                continue;
            }
            if (prof && j < 2) pc[2 + j] = clock64();
#ifdef CLB_LAB
            if (prof && j < 256) g_tc_trace[2][j] = clock64();
#endif
            const int valid = chunk_rows - j * TC_TILE - half * 64;       // rows of this half that exist
            const uint32_t row0 = (uint32_t)(j * TC_TILE + half * 64);
#ifdef CLB_LAB
            long long* tr = (prof && j < 256) ? &g_tc_trace[4][j] : nullptr;
            if (valid >= 64) tc_epilogue_half_tile<false>(tbase, row0, 64, bar_tempty + hb, lane, st, tr);
            else             tc_epilogue_half_tile<true>(tbase, row0, valid, bar_tempty + hb, lane, st, tr);
            if (prof && j < 256) g_tc_trace[3][j] = clock64();
#else
            if (valid >= 64) tc_epilogue_half_tile<false>(tbase, row0, 64, bar_tempty + hb, lane, st);
            else             tc_epilogue_half_tile<true>(tbase, row0, valid, bar_tempty + hb, lane, st);
#endif
        }
        // merge the two column halves of each query by packed key.  The staging array aliases ring stage 0, which the
        // last tile's MMAs of the OTHER query tile may still be reading: wait until every epilogue warp has seen its last tile.
        if (prof) pc[4] = clock64();
        asm volatile("bar.sync 1, %0;" ::"n"(TC4_EPI_WARPS * 32) : "memory");
        const uint2 mine = tc_statef_keys(st);
        if (half == 1) smem_res[m * TC_TILE + row] = mine;
        asm volatile("bar.sync 1, %0;" ::"n"(TC4_EPI_WARPS * 32) : "memory");
        if (half == 0 && qi < nQ) {
            const uint2 o = smem_res[m * TC_TILE + row];
            const uint32_t b = min(mine.x, o.x);
            const uint32_t sec = min(max(mine.x, o.x), min(mine.y, o.y));
            partial_chunk[qi] = make_uint2(b, sec);
        }
        if (LAB && prof) {
            pc[5] = clock64();
            printf("tc4 phases (cycles since the epilogue warp started): queries in TMEM %lld, first accumulator %lld, second %lld, sweep done %lld, merged %lld\n",
                   pc[1] - pc[0], pc[2] - pc[0], pc[3] - pc[0], pc[4] - pc[0], pc[5] - pc[0]);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
    }
    if (LAB && (dbg == 7 || (dbg >= 10 && dbg != 20)) && tid == 0 && qblock == 0 && tile0 == 0) {     // real SM clock while the sweep runs: cycles / globaltimer ns
        unsigned long long t1; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        const long long c1 = clock64();
        printf("tc4 qblock %d: %lld cycles in %llu ns = %.0f MHz, %d tiles, %.1f cycles/tile; barriers + TMEM allocation %lld cycles\n", qblock, c1 - dbg_c0, t1 - dbg_t0,
               1e3 * (double)(c1 - dbg_c0) / (double)(t1 - dbg_t0), ntiles, (double)(c1 - dbg_c0) / ntiles, dbg_c1 - dbg_c0);
    }
}

__global__ void __launch_bounds__(TC4_THREADS, 1)
match_tc4_kernel(const uint32_t* __restrict__ Q, int nQ, const uint8_t* __restrict__ Tx, int nT, int tiles_per_chunk,
                 uint2* __restrict__ partial) {
    const int total_tiles = (nT + TC_TILE - 1) / TC_TILE;
    const int tile0 = blockIdx.y * tiles_per_chunk;
    const int ntiles = min(tiles_per_chunk, total_tiles - tile0);
    const int chunk_rows = min(ntiles * TC_TILE, nT - tile0 * TC_TILE);
    tc4_cta<false>(Q, nQ, Tx, blockIdx.x, tile0, ntiles, chunk_rows, partial + (size_t)blockIdx.y * nQ, 0);
}
#ifdef CLB_LAB      // the same CTA with the timing modes compiled in (clb_set_param "match_tc_dbg" != 0)
__global__ void __launch_bounds__(TC4_THREADS, 1)
match_tc4_lab_kernel(const uint32_t* __restrict__ Q, int nQ, const uint8_t* __restrict__ Tx, int nT, int tiles_per_chunk,
                     uint2* __restrict__ partial, int dbg) {
    const int total_tiles = (nT + TC_TILE - 1) / TC_TILE;
    const int tile0 = blockIdx.y * tiles_per_chunk;
    const int ntiles = min(tiles_per_chunk, total_tiles - tile0);
    const int chunk_rows = min(ntiles * TC_TILE, nT - tile0 * TC_TILE);
    tc4_cta<true>(Q, nQ, Tx, blockIdx.x, tile0, ntiles, chunk_rows, partial + (size_t)blockIdx.y * nQ, dbg);
}
#endif

// Batched form: one CTA per job of a device-resident job list (all pairs of a pair list in one launch).
__global__ void __launch_bounds__(TC4_THREADS, 1)
match_tc4_jobs_kernel(const TcJob* __restrict__ jobs) {
    const TcJob j = jobs[blockIdx.x];
    tc4_cta<false>(j.Q, j.nQ, j.Tx, j.qblock, j.tile0, j.ntiles, j.chunk_rows, j.partial_chunk, 0);
}

// ---- measured tensor-pipe peak (clb_microbench_mma) -------------------------------------------------------------------
// A bare stream of the sweep's own instruction -- tcgen05.mma kind::mxf4.block_scale M128 N128 K64, A from TMEM, B from
// shared memory, unit scales, +-1 E2M1 operands -- on every SM: no TMA feed, no epilogue, no hand-off.  What it sustains is
// the roofline denominator of match_tc4_kernel (MEASURED_PEAKS.json has no fp4 figure).
__global__ void __launch_bounds__(128, 1)
tc_mma_bench_kernel(int groups, long long* __restrict__ cycles_out, unsigned long long* __restrict__ ns_out) {
    extern __shared__ uint8_t tc_smem_raw[];
    const uint32_t raw = tc_smem_u32(tc_smem_raw);
    const uint32_t sB = (raw + 1023u) & ~1023u;                   // two K-blocks of 16 KB
    const uint32_t sBar = sB + 2 * TC_KB_BYTES;
    const uint32_t bar_done = sBar, tmem_slot = sBar + 8;
    uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(tc_smem_raw + (tmem_slot - raw));
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        tc_mbar_init(bar_done, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(tmem_slot) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // +-1 nibbles, pseudo-random like real descriptors (the data decides the power drawn, and with it the clock)
    uint32_t* bw = reinterpret_cast<uint32_t*>(tc_smem_raw + (sB - raw));
    for (int i = tid; i < 2 * TC_KB_BYTES / 4; i += 128) {
        uint32_t h = (uint32_t)i * 2654435761u + blockIdx.x * 40503u; h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
        bw[i] = 0x22222222u | (h & 0x88888888u);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot_ptr;
    const uint32_t tm_sf = tmem + 256u, tm_a = tmem + 320u;
    {
        const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
        uint32_t r[32];
#pragma unroll
        for (int c = 0; c < 32; ++c) r[c] = 0x7F7F7F7Fu;
        tc_st32(lane_base + 256u, r); tc_st32(lane_base + 288u, r);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
#pragma unroll
            for (int c = 0; c < 32; ++c) { uint32_t h = (uint32_t)(tid * 131 + q * 32 + c) * 2654435761u; h ^= h >> 15; r[c] = 0x22222222u | (h & 0x88888888u); }
            tc_st32(lane_base + 320u + (uint32_t)(q * 32), r);
        }
        tc_wait_st();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 0) {
        const uint64_t desc_hi = ((uint64_t)(1024u >> 4) << 32) | (1ull << 46) | (2ull << 61);
        long long c0 = clock64();
        unsigned long long t0; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        for (int j = 0; j < groups; ++j) {
#pragma unroll
            for (int g = 0; g < 4; ++g) {          // one "DB tile" of the sweep: 2 accumulators x 2 K-blocks x 4 instructions
                if (tc_elect_one()) {
                    const uint64_t b0 = desc_hi | (uint64_t)(((sB + (g & 1) * TC_KB_BYTES) & 0x3FFFFu) >> 4);
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        tc_mma_mxf4_ts(tmem + (uint32_t)((g >> 1) * TC_TILE), tm_a + (uint32_t)((g >> 1) * 64 + (g & 1) * 32 + k * 8), b0 + (uint64_t)(k * 2),
                                       (uint32_t)(((g & 1) | k) != 0), tm_sf, tm_sf + 32u);
                }
                __syncwarp();
            }
        }
        if (tc_elect_one()) tc_commit(bar_done);
        __syncwarp();
        tc_mbar_wait(bar_done, 0);
        const long long c1 = clock64();
        unsigned long long t1; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        if (lane == 0) { cycles_out[blockIdx.x] = c1 - c0; ns_out[blockIdx.x] = t1 - t0; }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
    }
}

// out[0] cycles per MMA (median SM)  out[1] MAC per clock per SM  out[2] SM MHz while it ran (cycles / globaltimer)
// out[3] chip TFLOP/s by CUDA events over the whole launch  out[4] SM count  out[5] MMAs per SM
int microbench_mma(double* out6) {
    int dev = 0, sms = 148;
    CLB_CUDA(cudaGetDevice(&dev));
    CLB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int smem = 2 * TC_KB_BYTES + 64 + 1024;
    CLB_CUDA(cudaFuncSetAttribute(tc_mma_bench_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    long long* d_cyc = nullptr; unsigned long long* d_ns = nullptr;
    CLB_CUDA(cudaMalloc(&d_cyc, sms * sizeof(long long)));
    CLB_CUDA(cudaMalloc(&d_ns, sms * sizeof(unsigned long long)));
    cudaEvent_t e0, e1;
    CLB_CUDA(cudaEventCreate(&e0)); CLB_CUDA(cudaEventCreate(&e1));
    const int groups = 20000;                    // 320 000 MMAs per SM: ~10 ms, long enough for the clock to settle
    float ms = 0.f;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0, 0);
        tc_mma_bench_kernel<<<sms, 128, smem>>>(groups, d_cyc, d_ns);
        cudaEventRecord(e1, 0);
        g_launches.fetch_add(1, std::memory_order_relaxed);
        cudaError_t e = cudaEventSynchronize(e1);
        if (e != cudaSuccess) { cudaFree(d_cyc); cudaFree(d_ns); return set_cuda_error(e, "tc_mma_bench_kernel", __FILE__, __LINE__); }
        cudaEventElapsedTime(&ms, e0, e1);
    }
    std::vector<long long> cyc(sms); std::vector<unsigned long long> ns(sms);
    cudaMemcpy(cyc.data(), d_cyc, sms * sizeof(long long), cudaMemcpyDeviceToHost);
    cudaMemcpy(ns.data(), d_ns, sms * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
    cudaFree(d_cyc); cudaFree(d_ns); cudaEventDestroy(e0); cudaEventDestroy(e1);
    std::vector<double> per(sms), mhz(sms);
    const double mmas = 16.0 * groups;
    for (int i = 0; i < sms; ++i) { per[i] = (double)cyc[i] / mmas; mhz[i] = ns[i] ? 1e3 * (double)cyc[i] / (double)ns[i] : 0.0; }
    std::sort(per.begin(), per.end()); std::sort(mhz.begin(), mhz.end());
    const double mac = (double)TC_TILE * TC_TILE * 64.0;
    out6[0] = per[sms / 2]; out6[1] = mac / per[sms / 2]; out6[2] = mhz[sms / 2];
    out6[3] = 2.0 * mac * mmas * sms / (ms * 1e-3) / 1e12; out6[4] = sms; out6[5] = mmas;
    return CLB_OK;
}

#ifdef CLB_LAB
int tc_lab_tile_clocks(long long* out1024) {
    CLB_CUDA(cudaDeviceSynchronize());
    CLB_CUDA(cudaMemcpyFromSymbol(out1024, g_tc_trace, sizeof(long long) * 6 * 256));
    return CLB_OK;
}
#endif

// ---- host side --------------------------------------------------------------------------------
bool tc_wanted(int nQ, int nT) {
    const int mode = g_match_tc;
    if (mode == 0 || nQ <= 0 || nT <= 0) return false;
    if (mode > 0) return true;
    return (long long)nQ * (long long)nT >= (1ll << 24);    // below that the expansion passes dominate
}

int tc_version() { return g_match_tc_ver; }
size_t tc_q_bytes(int nQ) { return (size_t)((nQ + TC_QROWS - 1) / TC_QROWS) * (tc_version() >= 3 ? TC3_SMEM_A : TC_SMEM_A); }
size_t tc_t_bytes(int nT) { return (size_t)((nT + TC_TILE - 1) / TC_TILE) * (tc_version() >= 3 ? TC3_TILE_BYTES : TC_TILE_BYTES); }

int launch_expand(const uint64_t* d_rows, int n, size_t out_bytes, void* d_out, cudaStream_t s) {
    NvtxRange nvtx_range("clb:expand");
    if (out_bytes == 0) return CLB_OK;
    const size_t chunks = out_bytes / 16;
    const size_t blocks = (chunks + 255) / 256;
    if (blocks > 0x7FFFFFFFull) { snprintf(g_err, sizeof(g_err), "descriptor set too large to expand"); return CLB_EINVAL; }
    timer("expand").begin(s);
#ifdef CLB_LAB
    if (tc_version() < 3)
        expand_kernel<<<(unsigned)blocks, 256, 0, s>>>(reinterpret_cast<const uint32_t*>(d_rows), n, chunks, reinterpret_cast<uint4*>(d_out));
    else
#endif
        expand4_kernel<<<(unsigned)blocks, 256, 0, s>>>(reinterpret_cast<const uint32_t*>(d_rows), n, chunks, reinterpret_cast<uint4*>(d_out));
    timer("expand").end(s);
    CLB_LAUNCHED();
    return CLB_OK;
}

// CTA time model (us) fitted to one-wave sweeps (tools/tc_phases.py: 8 tiles 22.9 us, 32: 45.5, 84: 84.2, 160: 135.3): ~15 us of
// launch + prologue (TMEM allocation, query expansion, first operands) + tail, 0.65 us per DB tile in steady state, and 0.29 us
// more per tile over the first ~50 tiles while the per-query thresholds are still open and most groups take the update path.
// (Before the float-domain epilogue state the same fit gave 28 us + 0.60 + 0.157 over the first 200 tiles.)
double tc_cta_cost_us(int tiles) {
    return 15.4 + 0.65 * (double)tiles + 0.29 * (double)(tiles < 50 ? tiles : 50);
}

// choose the DB chunk count: fewest chunks whose CTA count fills whole waves of `sm_count` SMs
void tc_plan(int nQ, int nT, int sm_count, int* qblocks_out, int* nchunks_out, int* tpc_out) {
    const int qblocks = (nQ + TC_QROWS - 1) / TC_QROWS;
    const int total_tiles = (nT + TC_TILE - 1) / TC_TILE;
    const int min_chunks = (total_tiles + TC_MAX_TILES_PER_CHUNK - 1) / TC_MAX_TILES_PER_CHUNK;
    int max_chunks = total_tiles / 4;                       // >= 4 tiles (512 rows) per chunk
    if (max_chunks > 256) max_chunks = 256;
    if (max_chunks < min_chunks) max_chunks = min_chunks;
    if (max_chunks < 1) max_chunks = 1;
    int best_c = min_chunks > 0 ? min_chunks : 1;
    const int forced = g_match_tc_chunks;
    if (forced > 0) {
        best_c = forced < max_chunks ? forced : max_chunks;
        if (best_c < min_chunks) best_c = min_chunks;
    } else {
        // the busiest SM runs `waves` CTAs of tc_cta_cost_us(tpc) back to back
        double best_t = 1e30;
        for (int c = best_c; c <= max_chunks; ++c) {
            const int tpc = (total_tiles + c - 1) / c;
            const int real_c = (total_tiles + tpc - 1) / tpc;
            const long long ctas = (long long)qblocks * real_c;
            const long long waves = (ctas + sm_count - 1) / sm_count;
            const double t = (double)waves * tc_cta_cost_us(tpc);
            if (t < best_t * 0.98) { best_t = t; best_c = c; }
        }
    }
    const int tpc = (total_tiles + best_c - 1) / best_c;
    *qblocks_out = qblocks;
    *tpc_out = tpc > 0 ? tpc : 1;
    *nchunks_out = total_tiles > 0 ? (total_tiles + *tpc_out - 1) / *tpc_out : 0;
}

bool tc_needs_qx() { return tc_version() == 1 || tc_version() == 3; }

// ---- batched pair lists: always the v4 kernel on E2M1 images, whatever match_tc_ver says -------
size_t tc4_t_bytes(int nT) { return (size_t)((nT + TC_TILE - 1) / TC_TILE) * TC3_TILE_BYTES; }

int launch_expand4(const uint64_t* d_rows, int n, void* d_out, cudaStream_t s) {
    NvtxRange nvtx_range("clb:expand");
    const size_t chunks = tc4_t_bytes(n) / 16;
    if (chunks == 0) return CLB_OK;
    const size_t blocks = (chunks + 255) / 256;
    if (blocks > 0x7FFFFFFFull) { snprintf(g_err, sizeof(g_err), "descriptor set too large to expand"); return CLB_EINVAL; }
    timer("expand").begin(s);
    expand4_kernel<<<(unsigned)blocks, 256, 0, s>>>(reinterpret_cast<const uint32_t*>(d_rows), n, chunks, reinterpret_cast<uint4*>(d_out));
    timer("expand").end(s);
    CLB_LAUNCHED();
    return CLB_OK;
}

// One chunk length (in DB tiles) for every pair of the list, chosen with the same CTA cost model as tc_plan: the
// CTA count is summed over the pairs, so small pairs fill each other's waves.
int tc_plan_pairs(const int* nQ, const int* nT, int npairs, int sm_count, double* t_us_out) {
    int max_tiles = 1;
    for (int p = 0; p < npairs; ++p) {
        const int t = (nT[p] + TC_TILE - 1) / TC_TILE;
        if (t > max_tiles) max_tiles = t;
    }
    double best_t = 1e30;
    int best_tpc = max_tiles < TC_MAX_TILES_PER_CHUNK ? max_tiles : TC_MAX_TILES_PER_CHUNK;
    for (int c = 1; c <= 256; ++c) {
        int tpc = (max_tiles + c - 1) / c;
        if (tpc > TC_MAX_TILES_PER_CHUNK) continue;
        if (tpc < 4 && c > 1) break;
        long long ctas = 0;
        for (int p = 0; p < npairs; ++p) {
            if (nQ[p] <= 0 || nT[p] <= 0) continue;
            const long long tiles = (nT[p] + TC_TILE - 1) / TC_TILE;
            ctas += (long long)((nQ[p] + TC_QROWS - 1) / TC_QROWS) * ((tiles + tpc - 1) / tpc);
        }
        const long long waves = (ctas + sm_count - 1) / sm_count;
        const double t = (double)waves * tc_cta_cost_us(tpc);
        if (t < best_t * 0.98) { best_t = t; best_tpc = tpc; }
    }
    if (t_us_out) *t_us_out = best_t;
    return best_tpc > 0 ? best_tpc : 1;
}

// balanced layout: `units` DB tiles in `strips` (view, query block) strips, cut into one contiguous range per SM (more when a range
// would exceed what a packed key can index); an SM then runs its range as 1 + (strips that straddle a border) jobs
double tc_balanced_cost_us(long long units, long long strips, int sm_count, long long* parts_out) {
    long long parts = sm_count;
    while (units / parts > 30000) parts += sm_count;
    if (parts_out) *parts_out = parts;
    if (units <= 0) return 0.0;
    const long long quota = (units + parts - 1) / parts;
    const double extra_jobs = (double)(strips < parts ? strips : parts) / (double)parts;
    return 1.03 * (tc_cta_cost_us((int)quota) + 12.0 * extra_jobs);   // a job that starts mid-range: prologue + part of the warm-up again
}

int launch_match_tc_jobs(const TcJob* d_jobs, int njobs, cudaStream_t s) {
    NvtxRange nvtx_range("clb:match");
    if (njobs <= 0) return CLB_OK;
    static bool attr_done[64] = {false};
    int dev = 0;
    CLB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_done[dev]) {
        CLB_CUDA(cudaFuncSetAttribute(match_tc4_jobs_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TC4_SMEM_TOTAL));
        if (dev >= 0 && dev < 64) attr_done[dev] = true;
    }
    timer("match").begin(s);
    match_tc4_jobs_kernel<<<njobs, TC4_THREADS, TC4_SMEM_TOTAL, s>>>(d_jobs);
    timer("match").end(s);
    CLB_LAUNCHED();
    return CLB_OK;
}

int launch_match_tc_kernel(const uint64_t* dQ, const void* Qx, int nQ, const void* Tx, int nT, int qblocks, int nchunks, int tpc,
                           uint2* part, cudaStream_t s) {
    static bool attr_done[64] = {false};
    int dev = 0;
    CLB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_done[dev]) {
        CLB_CUDA(cudaFuncSetAttribute(match_tc4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TC4_SMEM_TOTAL));
#ifdef CLB_LAB
        CLB_CUDA(cudaFuncSetAttribute(match_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_TOTAL));
        CLB_CUDA(cudaFuncSetAttribute(match_tc2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TC2_SMEM_TOTAL));
        CLB_CUDA(cudaFuncSetAttribute(match_tc3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TC3_SMEM_TOTAL));
        CLB_CUDA(cudaFuncSetAttribute(match_tc4_lab_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TC4_SMEM_TOTAL));
#endif
        if (dev >= 0 && dev < 64) attr_done[dev] = true;
    }
    dim3 grid(qblocks, nchunks);
    timer("match").begin(s);
#ifdef CLB_LAB
    const int ver = tc_version(), dbg = g_match_tc_dbg;
    if (ver == 4 && dbg != 0)
        match_tc4_lab_kernel<<<grid, TC4_THREADS, TC4_SMEM_TOTAL, s>>>(reinterpret_cast<const uint32_t*>(dQ), nQ,
                                                                     reinterpret_cast<const uint8_t*>(Tx), nT, tpc, part, dbg);
    else if (ver == 3 && Qx)
        match_tc3_kernel<<<grid, TC_THREADS, TC3_SMEM_TOTAL, s>>>(reinterpret_cast<const uint8_t*>(Qx), nQ,
                                                                reinterpret_cast<const uint8_t*>(Tx), nT, tpc, part, dbg);
    else if (ver == 1 && Qx)
        match_tc_kernel<<<grid, TC_THREADS, TC_SMEM_TOTAL, s>>>(reinterpret_cast<const uint8_t*>(Qx), nQ,
                                                              reinterpret_cast<const uint8_t*>(Tx), nT, tpc, part, dbg);
    else if (ver == 2 || ver == 3 || ver == 1)
        match_tc2_kernel<<<grid, TC_THREADS, TC2_SMEM_TOTAL, s>>>(reinterpret_cast<const uint32_t*>(dQ), nQ,
                                                                reinterpret_cast<const uint8_t*>(Tx), nT, tpc, part, dbg);
    else
#endif
        match_tc4_kernel<<<grid, TC4_THREADS, TC4_SMEM_TOTAL, s>>>(reinterpret_cast<const uint32_t*>(dQ), nQ,
                                                                 reinterpret_cast<const uint8_t*>(Tx), nT, tpc, part);
    (void)Qx;
    timer("match").end(s);
    CLB_LAUNCHED();
    return CLB_OK;
}

}  // namespace clb
