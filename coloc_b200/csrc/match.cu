// match.cu -- brute-force 512-bit Hamming 2-NN on the INT/POPC pipes of sm_100a.
//
// Replaces CUDAK2NN (reference src/CUDAK2NN.cu:42-81) and its host wrapper
// GPUMatcher::computeMatches (include/coloc/GPUMatcher.hpp:180-226).  Semantics kept bit-exact:
// rows scanned in ascending order, first row attaining the minimum wins, `second` is the second
// smallest distance (== best on ties), match iff second - best > threshold (CUDAK2NN.cu:67-75).
//
// Design (not a port): one query per lane held in registers (RQ queries per thread), DB rows
// staged into shared memory by 1-D TMA bulk copies (cp.async.bulk + mbarrier, 4-stage ring) and
// read back as warp-broadcast LDS.128; the 16-word XOR stream is compressed by a carry-save
// adder tree (LOP3 on the ALU pipe) before the POPCs so the narrow POPC pipe sees 8 (or 5/11/16)
// instead of 16 ops per comparison; (distance,index) packed into one u32 key so best/second
// tracking is 3 min/max ops.  The DB is cut into chunks across blockIdx.y so small query sets
// still fill 148 SMs; a merge kernel folds the per-chunk (best,second) pairs in ascending chunk
// order, which preserves first-min-wins, and applies the margin (or ratio) test.
#include "common.cuh"

namespace clb {

static constexpr int MT_THREADS = 256;
static constexpr int MT_TILE_ROWS = 64;             // 64 rows x 64 B = 4 KB per stage
static constexpr int MT_STAGES = 4;
static constexpr int MT_KEY_SHIFT = 22;             // key = dist << 22 | row-in-chunk
static constexpr uint32_t MT_IDX_MASK = (1u << MT_KEY_SHIFT) - 1;
static constexpr int MT_MAX_CHUNK_ROWS = 1 << MT_KEY_SHIFT;

std::atomic<int> g_match_rq{0};      // 0 = heuristic
std::atomic<int> g_match_chunks{0};  // 0 = heuristic
std::atomic<int> g_last_match_path{0};   // 0 = POPC kernel, 4 = tensor-core kernel (what the last launch_match used)
#ifdef CLB_LAB
std::atomic<int> g_match_hs{2};      // carry-save depth: 0 none(16 popc) 1 (11) 2 (8) 3 (5)
#else
static constexpr int g_match_hs = 2; // the product build carries the 8-POPC tree only
#endif

// ---- PTX helpers ----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\t"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    } while (!ok);
}
// 1-D TMA bulk copy global -> shared, completion on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void tma_load_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ uint32_t lop3_xor3(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r; asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r;
}
__device__ __forceinline__ uint32_t lop3_maj(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r; asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r;
}
#define CSA(hi, lo, a, b, c) do { lo = lop3_xor3(a, b, c); hi = lop3_maj(a, b, c); } while (0)

// popcount of the 512-bit XOR of q and t; HS selects how deep the carry-save tree goes
template <int HS>
__device__ __forceinline__ uint32_t ham512(const uint32_t (&q)[16], const uint32_t (&t)[16]) {
    uint32_t x[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = q[i] ^ t[i];
    if (HS == 0) {
        uint32_t d = 0;
#pragma unroll
        for (int i = 0; i < 16; ++i) d += __popc(x[i]);
        return d;
    }
    // level 1: five 3:2 compressors over x0..x14 -> ones s0..s4 (+x15), twos c0..c4
    uint32_t s0, s1, s2, s3, s4, c0, c1, c2, c3, c4;
    CSA(c0, s0, x[0], x[1], x[2]);
    CSA(c1, s1, x[3], x[4], x[5]);
    CSA(c2, s2, x[6], x[7], x[8]);
    CSA(c3, s3, x[9], x[10], x[11]);
    CSA(c4, s4, x[12], x[13], x[14]);
    if (HS == 1) {   // 11 popc
        uint32_t ones = __popc(s0) + __popc(s1) + __popc(s2) + __popc(s3) + __popc(s4) + __popc(x[15]);
        uint32_t twos = __popc(c0) + __popc(c1) + __popc(c2) + __popc(c3) + __popc(c4);
        return ones + 2 * twos;
    }
    uint32_t o0, o1, c5, c6, t0, f0;
    CSA(c5, o0, s0, s1, s2);
    CSA(c6, o1, s3, s4, x[15]);
    CSA(f0, t0, c0, c1, c2);
    if (HS == 2) {   // 8 popc
        uint32_t ones = __popc(o0) + __popc(o1);
        uint32_t twos = __popc(t0) + __popc(c3) + __popc(c4) + __popc(c5) + __popc(c6);
        return ones + 2 * twos + 4 * __popc(f0);
    }
    // HS == 3: 5 popc (o0,o1 stay; twos t0,c3..c6 -> 1 two + 2 fours; fours f0,f1,f2 -> 1 four + 1 eight)
    uint32_t f1, f2, t1, t2, e0, f3;
    CSA(f1, t1, t0, c3, c4);
    CSA(f2, t2, t1, c5, c6);
    CSA(e0, f3, f0, f1, f2);
    uint32_t ones = __popc(o0) + __popc(o1);
    return ones + 2 * __popc(t2) + 4 * __popc(f3) + 8 * __popc(e0);
}

// ---- main kernel ----------------------------------------------------------------------------
// grid = (query blocks, DB chunks).  partial[chunk][q] = (best_key, second_key), 0xFFFFFFFF = none.
template <int RQ, int HS>
__global__ void __launch_bounds__(MT_THREADS, RQ >= 4 ? 2 : 3)
match_kernel(const uint4* __restrict__ Q, int nQ, const uint4* __restrict__ T, int nT, int rows_per_chunk,
             uint2* __restrict__ partial) {
    __shared__ __align__(128) uint4 tile[MT_STAGES][MT_TILE_ROWS * 4];
    __shared__ __align__(8) uint64_t full_bar[MT_STAGES];
    __shared__ __align__(8) uint64_t empty_bar[MT_STAGES];

    const int tid = threadIdx.x;
    const int chunk = blockIdx.y;
    const int row0 = chunk * rows_per_chunk;
    const int rows = min(rows_per_chunk, nT - row0);
    const int ntiles = (rows + MT_TILE_ROWS - 1) / MT_TILE_ROWS;
    const uint4* Tc = T + (size_t)row0 * 4;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < MT_STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], MT_THREADS / 32); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    auto issue = [&](int t) {   // thread 0 only
        const int s = t % MT_STAGES;
        const int r = min(MT_TILE_ROWS, rows - t * MT_TILE_ROWS);
        mbar_expect_tx(&full_bar[s], (uint32_t)r * 64u);
        tma_load_1d(&tile[s][0], Tc + (size_t)t * MT_TILE_ROWS * 4, (uint32_t)r * 64u, &full_bar[s]);
    };
    if (tid == 0) {
        for (int t = 0; t < MT_STAGES - 1 && t < ntiles; ++t) issue(t);
    }

    // queries: one 64-byte row per lane per j, kept in registers for the whole sweep
    uint32_t q[RQ][16];
    const int qbase = blockIdx.x * (MT_THREADS * RQ) + tid;
#pragma unroll
    for (int j = 0; j < RQ; ++j) {
        const int qi = min(qbase + j * MT_THREADS, nQ - 1);
        const uint4* src = Q + (size_t)qi * 4;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint4 v = __ldg(src + k);
            q[j][4 * k + 0] = v.x; q[j][4 * k + 1] = v.y; q[j][4 * k + 2] = v.z; q[j][4 * k + 3] = v.w;
        }
    }
    uint32_t best[RQ], second[RQ];
#pragma unroll
    for (int j = 0; j < RQ; ++j) { best[j] = 0xFFFFFFFFu; second[j] = 0xFFFFFFFFu; }

    for (int it = 0; it < ntiles; ++it) {
        const int s = it % MT_STAGES;
        if (tid == 0) {
            const int nt = it + MT_STAGES - 1;          // refill the stage tile it-1 used
            if (nt < ntiles) {
                if (it > 0) mbar_wait(&empty_bar[(it - 1) % MT_STAGES], (uint32_t)(((it - 1) / MT_STAGES) & 1));
                issue(nt);
            }
        }
        mbar_wait(&full_bar[s], (uint32_t)((it / MT_STAGES) & 1));
        const int r_end = min(MT_TILE_ROWS, rows - it * MT_TILE_ROWS);
        const uint4* tp = &tile[s][0];
        uint32_t key_lo = (uint32_t)(it * MT_TILE_ROWS);
#pragma unroll 2
        for (int r = 0; r < r_end; ++r, ++key_lo) {
            uint32_t t[16];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint4 v = tp[r * 4 + k];            // same address in every lane: broadcast
                t[4 * k + 0] = v.x; t[4 * k + 1] = v.y; t[4 * k + 2] = v.z; t[4 * k + 3] = v.w;
            }
#pragma unroll
            for (int j = 0; j < RQ; ++j) {
                const uint32_t d = ham512<HS>(q[j], t);
                const uint32_t key = (d << MT_KEY_SHIFT) + key_lo;
                const uint32_t m = max(key, best[j]);
                second[j] = min(second[j], m);
                best[j] = min(best[j], key);
            }
        }
        __syncwarp();
        if ((tid & 31) == 0) mbar_arrive(&empty_bar[s]);
    }

#pragma unroll
    for (int j = 0; j < RQ; ++j) {
        const int qi = qbase + j * MT_THREADS;
        if (qi < nQ) partial[(size_t)chunk * nQ + qi] = make_uint2(best[j], second[j]);
    }
}

// ---- merge: fold per-chunk pairs in ascending chunk order, apply the decision -------------------
__device__ __forceinline__ void feed(int d, int idx, int& best, int& best_i, int& second) {
    // CUDAK2NN.cu:67-72 recurrence
    second = min(d, second);
    if (d < best) { second = best; best_i = idx; best = d; }
}

__global__ void __launch_bounds__(256)
merge_chunks_kernel(const uint2* __restrict__ partial, int nchunks, int rows_per_chunk, int nQ, int nT,
                    long long t_base, int threshold, float ratio,
                    int32_t* __restrict__ match, int32_t* __restrict__ best_out, int32_t* __restrict__ second_out,
                    clb_partial* __restrict__ part_out) {
    const int qi = blockIdx.x * blockDim.x + threadIdx.x;
    if (qi >= nQ) return;
    int best = CLB_BEST_NONE, second = CLB_SECOND_NONE, best_i = -1;
    for (int c = 0; c < nchunks; ++c) {
        const uint2 p = partial[(size_t)c * nQ + qi];
        if (p.x == 0xFFFFFFFFu) continue;
        feed((int)(p.x >> MT_KEY_SHIFT), c * rows_per_chunk + (int)(p.x & MT_IDX_MASK), best, best_i, second);
        feed(p.y == 0xFFFFFFFFu ? CLB_BEST_NONE : (int)(p.y >> MT_KEY_SHIFT), -1, best, best_i, second);
    }
    if (part_out) {
        clb_partial o; o.best_d = best; o.best_i = best_i < 0 ? -1 : (int32_t)(t_base + best_i); o.second_d = second; o._pad = 0;
        part_out[qi] = o;
    }
    if (match) {
        bool ok;
        if (ratio > 0.f) ok = nT >= 2 && (float)best < ratio * (float)second;   // matching_filters.hpp:56
        else             ok = nT > 0 && (second - best > threshold);            // CUDAK2NN.cu:75
        match[qi] = ok ? best_i : CLB_NO_MATCH;
    }
    if (best_out) best_out[qi] = best;
    if (second_out) second_out[qi] = second;
}

__global__ void __launch_bounds__(256)
merge_partials_kernel(const clb_partial* __restrict__ parts, int nparts, int nQ, int threshold,
                      int32_t* __restrict__ match, int32_t* __restrict__ best_out, int32_t* __restrict__ second_out) {
    const int qi = blockIdx.x * blockDim.x + threadIdx.x;
    if (qi >= nQ) return;
    int best = CLB_BEST_NONE, second = CLB_SECOND_NONE, best_i = -1;
    bool any = false;
    for (int p = 0; p < nparts; ++p) {
        const clb_partial v = parts[(size_t)p * nQ + qi];
        if (v.best_i < 0) continue;
        any = true;
        feed(v.best_d, v.best_i, best, best_i, second);
        feed(v.second_d, -1, best, best_i, second);
    }
    if (match) match[qi] = (any && second - best > threshold) ? best_i : CLB_NO_MATCH;
    if (best_out) best_out[qi] = best;
    if (second_out) second_out[qi] = second;
}

// several partial arrays [part][nQ] (ascending t_base) folded into one partial array: the slices of a streamed DB
__global__ void __launch_bounds__(256)
fold_partials_kernel(const clb_partial* __restrict__ parts, int nparts, int nQ, clb_partial* __restrict__ out) {
    const int qi = blockIdx.x * blockDim.x + threadIdx.x;
    if (qi >= nQ) return;
    int best = CLB_BEST_NONE, second = CLB_SECOND_NONE, best_i = -1;
    for (int p = 0; p < nparts; ++p) {
        const clb_partial v = parts[(size_t)p * nQ + qi];
        if (v.best_i < 0) continue;
        feed(v.best_d, v.best_i, best, best_i, second);
        feed(v.second_d, -1, best, best_i, second);
    }
    clb_partial o; o.best_d = best; o.best_i = best_i; o.second_d = second; o._pad = 0;
    out[qi] = o;
}

// ---- host side ------------------------------------------------------------------------------
struct MatchPlan { int rq, qblocks, nchunks, rpc; };

static MatchPlan plan_match(int nQ, int nT, int sm_count) {
    MatchPlan p;
    const int rq_knob = g_match_rq, chunks_knob = g_match_chunks;
    p.rq = rq_knob ? rq_knob : (nQ >= 2048 ? 4 : (nQ >= 512 ? 2 : 1));
    p.qblocks = (nQ + MT_THREADS * p.rq - 1) / (MT_THREADS * p.rq);
    const int per_sm = p.rq >= 4 ? 2 : 3;
    const int target = sm_count * per_sm * 3;                       // ~3 waves of resident CTAs
    int nchunks = chunks_knob ? chunks_knob : (target + p.qblocks - 1) / p.qblocks;
    const int max_chunks = (nT + 4 * MT_TILE_ROWS - 1) / (4 * MT_TILE_ROWS);   // >= 256 rows per chunk
    if (nchunks > max_chunks) nchunks = max_chunks;
    const int min_chunks = (nT + MT_MAX_CHUNK_ROWS - 1) / MT_MAX_CHUNK_ROWS;
    if (nchunks < min_chunks) nchunks = min_chunks;
    if (nchunks < 1) nchunks = 1;
    int rpc = (nT + nchunks - 1) / nchunks;
    rpc = (rpc + MT_TILE_ROWS - 1) / MT_TILE_ROWS * MT_TILE_ROWS;
    if (rpc < MT_TILE_ROWS) rpc = MT_TILE_ROWS;
    p.rpc = rpc;
    p.nchunks = nT > 0 ? (nT + rpc - 1) / rpc : 0;
    return p;
}

size_t match_scratch_bytes(int nQ, int nT) {
    // worst case over the knobs: chunks are at least 256 rows
    size_t max_chunks = (size_t)(nT + 4 * MT_TILE_ROWS - 1) / (4 * MT_TILE_ROWS) + 1;
    if (max_chunks > 4096) max_chunks = 4096;
    if (max_chunks < 4) max_chunks = 4;
    return max_chunks * (size_t)(nQ > 0 ? nQ : 1) * sizeof(uint2);
}

template <int RQ>
static void launch_rq(int hs, dim3 grid, cudaStream_t s, const uint4* Q, int nQ, const uint4* T, int nT, int rpc, uint2* part) {
    switch (hs) {
#ifdef CLB_LAB
        case 0: match_kernel<RQ, 0><<<grid, MT_THREADS, 0, s>>>(Q, nQ, T, nT, rpc, part); break;
        case 1: match_kernel<RQ, 1><<<grid, MT_THREADS, 0, s>>>(Q, nQ, T, nT, rpc, part); break;
        case 3: match_kernel<RQ, 3><<<grid, MT_THREADS, 0, s>>>(Q, nQ, T, nT, rpc, part); break;
#endif
        default: match_kernel<RQ, 2><<<grid, MT_THREADS, 0, s>>>(Q, nQ, T, nT, rpc, part); break;
    }
}

int launch_match(const uint64_t* dQ, int nQ, const uint64_t* dT, int nT, int64_t t_base,
                 void* scratch, size_t scratch_bytes, int sm_count,
                 int threshold, float ratio, int32_t* d_match, int32_t* d_best, int32_t* d_second,
                 clb_partial* d_partial, cudaStream_t s, const void* Qx, const void* Tx) {
    NvtxRange nvtx_range("clb:match");
    if (nQ <= 0) return CLB_OK;
    if (((uintptr_t)dQ & 15) || ((uintptr_t)dT & 15)) { snprintf(g_err, sizeof(g_err), "descriptor pointers must be 16-byte aligned"); return CLB_EINVAL; }
    if (Tx && nT > 0) {
        // tensor-core sweep (match_tc.cu) feeding the same per-chunk key pairs into the same merge
        int qblocks, nchunks, tpc;
        tc_plan(nQ, nT, sm_count, &qblocks, &nchunks, &tpc);
        if ((size_t)nchunks * nQ * sizeof(uint2) > scratch_bytes) { snprintf(g_err, sizeof(g_err), "match scratch too small"); return CLB_EINVAL; }
        if (nchunks > 65535) { snprintf(g_err, sizeof(g_err), "too many DB chunks"); return CLB_EINVAL; }
        uint2* part = reinterpret_cast<uint2*>(scratch);
        g_last_match_path = (tc_version() == 4 || Qx) ? tc_version() : 2;
        int rc = launch_match_tc_kernel(dQ, Qx, nQ, Tx, nT, qblocks, nchunks, tpc, part, s);
        if (rc) return rc;
        timer("merge").begin(s);
        merge_chunks_kernel<<<(nQ + 255) / 256, 256, 0, s>>>(part, nchunks, tpc * 128, nQ, nT, (long long)t_base, threshold, ratio,
                                                            d_match, d_best, d_second, d_partial);
        timer("merge").end(s);
        CLB_LAUNCHED();
        return CLB_OK;
    }
    g_last_match_path = 0;
    MatchPlan p = plan_match(nQ, nT, sm_count);
    if ((size_t)p.nchunks * nQ * sizeof(uint2) > scratch_bytes) { snprintf(g_err, sizeof(g_err), "match scratch too small"); return CLB_EINVAL; }
    if (p.nchunks > 65535) { snprintf(g_err, sizeof(g_err), "too many DB chunks"); return CLB_EINVAL; }
    uint2* part = reinterpret_cast<uint2*>(scratch);
    if (p.nchunks > 0) {
        dim3 grid(p.qblocks, p.nchunks);
        const uint4* Q4 = reinterpret_cast<const uint4*>(dQ);
        const uint4* T4 = reinterpret_cast<const uint4*>(dT);
        timer("match").begin(s);
        if (p.rq >= 4) launch_rq<4>(g_match_hs, grid, s, Q4, nQ, T4, nT, p.rpc, part);
        else if (p.rq == 2) launch_rq<2>(g_match_hs, grid, s, Q4, nQ, T4, nT, p.rpc, part);
        else launch_rq<1>(g_match_hs, grid, s, Q4, nQ, T4, nT, p.rpc, part);
        timer("match").end(s);
        CLB_LAUNCHED();
    }
    timer("merge").begin(s);
    merge_chunks_kernel<<<(nQ + 255) / 256, 256, 0, s>>>(part, p.nchunks, p.rpc, nQ, nT, (long long)t_base, threshold, ratio,
                                                        d_match, d_best, d_second, d_partial);
    timer("merge").end(s);
    CLB_LAUNCHED();
    return CLB_OK;
}

// batched merge_chunks_kernel: blockIdx.y = pair of the list; pieces are uniform chunks or the per-query-block pieces of a
// balanced job list
__global__ void __launch_bounds__(256)
merge_jobs_kernel(const MergeJob* __restrict__ jobs, int threshold, float ratio) {
    const MergeJob j = jobs[blockIdx.y];
    const int qi = blockIdx.x * blockDim.x + threadIdx.x;
    if (qi >= j.nQ) return;
    const int qb = qi >> 8;
    const int np = j.pieces ? j.pieces[qb] : j.nchunks;
    int best = CLB_BEST_NONE, second = CLB_SECOND_NONE, best_i = -1;
    for (int c = 0; c < np; ++c) {
        const uint2 p = j.partial[(size_t)c * j.nQ + qi];
        if (p.x == 0xFFFFFFFFu) continue;
        const int base = j.pieces ? j.base_rows[qb * j.maxp + c] : c * j.rows_per_chunk;
        feed((int)(p.x >> MT_KEY_SHIFT), base + (int)(p.x & MT_IDX_MASK), best, best_i, second);
        feed(p.y == 0xFFFFFFFFu ? CLB_BEST_NONE : (int)(p.y >> MT_KEY_SHIFT), -1, best, best_i, second);
    }
    if (j.part_out) {
        clb_partial o; o.best_d = best; o.best_i = best_i < 0 ? -1 : (int32_t)(j.t_base + best_i); o.second_d = second; o._pad = 0;
        j.part_out[qi] = o;
    }
    if (j.match) {
        bool ok;
        if (ratio > 0.f) ok = j.nT >= 2 && (float)best < ratio * (float)second;
        else             ok = j.nT > 0 && (second - best > threshold);
        j.match[qi] = ok ? best_i : CLB_NO_MATCH;
    }
    if (j.best) j.best[qi] = best;
    if (j.second) j.second[qi] = second;
}

int launch_merge_jobs(const MergeJob* d_jobs, int njobs, int max_nQ, int threshold, float ratio, cudaStream_t s) {
    NvtxRange nvtx_range("clb:merge");
    if (njobs <= 0 || max_nQ <= 0) return CLB_OK;
    if (njobs > 65535) { snprintf(g_err, sizeof(g_err), "too many pairs in one list"); return CLB_EINVAL; }
    dim3 grid((max_nQ + 255) / 256, njobs);
    timer("merge").begin(s);
    merge_jobs_kernel<<<grid, 256, 0, s>>>(d_jobs, threshold, ratio);
    timer("merge").end(s);
    CLB_LAUNCHED();
    return CLB_OK;
}

int launch_fold_partials(const clb_partial* parts, int nparts, int nQ, clb_partial* out, cudaStream_t s) {
    if (nQ <= 0) return CLB_OK;
    fold_partials_kernel<<<(nQ + 255) / 256, 256, 0, s>>>(parts, nparts, nQ, out);
    CLB_LAUNCHED();
    return CLB_OK;
}

int launch_merge_partials(const clb_partial* parts, int nparts, int nQ, int threshold,
                          int32_t* d_match, int32_t* d_best, int32_t* d_second, cudaStream_t s) {
    NvtxRange nvtx_range("clb:merge");
    if (nQ <= 0) return CLB_OK;
    merge_partials_kernel<<<(nQ + 255) / 256, 256, 0, s>>>(parts, nparts, nQ, threshold, d_match, d_best, d_second);
    CLB_LAUNCHED();
    return CLB_OK;
}

// ---- allocation-free entry point (clb_match_2nn_noalloc) ------------------------------------------------------------------
// For callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it): the POPC sweep
// (match_kernel<RQ, 2>) + merge_chunks_kernel with the per-chunk (best, second) key pairs -- the only scratch this path needs --
// in a file-scope device array instead of the matcher object's cudaMalloc'ed buffer.  4 MiB, held by every GPU that loads the
// library until the process exits.  In host code `g_noalloc_scratch` is only a stand-in for the symbol: the device address comes
// from cudaGetSymbolAddress, once per device with that device current, and is what the kernels are given.
static constexpr size_t NOALLOC_SCRATCH_BYTES = 4u << 20;
static constexpr size_t NOALLOC_ENTRIES = NOALLOC_SCRATCH_BYTES / sizeof(uint2);       // 524 288 (chunk, query) slots
static constexpr int NOALLOC_MAX_Q = 32768;                                            // leaves >= 16 chunks at the largest nQ
static constexpr long long NOALLOC_MAX_T = 16ll * MT_MAX_CHUNK_ROWS;                   // 2^26 rows: 16 chunks of 2^22 rows at most
static __device__ __align__(256) uint2 g_noalloc_scratch[NOALLOC_ENTRIES];

static uint2* g_noalloc_ptr[64] = {nullptr};      // per device: address of the array ON that device
static int g_noalloc_sms[64] = {0};

int match_2nn_noalloc(const uint64_t* dQ, int nQ, const uint64_t* dT, int nT, int threshold,
                      int32_t* d_match, int32_t* d_best, int32_t* d_second, cudaStream_t s) {
    if (nQ < 0 || nT < 0 || (nQ > 0 && (!dQ || !d_match)) || (nT > 0 && !dT)) { snprintf(g_err, sizeof(g_err), "bad descriptor set"); return CLB_EINVAL; }
    if (nQ > NOALLOC_MAX_Q || (long long)nT > NOALLOC_MAX_T) {
        snprintf(g_err, sizeof(g_err), "clb_match_2nn_noalloc: %d queries x %d rows exceed the static scratch (at most %d queries, %lld rows); "
                 "use clb_match_2nn_dev", nQ, nT, NOALLOC_MAX_Q, NOALLOC_MAX_T);
        return CLB_EINVAL;
    }
    if (nQ == 0) return CLB_OK;
    if (((uintptr_t)dQ & 15) || ((uintptr_t)dT & 15)) { snprintf(g_err, sizeof(g_err), "descriptor pointers must be 16-byte aligned"); return CLB_EINVAL; }
    int dev = 0;
    CLB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) { snprintf(g_err, sizeof(g_err), "device index %d out of range", dev); return CLB_EINVAL; }
    if (!g_noalloc_ptr[dev]) {                     // set-up, once per device, with that device current
        uint2* ptr = nullptr;
        CLB_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_noalloc_scratch));
        int sms = 0;
        CLB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        g_noalloc_sms[dev] = sms > 0 ? sms : 148;
        g_noalloc_ptr[dev] = ptr;
    }
    uint2* part = g_noalloc_ptr[dev];
    // chunk plan of the POPC kernel, with the chunk count capped by what the static array holds for nQ queries
    MatchPlan p = plan_match(nQ, nT, g_noalloc_sms[dev]);
    const int fit = (int)(NOALLOC_ENTRIES / (size_t)nQ);                    // >= 16
    if (p.nchunks > fit) {
        int rpc = (nT + fit - 1) / fit;
        rpc = (rpc + MT_TILE_ROWS - 1) / MT_TILE_ROWS * MT_TILE_ROWS;
        p.rpc = rpc; p.nchunks = (nT + rpc - 1) / rpc;
    }
    if (p.rpc > MT_MAX_CHUNK_ROWS || (size_t)p.nchunks * (size_t)nQ > NOALLOC_ENTRIES || p.nchunks > 65535) {
        snprintf(g_err, sizeof(g_err), "clb_match_2nn_noalloc: no chunk plan fits the static scratch for %d x %d", nQ, nT); return CLB_EINVAL;
    }
    // every (chunk, query) slot the merge reads is written by the sweep of THIS call: nothing has to start at zero, and what
    // the previous call left in the array is never read
    if (p.nchunks > 0) {
        dim3 grid(p.qblocks, p.nchunks);
        const uint4* Q4 = reinterpret_cast<const uint4*>(dQ);
        const uint4* T4 = reinterpret_cast<const uint4*>(dT);
        if (p.rq >= 4) match_kernel<4, 2><<<grid, MT_THREADS, 0, s>>>(Q4, nQ, T4, nT, p.rpc, part);
        else if (p.rq == 2) match_kernel<2, 2><<<grid, MT_THREADS, 0, s>>>(Q4, nQ, T4, nT, p.rpc, part);
        else match_kernel<1, 2><<<grid, MT_THREADS, 0, s>>>(Q4, nQ, T4, nT, p.rpc, part);
        CLB_LAUNCHED();
    }
    merge_chunks_kernel<<<(nQ + 255) / 256, 256, 0, s>>>(part, p.nchunks, p.rpc, nQ, nT, 0ll, threshold, 0.f, d_match, d_best, d_second, nullptr);
    CLB_LAUNCHED();
    return CLB_OK;
}

}  // namespace clb

extern "C" int clb_match_2nn_noalloc(const uint64_t* dQ, int nQ, const uint64_t* dT, int nT, int threshold,
                                     int32_t* d_match_out, int32_t* d_best_out, int32_t* d_second_out, void* stream) {
    // NULL = the legacy default stream (there is no object whose stream it could mean); cudaStreamLegacy / PerThread pass through
    return clb::match_2nn_noalloc(dQ, nQ, dT, nT, threshold, d_match_out, d_best_out, d_second_out, (cudaStream_t)stream);
}
