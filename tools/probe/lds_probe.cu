// lds_probe.cu -- which lanes of a warp share a shared-memory wavefront on a 64-bit load (B200)?  One launch per address pattern,
// one warp, the same LDS.64 executed 1024 times; run under
//   ncu --metrics l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum,smsp__inst_executed_op_shared_ld.sum --csv
// and read wavefronts / 1024 per launch.  Patterns: a few synthetic ones, then every (warp, step, window, row) load of latch_kernel
// with the model's own prediction (half-warp rule) printed beside it.
// Build: nvcc -O2 -gencode arch=compute_100a,code=sm_100a -I include -I coloc_b200/csrc tools/probe/lds_probe.cu -o /tmp/lds_probe
#include <cstdio>
#include <cstdint>
#include <vector>
#include <map>
#include <set>
#include <algorithm>
#include <cuda_runtime.h>
#include "coloc_b200_latch_table.h"
#include "latch_sched.h"

__global__ void lds64_probe(const int* __restrict__ offs, int iters, unsigned long long* out) {
    extern __shared__ __align__(16) unsigned char sm[];
    for (int i = threadIdx.x; i < 45056 / 4; i += 32) reinterpret_cast<uint32_t*>(sm)[i] = (uint32_t)i * 2654435761u;
    __syncwarp();
    const int o = offs[threadIdx.x];
    uint32_t ax = 0, ay = 0;
    for (int it = 0; it < iters; ++it) {
        uint32_t vx, vy;
        asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(vx), "=r"(vy) : "r"((uint32_t)__cvta_generic_to_shared(sm + o)) : "memory");
        ax += vx; ay ^= vy;
    }
    out[threadIdx.x] = (unsigned long long)ax + ay;
}

static int model_halfwarp(const int* o) {      // wavefronts if lanes 0-15 / 16-31 are served separately, 32 banks of 4 bytes
    int tot = 0;
    for (int h = 0; h < 2; ++h) {
        std::map<int, std::set<int>> banks;
        for (int l = 16 * h; l < 16 * h + 16; ++l)
            for (int w = 0; w < 2; ++w) banks[(o[l] / 4 + w) % 32].insert(o[l] / 4 + w);
        size_t m = 0;
        for (auto& kv : banks) m = std::max(m, kv.second.size());
        tot += (int)m;
    }
    return tot;
}

int main() {
    std::vector<std::vector<int>> pats; std::vector<const char*> names;
    auto add = [&](const char* n, std::vector<int> v) { pats.push_back(v); names.push_back(n); };
    { std::vector<int> v(32, 64); add("broadcast", v); }
    { std::vector<int> v(32); for (int i = 0; i < 32; ++i) v[i] = 8 * i; add("contiguous 256 B", v); }
    { std::vector<int> v(32); for (int i = 0; i < 32; ++i) v[i] = 8 * (i & 15) + (i >> 4) * 1280; add("halves on the same banks, other row", v); }
    { std::vector<int> v(32); for (int i = 0; i < 32; ++i) v[i] = 8 * (i >> 1) + (i & 1) * 1280; add("even/odd lanes on the same banks", v); }
    { std::vector<int> v(32); for (int i = 0; i < 32; ++i) v[i] = 8 * (i >> 2); add("4 lanes per address, 64 B (schedule table load)", v); }
    { std::vector<int> v(32); for (int i = 0; i < 32; ++i) v[i] = 8 * (i & 7) + (i >> 3) * 1280; add("quarters on the same 16 banks", v); }
    { std::vector<int> v(32); for (int i = 0; i < 32; ++i) v[i] = 8 * (i & 7) + (i >> 3) * (1280 + 64); add("quarters: 0/2 and 1/3 share banks", v); }
    { std::vector<int> v(32); for (int i = 0; i < 32; ++i) v[i] = 16 * i; add("stride 16 B (2-way by address range)", v); }
    const int npre = (int)pats.size();
    const int STR = CLB_LATCH_SCHED_STRIDE, COPY = CLB_LATCH_SCHED_COPY;
    for (int warp = 0; warp < 8; ++warp)
        for (int step = 0; step < 8; ++step)
            for (int k = 0; k < 3; ++k)
                for (int rowadd = 0; rowadd < 2; ++rowadd) {
                    std::vector<int> v(32);
                    for (int lane = 0; lane < 32; ++lane) {
                        const int bit = clb_latch_order[warp * 64 + step * 8 + (lane >> 2)];
                        const int r = clb_latch_triplets[bit][2 * k], c = clb_latch_triplets[bit][2 * k + 1];
                        v[lane] = (c & 7) * COPY + r * STR + (c & ~7) + (lane & 3) * 2 * STR + rowadd * STR;
                    }
                    add("latch", v);
                }
    int* d_off; unsigned long long* d_out;
    cudaMalloc(&d_off, pats.size() * 32 * sizeof(int)); cudaMalloc(&d_out, 32 * 8);
    for (size_t p = 0; p < pats.size(); ++p) cudaMemcpy(d_off + 32 * p, pats[p].data(), 32 * sizeof(int), cudaMemcpyHostToDevice);
    cudaFuncSetAttribute(lds64_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 45056);
    long long model_sum = 0;
    for (size_t p = 0; p < pats.size(); ++p) {
        lds64_probe<<<1, 32, 45056>>>(d_off + 32 * p, 1024, d_out);
        const int m = model_halfwarp(pats[p].data());
        if ((int)p >= npre) model_sum += m;
        printf("launch %zu %s model %d\n", p, names[p], m);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("done: %s; latch patterns %zu, model wavefronts per keypoint %lld\n", cudaGetErrorString(e), pats.size() - npre, model_sum);
    return e != cudaSuccess;
}
