// microbench.cu -- pipe-throughput micro-benchmarks behind the match roofline:
// MEASURED_PEAKS.json has no POPC figure, so the denominator is measured here: independent POPC
// chains, independent LOP3 chains, and both interleaved, timed with clock64 inside the kernel.
#include "common.cuh"

namespace clb {

template <int MODE>   // 0 = POPC only, 1 = LOP3 only, 2 = 1 POPC : 2 LOP3 interleaved
__global__ void __launch_bounds__(1024)
pipe_kernel(uint32_t seed, int iters, uint32_t* out, long long* cycles) {
    uint32_t a[8], b[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { a[i] = seed * (threadIdx.x + 1) + i * 0x9E3779B9u; b[i] = a[i] ^ 0x5bd1e995u; }
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0 || MODE == 2) {
                // popc feeding an xor keeps the chain dependent per i but independent across i
                uint32_t p; asm volatile("popc.b32 %0, %1;" : "=r"(p) : "r"(a[i]));
                a[i] = p + a[i];   // IADD (ALU/FMA pipe), far below its limit
            }
            if (MODE == 1 || MODE == 2) {
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(b[i]) : "r"(a[(i + 1) & 7]), "r"(seed));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0xE8;" : "+r"(b[i]) : "r"(a[(i + 3) & 7]), "r"(seed));
            }
        }
    }
    const long long t1 = clock64();
    uint32_t acc = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc ^= a[i] ^ b[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

int microbench_pipes(double* out) {
    int dev = 0, sms = 0, khz = 0;
    CLB_CUDA(cudaGetDevice(&dev));
    CLB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    CLB_CUDA(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev));
    const int threads = 1024, blocks = sms * 2, iters = 4096;
    uint32_t* d_out; long long* d_cyc;
    CLB_CUDA(cudaMalloc(&d_out, (size_t)threads * blocks * 4));
    CLB_CUDA(cudaMalloc(&d_cyc, (size_t)blocks * 8));
    long long* h_cyc = new long long[blocks];
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    double res[3][2] = {{0}};   // [mode][cycles-based ops/clk/SM for popc, lop3]
    float ms_mode[3] = {0};
    for (int mode = 0; mode < 3; ++mode) {
        for (int rep = 0; rep < 3; ++rep) {   // last rep counts
            cudaEventRecord(e0);
            if (mode == 0) pipe_kernel<0><<<blocks, threads>>>(12345u, iters, d_out, d_cyc);
            else if (mode == 1) pipe_kernel<1><<<blocks, threads>>>(12345u, iters, d_out, d_cyc);
            else pipe_kernel<2><<<blocks, threads>>>(12345u, iters, d_out, d_cyc);
            cudaEventRecord(e1);
            g_launches.fetch_add(1);
            CLB_CUDA(cudaEventSynchronize(e1));
            cudaEventElapsedTime(&ms_mode[mode], e0, e1);
        }
        CLB_CUDA(cudaMemcpy(h_cyc, d_cyc, (size_t)blocks * 8, cudaMemcpyDeviceToHost));
        double mean = 0; for (int i = 0; i < blocks; ++i) mean += (double)h_cyc[i]; mean /= blocks;
        // 2 CTAs of 1024 threads per SM run concurrently; per CTA: iters*8 popc and iters*16 lop3 per thread
        const double popc_ops = (mode != 1) ? 2.0 * threads * iters * 8.0 : 0.0;
        const double lop_ops = (mode != 0) ? 2.0 * threads * iters * 16.0 : 0.0;
        res[mode][0] = popc_ops / mean; res[mode][1] = lop_ops / mean;
    }
    out[0] = res[0][0]; out[1] = res[1][1]; out[2] = res[2][0]; out[3] = res[2][1];
    // effective SM clock from wall time of the POPC run: cycles / time
    {
        double mean = 0; // reuse last (mode 2) cycles for the clock estimate
        for (int i = 0; i < blocks; ++i) mean += (double)h_cyc[i]; mean /= blocks;
        out[4] = mean / (ms_mode[2] * 1e-3) / 1e6;
    }
    out[5] = sms;
    // chip-level rates straight from wall time (no clock assumption): ops / event time
    out[6] = 2.0 * threads * iters * 8.0 * sms / (ms_mode[0] * 1e-3) / 1e9;    // G POPC32/s, POPC-only stream
    out[7] = 2.0 * threads * iters * 16.0 * sms / (ms_mode[1] * 1e-3) / 1e9;   // G LOP3/s, LOP3-only stream
    delete[] h_cyc; cudaFree(d_out); cudaFree(d_cyc); cudaEventDestroy(e0); cudaEventDestroy(e1);
    (void)khz;
    return CLB_OK;
}

}  // namespace clb
