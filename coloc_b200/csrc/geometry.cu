// geometry.cu -- residuals of two-view model hypotheses over a match list: the Errors() pass of the AC-RANSAC kernel
// adaptors (robust_estimator_ACRansacKernelAdaptator.hpp:93-99, :244-252), which RobustMatcher::filterFundamental / filterEssential /
// filterHomography (include/coloc/RobustMatcher.hpp:128-239) run once per hypothesis on the host, here for a batch of hypotheses.
//   kind 0  fundamental::kernel::EpipolarDistanceError           solver_fundamental_kernel.cpp:183-192
//   kind 1  fundamental::kernel::SymmetricEpipolarDistanceError  solver_fundamental_kernel.cpp:168-180
//   kind 2  fundamental::kernel::SampsonError                    solver_fundamental_kernel.cpp:156-166
//   kind 3  homography::kernel::AsymmetricError                  solver_homography_kernel.hpp:59-63
// fp64 throughout; the hypothesis solvers and the a-contrario model selection stay on the host (north_star).
#include "matcher_priv.cuh"

using namespace clb;

namespace {

template <int KIND>
__device__ __forceinline__ double residual(const double (&M)[9], double x0, double x1, double y0, double y1) {
    const double a0 = M[0] * x0 + M[1] * x1 + M[2];
    const double a1 = M[3] * x0 + M[4] * x1 + M[5];
    const double a2 = M[6] * x0 + M[7] * x1 + M[8];
    if (KIND == 3) {
        const double dx = y0 - a0 / a2, dy = y1 - a1 / a2;
        return dx * dx + dy * dy;
    }
    const double d = a0 * y0 + a1 * y1 + a2;
    const double na = a0 * a0 + a1 * a1;
    if (KIND == 0) return d * d / na;
    const double b0 = M[0] * y0 + M[3] * y1 + M[6];
    const double b1 = M[1] * y0 + M[4] * y1 + M[7];
    const double nb = b0 * b0 + b1 * b1;
    if (KIND == 1) return d * d * (1.0 / na + 1.0 / nb) / 4.0;
    return d * d / (na + nb);
}

// grid (ceil(N / 256), H): one thread per (hypothesis, match); the inlier count of a hypothesis is a ballot + one atomic per warp
template <int KIND>
__global__ void __launch_bounds__(256)
residuals_kernel(const double* __restrict__ models, const double2* __restrict__ x1, const double2* __restrict__ x2, int n,
                 double thr, double* __restrict__ res_out, int* __restrict__ inliers) {
    __shared__ double sM[9];
    const int h = blockIdx.y;
    if (threadIdx.x < 9) sM[threadIdx.x] = models[(size_t)h * 9 + threadIdx.x];
    __syncthreads();
    double M[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) M[k] = sM[k];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    bool in = false;
    if (i < n) {
        const double2 a = x1[i], b = x2[i];
        const double r = residual<KIND>(M, a.x, a.y, b.x, b.y);
        if (res_out) res_out[(size_t)h * n + i] = r;
        in = r < thr;
    }
    if (inliers) {
        const unsigned m = __ballot_sync(0xFFFFFFFFu, in);
        if ((threadIdx.x & 31) == 0 && m) atomicAdd(&inliers[h], __popc(m));
    }
}

int launch_residuals(int kind, const double* d_models, int H, const double* d_x1, const double* d_x2, int n, double thr,
                     double* d_res, int* d_inl, cudaStream_t s) {
    if (H <= 0 || n <= 0) return CLB_OK;
    if (H > 65535) { snprintf(g_err, sizeof(g_err), "more than 65535 hypotheses in one call"); return CLB_EINVAL; }
    if (d_inl) CLB_CUDA(cudaMemsetAsync(d_inl, 0, (size_t)H * sizeof(int), s));
    dim3 grid((n + 255) / 256, H);
    const double2* a = reinterpret_cast<const double2*>(d_x1);
    const double2* b = reinterpret_cast<const double2*>(d_x2);
    timer("residuals").begin(s);
    switch (kind) {
        case 0: residuals_kernel<0><<<grid, 256, 0, s>>>(d_models, a, b, n, thr, d_res, d_inl); break;
        case 1: residuals_kernel<1><<<grid, 256, 0, s>>>(d_models, a, b, n, thr, d_res, d_inl); break;
        case 2: residuals_kernel<2><<<grid, 256, 0, s>>>(d_models, a, b, n, thr, d_res, d_inl); break;
        case 3: residuals_kernel<3><<<grid, 256, 0, s>>>(d_models, a, b, n, thr, d_res, d_inl); break;
        default: snprintf(g_err, sizeof(g_err), "unknown residual kind %d", kind); return CLB_EINVAL;
    }
    timer("residuals").end(s);
    CLB_LAUNCHED();
    return CLB_OK;
}

}  // namespace

extern "C" {

int clb_model_residuals_dev(clb_matcher* m, int kind, const double* d_models, int n_models, const double* d_x1, const double* d_x2,
                            int n, double threshold, double* d_residuals_out, int32_t* d_inliers_out, void* stream) {
    if (!m || n_models < 0 || n < 0 || kind < 0 || kind > 3) { snprintf(g_err, sizeof(g_err), "bad residual request"); return CLB_EINVAL; }
    if ((n_models > 0 && !d_models) || (n > 0 && (!d_x1 || !d_x2))) return CLB_EINVAL;
    CLB_CUDA(cudaSetDevice(m->device));
    return launch_residuals(kind, d_models, n_models, d_x1, d_x2, n, threshold, d_residuals_out, d_inliers_out,
                            stream ? (cudaStream_t)stream : m->stream);
}

int clb_model_residuals(clb_matcher* m, int kind, const double* models, int n_models, const double* x1, const double* x2, int n,
                        double threshold, double* residuals_out, int32_t* inliers_out) {
    if (!m || n_models < 0 || n < 0 || kind < 0 || kind > 3) { snprintf(g_err, sizeof(g_err), "bad residual request"); return CLB_EINVAL; }
    if ((n_models > 0 && !models) || (n > 0 && (!x1 || !x2))) return CLB_EINVAL;
    if (n_models == 0 || n == 0) { if (inliers_out) for (int h = 0; h < n_models; ++h) inliers_out[h] = 0; return CLB_OK; }
    CLB_CUDA(cudaSetDevice(m->device));
    cudaStream_t s = m->stream;
    const size_t mb = (size_t)n_models * 72, xb = (size_t)n * 16, rb = residuals_out ? (size_t)n_models * n * 8 : 0;
    const size_t off_x1 = (mb + 255) & ~(size_t)255, off_x2 = off_x1 + ((xb + 255) & ~(size_t)255);
    const size_t off_in = off_x2 + ((xb + 255) & ~(size_t)255), off_r = off_in + (((size_t)n_models * 4 + 255) & ~(size_t)255);
    CLB_CUDA(m->geo.reserve(off_r + rb));
    char* base = static_cast<char*>(m->geo.p);
    CLB_CUDA(cudaMemcpyAsync(base, models, mb, cudaMemcpyHostToDevice, s));
    CLB_CUDA(cudaMemcpyAsync(base + off_x1, x1, xb, cudaMemcpyHostToDevice, s));
    CLB_CUDA(cudaMemcpyAsync(base + off_x2, x2, xb, cudaMemcpyHostToDevice, s));
    int rc = launch_residuals(kind, reinterpret_cast<double*>(base), n_models, reinterpret_cast<double*>(base + off_x1),
                              reinterpret_cast<double*>(base + off_x2), n, threshold,
                              residuals_out ? reinterpret_cast<double*>(base + off_r) : nullptr,
                              inliers_out ? reinterpret_cast<int*>(base + off_in) : nullptr, s);
    if (rc) return rc;
    if (residuals_out) CLB_CUDA(cudaMemcpyAsync(residuals_out, base + off_r, rb, cudaMemcpyDeviceToHost, s));
    if (inliers_out) CLB_CUDA(cudaMemcpyAsync(inliers_out, base + off_in, (size_t)n_models * 4, cudaMemcpyDeviceToHost, s));
    CLB_CUDA(cudaStreamSynchronize(s));
    return CLB_OK;
}

}  // extern "C"
