// common.cuh -- shared device helpers for the ADKF-IFT B200 kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>

#define ADKF_LAUNCH_CHECK()                         \
    do {                                            \
        cudaError_t e__ = cudaGetLastError();       \
        if (e__ != cudaSuccess) return (int)e__;    \
    } while (0)

// number of kernel launches issued through the C-ABI since load (host-side counter, see capi.cu)
extern "C" void adkf_count_launches(int n);

namespace adkf {

constexpr int kWarp = 32;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Block-wide sum of K per-thread partials; the cross-thread part runs in fp64.  Result is broadcast to
// every thread through `out` (shared, >= K doubles).  `scratch` is shared, >= K * 32 doubles.
// Contains __syncthreads(): must be called by all threads of the block.
template <int K>
__device__ __forceinline__ void block_sum_multi(const float (&part)[K], double* scratch, double* out) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        double v = warp_sum((double)part[k]);
        if (lane == 0) scratch[k * 32 + wid] = v;
    }
    __syncthreads();
    if (wid == 0) {
#pragma unroll
        for (int k = 0; k < K; ++k) {
            double v = (lane < nw) ? scratch[k * 32 + lane] : 0.0;
            v = warp_sum(v);
            if (lane == 0) out[k] = v;
        }
    }
    __syncthreads();
}

// gram.cu: mu[t][:] = mean of the support rows (or 0 when center == 0), nrm[row] = |z_row - mu_t|^2
void launch_center_stats(const float* Z, int d, int T, const int32_t* z_off, const int32_t* n_sup, int max_n,
                         float* mu_ws, float* nrm_ws, cudaStream_t s, int center = 1);
void launch_pairstat_diag(int stat, int T, const int32_t* z_off, const int64_t* d2_off, int max_n, const float* nrm,
                          float* D2, cudaStream_t s);

// pair statistic of two rows from their squared norms and dot product (stat: ADKF_STAT_* of adkf_b200.h)
//   0 centred squared distance (gpytorch sq_dist: clamp_min 0, exact zero diagonal)
//   1 dot product (diagonal = the row's squared norm)
//   2 Tanimoto similarity (x.y + eps) / (eps + |x|^2 + |y|^2 - x.y), eps = 1e-6 (gp_utils.py:106-117); diagonal = 1
__device__ __forceinline__ float pairstat_value(int stat, float ni, float nj, float dot, bool diag) {
    if (stat == 0) return diag ? 0.f : fmaxf(ni + nj - 2.f * dot, 0.f);
    if (stat == 1) return diag ? ni : dot;
    return diag ? 1.f : (dot + 1e-6f) / (1e-6f + ni + nj - dot);
}

__device__ __forceinline__ float softplusf(float x) { return (x > 20.f) ? x : log1pf(expf(x)); }
__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + expf(-x)); }

}  // namespace adkf
