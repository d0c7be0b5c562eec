// gp_ard.cu -- the two contractions over the feature dimension that --use-ard adds to the batched GP tail
// (ExactGPLayer(..., ard_num_dims=d): fs_mol/utils/gp_utils.py:14,27,30; fs_mol/models/adaptive_dkt.py:107-110,124).
//
// With one lengthscale per feature, l_k, the kernel acts on the WEIGHTED squared distance u_ab = sum_k w_k (z_ak - z_bk)^2,
// w_k = 1 / l_k^2.  Everything n x n (Cholesky, MLL, its derivatives w.r.t. u, the joint predictive NLL) is done by the existing
// one-CTA-per-task kernels on u with the lengthscale fixed to 1; what is new is
//   adkf_ard_wsqdist :  u_ab = sum_k c_k (z_ak - z_bk)^2 for per-task weights c [T,d] of either sign (w itself, or its
//                       directional derivative for a Hessian-vector product)
//   adkf_ard_pairquad:  q_k = sum_ab G_ab (z_ak - z_bk)^2 for a matrix G over the pairs (df/du, dF/du or their directional
//                       derivatives): the gradient / Hessian-vector product w.r.t. the d lengthscales is q_k times a per-k factor.
// A Hessian-vector product of the 2050-dimensional inner problem is wsqdist -> adkf_gp_ift mode 4 -> pairquad: the (d+2)^2
// Hessian of the reference (cauchy_hypergradient.py:41-64, dense torch.linalg.solve :136) is never formed.
#include "common.cuh"
#include "../../include/adkf_b200.h"

namespace adkf {

constexpr int kAT = 32;   // pair tile edge

// grid (ceil(R/32), ceil(R/32), T), 256 threads: thread (ty, tx) owns pairs (a0 + ty + 8 i, b0 + tx), i < 4
__global__ void __launch_bounds__(256)
ard_wsqdist_kernel(const float* __restrict__ Z, int d, const int32_t* __restrict__ z_off, const int32_t* __restrict__ n_sup,
                   const int64_t* __restrict__ d2_off, const float* __restrict__ c, int support_only, float* __restrict__ D2) {
    __shared__ float za[kAT][kAT + 1], zb[kAT][kAT + 1], cs[kAT];
    const int t = blockIdx.z;
    const int r0 = z_off[t], N = z_off[t + 1] - r0;
    const int R = support_only ? n_sup[t] : N;
    const int a0 = blockIdx.y * kAT, b0 = blockIdx.x * kAT;
    if (a0 >= R || b0 >= R || b0 > a0) return;                 // symmetric: lower tiles, mirrored on store
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    const float* ct = c + (size_t)t * d;
    for (int k0 = 0; k0 < d; k0 += kAT) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int r = ty + 8 * i, k = k0 + tx;
            za[r][tx] = (a0 + r < R && k < d) ? Z[(size_t)(r0 + a0 + r) * d + k] : 0.f;
            zb[r][tx] = (b0 + r < R && k < d) ? Z[(size_t)(r0 + b0 + r) * d + k] : 0.f;
        }
        if (threadIdx.x < kAT) cs[threadIdx.x] = (k0 + threadIdx.x < d) ? ct[k0 + threadIdx.x] : 0.f;
        __syncthreads();
#pragma unroll 8
        for (int k = 0; k < kAT; ++k) {
            const float zbk = zb[tx][k], ck = cs[k];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float df = za[ty + 8 * i][k] - zbk;
                acc[i] = fmaf(ck * df, df, acc[i]);
            }
        }
        __syncthreads();
    }
    float* Dt = D2 + d2_off[t];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int a = a0 + ty + 8 * i, b = b0 + tx;
        if (a < R && b < R) {
            const float v = (a == b) ? 0.f : acc[i];
            Dt[(size_t)a * N + b] = v;
            Dt[(size_t)b * N + a] = v;
        }
    }
}

// grid (ceil(d/64), T), 256 threads = 64 columns x 4 row groups; G is read through shared memory in 32 x 32 tiles.
//   q_k = sum_a r_a z_ak^2 + sum_b c_b z_bk^2 - 2 sum_ab G_ab z_ak z_bk      (r / c = row / column sums of G)
__global__ void __launch_bounds__(256)
ard_pairquad_kernel(const float* __restrict__ Z, int d, const int32_t* __restrict__ z_off, const int32_t* __restrict__ n_sup,
                    const int64_t* __restrict__ d2_off, const float* __restrict__ G, int support_only, float* __restrict__ q) {
    __shared__ float gs[kAT][kAT + 1];
    __shared__ float zbs[kAT][64];
    __shared__ float part[4][64];
    const int t = blockIdx.y;
    const int r0 = z_off[t], N = z_off[t + 1] - r0;
    const int R = support_only ? n_sup[t] : N;
    const int col = threadIdx.x & 63, grp = threadIdx.x >> 6;          // 4 groups of rows a
    const int k = blockIdx.x * 64 + col;
    const float* Gt = G + d2_off[t];
    float acc = 0.f;
    for (int a0 = 0; a0 < R; a0 += kAT) {
        for (int b0 = 0; b0 < R; b0 += kAT) {
            for (int idx = threadIdx.x; idx < kAT * kAT; idx += 256) {
                const int i = idx >> 5, j = idx & 31;
                gs[i][j] = (a0 + i < R && b0 + j < R) ? Gt[(size_t)(a0 + i) * N + b0 + j] : 0.f;
            }
            for (int idx = threadIdx.x; idx < kAT * 64; idx += 256) {
                const int j = idx >> 6, cc = idx & 63;
                zbs[j][cc] = (b0 + j < R && blockIdx.x * 64 + cc < d) ? Z[(size_t)(r0 + b0 + j) * d + blockIdx.x * 64 + cc] : 0.f;
            }
            __syncthreads();
            if (k < d) {
                for (int i = grp; i < kAT; i += 4) {
                    if (a0 + i >= R) break;
                    const float zak = Z[(size_t)(r0 + a0 + i) * d + k];
                    float s1 = 0.f;
#pragma unroll 8
                    for (int j = 0; j < kAT; ++j) {
                        const float df = zak - zbs[j][col];
                        s1 = fmaf(gs[i][j] * df, df, s1);
                    }
                    acc += s1;
                }
            }
            __syncthreads();
        }
    }
    part[grp][col] = acc;
    __syncthreads();
    if (grp == 0 && k < d) q[(size_t)t * d + k] = part[0][col] + part[1][col] + part[2][col] + part[3][col];
}

}  // namespace adkf

using namespace adkf;

extern "C" int adkf_ard_wsqdist(const float* Z, int32_t d, int32_t T, const int32_t* z_off, const int32_t* n_sup,
                                const int64_t* d2_off, int32_t max_n, const float* c, int32_t support_only, float* D2, void* stream) {
    if (T <= 0 || max_n <= 0) return 0;
    if (!Z || !z_off || !n_sup || !d2_off || !c || !D2 || d <= 0) return ADKF_ERR_BADARG;
    const int nt = (max_n + kAT - 1) / kAT;
    ard_wsqdist_kernel<<<dim3(nt, nt, T), 256, 0, (cudaStream_t)stream>>>(Z, d, z_off, n_sup, d2_off, c, support_only, D2);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_ard_pairquad(const float* Z, int32_t d, int32_t T, const int32_t* z_off, const int32_t* n_sup,
                                 const int64_t* d2_off, const float* G, int32_t support_only, float* q, void* stream) {
    if (T <= 0) return 0;
    if (!Z || !z_off || !n_sup || !d2_off || !G || !q || d <= 0) return ADKF_ERR_BADARG;
    ard_pairquad_kernel<<<dim3((d + 63) / 64, T), 256, 0, (cudaStream_t)stream>>>(Z, d, z_off, n_sup, d2_off, G, support_only, q);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}
