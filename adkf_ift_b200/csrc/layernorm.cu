// layernorm.cu -- nn.LayerNorm(128) of the GNN block (`boom_norm_layer`, fs_mol/modules/gnn.py:462, applied at :504) forward
// and backward.  Rows are 128 floats = 512 B: one warp per row, one float4 per lane, the two row statistics by warp
// shuffles, so every byte of x / y / dy / dx moves exactly once with 512-byte coalesced accesses.  HBM-bound:
//   forward   4*128*V read + 4*128*V write (+ 8*V statistics)           68.5 + 68.5 MB at V = 133,728
//   backward  2 * 4*128*V read + 4*128*V write                           205 MB
// (torch's generic row-per-block kernel takes 0.20 ms / 0.13 ms for these at V = 133,728, 6-10x the HBM time.)
// dgamma / dbeta are deterministic: per-block partials in registers -> shared memory -> workspace, fixed-order finish.
#include "common.cuh"
#include "../../include/adkf_b200.h"

namespace adkf {

constexpr int kLnThreads = 256;
constexpr int kLnWarps = kLnThreads / 32;

__device__ __forceinline__ float ln_warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__global__ void __launch_bounds__(kLnThreads)
layernorm128_fwd_kernel(const float* __restrict__ x, int64_t M, const float* __restrict__ gamma, const float* __restrict__ beta,
                        float eps, float* __restrict__ y, float* __restrict__ mean, float* __restrict__ rstd) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * kLnWarps + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * kLnWarps;
    const float4 g = reinterpret_cast<const float4*>(gamma)[lane];
    const float4 b = reinterpret_cast<const float4*>(beta)[lane];
    for (int64_t r = warp; r < M; r += nwarps) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(x + r * 128) + lane);
        const float mu = ln_warp_sum((v.x + v.y) + (v.z + v.w)) * (1.f / 128.f);
        const float d0 = v.x - mu, d1 = v.y - mu, d2 = v.z - mu, d3 = v.w - mu;
        const float var = ln_warp_sum((d0 * d0 + d1 * d1) + (d2 * d2 + d3 * d3)) * (1.f / 128.f);
        const float rs = rsqrtf(var + eps);
        float4 o;
        o.x = d0 * rs * g.x + b.x;
        o.y = d1 * rs * g.y + b.y;
        o.z = d2 * rs * g.z + b.z;
        o.w = d3 * rs * g.w + b.w;
        reinterpret_cast<float4*>(y + r * 128)[lane] = o;
        if (lane == 0) {
            mean[r] = mu;
            rstd[r] = rs;
        }
    }
}

// block b owns rows [b * rows_per_block, ...): dx per row, dgamma / dbeta partials per block -> partial[b][0:128 | 128:256]
__global__ void __launch_bounds__(kLnThreads)
layernorm128_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ mean,
                        const float* __restrict__ rstd, const float* __restrict__ gamma, int64_t M, int rows_per_block,
                        float* __restrict__ dx, float* __restrict__ partial) {
    __shared__ float4 red[kLnWarps][64];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
    const int64_t r1 = min(M, r0 + (int64_t)rows_per_block);
    const float4 g = reinterpret_cast<const float4*>(gamma)[lane];
    float4 ag = make_float4(0.f, 0.f, 0.f, 0.f), ab = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int64_t r = r0 + w; r < r1; r += kLnWarps) {
        const float4 d = __ldg(reinterpret_cast<const float4*>(dy + r * 128) + lane);
        const float4 v = __ldg(reinterpret_cast<const float4*>(x + r * 128) + lane);
        const float mu = mean[r], rs = rstd[r];
        const float h0 = (v.x - mu) * rs, h1 = (v.y - mu) * rs, h2 = (v.z - mu) * rs, h3 = (v.w - mu) * rs;
        const float g0 = d.x * g.x, g1 = d.y * g.y, g2 = d.z * g.z, g3 = d.w * g.w;
        const float m1 = ln_warp_sum((g0 + g1) + (g2 + g3)) * (1.f / 128.f);
        const float m2 = ln_warp_sum((g0 * h0 + g1 * h1) + (g2 * h2 + g3 * h3)) * (1.f / 128.f);
        float4 o;
        o.x = rs * (g0 - m1 - h0 * m2);
        o.y = rs * (g1 - m1 - h1 * m2);
        o.z = rs * (g2 - m1 - h2 * m2);
        o.w = rs * (g3 - m1 - h3 * m2);
        reinterpret_cast<float4*>(dx + r * 128)[lane] = o;
        ag.x += d.x * h0; ag.y += d.y * h1; ag.z += d.z * h2; ag.w += d.w * h3;
        ab.x += d.x; ab.y += d.y; ab.z += d.z; ab.w += d.w;
    }
    red[w][lane] = ag;
    red[w][32 + lane] = ab;
    __syncthreads();
    if (threadIdx.x < 64) {
        float4 s = red[0][threadIdx.x];
#pragma unroll
        for (int k = 1; k < kLnWarps; ++k) {
            const float4 t = red[k][threadIdx.x];
            s.x += t.x; s.y += t.y; s.z += t.z; s.w += t.w;
        }
        reinterpret_cast<float4*>(partial + (size_t)blockIdx.x * 256)[threadIdx.x] = s;
    }
}

// out[c] = sum_b partial[b][c], c < 256, fixed order: thread (ty, tx) sums b = ty, ty + 8, ... of column tx of its 32-column block
__global__ void __launch_bounds__(256)
layernorm128_finish_kernel(const float* __restrict__ partial, int nblk, float* __restrict__ dgamma, float* __restrict__ dbeta) {
    __shared__ float red[8][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + tx;
    float s = 0.f;
    for (int b = ty; b < nblk; b += 8) s += partial[(size_t)b * 256 + c];
    red[ty][tx] = s;
    __syncthreads();
    if (ty == 0) {
        float t = red[0][tx];
#pragma unroll
        for (int k = 1; k < 8; ++k) t += red[k][tx];
        if (c < 128) dgamma[c] = t;
        else dbeta[c - 128] = t;
    }
}

}  // namespace adkf

using namespace adkf;

static int ln_bwd_blocks(int64_t M) {
    const int64_t want = (M + 63) / 64;
    return (int)(want < 1 ? 1 : (want > 148 * 8 ? 148 * 8 : want));
}

extern "C" int64_t adkf_layernorm128_bwd_ws_floats(int64_t M) { return (int64_t)ln_bwd_blocks(M) * 256; }

extern "C" int adkf_layernorm128_fwd(const float* x, int64_t M, const float* gamma, const float* beta, float eps, float* y,
                                     float* mean, float* rstd, void* stream) {
    if (M <= 0) return 0;
    if (!x || !gamma || !beta || !y || !mean || !rstd) return ADKF_ERR_BADARG;
    if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(gamma) |
         reinterpret_cast<uintptr_t>(beta)) & 15)
        return ADKF_ERR_BADARG;
    (void)cudaGetLastError();
    const int64_t want = (M + kLnWarps - 1) / kLnWarps;
    const int grid = (int)(want > 148 * 8 ? 148 * 8 : want);
    layernorm128_fwd_kernel<<<grid, kLnThreads, 0, (cudaStream_t)stream>>>(x, M, gamma, beta, eps, y, mean, rstd);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_layernorm128_bwd(const float* dy, const float* x, const float* mean, const float* rstd, const float* gamma,
                                     int64_t M, float* dx, float* dgamma, float* dbeta, float* ws, int64_t ws_floats,
                                     void* stream) {
    if (!dgamma || !dbeta) return ADKF_ERR_BADARG;
    cudaStream_t s = (cudaStream_t)stream;
    if (M <= 0) {
        cudaMemsetAsync(dgamma, 0, 128 * sizeof(float), s);
        cudaMemsetAsync(dbeta, 0, 128 * sizeof(float), s);
        return 0;
    }
    if (!dy || !x || !mean || !rstd || !gamma || !dx || !ws) return ADKF_ERR_BADARG;
    if ((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(dx) |
         reinterpret_cast<uintptr_t>(gamma) | reinterpret_cast<uintptr_t>(ws)) & 15)
        return ADKF_ERR_BADARG;
    const int nblk = ln_bwd_blocks(M);
    if (ws_floats < (int64_t)nblk * 256) return ADKF_ERR_BADARG;
    (void)cudaGetLastError();
    const int rows_per_block = (int)((M + nblk - 1) / nblk);
    layernorm128_bwd_kernel<<<nblk, kLnThreads, 0, s>>>(dy, x, mean, rstd, gamma, M, rows_per_block, dx, ws);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    layernorm128_finish_kernel<<<8, 256, 0, s>>>(ws, nblk, dgamma, dbeta);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}
