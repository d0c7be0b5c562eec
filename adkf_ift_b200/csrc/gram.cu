// gram.cu -- per-task squared-distance (Gram) matrices of the GP inputs and their backward (sm_100a).
//
// Forward replaces gpytorch's sq_dist inside MaternKernel / RBFKernel (call sites
// fs_mol/utils/gp_utils.py:28,31): inputs centred by the mean of the SUPPORT rows, quadratic expansion
// |a|^2 - 2 a.b + |b|^2, clamp_min(0), exact zero diagonal.  One launch covers all T tasks; the joint
// (support+query) matrix is produced so that the fit (ss block), the prediction (qs block) and the
// predictive joint NLL of the IFT stage (whole matrix) read the same buffer.
// Backward: anything that depends on Z only through D2 has dZ_a = 4 sum_b Wm_ab (z_a - z_b) for the
// symmetric weight matrix Wm = dLoss/dD2 -- one batched [N x N] x [N x d] product.
//
// Both are FP32 CUDA-core tiled GEMMs (64x64 tile, 4x4 per thread): the contraction length is d = 2048 and
// fp32-accurate distances are required for the 1e-4 parity target (SURVEY.md section 7 hard part (b)).
#include "common.cuh"
#include "../../include/adkf_b200.h"

namespace adkf {

// rows of stat != SQDIST that the tile kernels skip (query x query block in meta-test mode) still need their diagonal:
// the predictive variance reads k(x*, x*) from it.  dot: |z|^2, Tanimoto: 1.
__global__ void pairstat_diag_kernel(int stat, const int32_t* __restrict__ z_off, const int64_t* __restrict__ d2_off,
                                     const float* __restrict__ nrm, float* __restrict__ D2) {
    const int t = blockIdx.y;
    const int r0 = z_off[t], N = z_off[t + 1] - r0;
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < N; r += gridDim.x * blockDim.x)
        D2[d2_off[t] + (size_t)r * N + r] = (stat == ADKF_STAT_DOT) ? nrm[r0 + r] : 1.f;
}

constexpr int kTile = 64;
constexpr int kBK = 16;
constexpr int kPad = 4;

// mu[t][c] = mean over the support rows of task t
__global__ void task_mean_kernel(const float* __restrict__ Z, int d, const int32_t* __restrict__ z_off,
                                 const int32_t* __restrict__ n_sup, float* __restrict__ mu) {
    const int t = blockIdx.y;
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= d) return;
    const int r0 = z_off[t], n = n_sup[t];
    float acc = 0.f;
    for (int r = 0; r < n; ++r) acc += Z[(size_t)(r0 + r) * d + c];
    mu[(size_t)t * d + c] = acc / (float)max(n, 1);
}

// nrm[row] = |z_row - mu_t|^2, one warp per row
__global__ void row_norm_kernel(const float* __restrict__ Z, int d, const int32_t* __restrict__ z_off,
                                const float* __restrict__ mu, float* __restrict__ nrm) {
    const int t = blockIdx.y;
    const int r0 = z_off[t], N = z_off[t + 1] - r0;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const float* m = mu + (size_t)t * d;
    for (int r = blockIdx.x * nw + wid; r < N; r += gridDim.x * nw) {
        const float* z = Z + (size_t)(r0 + r) * d;
        float acc = 0.f;
        for (int c = lane; c < d; c += 32) { const float v = z[c] - m[c]; acc += v * v; }
        acc = warp_sum(acc);
        if (lane == 0) nrm[r0 + r] = acc;
    }
}

// one 64x64 tile of D2 for task blockIdx.z; tiles above the diagonal are mirrored from below
__global__ void __launch_bounds__(256)
sqdist_tile_kernel(const float* __restrict__ Z, int d, const int32_t* __restrict__ z_off,
                   const int32_t* __restrict__ n_sup, const int64_t* __restrict__ d2_off, int qq_block,
                   const float* __restrict__ mu, const float* __restrict__ nrm, float* __restrict__ D2, int stat) {
    const int t = blockIdx.z;
    const int r0 = z_off[t], N = z_off[t + 1] - r0, n = n_sup[t];
    const int ti = blockIdx.y, tj = blockIdx.x;
    if (tj > ti) return;
    const int row0 = ti * kTile, col0 = tj * kTile;
    if (row0 >= N || col0 >= N) return;
    if (!qq_block && col0 >= n) return;  // (query x query) tile: not needed for meta-test
    __shared__ float As[kBK][kTile + kPad];
    __shared__ float Bs[kBK][kTile + kPad];
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const int lr = tid >> 2, lk = (tid & 3) * 4;  // loader: row within tile, k offset
    const float* m = mu + (size_t)t * d;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    const bool arow_ok = row0 + lr < N, brow_ok = col0 + lr < N;
    const float* za = Z + (size_t)(r0 + row0 + lr) * d;
    const float* zb = Z + (size_t)(r0 + col0 + lr) * d;
    for (int k0 = 0; k0 < d; k0 += kBK) {
        float a4[4], b4[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int k = k0 + lk + q;
            const float mk = (k < d) ? m[k] : 0.f;
            a4[q] = (arow_ok && k < d) ? za[k] - mk : 0.f;
            b4[q] = (brow_ok && k < d) ? zb[k] - mk : 0.f;
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) { As[lk + q][lr] = a4[q]; Bs[lk + q][lr] = b4[q]; }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kBK; ++k) {
            const float4 av = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
            const float4 bv = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
            const float a[4] = {av.x, av.y, av.z, av.w}, b[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
    }
    float* Dt = D2 + d2_off[t];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int r = row0 + ty * 4 + i;
        if (r >= N) continue;
        const float ni = nrm[r0 + r];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int c = col0 + tx * 4 + j;
            if (c >= N) continue;
            const float v = pairstat_value(stat, ni, nrm[r0 + c], acc[i][j], r == c);
            Dt[(size_t)r * N + c] = v;
            if (ti != tj) Dt[(size_t)c * N + r] = v;
        }
    }
}

// dZ tile: rows a in [row0, row0+64), feature columns c in [c0, c0+64)
__global__ void __launch_bounds__(256)
sqdist_bwd_kernel(const float* __restrict__ Z, int d, const int32_t* __restrict__ z_off,
                  const int64_t* __restrict__ d2_off, const float* __restrict__ Wm, const float* __restrict__ mu,
                  float scale, float* __restrict__ dZ, int stat) {
    const int t = blockIdx.z;
    const int r0 = z_off[t], N = z_off[t + 1] - r0;
    const int row0 = blockIdx.y * kTile, c0 = blockIdx.x * kTile;
    if (row0 >= N) return;
    __shared__ float As[kBK][kTile + kPad];   // W[a][b] stored [b][a]
    __shared__ float Bs[kBK][kTile + kPad];   // Zc[b][c]
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const float* Wt = Wm + d2_off[t];
    const float* m = mu + (size_t)t * d;
    float acc[4][4], rs[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    const int la = tid >> 2, lb = (tid & 3) * 4;   // W loader: row a, 4 consecutive b
    const int zb = tid >> 4, zc = (tid & 15) * 4;  // Z loader: row b, 4 consecutive c
    for (int b0 = 0; b0 < N; b0 += kBK) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int a = row0 + la, b = b0 + lb + q;
            As[lb + q][la] = (a < N && b < N) ? Wt[(size_t)a * N + b] : 0.f;
        }
        {
            const int b = b0 + zb;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int c = c0 + zc + q;
                Bs[zb][zc + q] = (b < N && c < d) ? Z[(size_t)(r0 + b) * d + c] - m[c] : 0.f;
            }
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kBK; ++k) {
            const float4 av = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
            const float4 bv = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
            const float a[4] = {av.x, av.y, av.z, av.w}, b[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                rs[i] += a[i];
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int a = row0 + ty * 4 + i;
        if (a >= N) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int c = c0 + tx * 4 + j;
            if (c >= d) continue;
            const float zac = Z[(size_t)(r0 + a) * d + c] - m[c];
            // sqdist: du_ab = 2 (z_a - z_b).(dz_a - dz_b);  dot: du_ab = z_b.dz_a + z_a.dz_b   (Wm symmetric)
            dZ[(size_t)(r0 + a) * d + c] = (stat == ADKF_STAT_SQDIST) ? scale * 4.f * (rs[i] * zac - acc[i][j])
                                                                      : scale * 2.f * acc[i][j];
        }
    }
}

}  // namespace adkf

using namespace adkf;

// centring statistics shared with the tensor-core Gram (gemm_tf32x3.cu); center = 0 -> mu = 0, nrm = |z|^2
void adkf::launch_center_stats(const float* Z, int d, int T, const int32_t* z_off, const int32_t* n_sup, int max_n,
                               float* mu_ws, float* nrm_ws, cudaStream_t s, int center) {
    if (center) task_mean_kernel<<<dim3((d + 255) / 256, T), 256, 0, s>>>(Z, d, z_off, n_sup, mu_ws);
    else cudaMemsetAsync(mu_ws, 0, sizeof(float) * (size_t)T * d, s);
    row_norm_kernel<<<dim3((max_n + 7) / 8, T), 256, 0, s>>>(Z, d, z_off, mu_ws, nrm_ws);
}

void adkf::launch_pairstat_diag(int stat, int T, const int32_t* z_off, const int64_t* d2_off, int max_n, const float* nrm,
                                float* D2, cudaStream_t s) {
    pairstat_diag_kernel<<<dim3((max_n + 255) / 256, T), 256, 0, s>>>(stat, z_off, d2_off, nrm, D2);
}

extern "C" int adkf_gram_pairstat(int32_t stat, const float* Z, int32_t d, int32_t T, const int32_t* z_off,
                                  const int32_t* n_sup, const int64_t* d2_off, int32_t max_n, int32_t qq_block,
                                  float* mu_ws, float* nrm_ws, float* D2, void* stream) {
    if (T <= 0 || max_n <= 0) return 0;
    if (!Z || !z_off || !n_sup || !d2_off || !mu_ws || !nrm_ws || !D2 || d <= 0) return ADKF_ERR_BADARG;
    if (stat < ADKF_STAT_SQDIST || stat > ADKF_STAT_TANIMOTO) return ADKF_ERR_BADARG;
    cudaStream_t s = (cudaStream_t)stream;
    launch_center_stats(Z, d, T, z_off, n_sup, max_n, mu_ws, nrm_ws, s, stat == ADKF_STAT_SQDIST);
    const int nt = (max_n + kTile - 1) / kTile;
    sqdist_tile_kernel<<<dim3(nt, nt, T), 256, 0, s>>>(Z, d, z_off, n_sup, d2_off, qq_block, mu_ws, nrm_ws, D2, stat);
    if (stat != ADKF_STAT_SQDIST && !qq_block) launch_pairstat_diag(stat, T, z_off, d2_off, max_n, nrm_ws, D2, s);
    adkf_count_launches(3);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_gram_sqdist(const float* Z, int32_t d, int32_t T, const int32_t* z_off,
                                const int32_t* n_sup, const int64_t* d2_off, int32_t max_n, int32_t qq_block,
                                float* mu_ws, float* nrm_ws, float* D2, void* stream) {
    return adkf_gram_pairstat(ADKF_STAT_SQDIST, Z, d, T, z_off, n_sup, d2_off, max_n, qq_block, mu_ws, nrm_ws, D2, stream);
}

extern "C" int adkf_gram_pairstat_bwd(int32_t stat, const float* Z, int32_t d, int32_t T, const int32_t* z_off,
                                      const int64_t* d2_off, int32_t max_n, const float* Wm, const float* mu_ws,
                                      float scale, float* dZ, void* stream) {
    if (T <= 0 || max_n <= 0) return 0;
    if (!Z || !z_off || !d2_off || !Wm || !mu_ws || !dZ || d <= 0) return ADKF_ERR_BADARG;
    if (stat != ADKF_STAT_SQDIST && stat != ADKF_STAT_DOT) return ADKF_ERR_BADARG;   // Tanimoto inputs are fixed fingerprints
    cudaStream_t s = (cudaStream_t)stream;
    sqdist_bwd_kernel<<<dim3((d + kTile - 1) / kTile, (max_n + kTile - 1) / kTile, T), 256, 0, s>>>(
        Z, d, z_off, d2_off, Wm, mu_ws, scale, dZ, stat);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_gram_sqdist_bwd(const float* Z, int32_t d, int32_t T, const int32_t* z_off,
                                    const int64_t* d2_off, int32_t max_n, const float* Wm, const float* mu_ws,
                                    float scale, float* dZ, void* stream) {
    return adkf_gram_pairstat_bwd(ADKF_STAT_SQDIST, Z, d, T, z_off, d2_off, max_n, Wm, mu_ws, scale, dZ, stream);
}
